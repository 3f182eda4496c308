/*
 * mcmcpp_b200/csrc/builtin_movers.cu -- internal: binds the built-in Calculators (include/mcmcpp_b200/targets.h) to the
 * launchers of Mover::DifferentialEvolution and Mover::WalkMove (mcmcpp_sampler_create_builtin_mover).  A translation unit of
 * its own so that it compiles beside builtin_targets.cu (the two carry most of the library's template instantiations).
 */
#include <cmath>
#include <cstring>
#include "builtin_targets.h"
#include "mcmcpp_b200/diff_evo_kernels.cuh"
#include "mcmcpp_b200/walk_move_kernels.cuh"
#include "mcmcpp_b200/targets.h"

namespace
{
template <class Calc, int FixedD>
int bindDiffEvo(const Calc& c, int D, const MCMC::Device::DiffEvoParams& mv, mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator,
                mcmcpp_fused_fn* fused)
{
    if (FixedD > 0 && D != FixedD) return MCMCPP_ERR_ARG;
    MCMC::Device::DiffEvoBlob<Calc> blob = {c, mv};
    calculator->resize(sizeof blob);
    std::memcpy(calculator->data(), &blob, sizeof blob);
    *launcher = &MCMC::Device::DiffEvoLauncher<Calc>::launch;
    if (fused) *fused = D <= MCMC::Device::ThreadKernelMaxDim ? &MCMC::Device::DiffEvoFusedLauncher<Calc, FixedD>::launch : nullptr;
    return MCMCPP_OK;
}
}

namespace
{
template <class Calc, int FixedD>
int bindWalk(const Calc& c, int D, int numPoints, mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator, mcmcpp_fused_fn* fused)
{
    if (FixedD > 0 && D != FixedD) return MCMCPP_ERR_ARG;
    MCMC::Device::WalkMoveBlob<Calc> blob = {c, numPoints};
    calculator->resize(sizeof blob);
    std::memcpy(calculator->data(), &blob, sizeof blob);
    *launcher = &MCMC::Device::WalkMoveLauncher<Calc>::launch;
    if (fused) *fused = D <= MCMC::Device::ThreadKernelMaxDim ? &MCMC::Device::WalkMoveFusedLauncher<Calc, FixedD>::launch : nullptr;
    return MCMCPP_OK;
}
}

/* DifferentialEvolution: moverParams = {gamma (<= 0: 2.38/sqrt(2D)), rand (<= 0: 1e-4)};  WalkMove: moverParams = {numSamples} */
int mcmcpp_builtin_bind_mover(int moverKind, int kind, int W, int D, const double* moverParams, size_t nMover, const double* hostParams,
                              size_t nValues, const double* devParams, mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator,
                              mcmcpp_fused_fn* fused)
{
    using namespace MCMC::Targets;
    if (fused) *fused = nullptr;
    if (moverKind == MCMCPP_MOVER_WALK) {
        if (nMover < 1 || !moverParams) return MCMCPP_ERR_ARG;
        const int pts = (int)moverParams[0];
        if (pts < 1 || pts > MCMC::Device::WalkMoveMaxPoints || pts > W / 2) return MCMCPP_ERR_ARG;     /* WalkMove.h:119 asserts numPoints <= numWalkers */
        switch (kind) {
        case MCMCPP_TARGET_SKEWED_GAUSS_2D:
            if (nValues != 1) return MCMCPP_ERR_ARG;
            return bindWalk<SkewedGaussianTwoDim<double>, 2>(SkewedGaussianTwoDim<double>(hostParams[0]), D, pts, launcher, calculator, fused);
        case MCMCPP_TARGET_ROSENBROCK_2D: return bindWalk<RosenbrockTwoDim<double>, 2>(RosenbrockTwoDim<double>(), D, pts, launcher, calculator, fused);
        case MCMCPP_TARGET_DIAG_GAUSS:
            if (nValues != (size_t)D) return MCMCPP_ERR_ARG;
            return bindWalk<DiagGaussian<double>, 0>(DiagGaussian<double>(devParams, D), D, pts, launcher, calculator, fused);
        case MCMCPP_TARGET_DENSE_GAUSS:
            if (nValues != (size_t)D * (size_t)D) return MCMCPP_ERR_ARG;
            return bindWalk<DenseGaussian<double>, 0>(DenseGaussian<double>(devParams, D), D, pts, launcher, calculator, fused);
        default: return MCMCPP_ERR_UNSUPPORTED;
        }
    }
    if (moverKind != MCMCPP_MOVER_DIFF_EVOLUTION) return MCMCPP_ERR_UNSUPPORTED;
    MCMC::Device::DiffEvoParams mv;
    mv.gamma = (nMover > 0 && moverParams[0] > 0.0) ? moverParams[0] : 2.38 / std::sqrt((double)(2 * D));   /* DifferentialEvolution.h:58 */
    const double rnd = (nMover > 1 && moverParams[1] > 0.0) ? moverParams[1] : 1.0e-4;                      /* :83,150-151 */
    mv.lowEdge = -rnd;
    mv.width = 2.0 * rnd;
    switch (kind) {
    case MCMCPP_TARGET_SKEWED_GAUSS_2D:
        if (nValues != 1) return MCMCPP_ERR_ARG;
        return bindDiffEvo<SkewedGaussianTwoDim<double>, 2>(SkewedGaussianTwoDim<double>(hostParams[0]), D, mv, launcher, calculator, fused);
    case MCMCPP_TARGET_ROSENBROCK_2D:
        return bindDiffEvo<RosenbrockTwoDim<double>, 2>(RosenbrockTwoDim<double>(), D, mv, launcher, calculator, fused);
    case MCMCPP_TARGET_DIAG_GAUSS:
        if (nValues != (size_t)D) return MCMCPP_ERR_ARG;
        return bindDiffEvo<DiagGaussian<double>, 0>(DiagGaussian<double>(devParams, D), D, mv, launcher, calculator, fused);
    case MCMCPP_TARGET_DENSE_GAUSS:
        if (nValues != (size_t)D * (size_t)D) return MCMCPP_ERR_ARG;
        return bindDiffEvo<DenseGaussian<double>, 0>(DenseGaussian<double>(devParams, D), D, mv, launcher, calculator, fused);
    default: return MCMCPP_ERR_UNSUPPORTED;
    }
}
