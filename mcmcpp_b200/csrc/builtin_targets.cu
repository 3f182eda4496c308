/*
 * mcmcpp_b200/csrc/builtin_targets.cu -- half-step launchers for the built-in Calculators
 * (BASELINE.json configs; include/mcmcpp_b200/targets.h).  Each is an instantiation of the same
 * templates a user's CUDA translation unit instantiates for its own Calculator
 * (MCMC::Device::HalfStepLauncher<Calc>, include/mcmcpp_b200/stretch_kernels.cuh).
 */
#include <cstring>
#include "builtin_targets.h"
#include "mcmcpp_b200/stretch_kernels.cuh"
#include "mcmcpp_b200/dim_shard_kernels.cuh"
#include "mcmcpp_b200/diff_evo_kernels.cuh"
#include "mcmcpp_b200/walk_move_kernels.cuh"
#include "mcmcpp_b200/targets.h"

namespace
{
template <class Calc>
void storeCalc(const Calc& c, std::vector<char>* blob)
{
    blob->resize(sizeof(Calc));
    std::memcpy(blob->data(), &c, sizeof(Calc));
}
}

int mcmcpp_builtin_bind(int kind, int W, int D, const double* hostParams, size_t nValues, const double* devParams,
                        mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator, void** aux, int shardMode, int dimBegin, int dimCount)
{
    using namespace MCMC::Targets;
    *aux = nullptr;
    if (shardMode == MCMCPP_SHARD_DIM) {
        /* parameter-dimension sharding: separable Calculators only; the functor views this rank's columns */
        if (kind != MCMCPP_TARGET_DIAG_GAUSS || nValues != (size_t)D) return MCMCPP_ERR_UNSUPPORTED;
        storeCalc(DiagGaussian<double>(devParams + dimBegin, dimCount), calculator);
        *launcher = &MCMC::Device::DimShardLauncher<DiagGaussian<double> >::launch;
        return MCMCPP_OK;
    }
    switch (kind) {
    case MCMCPP_TARGET_SKEWED_GAUSS_2D: {
        if (D != 2 || nValues != 1) return MCMCPP_ERR_ARG;
        storeCalc(SkewedGaussianTwoDim<double>(hostParams[0]), calculator);
        *launcher = &MCMC::Device::HalfStepLauncher<SkewedGaussianTwoDim<double>, 2>::launch;
        return MCMCPP_OK;
    }
    case MCMCPP_TARGET_ROSENBROCK_2D: {
        if (D != 2) return MCMCPP_ERR_ARG;
        storeCalc(RosenbrockTwoDim<double>(), calculator);
        *launcher = &MCMC::Device::HalfStepLauncher<RosenbrockTwoDim<double>, 2>::launch;
        return MCMCPP_OK;
    }
    case MCMCPP_TARGET_DIAG_GAUSS: {
        if (nValues != (size_t)D) return MCMCPP_ERR_ARG;
        storeCalc(DiagGaussian<double>(devParams, D), calculator);
        *launcher = &MCMC::Device::HalfStepLauncher<DiagGaussian<double> >::launch;
        return MCMCPP_OK;
    }
    case MCMCPP_TARGET_DENSE_GAUSS: {
        if (nValues != (size_t)D * (size_t)D) return MCMCPP_ERR_ARG;
        if (mcmcpp_dense_bind(D, devParams, launcher, calculator)) return MCMCPP_OK;
        storeCalc(DenseGaussian<double>(devParams, D), calculator);
        *launcher = &MCMC::Device::HalfStepLauncher<DenseGaussian<double> >::launch;
        return MCMCPP_OK;
    }
    case MCMCPP_TARGET_LINREG: {
        if (nValues < 2) return MCMCPP_ERR_ARG;
        const long M = (long)hostParams[0];
        if (M <= 0 || nValues != 2 + (size_t)M * (size_t)D + (size_t)M) return MCMCPP_ERR_ARG;
        const double* X = devParams + 2;
        if (mcmcpp_linreg_bind(W, D, M, hostParams[1], X, X + (size_t)M * D, launcher, calculator, aux)) return MCMCPP_OK;
        if (*aux) { mcmcpp_linreg_release(*aux); *aux = nullptr; }
        storeCalc(LinearRegression<double>(X, X + (size_t)M * D, M, D, hostParams[1]), calculator);
        *launcher = &MCMC::Device::HalfStepLauncher<LinearRegression<double> >::launch;
        return MCMCPP_OK;
    }
    default: return MCMCPP_ERR_ARG;
    }
}

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

mcmcpp_fused_fn mcmcpp_builtin_fused(int kind, int D, int shardMode)
{
    using namespace MCMC::Targets;
    if (shardMode != MCMCPP_SHARD_NONE || D > MCMC::Device::ThreadKernelMaxDim) return nullptr;
    switch (kind) {
    case MCMCPP_TARGET_SKEWED_GAUSS_2D: return &MCMC::Device::FusedLauncher<SkewedGaussianTwoDim<double>, 2>::launch;
    case MCMCPP_TARGET_ROSENBROCK_2D: return &MCMC::Device::FusedLauncher<RosenbrockTwoDim<double>, 2>::launch;
    case MCMCPP_TARGET_DIAG_GAUSS: return &MCMC::Device::FusedLauncher<DiagGaussian<double> >::launch;
    case MCMCPP_TARGET_DENSE_GAUSS: return &MCMC::Device::FusedLauncher<DenseGaussian<double> >::launch;   /* D <= 8: generic calculator blob */
    default: return nullptr;
    }
}

mcmcpp_group_fn mcmcpp_builtin_group(int kind, int shardMode)
{
    using namespace MCMC::Targets;
    if (shardMode == MCMCPP_SHARD_DIM && kind == MCMCPP_TARGET_DIAG_GAUSS) return &MCMC::Device::DimShardLauncher<DiagGaussian<double> >::launchGroup;
    return nullptr;
}

void mcmcpp_builtin_release(void* aux) { mcmcpp_linreg_release(aux); }   /* the only target with scratch */
