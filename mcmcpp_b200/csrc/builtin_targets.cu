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
