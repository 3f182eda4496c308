/*
 * mcmcpp_b200/csrc/builtin_targets.h -- internal: binds a built-in Calculator (include/mcmcpp_b200/targets.h)
 * to its half-step launcher.  Not part of the C-ABI.
 */
#ifndef MCMCPP_B200_BUILTIN_TARGETS_H
#define MCMCPP_B200_BUILTIN_TARGETS_H
#include <cstddef>
#include <vector>
#include "mcmcpp_b200/capi.h"

/* hostParams/nValues: the caller's parameter vector; devParams: its device copy (viewed by the functor).
 * Fills *launcher and the by-value calculator blob; *aux receives optional scratch owned by the target. */
int mcmcpp_builtin_bind(int kind, int W, int D, const double* hostParams, size_t nValues, const double* devParams,
                        mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator, void** aux, int shardMode, int dimBegin, int dimCount);
int mcmcpp_builtin_bind_mover(int moverKind, int kind, int W, int D, const double* moverParams, size_t nMover, const double* hostParams,
                              size_t nValues, const double* devParams, mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator,
                              mcmcpp_fused_fn* fused);     /* *fused: the mover's small-ensemble launcher (D <= 8), or NULL */
/* the fused small-ensemble launcher that goes with mcmcpp_builtin_bind's generic launcher for this target, or NULL */
mcmcpp_fused_fn mcmcpp_builtin_fused(int kind, int D, int shardMode);
/* the rank-group launcher (all ranks of a DIM-sharded ensemble in one cooperative launch) for this target, or NULL */
mcmcpp_group_fn mcmcpp_builtin_group(int kind, int shardMode);
void mcmcpp_builtin_release(void* aux);
/* dense_gauss.cu: FP64 tensor-core half-step for the dense Gaussian; false = use the generic launcher */
bool mcmcpp_dense_bind(int D, const double* devP, mcmcpp_halfstep_fn* launcher, std::vector<char>* calculator);
/* linreg.cu: FP64 tensor-core GEMM half-step for the linear-regression target; false = use the generic launcher */
bool mcmcpp_linreg_bind(int W, int D, long M, double priorVar, const double* devX, const double* devY, mcmcpp_halfstep_fn* launcher,
                        std::vector<char>* calculator, void** aux);
void mcmcpp_linreg_release(void* aux);
#endif
