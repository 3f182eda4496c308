/*
 * mcmcpp_b200/csrc/analysis.h -- internal: device-side accumulator behind the covariance entry points of capi.h
 * (MCMC::Analysis::CovarianceMatrix, MCMCpp/Analysis/CovarianceMatrix.h).  Not part of the C-ABI.
 */
#ifndef MCMCPP_B200_ANALYSIS_H
#define MCMCPP_B200_ANALYSIS_H
#include <cuda_runtime.h>

struct mcmcpp_cov_accum;

/* D <= 128.  Returns nullptr when unsupported or out of memory. */
mcmcpp_cov_accum* mcmcpp_cov_create(int D);
void mcmcpp_cov_destroy(mcmcpp_cov_accum* c);
/* zeroStorage (CovarianceMatrix.h:247-264) */
int mcmcpp_cov_zero(mcmcpp_cov_accum* c, cudaStream_t stream);
/* processPoint (CovarianceMatrix.h:201-214) for nPoints contiguous rows of D doubles in device memory */
int mcmcpp_cov_accumulate(mcmcpp_cov_accum* c, const double* devRows, long long nPoints, cudaStream_t stream);
/* finalizeMatrix (CovarianceMatrix.h:143-185); host outputs mean[D], cov[D*D], corr[D*D] (any may be NULL) */
int mcmcpp_cov_finalize(mcmcpp_cov_accum* c, long long nPoints, double* mean, double* cov, double* corr, cudaStream_t stream);
/* AutoCorrCalc::calcAutoCorrTimes over a DEVICE chain [steps][W][D] (all walkers) -> host iat[D]; see mcmcpp_autocorr_times */
int mcmcpp_autocorr_device(const double* chain_dev, long long steps, int W, int D, int window_scaling, int reference_buffer_reuse, double* iat_out);
/* sets the message mcmcpp_last_error returns (runtime.cu) and hands `code` back */
int mcmcpp_internal_fail(int code, const char* what);
#endif
