/*
 * oracle/ref_driver.cpp -- TEST INFRASTRUCTURE (oracle), not product code.
 *
 * Thin extern "C" driver around the UNMODIFIED reference headers under
 * /root/reference/MCMCpp (never copied into this repo).  Built by oracle/Makefile
 * into oracle/_ref/ (git-ignored):
 *   libmcmcpp_ref.so        reference headers as they lie (BlockSize 10000)  -> parity + goldens
 *   libmcmcpp_ref_small.so  shadow copy whose only change is Chain::Detail::BlockSize
 *                           (Chain/ChainBlock.h:31) so W >= ~2e4 can be allocated -> CPU timing
 * Single TU on purpose: MCMCpp/Utility/Misc.h:34-67 defines non-inline functions.
 */
#include <chrono>
#include <cmath>
#include <cstring>
#include <random>
#include <thread>
#include <vector>
#include "EnsembleSampler.h"              /* MCMCpp/EnsembleSampler.h */
#include "ParallelEnsembleSampler.h"      /* MCMCpp/ParallelEnsembleSampler.h */
#include "Movers/StretchMove.h"           /* MCMCpp/Movers/StretchMove.h */
#include "Movers/DifferentialEvolution.h" /* MCMCpp/Movers/DifferentialEvolution.h */
#include "Movers/WalkMove.h"              /* MCMCpp/Movers/WalkMove.h */
#include "Analysis/CovarianceMatrix.h"    /* MCMCpp/Analysis/CovarianceMatrix.h */
#include "Analysis/AutoCorrCalc.h"        /* MCMCpp/Analysis/AutoCorrCalc.h */
#include "Common/SkewedGaussian.h"        /* test/sequential/SkewedGaussian/Common/SkewedGaussian.h */
#include "mcmcpp_b200/targets.h"          /* HD functors for the synthetic configs (kinds 1-4) */

namespace {

/* Run `nsteps` stored steps on a sequential reference sampler; copy chain rows out. */
template <class Calc>
int runSequential(const Calc& calc, int W, int D, int seed, double* initPos, double* initAux,
                  int nsteps, int interval, double* chainOut,
                  unsigned long long* acc, unsigned long long* tot)
{
    typedef MCMC::Mover::StretchMove<double, Calc> Mover;
    Mover mover(D, seed, calc);
    /* chain cap large enough for nsteps+1 rows in 10000-step blocks */
    unsigned long long blocks = (unsigned long long)(nsteps + 1) / MCMC::Chain::Detail::BlockSize + 2ULL;
    unsigned long long maxBytes = blocks * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
    if (interval > 1) sampler.setSlicingMode(true, interval);
    sampler.setInitialWalkerPos(initPos, initAux);
    bool ok = sampler.runMCMC(nsteps);
    *acc = sampler.getAcceptedSteps();
    *tot = sampler.getTotalSteps();
    if (chainOut) {
        auto it = sampler.getStepIttBegin();
        auto end = sampler.getStepIttEnd();
        size_t row = (size_t)W * D;
        int s = 0;
        for (; it != end && s <= nsteps; ++it, ++s) std::memcpy(chainOut + s * row, *it, row * sizeof(double));
    }
    return ok ? sampler.getStoredSteps() : -sampler.getStoredSteps();
}

/* Sequential reference sampler with the DifferentialEvolution mover (Movers/DifferentialEvolution.h:113-142). */
template <class Calc>
int runDiffEvo(const Calc& calc, int W, int D, int seed, double* initPos, double* initAux, int nsteps, double* chainOut,
               unsigned long long* acc, unsigned long long* tot)
{
    typedef MCMC::Mover::DifferentialEvolution<double, Calc> Mover;
    Mover mover(D, seed, calc);
    unsigned long long blocks = (unsigned long long)(nsteps + 1) / MCMC::Chain::Detail::BlockSize + 2ULL;
    unsigned long long maxBytes = blocks * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
    sampler.setInitialWalkerPos(initPos, initAux);
    sampler.runMCMC(nsteps);
    *acc = sampler.getAcceptedSteps();
    *tot = sampler.getTotalSteps();
    auto it = sampler.getStepIttBegin();
    auto end = sampler.getStepIttEnd();
    size_t row = (size_t)W * D;
    int s = 0;
    for (; it != end && s <= nsteps; ++it, ++s) std::memcpy(chainOut + s * row, *it, row * sizeof(double));
    return sampler.getStoredSteps();
}

/* Sequential reference sampler with the WalkMove mover (Movers/WalkMove.h:95-186), numSamples partner points. */
template <class Calc>
int runWalk(const Calc& calc, int W, int D, int seed, int numSamples, double* initPos, double* initAux, int nsteps, double* chainOut,
            unsigned long long* acc, unsigned long long* tot)
{
    typedef MCMC::Mover::WalkMove<double, Calc> Mover;
    Mover mover(D, seed, calc, numSamples);
    unsigned long long blocks = (unsigned long long)(nsteps + 1) / MCMC::Chain::Detail::BlockSize + 2ULL;
    unsigned long long maxBytes = blocks * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
    sampler.setInitialWalkerPos(initPos, initAux);
    sampler.runMCMC(nsteps);
    *acc = sampler.getAcceptedSteps();
    *tot = sampler.getTotalSteps();
    auto it = sampler.getStepIttBegin();
    auto end = sampler.getStepIttEnd();
    size_t row = (size_t)W * D;
    int s = 0;
    for (; it != end && s <= nsteps; ++it, ++s) std::memcpy(chainOut + s * row, *it, row * sizeof(double));
    return sampler.getStoredSteps();
}

/* Same run, then the reference's own CovarianceMatrix over the stored chain (Analysis/CovarianceMatrix.h:120-141). */
template <class Calc>
int runCovariance(const Calc& calc, int W, int D, int seed, double* initPos, double* initAux, int nsteps, int covInterval,
                  double* chainOut, double* covOut, double* corrOut)
{
    typedef MCMC::Mover::StretchMove<double, Calc> Mover;
    Mover mover(D, seed, calc);
    unsigned long long blocks = (unsigned long long)(nsteps + 1) / MCMC::Chain::Detail::BlockSize + 2ULL;
    unsigned long long maxBytes = blocks * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
    sampler.setInitialWalkerPos(initPos, initAux);
    sampler.runMCMC(nsteps);
    MCMC::Analysis::CovarianceMatrix<double> cm(D, W);
    cm.calculateCovar(sampler.getStepIttBegin(), sampler.getStepIttEnd(), covInterval);
    for (int i = 0; i < D; ++i)
        for (int j = 0; j < D; ++j) {
            covOut[i * D + j] = cm.getCovarianceMatrixElement(i, j);
            corrOut[i * D + j] = cm.getCorrelationMatrixElement(i, j);
        }
    auto it = sampler.getStepIttBegin();
    auto end = sampler.getStepIttEnd();
    size_t row = (size_t)W * D;
    int s = 0;
    for (; it != end && s <= nsteps; ++it, ++s) std::memcpy(chainOut + s * row, *it, row * sizeof(double));
    return sampler.getStoredSteps();
}

/* Same run, then the reference's own AutoCorrCalc over the stored chain, all walkers (Analysis/AutoCorrCalc.h:150-168). */
template <class Calc>
int runAutoCorr(const Calc& calc, int W, int D, int seed, double* initPos, double* initAux, int nsteps, double* chainOut, double* iatOut)
{
    typedef MCMC::Mover::StretchMove<double, Calc> Mover;
    Mover mover(D, seed, calc);
    unsigned long long blocks = (unsigned long long)(nsteps + 1) / MCMC::Chain::Detail::BlockSize + 2ULL;
    unsigned long long maxBytes = blocks * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
    sampler.setInitialWalkerPos(initPos, initAux);
    sampler.runMCMC(nsteps);
    MCMC::Analysis::AutoCorrCalc<double> ac(D, W);
    auto b = sampler.getStepIttBegin();
    auto e = sampler.getStepIttEnd();
    ac.calcAutoCorrTimes(b, e, sampler.getStoredSteps());
    for (int i = 0; i < D; ++i) iatOut[i] = ac.retrieveAutoCorrelationTime(i);
    auto it = sampler.getStepIttBegin();
    size_t row = (size_t)W * D;
    int s = 0;
    for (; it != e && s <= nsteps; ++it, ++s) std::memcpy(chainOut + s * row, *it, row * sizeof(double));
    return sampler.getStoredSteps();
}

/* Throughput run: steps are issued in bursts that fit one chain block, reset() in between
 * (EnsembleSampler.h:312-322 keeps walker positions).  threads==0 -> sequential sampler. */
template <class Calc>
double runTimed(const Calc& calc, int W, int D, int seed, double* initPos, double* initAux,
                int nsteps, int threads, unsigned long long* acc, unsigned long long* tot)
{
    typedef MCMC::Mover::StretchMove<double, Calc> Mover;
    Mover mover(D, seed, calc);
    const int burst = MCMC::Chain::Detail::BlockSize - 1;
    unsigned long long maxBytes = 2ULL * sizeof(double) * D * MCMC::Chain::Detail::BlockSize * (unsigned long long)W;
    double secs = 0.0;
    unsigned long long a = 0, t = 0;
    if (threads <= 0) {
        MCMC::EnsembleSampler<double, Mover> sampler(seed, W, D, mover, maxBytes);
        sampler.setInitialWalkerPos(initPos, initAux);
        for (int done = 0; done < nsteps;) {
            int n = (nsteps - done < burst) ? nsteps - done : burst;
            sampler.reset();
            auto t0 = std::chrono::steady_clock::now();
            sampler.runMCMC(n);
            secs += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            a += sampler.getAcceptedSteps(); t += sampler.getTotalSteps();
            done += n;
        }
    } else {
        MCMC::ParallelEnsembleSampler<double, Mover, MCMC::Utility::NoAction<double>, true> sampler(seed, threads, W, D, mover, maxBytes);
        sampler.setInitialWalkerPos(initPos, initAux);
        for (int done = 0; done < nsteps;) {
            int n = (nsteps - done < burst) ? nsteps - done : burst;
            sampler.reset();
            auto t0 = std::chrono::steady_clock::now();
            sampler.runMCMC(n);
            secs += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            /* the parallel sampler only exposes the fraction (ParallelEnsembleSampler.h:171) */
            unsigned long long tt = (unsigned long long)W * n;
            a += (unsigned long long)(sampler.getAcceptanceFraction() * (double)tt + 0.5); t += tt;
            done += n;
        }
    }
    *acc = a; *tot = t;
    return secs;
}

/*
 * Restatement driver for the PER_WALKER32 rng discipline: the reference's facade (Utility/MultiSampler.h:54-99) with the
 * vendored 32-bit engine -- `pcg32 engine` (pcg_random.hpp:1568) where MultiSampler.h:120 has pcg64 -- and the SAME
 * libstdc++ adapters and the reference's own GwDistribution; one facade per walker, engine.seed(seed, walker).
 * The update is StretchMove::updateWalker's arithmetic (Movers/StretchMove.h:100-123) on flat arrays.
 */
struct Facade32
{
    pcg32 engine;
    std::uniform_real_distribution<double> realDist;
    std::exponential_distribution<double> expDist;
    MCMC::Utility::GwDistribution<double, 2, 1> customDist;
    void setPrng(unsigned long long seed, unsigned long long stream) { engine.seed(seed, stream); }
    double getCustomSample() { return customDist(realDist(engine)); }
    double getNegExponentialReal() { return -expDist(engine); }
    int getNonOffSetInt(int max) { return engine(max); }
};

template <class Calc>
int runPerWalker32(Calc calc, int W, int D, int seed, double* pos, double* logp, int nsteps, int interval, double* chainOut,
                   unsigned long long* acc, unsigned long long* tot)
{
    const int K = W / 2;
    std::vector<Facade32> prng((size_t)W);
    for (int w = 0; w < W; ++w) prng[(size_t)w].setPrng((unsigned long long)seed, (unsigned long long)w);
    std::vector<double> prop((size_t)D);
    unsigned long long a = (unsigned long long)W, t = (unsigned long long)W;      /* setFirstPoint counts (Walker.h:76,162-170) */
    size_t row = (size_t)W * D;
    if (chainOut) std::memcpy(chainOut, pos, row * sizeof(double));
    for (int s = 0; s < nsteps; ++s) {
        for (int sub = 0; sub < interval; ++sub) {
            for (int h = 0; h < 2; ++h) {
                for (int i = 0; i < K; ++i) {
                    const int w = h * K + i;
                    Facade32& g = prng[(size_t)w];
                    const double* sel = pos + (size_t)((1 - h) * K + g.getNonOffSetInt(K)) * D;
                    double* cur = pos + (size_t)w * D;
                    double scale = g.getCustomSample();
                    for (int d = 0; d < D; ++d) prop[(size_t)d] = sel[d] + scale * (cur[d] - sel[d]);
                    double probScaling = std::log(scale) * static_cast<double>(D - 1);
                    double newProb = calc.calcLogPostProb(prop.data());
                    double probDiff = probScaling + newProb - logp[w];
                    if (g.getNegExponentialReal() < probDiff) { std::memcpy(cur, prop.data(), sizeof(double) * (size_t)D); logp[w] = newProb; ++a; }
                    ++t;
                }
            }
        }
        if (chainOut) std::memcpy(chainOut + (size_t)(s + 1) * row, pos, row * sizeof(double));
    }
    *acc = a; *tot = t;
    return nsteps + 1;
}

} // namespace

/* kind / params encoding is the one of oracle/oracle.c: 0 {eps}; 1 {}; 2 iv[D];
 * 3 P[D*D]; 4 {M, priorVar, X[M*D], y[M]} */
#define DISPATCH(CALL)                                                                             \
    switch (kind) {                                                                                \
    case 0: { SkewedGaussianTwoDim<double> c(p[0]); return CALL(c); }                              \
    case 1: { MCMC::Targets::RosenbrockTwoDim<double> c; return CALL(c); }                         \
    case 2: { MCMC::Targets::DiagGaussian<double> c(p, D); return CALL(c); }                       \
    case 3: { MCMC::Targets::DenseGaussian<double> c(p, D); return CALL(c); }                      \
    case 4: { long M = (long)p[0]; const double* X = p + 2;                                        \
              MCMC::Targets::LinearRegression<double> c(X, X + (size_t)M * D, M, D, p[1]); return CALL(c); } \
    }

extern "C" {

int ref_block_size() { return MCMC::Chain::Detail::BlockSize; }
int ref_hardware_concurrency() { return (int)std::thread::hardware_concurrency(); }

/* log posterior through the functor the reference mover would call (StretchMove.h:111) */
double ref_logp(int kind, int D, const double* p, double* x)
{
#define CALL_LOGP(c) c.calcLogPostProb(x)
    DISPATCH(CALL_LOGP)
    return 0.0;
}

/* Sequential EnsembleSampler (EnsembleSampler.h:284-310).  chainOut: (nsteps+1) rows of W*D
 * (row 0 = initial positions, EnsembleSampler.h:220-230) or NULL.  Returns stored steps
 * (negative if runMCMC returned false). */
int ref_stretch_run(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux,
                    int nsteps, int interval, double* chainOut, unsigned long long* acc, unsigned long long* tot)
{
#define CALL_RUN(c) runSequential(c, W, D, seed, initPos, initAux, nsteps, interval, chainOut, acc, tot)
    DISPATCH(CALL_RUN)
    return 0;
}

/* PER_WALKER32: vendored pcg32 + libstdc++ adapters + reference GwDistribution, one facade per walker.
 * initPos/initAux are updated in place; chainOut: (nsteps+1) rows (row 0 = initial positions) or NULL. */
int ref_stretch_run_pcg32(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux,
                          int nsteps, int interval, double* chainOut, unsigned long long* acc, unsigned long long* tot)
{
#define CALL_RUN32(c) runPerWalker32(c, W, D, seed, initPos, initAux, nsteps, interval, chainOut, acc, tot)
    DISPATCH(CALL_RUN32)
    return 0;
}

/* reference CovarianceMatrix::calculateCovar over the chain of a sequential run (chainOut: nsteps+1 rows) */
int ref_covariance(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux, int nsteps,
                   int covInterval, double* chainOut, double* covOut, double* corrOut)
{
#define CALL_COV(c) runCovariance(c, W, D, seed, initPos, initAux, nsteps, covInterval, chainOut, covOut, corrOut)
    DISPATCH(CALL_COV)
    return 0;
}

/* reference AutoCorrCalc::calcAutoCorrTimes over the chain of a sequential run (chainOut: nsteps+1 rows) */
int ref_autocorr(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux, int nsteps, double* chainOut, double* iatOut)
{
#define CALL_AC(c) runAutoCorr(c, W, D, seed, initPos, initAux, nsteps, chainOut, iatOut)
    DISPATCH(CALL_AC)
    return 0;
}

/* sequential EnsembleSampler<DifferentialEvolution> (chainOut: nsteps+1 rows) */
int ref_de_run(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux, int nsteps, double* chainOut,
               unsigned long long* acc, unsigned long long* tot)
{
#define CALL_DE(c) runDiffEvo(c, W, D, seed, initPos, initAux, nsteps, chainOut, acc, tot)
    DISPATCH(CALL_DE)
    return 0;
}

/* sequential EnsembleSampler<WalkMove> (chainOut: nsteps+1 rows) */
int ref_walk_run(int kind, const double* p, int W, int D, int seed, int numSamples, double* initPos, double* initAux, int nsteps,
                 double* chainOut, unsigned long long* acc, unsigned long long* tot)
{
#define CALL_WALK(c) runWalk(c, W, D, seed, numSamples, initPos, initAux, nsteps, chainOut, acc, tot)
    DISPATCH(CALL_WALK)
    return 0;
}

/* Timed run; threads==0 -> EnsembleSampler on one core, else ParallelEnsembleSampler<..., true>
 * with `threads` workers (ParallelEnsembleSampler.h:228-291).  Returns seconds inside runMCMC. */
double ref_stretch_time(int kind, const double* p, int W, int D, int seed, double* initPos, double* initAux,
                        int nsteps, int threads, unsigned long long* acc, unsigned long long* tot)
{
#define CALL_TIME(c) runTimed(c, W, D, seed, initPos, initAux, nsteps, threads, acc, tot)
    DISPATCH(CALL_TIME)
    return -1.0;
}

/* Test-side initial positions exactly as the survey KAT / reference tests make them:
 * pcg32 eng(seed) + std::normal_distribution<double>(0,1), filled in order
 * (test/sequential/SkewedGaussian/StretchMove/src/main.cpp:146). */
void ref_init_normal(unsigned long long seed, double* out, long n)
{
    pcg32 eng(seed);
    std::normal_distribution<double> nd(0.0, 1.0);
    for (long i = 0; i < n; ++i) out[i] = nd(eng);
}

/* raw MultiSampler draws in StretchMove order (j, z, log(1-u)) for n updates
 * (Utility/MultiSampler.h:60,86,99) */
void ref_draws(long long seed, long long stream, int K, long n, int* j, double* z, double* e)
{
    MCMC::Utility::MultiSampler<double, MCMC::Utility::GwDistribution<double, 2, 1> > ms;
    ms.setPrng(seed, stream);
    for (long i = 0; i < n; ++i) { j[i] = ms.getNonOffSetInt(K); z[i] = ms.getCustomSample(); e[i] = ms.getNegExponentialReal(); }
}

/* raw pcg64 / pcg32 outputs */
void ref_pcg64_raw(unsigned long long seed, unsigned long long stream, unsigned long long skip, unsigned long long* out, long n)
{
    pcg64 g(seed, stream);
    g.advance(skip);
    for (long i = 0; i < n; ++i) out[i] = g();
}
void ref_pcg32_raw(unsigned long long seed, unsigned long long stream, unsigned int* out, long n)
{
    pcg32 g(seed, stream);
    for (long i = 0; i < n; ++i) out[i] = g();
}

} /* extern "C" */
