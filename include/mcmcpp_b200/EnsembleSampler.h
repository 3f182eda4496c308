/*
 * include/mcmcpp_b200/EnsembleSampler.h -- C++ drop-in for the reference's sampler front-ends on the
 * ensemble-update path.  Same namespaces, class names, template parameters and member signatures as
 *   MCMC::EnsembleSampler            MCMCpp/EnsembleSampler.h:39-197
 *   MCMC::ParallelEnsembleSampler    MCMCpp/ParallelEnsembleSampler.h:78-226
 *   MCMC::Mover::StretchMove         MCMCpp/Movers/StretchMove.h:40-131
 *   MCMC::Mover::DifferentialEvolution  MCMCpp/Movers/DifferentialEvolution.h:41-153
 *   MCMC::Mover::WalkMove            MCMCpp/Movers/WalkMove.h:42-197
 *   MCMC::Utility::GwDistribution    MCMCpp/Utility/GwDistribution.h:33-61
 *   MCMC::Utility::NoAction          MCMCpp/Utility/NoAction.h:34-46
 * so a call site such as test/sequential/SkewedGaussian/StretchMove/src/main.cpp:22-139 compiles unchanged when
 * built with nvcc (the Calculator needs __host__ __device__ members; pointers inside it must be device pointers).
 * Everything forwards to the C-ABI (capi.h); the update kernels for the user's Calculator are instantiated HERE, in
 * the user's translation unit.
 *
 * The host chain -- two back-ends, chosen at compile time:
 *   * the reference's headers are on the include path (__has_include("Chain/Chain.h")): the sampler owns a real
 *     MCMC::Chain::Chain<ParamType> (Chain/Chain.h:54-174) and fills its ChainBlock list from the runtime's pinned rows
 *     (a bounded staging window, mcmcpp_sampler_chain_consume).  getStepIttBegin/End, getParamSetIttBegin/End return
 *     the REFERENCE's own iterators, sliceAndBurnChain is the reference's Chain::resetChainForSubSampling, the byte cap
 *     has the reference's whole-block granularity (Chain.h:178,204-236): the reference's Analysis classes
 *     (AutoCorrCalc, CovarianceMatrix, CornerHistograms, PercentileAndMaximumFinder) run on it unchanged.  The
 *     reference's Utility/GwDistribution.h and Utility/NoAction.h are used as they are.
 *   * stand-alone (no reference headers, or -DMCMCPP_B200_STANDALONE_CHAIN): flat iterators with the same operator
 *     surface over the runtime's contiguous pinned rows [step][walker][param].
 * include/mcmcpp_b200/dropin/ holds forwarding headers named like the reference's (EnsembleSampler.h,
 * ParallelEnsembleSampler.h, Movers/StretchMove.h, ...): put it on the include path BEFORE the reference's MCMCpp
 * directory and the reference's own programs build against the device sampler with no source edit.
 * Compile:  nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -I<repo>/include [-I<repo>/include/mcmcpp_b200/dropin -I<MCMCpp>]
 *           user.cu -L<repo>/mcmcpp_b200 -lmcmcpp_b200
 */
#ifndef MCMCPP_B200_ENSEMBLESAMPLER_H
#define MCMCPP_B200_ENSEMBLESAMPLER_H

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include "mcmcpp_b200/capi.h"
#include "mcmcpp_b200/stretch_kernels.cuh"
#include "mcmcpp_b200/diff_evo_kernels.cuh"
#include "mcmcpp_b200/walk_move_kernels.cuh"

#if !defined(MCMCPP_B200_STANDALONE_CHAIN) && defined(__has_include)
#if __has_include("Chain/Chain.h") && __has_include("Utility/GwDistribution.h") && __has_include("Utility/NoAction.h")
#define MCMCPP_B200_REFERENCE_CHAIN 1
#endif
#endif

#ifdef MCMCPP_B200_REFERENCE_CHAIN
#include "Chain/Chain.h"                 /* the reference's own chain, blocks and iterators */
#include "Utility/GwDistribution.h"
#include "Utility/NoAction.h"
namespace MCMC
{
#else
namespace MCMC
{

namespace Utility
{
/* g(z) ~ 1/sqrt(z) on [1/a, a], a = AlphaNum/AlphaDenom, by inverse CDF (GwDistribution.h:45-58) */
template <class ParamType, int AlphaNum = 2, int AlphaDenom = 1>
struct GwDistribution
{
    static const int alphaNum = AlphaNum;
    static const int alphaDenom = AlphaDenom;
    ParamType operator()(ParamType u) const
    {
        const ParamType alpha = static_cast<ParamType>(AlphaNum) / static_cast<ParamType>(AlphaDenom);
        const ParamType invSqrt = static_cast<ParamType>(1) / std::sqrt(alpha);
        const ParamType t = (std::sqrt(alpha) - invSqrt) * u + invSqrt;
        return t * t;
    }
};
}

namespace Chain
{
/* Read-side views over the contiguous host chain [step][walker][param] the runtime fills; same operator surface
 * as the reference iterators the tests and Analysis classes use. */
template <class ParamType>
class ChainStepIterator
{
public:
    ChainStepIterator() : row(nullptr), stride(0) {}
    ChainStepIterator(ParamType* r, long long cellsPerStep) : row(r), stride(cellsPerStep) {}
    ParamType* operator*() const { return row; }
    ChainStepIterator& operator++() { row += stride; return *this; }
    ChainStepIterator operator++(int) { ChainStepIterator t(*this); row += stride; return t; }
    ChainStepIterator& operator--() { row -= stride; return *this; }
    ChainStepIterator operator--(int) { ChainStepIterator t(*this); row -= stride; return t; }
    ChainStepIterator& operator+=(int n) { row += stride * n; return *this; }
    ChainStepIterator& operator-=(int n) { row -= stride * n; return *this; }
    bool operator==(const ChainStepIterator& o) const { return row == o.row; }
    bool operator!=(const ChainStepIterator& o) const { return row != o.row; }
private:
    ParamType* row;
    long long stride;
};

template <class ParamType>
class ChainPsetIterator
{
public:
    ChainPsetIterator() : row(nullptr), stride(0) {}
    ChainPsetIterator(ParamType* r, int cellsPerWalker) : row(r), stride(cellsPerWalker) {}
    ParamType* operator*() const { return row; }
    ChainPsetIterator& operator++() { row += stride; return *this; }
    ChainPsetIterator operator++(int) { ChainPsetIterator t(*this); row += stride; return t; }
    ChainPsetIterator& operator--() { row -= stride; return *this; }
    bool operator==(const ChainPsetIterator& o) const { return row == o.row; }
    bool operator!=(const ChainPsetIterator& o) const { return row != o.row; }
private:
    ParamType* row;
    int stride;
};
}

namespace Utility
{
template <class ParamType>
struct NoAction
{
    typedef Chain::ChainStepIterator<ParamType> StepItt;
    void performAction(const StepItt&, const StepItt&) {}
};
}
#endif

namespace Detail
{
/* AlphaNum / AlphaDenom of a GwDistribution<ParamType, AlphaNum, AlphaDenom> (ours or the reference's: same template) */
template <class Dist> struct GwAlpha;
template <class P, int N, int D> struct GwAlpha<Utility::GwDistribution<P, N, D> > { static const int num = N, denom = D; };
}

namespace Mover
{
/* Holds what the reference mover holds that matters on the device path: the parameter count and a copy of the
 * Calculator (StretchMove.h:62-66).  As in the reference the prngInit argument does not seed the run: the sampler
 * re-seeds its copy with (randSeed, stream) (StretchMove.h:74-78, EnsembleSampler.h:217). */
template <class ParamType, class Calculator, class CustomDistribution = Utility::GwDistribution<ParamType, 2, 1> >
class StretchMove
{
public:
    typedef Calculator CalculatorType;
    typedef CustomDistribution DistributionType;
    StretchMove(int numParams, long long prngInit, const Calculator& orig) : paramCount(numParams), seed(prngInit), stream(0), calc(orig) {}
    void setPrng(long long s, long long st) { seed = s; stream = st; }
    int getParamCount() const { return paramCount; }
    const Calculator& getCalculator() const { return calc; }
    /* the kernels for THIS Calculator are instantiated here, in the user's translation unit */
    int createSampler(mcmcpp_config& cfg, mcmcpp_sampler** out) const
    {
        Calculator c = calc;
        int rc = mcmcpp_sampler_create(&cfg, &Device::HalfStepLauncher<Calculator>::launch, &c, sizeof(Calculator), out);
        /* small ensembles (the reference's own test sizes): the whole runMCMC loop in one launch */
        if (rc == MCMCPP_OK && cfg.num_params <= Device::ThreadKernelMaxDim)
            rc = mcmcpp_sampler_set_fused_launcher(*out, &Device::FusedLauncher<Calculator>::launch);
        return rc;
    }
private:
    int paramCount;
    long long seed, stream;
    Calculator calc;
};

/* MCMCpp/Movers/DifferentialEvolution.h:41-153.  The number of draws per update varies (the second partner is redrawn
 * until it differs, :117-120), so this mover always runs on one pcg64 stream per walker, whichever front-end holds it. */
template <class ParamType, class Calculator>
class DifferentialEvolution
{
public:
    typedef Calculator CalculatorType;
    typedef Utility::GwDistribution<ParamType, 2, 1> DistributionType;      /* unused by this mover */
    DifferentialEvolution(int numParams, long long prngInit, const Calculator& orig)
        : paramCount(numParams), seed(prngInit), stream(0), calc(orig),
          gamma(2.38 / std::sqrt(2.0 * static_cast<ParamType>(numParams))), smallRandLowEdge(-1.0e-4), smallRandWidth(2.0e-4) {}
    void overrideGamma(ParamType newGamma) { gamma = newGamma; }                                            /* :88 */
    void overrideRandBounds(ParamType newRand) { smallRandLowEdge = -newRand; smallRandWidth = 2 * newRand; }   /* :96 */
    void setPrng(long long s, long long st) { seed = s; stream = st; }
    int getParamCount() const { return paramCount; }
    const Calculator& getCalculator() const { return calc; }
    int createSampler(mcmcpp_config& cfg, mcmcpp_sampler** out) const
    {
        Device::DiffEvoBlob<Calculator> blob = {calc, {gamma, smallRandLowEdge, smallRandWidth}};
        cfg.rng_mode = MCMCPP_RNG_PER_WALKER;
        int rc = mcmcpp_sampler_create(&cfg, &Device::DiffEvoLauncher<Calculator>::launch, &blob, sizeof(blob), out);
        if (rc == MCMCPP_OK && paramCount <= Device::ThreadKernelMaxDim)     /* small ensembles: the whole runMCMC call in one launch */
            rc = mcmcpp_sampler_set_fused_launcher(*out, &Device::DiffEvoFusedLauncher<Calculator>::launch);
        return rc;
    }
private:
    int paramCount;
    long long seed, stream;
    Calculator calc;
    ParamType gamma, smallRandLowEdge, smallRandWidth;
};

/* MCMCpp/Movers/WalkMove.h:42-197.  Per-walker streams only (selection sampling draws a variable number of uniforms);
 * the saved second value of std::normal_distribution is per-walker state kept by the runtime (mover_state_doubles). */
template <class ParamType, class Calculator>
class WalkMove
{
public:
    typedef Calculator CalculatorType;
    typedef Utility::GwDistribution<ParamType, 2, 1> DistributionType;      /* unused by this mover */
    WalkMove(int numParams, long long prngInit, const Calculator& orig, int numSamples)
        : paramCount(numParams), numPoints(numSamples), seed(prngInit), stream(0), calc(orig) {}
    void setPrng(long long s, long long st) { seed = s; stream = st; }
    int getParamCount() const { return paramCount; }
    const Calculator& getCalculator() const { return calc; }
    int createSampler(mcmcpp_config& cfg, mcmcpp_sampler** out) const
    {
        Device::WalkMoveBlob<Calculator> blob = {calc, numPoints};
        cfg.rng_mode = MCMCPP_RNG_PER_WALKER;
        cfg.mover_state_doubles = 2;
        int rc = mcmcpp_sampler_create(&cfg, &Device::WalkMoveLauncher<Calculator>::launch, &blob, sizeof(blob), out);
        if (rc == MCMCPP_OK && paramCount <= Device::ThreadKernelMaxDim)     /* small ensembles: the whole runMCMC call in one launch */
            rc = mcmcpp_sampler_set_fused_launcher(*out, &Device::WalkMoveFusedLauncher<Calculator>::launch);
        return rc;
    }
private:
    int paramCount, numPoints;
    long long seed, stream;
    Calculator calc;
};
}

namespace Detail
{
inline void checkStatus(int rc, const char* what)
{
    /* the reference has no exceptions: hard errors are asserts (EnsembleSampler.h:207-208,327,335-336) */
    if (rc < 0) {
        std::fprintf(stderr, "mcmcpp_b200: %s failed (%d): %s\n", what, rc, mcmcpp_last_error());
        std::abort();
    }
}

/* Shared implementation of both front-ends; RngMode is the only thing that differs. */
template <class ParamType, class Mover, class PostStepAction>
class DeviceSamplerBase
{
public:
    typedef Chain::ChainPsetIterator<ParamType> PsetItt;
    typedef Chain::ChainStepIterator<ParamType> StepItt;
    typedef typename Mover::CalculatorType Calculator;
    typedef typename Mover::DistributionType Distribution;

    DeviceSamplerBase(int randSeed, int rngMode, int numWalker, int numParameter, const Mover& move,
                      unsigned long long maxChainSizeBytes, PostStepAction* stepAct)
        : stepAction(stepAct), moveProposer(move), numParams(numParameter), numWalkers(numWalker), handle(nullptr)
#ifdef MCMCPP_B200_REFERENCE_CHAIN
          , markovChain(numWalker, numParameter, maxChainSizeBytes)
#endif
    {
        static_assert(sizeof(ParamType) == sizeof(double), "the B200 path computes in FP64 (ParamType = double)");
        assert(numWalkers % 2 == 0);
        assert(numWalkers > (2 * numParams));
        mcmcpp_config cfg;
        cfg.num_walkers = numWalker; cfg.num_params = numParameter; cfg.seed = randSeed; cfg.rng_mode = rngMode;
        cfg.alpha_num = GwAlpha<Distribution>::num; cfg.alpha_denom = GwAlpha<Distribution>::denom;
        cfg.device = 0; cfg.rank = 0; cfg.world = 1; cfg.shard_mode = MCMCPP_SHARD_NONE; cfg.trace_updates = 0;
        cfg.max_chain_bytes = maxChainSizeBytes; cfg.mover_state_doubles = 0; cfg.chain_block_steps = 0; cfg.device_chain_bytes = 0;
#ifdef MCMCPP_B200_REFERENCE_CHAIN
        /* the runtime's pinned rows are a staging window in front of the reference chain: ~64 MB of rows, at least 2 */
        const unsigned long long rowBytes = sizeof(ParamType) * (unsigned long long)numWalker * (unsigned long long)numParameter;
        windowRows = (long long)std::max<unsigned long long>(2ULL, std::min<unsigned long long>(16384ULL, (64ULL << 20) / rowBytes));
        cfg.max_chain_bytes = (unsigned long long)(windowRows + 1) * rowBytes;
        /* capacity in steps, as Chain::Chain computes it: whole blocks, and the first block always exists (Chain.h:178-186) */
        const long long maxBlocks = (long long)(maxChainSizeBytes / (sizeof(ParamType) * numParameter * Chain::Detail::BlockSize * numWalker));
        capSteps = std::max<long long>(1, maxBlocks) * (long long)Chain::Detail::BlockSize;
#endif
        if (const char* dev = std::getenv("MCMCPP_B200_DEVICE")) cfg.device = std::atoi(dev);
        /* the mover picks its launcher (and, for movers with a variable draw count, per-walker streams) */
        checkStatus(moveProposer.createSampler(cfg, &handle), "EnsembleSampler()");
    }
    ~DeviceSamplerBase() { mcmcpp_sampler_destroy(handle); }
    DeviceSamplerBase(const DeviceSamplerBase&) = delete;
    DeviceSamplerBase& operator=(const DeviceSamplerBase&) = delete;

    void setInitialWalkerPos(ParamType* positions, ParamType* auxValues)
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        checkStatus(mcmcpp_sampler_set_initial(handle, positions, auxValues), "setInitialWalkerPos");
        drainRows(1);
    }
    void storeCurrentWalkerPositions()
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        checkStatus(mcmcpp_sampler_store_current(handle), "storeCurrentWalkerPositions");
        drainRows(1);
    }
    bool runMCMC(int numSteps)
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        const int perStored = subSampling ? subSamplingInterval : 1;
        int done = 0;
        while (done < numSteps) {
            /* With an action the reference calls it after EVERY performStep, saved or not, before the step just made is
             * counted (EnsembleSampler.h:289-308,356-359): one stored step per device call then, so the rows are current. */
            int chunk = stepAction != nullptr ? 1 : numSteps - done;
#ifdef MCMCPP_B200_REFERENCE_CHAIN
            const long long room = capSteps - (long long)markovChain.getStoredStepCount();
            if (room <= 0) return false;                        /* the reference would write past its last block here */
            chunk = (int)std::min<long long>(chunk, std::min(windowRows, room));
#endif
            int rc = mcmcpp_sampler_run(handle, chunk, nullptr);
            checkStatus(rc, "runMCMC");
            if (stepAction != nullptr) {
                for (int j = 0; j < perStored; ++j) stepAction->performAction(getStepIttBegin(), actionEnd());
            }
            const bool full = drainRows(chunk) || rc == MCMCPP_CHAIN_FULL;
            done += chunk;
            if (full) return false;                             /* incrementChainStep() == EndOfChain (EnsembleSampler.h:293,306) */
        }
        return true;
    }
    void reset()
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        checkStatus(mcmcpp_sampler_reset(handle), "reset");
#ifdef MCMCPP_B200_REFERENCE_CHAIN
        storedSteps = 0;
        markovChain.resetChain();
#endif
    }
    ParamType getAcceptanceFraction()
    {
        return static_cast<ParamType>(getAcceptedSteps()) / static_cast<ParamType>(getTotalSteps());
    }
    unsigned long long getAcceptedSteps()
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        unsigned long long a = 0, t = 0;
        checkStatus(mcmcpp_sampler_get_counts(handle, &a, &t), "getAcceptedSteps");
        return a;
    }
    unsigned long long getTotalSteps()
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        unsigned long long a = 0, t = 0;
        checkStatus(mcmcpp_sampler_get_counts(handle, &a, &t), "getTotalSteps");
        return t;
    }
    void sliceAndBurnChain(int slicingInterval, int burnIn)
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        assert(slicingInterval > 0);
        assert(burnIn >= 0);
#ifdef MCMCPP_B200_REFERENCE_CHAIN
        markovChain.resetChainForSubSampling(burnIn, slicingInterval);      /* EnsembleSampler.h:332-339 */
        storedSteps = markovChain.getStoredStepCount();
#else
        checkStatus(mcmcpp_sampler_slice_and_burn(handle, slicingInterval, burnIn), "sliceAndBurnChain");
#endif
    }

#ifdef MCMCPP_B200_REFERENCE_CHAIN
    int getStoredSteps() { std::lock_guard<std::recursive_mutex> lock(samplerMutex); return storedSteps; }
    PsetItt getParamSetIttBegin() { std::lock_guard<std::recursive_mutex> lock(samplerMutex); return markovChain.getPsetIteratorBegin(); }
    PsetItt getParamSetIttEnd() { std::lock_guard<std::recursive_mutex> lock(samplerMutex); return markovChain.getPsetIteratorEnd(); }
    StepItt getStepIttBegin() { std::lock_guard<std::recursive_mutex> lock(samplerMutex); return markovChain.getStepIteratorBegin(); }
    StepItt getStepIttEnd() { std::lock_guard<std::recursive_mutex> lock(samplerMutex); return markovChain.getStepIteratorEnd(); }
#else
    int getStoredSteps() { return mcmcpp_sampler_get_stored_steps(handle); }
    PsetItt getParamSetIttBegin() { return PsetItt(rows(), numParams); }
    PsetItt getParamSetIttEnd() { return PsetItt(rows() + cellsStored(), numParams); }
    StepItt getStepIttBegin() { return StepItt(rows(), (long long)numWalkers * numParams); }
    StepItt getStepIttEnd() { return StepItt(rows() + cellsStored(), (long long)numWalkers * numParams); }
#endif

    /* not in the reference: current walker state back to the host (Walker::getCurrState / getCurrAuxData) */
    void getWalkerState(ParamType* positions, ParamType* auxValues)
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        checkStatus(mcmcpp_sampler_get_state(handle, positions, auxValues), "getWalkerState");
    }
    mcmcpp_sampler* nativeHandle() { return handle; }

protected:
    void setSubSampling(bool use, int interval)
    {
        std::lock_guard<std::recursive_mutex> lock(samplerMutex);
        assert(interval > 0);
        subSampling = use; subSamplingInterval = interval;
        checkStatus(mcmcpp_sampler_set_slicing(handle, use ? 1 : 0, interval), "setSlicingMode");
    }
    ParamType* rows()
    {
        const double* r = nullptr; long long n = 0;
        mcmcpp_sampler_chain(handle, &r, &n);
        return const_cast<ParamType*>(reinterpret_cast<const ParamType*>(r));
    }
    long long cellsStored()
    {
        const double* r = nullptr; long long n = 0;
        mcmcpp_sampler_chain(handle, &r, &n);
        return n * (long long)numWalkers * numParams;
    }
#ifdef MCMCPP_B200_REFERENCE_CHAIN
    /* the step just made is in the staging window, not yet in the chain: the range an action sees ends before it, as in the
     * reference, where the step is not counted until after the action (EnsembleSampler.h:291-293) */
    StepItt actionEnd() { return markovChain.getStepIteratorEnd(); }
    /* moves the `n` rows waiting in the runtime's staging window into the reference chain: the write side of
     * ChainBlock::storeWalker for every walker at once (ChainBlock.h:125-131), then Chain::incrementChainStep (Chain.h:204-236).
     * Returns true when the chain reported EndOfChain. */
    bool drainRows(long long n)
    {
        const double* staged = nullptr; long long have = 0;
        mcmcpp_sampler_chain(handle, &staged, &have);
        n = std::min(n, have);
        const size_t cells = (size_t)numWalkers * (size_t)numParams;
        bool full = false;
        for (long long r = 0; r < n && !full; ++r) {
            StepItt slot = markovChain.getStepIteratorEnd();          /* first empty step of the current block */
            std::memcpy(*slot, staged + cells * (size_t)r, sizeof(ParamType) * cells);
            ++storedSteps;
            full = Chain::IncrementStatus::EndOfChain == markovChain.incrementChainStep();
        }
        checkStatus(mcmcpp_sampler_chain_consume(handle, have), "runMCMC (chain hand-over)");
        return full;
    }
#else
    StepItt actionEnd() { return getStepIttEnd() -= 1; }
    bool drainRows(long long) { return false; }
#endif
    PostStepAction* stepAction;
    Mover moveProposer;
    int numParams, numWalkers;
    bool subSampling = false;
    int subSamplingInterval = 1;
    mcmcpp_sampler* handle;
    /* ParallelEnsembleSampler's contract: every public method takes the sampler's mutex (ParallelEnsembleSampler.h:184-210,
     * 236-347).  Recursive because a PostStepAction may call the getters from inside runMCMC. */
    std::recursive_mutex samplerMutex;
#ifdef MCMCPP_B200_REFERENCE_CHAIN
    Chain::Chain<ParamType> markovChain;
    long long windowRows = 2, capSteps = 0;
    int storedSteps = 0;
#endif
};
}

/* MCMCpp/EnsembleSampler.h:39-197 */
template <class ParamType, class Mover, class PostStepAction = Utility::NoAction<ParamType> >
class EnsembleSampler : public Detail::DeviceSamplerBase<ParamType, Mover, PostStepAction>
{
    typedef Detail::DeviceSamplerBase<ParamType, Mover, PostStepAction> Base;
public:
    EnsembleSampler(int randSeed, int numWalker, int numParameter, const Mover& move,
                    unsigned long long maxChainSizeBytes = 2147483648ULL, PostStepAction* stepAct = nullptr)
        : Base(randSeed, MCMCPP_RNG_SEQUENTIAL, numWalker, numParameter, move, maxChainSizeBytes, stepAct) {}
    void setSlicingMode(bool useSlicing = false, int slicingInterval = 1) { this->setSubSampling(useSlicing, slicingInterval); }
};

/* MCMCpp/ParallelEnsembleSampler.h:78-226.  One thread reproduces the sequential sampler bit for bit (as the
 * reference does); more threads select one RNG stream per walker (the reference is non-deterministic there,
 * ParallelEnsembleSampler.h:71-76).  UseSpinLocks has no meaning on the device and is accepted for source compatibility. */
template <class ParamType, class Mover, class PostStepAction = Utility::NoAction<ParamType>, bool UseSpinLocks = true>
class ParallelEnsembleSampler : public Detail::DeviceSamplerBase<ParamType, Mover, PostStepAction>
{
    typedef Detail::DeviceSamplerBase<ParamType, Mover, PostStepAction> Base;
public:
    ParallelEnsembleSampler(int randSeed, int threadCount, int numWalker, int numParameter, const Mover& move,
                            unsigned long long maxChainSizeBytes = 2147483648ULL, PostStepAction* stepAct = nullptr)
        : Base(randSeed, threadCount <= 1 ? MCMCPP_RNG_SEQUENTIAL : MCMCPP_RNG_PER_WALKER, numWalker, numParameter, move,
               maxChainSizeBytes, stepAct) {}
    /* ParallelEnsembleSampler.h:178,322-330: sets the interval for later runs and thins/burns what is already stored.
     * Later runs follow the sequential sampler's subsampling (interval-1 unsaved steps, then one saved); the parallel
     * controller's own subsampling is broken. */
    void setSamplingMode(int subSamplingInt = 1, int burnIn = 0)
    {
        assert(subSamplingInt > 0);
        assert(burnIn >= 0);
        this->setSubSampling(subSamplingInt > 1, subSamplingInt);
        this->sliceAndBurnChain(subSamplingInt, burnIn);
    }
    void setSlicingMode(bool useSlicing = false, int slicingInterval = 1) { this->setSubSampling(useSlicing, slicingInterval); }
};

}
#endif
