/*
 * examples/skewed_gaussian.cu -- the reference's own stretch-move test program
 * (test/sequential/SkewedGaussian/StretchMove/src/main.cpp:22-80) as a user of the drop-in headers: a USER-defined
 * Calculator (not one of the library's built-ins) given a __host__ __device__ calcLogPostProb, the reference's class
 * names and call sequence, kernels instantiated in this translation unit.
 * `--mover de|walk` plugs MCMC::Mover::DifferentialEvolution / WalkMove into the same sampler template.
 * Prints acceptance counts, means and covariance; `--dump N` prints walker 0 of the last stored step in hex floats so
 * tests can compare with the oracle bit for bit.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include "mcmcpp_b200/EnsembleSampler.h"

/* same arithmetic as test/sequential/SkewedGaussian/Common/SkewedGaussian.h:52-57, restated with HD qualifiers */
template <class ParamType>
class UserSkewedGaussian
{
public:
    __host__ __device__ UserSkewedGaussian(const ParamType& eps) : epsilon(eps) {}
    __host__ __device__ ParamType calcLogPostProb(ParamType* p)
    {
        ParamType t1 = (p[0] / static_cast<ParamType>(2) - p[1]);
        ParamType t2 = (p[0] / static_cast<ParamType>(2) + p[1]);
        return ((((t1 * t1) / epsilon) + (t2 * t2)) / static_cast<ParamType>(-2));
    }
private:
    ParamType epsilon;
};

using namespace MCMC;

/* a PostStepAction (concept: EnsembleSampler.h:52-56): counts its calls and remembers the range it was last shown */
struct CountingAction
{
    typedef Chain::ChainStepIterator<double> StepItt;
    long long calls = 0, lastSteps = 0;
    double lastFirstCell = 0.0;
    void performAction(const StepItt& start, const StepItt& end)
    {
        ++calls; lastSteps = 0;
        for (StepItt itt = start; itt != end; ++itt) { ++lastSteps; lastFirstCell = (*itt)[0]; }
    }
};

int main(int argc, char** argv)
{
    typedef Mover::StretchMove<double, UserSkewedGaussian<double> > Walker;
    typedef Mover::DifferentialEvolution<double, UserSkewedGaussian<double> > DeWalker;      /* --mover de */
    typedef Mover::WalkMove<double, UserSkewedGaussian<double> > WalkWalker;                 /* --mover walk [--points S] */
    const int runNumber = 0;
    int numWalkers = 320, numSteps = 2000, slicing = 3, threads = 0;
    const int numParams = 2;
    const double eps = 0.13;
    bool dump = false, withAction = false;
    const char* moverName = "stretch";
    int points = 3;
    for (int i = 1; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--walkers") && i + 1 < argc) numWalkers = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--steps") && i + 1 < argc) numSteps = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--slicing") && i + 1 < argc) slicing = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--threads") && i + 1 < argc) threads = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--mover") && i + 1 < argc) moverName = argv[++i];
        else if (!std::strcmp(argv[i], "--points") && i + 1 < argc) points = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--dump")) dump = true;
        else if (!std::strcmp(argv[i], "--action")) withAction = true;
    }
    UserSkewedGaussian<double> likelihood(eps);
    Walker mover(numParams, runNumber, likelihood);

    /* deterministic initial ensemble the Python tests can regenerate: x = sin/cos lattice, aux = logp */
    double* initVals = new double[numWalkers * numParams];
    double* auxVals = new double[numWalkers];
    for (int w = 0; w < numWalkers; ++w) {
        initVals[2 * w] = 0.001 * (double)((w * 7919) % 2003) - 1.0;
        initVals[2 * w + 1] = 0.002 * (double)((w * 104729) % 1009) - 1.0;
        auxVals[w] = likelihood.calcLogPostProb(initVals + 2 * w);
    }

    unsigned long long acc = 0, tot = 0;
    int stored = 0;
    double last[2] = {0, 0}, mean[2] = {0, 0}, cov[3] = {0, 0, 0};
    auto analyse = [&](auto& sampler) {
        sampler.setInitialWalkerPos(initVals, auxVals);
        bool status = sampler.runMCMC(numSteps);
        std::printf("status %s\n", status ? "complete" : "chain full");
        sampler.sliceAndBurnChain(1, 20);
        acc = sampler.getAcceptedSteps(); tot = sampler.getTotalSteps(); stored = sampler.getStoredSteps();
        long long n = 0;
        auto end = sampler.getParamSetIttEnd();
        for (auto itt = sampler.getParamSetIttBegin(); itt != end; ++itt) { mean[0] += (*itt)[0]; mean[1] += (*itt)[1]; ++n; }
        mean[0] /= n; mean[1] /= n;
        for (auto itt = sampler.getParamSetIttBegin(); itt != end; ++itt) {
            double a = (*itt)[0] - mean[0], b = (*itt)[1] - mean[1];
            cov[0] += a * a; cov[1] += a * b; cov[2] += b * b;
        }
        cov[0] /= n; cov[1] /= n; cov[2] /= n;
        auto lastStep = sampler.getStepIttEnd();
        --lastStep;
        last[0] = (*lastStep)[0]; last[1] = (*lastStep)[1];
        std::printf("fraction %.12f\n", sampler.getAcceptanceFraction());
    };
    if (withAction) {
        CountingAction action;
        EnsembleSampler<double, Walker, CountingAction> sampler(runNumber, numWalkers, numParams, mover, 2147483648ULL, &action);
        if (slicing > 1) sampler.setSlicingMode(true, slicing);
        analyse(sampler);
        std::printf("action calls %lld last_range %lld last_first %a\n", action.calls, action.lastSteps, action.lastFirstCell);
    } else if (!std::strcmp(moverName, "de")) {
        DeWalker deMover(numParams, runNumber, likelihood);
        EnsembleSampler<double, DeWalker> sampler(runNumber, numWalkers, numParams, deMover);
        if (slicing > 1) sampler.setSlicingMode(true, slicing);
        analyse(sampler);
    } else if (!std::strcmp(moverName, "walk")) {
        WalkWalker walkMover(numParams, runNumber, likelihood, points);
        EnsembleSampler<double, WalkWalker> sampler(runNumber, numWalkers, numParams, walkMover);
        if (slicing > 1) sampler.setSlicingMode(true, slicing);
        analyse(sampler);
    } else if (threads > 0) {
        ParallelEnsembleSampler<double, Walker> sampler(runNumber, threads, numWalkers, numParams, mover);
        if (slicing > 1) sampler.setSamplingMode(slicing, 0);
        analyse(sampler);
    } else {
        EnsembleSampler<double, Walker> sampler(runNumber, numWalkers, numParams, mover);
        if (slicing > 1) sampler.setSlicingMode(true, slicing);
        analyse(sampler);
    }
    delete[] initVals;
    delete[] auxVals;
    std::printf("accepted %llu total %llu stored %d\n", acc, tot, stored);
    std::printf("mean %.6f %.6f cov %.6f %.6f %.6f (exact 1.13 0.435 0.2825)\n", mean[0], mean[1], cov[0], cov[1], cov[2]);
    if (dump) std::printf("last %a %a\n", last[0], last[1]);
    return 0;
}
