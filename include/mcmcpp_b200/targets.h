/*
 * include/mcmcpp_b200/targets.h -- log-posterior Calculators of the BASELINE.json
 * configurations, in __host__ __device__ form.
 *
 * Each class satisfies the reference's Calculator concept
 * (MCMCpp/Movers/StretchMove.h:47-53: `ParamType calcLogPostProb(ParamType*)`,
 * copy-constructible), so the SAME source is plugged into
 *   - the unmodified reference `MCMC::Mover::StretchMove<double, Calc>` on the host
 *     (oracle/ref_driver.cpp, CPU baseline), and
 *   - the sm_100a half-step kernels (mcmcpp_b200/csrc/stretch_kernels.cuh).
 * Functors with array state are non-owning VIEWS: the pointers are host
 * pointers on the host side and device pointers inside kernels.
 *
 * Arithmetic order is part of each functor's definition (parity with the
 * sequential reference is bit-exact only if host and device add in the same
 * order); nothing here may be FMA-contracted (host: -ffp-contract=off,
 * device: -fmad=false).
 */
#ifndef MCMCPP_B200_TARGETS_H
#define MCMCPP_B200_TARGETS_H

#if defined(__CUDACC__)
#define MCMCPP_HD __host__ __device__ __forceinline__
#else
#define MCMCPP_HD inline
#endif

namespace MCMC
{
namespace Targets
{

/* Same target as the reference's only Calculator,
 * test/sequential/SkewedGaussian/Common/SkewedGaussian.h:52-57 (restated, HD). */
template <class ParamType>
class SkewedGaussianTwoDim
{
public:
    MCMCPP_HD SkewedGaussianTwoDim(const ParamType& eps) : epsilon(eps) {}
    MCMCPP_HD ParamType calcLogPostProb(ParamType* paramSet) const
    {
        ParamType temp1 = (paramSet[0] / static_cast<ParamType>(2) - paramSet[1]);
        ParamType temp2 = (paramSet[0] / static_cast<ParamType>(2) + paramSet[1]);
        return ((((temp1 * temp1) / epsilon) + (temp2 * temp2)) / static_cast<ParamType>(-2));
    }
    ParamType epsilon;
};

namespace Detail
{
/* v / -20 as an IEEE division, any type */
template <class ParamType>
MCMCPP_HD ParamType overMinusTwenty(ParamType v) { return v / static_cast<ParamType>(-20); }

/* ... and for double on the device without the generic division sequence (a MUFU, four Newton DFMAs and a range check to divide
 * by a constant): q = RN(v c), r = v + 20 q (exact in one FMA), RN(q + r c) with c = RN(1/-20) is the correctly rounded
 * quotient (Markstein), i.e. the very bits of v / -20.0 wherever nothing leaves the normal range; outside it the division itself. */
MCMCPP_HD double overMinusTwenty(double v)
{
#if defined(__CUDA_ARCH__)
    const unsigned h = (unsigned)__double2hiint(v) & 0x7fffffffu;
    if (h - 0x07000000u < 0x70000000u) {                 /* 2^-911 <= |v| < 2^881 */
        const double c = -0x1.999999999999ap-5;
        const double q = __dmul_rn(v, c);
        return fma(fma(20.0, q, v), c, q);
    }
#endif
    return v / -20.0;
}
}

/* BASELINE config 2: Goodman-Weare banana, -(100 (y-x^2)^2 + (1-x)^2)/20 */
template <class ParamType>
class RosenbrockTwoDim
{
public:
    MCMCPP_HD RosenbrockTwoDim() {}
    MCMCPP_HD ParamType calcLogPostProb(ParamType* paramSet) const
    {
        ParamType t1 = paramSet[1] - paramSet[0] * paramSet[0];
        ParamType t2 = static_cast<ParamType>(1) - paramSet[0];
        return Detail::overMinusTwenty((static_cast<ParamType>(100) * (t1 * t1)) + (t2 * t2));
    }
};

/* Canonical 64-way strided sum finished by an xor-butterfly: partial[d%64]
 * accumulates in increasing d; s[l]=p[2l]+p[2l+1]; s[l]+=s[l^off], off=16..1.
 * One warp whose lane l owns elements 64k+2l, 64k+2l+1 reproduces this order
 * exactly with __shfl_xor_sync. */
template <class ParamType>
MCMCPP_HD ParamType canonical64Finish(const ParamType* p)
{
    ParamType s[32], t[32];
    for (int l = 0; l < 32; ++l) s[l] = p[2 * l] + p[2 * l + 1];
    for (int off = 16; off >= 1; off >>= 1) {
        for (int l = 0; l < 32; ++l) t[l] = s[l] + s[l ^ off];
        for (int l = 0; l < 32; ++l) s[l] = t[l];
    }
    return s[0];
}

/* BASELINE config 4: -1/2 sum_d x_d^2 iv_d (diagonal Gaussian, inverse variances iv). */
template <class ParamType>
class DiagGaussian
{
public:
    MCMCPP_HD DiagGaussian(const ParamType* invVar, int numParams) : iv(invVar), D(numParams) {}
    MCMCPP_HD ParamType calcLogPostProb(ParamType* x) const
    {
        ParamType p[64];
        for (int l = 0; l < 64; ++l) p[l] = static_cast<ParamType>(0);
        for (int d = 0; d < D; ++d) p[d & 63] += (x[d] * x[d]) * iv[d];
        return static_cast<ParamType>(-0.5) * canonical64Finish(p);
    }
    /* element term, used by the warp-cooperative kernels (same arithmetic) */
    MCMCPP_HD ParamType term(ParamType xd, int d) const { return (xd * xd) * iv[d]; }
    MCMCPP_HD ParamType finish(ParamType sum) const { return static_cast<ParamType>(-0.5) * sum; }
    const ParamType* iv;
    int D;
};

/* BASELINE config 3: -1/2 x^T P x with dense precision P (row-major D x D). */
template <class ParamType>
class DenseGaussian
{
public:
    MCMCPP_HD DenseGaussian(const ParamType* precision, int numParams) : P(precision), D(numParams) {}
    MCMCPP_HD ParamType calcLogPostProb(ParamType* x) const
    {
        ParamType q = static_cast<ParamType>(0);
        for (int i = 0; i < D; ++i) {
            ParamType y = static_cast<ParamType>(0);
            for (int j = 0; j < D; ++j) y += P[(size_t)i * D + j] * x[j];
            q += x[i] * y;
        }
        return static_cast<ParamType>(-0.5) * q;
    }
    const ParamType* P;
    int D;
};

/* BASELINE config 5: Bayesian linear regression,
 * -1/2 ||y - X th||^2 - 1/2 ||th||^2 / priorVar, X row-major M x D. */
template <class ParamType>
class LinearRegression
{
public:
    MCMCPP_HD LinearRegression(const ParamType* Xmat, const ParamType* yvec, long numData, int numParams, ParamType priorVariance)
        : X(Xmat), y(yvec), M(numData), D(numParams), priorVar(priorVariance) {}
    MCMCPP_HD ParamType calcLogPostProb(ParamType* th) const
    {
        ParamType ss = static_cast<ParamType>(0), tt = static_cast<ParamType>(0);
        for (long m = 0; m < M; ++m) {
            ParamType r = static_cast<ParamType>(0);
            for (int j = 0; j < D; ++j) r += X[(size_t)m * D + j] * th[j];
            r = y[m] - r;
            ss += r * r;
        }
        for (int j = 0; j < D; ++j) tt += th[j] * th[j];
        return static_cast<ParamType>(-0.5) * ss - static_cast<ParamType>(0.5) * (tt / priorVar);
    }
    const ParamType* X;
    const ParamType* y;
    long M;
    int D;
    ParamType priorVar;
};

}
}
#endif
