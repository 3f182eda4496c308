/*
 * include/mcmcpp_b200/stretch_kernels.cuh -- the red/black half-ensemble stretch move as sm_100a
 * kernels, templated on the user's Calculator (the reference's log-posterior plug-in concept,
 * MCMCpp/Movers/StretchMove.h:47-53, in __host__ __device__ form).
 *
 * One launch = one colour of EnsembleSampler::performStep (MCMCpp/EnsembleSampler.h:345-354):
 * every owned walker of that colour runs StretchMove::updateWalker (Movers/StretchMove.h:100-123)
 * against the frozen other colour:
 *     j  = bounded pcg64 draw            (Utility/MultiSampler.h:99)
 *     z  = g(z) inverse CDF              (Utility/MultiSampler.h:60, Utility/GwDistribution.h:58)
 *     prop[d] = sel[d] + z*(cur[d]-sel[d])
 *     q  = log(z)*(D-1) + logp(prop) - logp(cur)
 *     accept iff log(1-u) < q            (Utility/MultiSampler.h:86)
 *     accept: state <- prop, aux <- logp(prop), ++accepted   (Walker/Walker.h:172-179)
 *     stored step: chain row <- current state, accepted or not (Walker/Walker.h:105, Chain/ChainBlock.h:125-131)
 *
 * Two kernels:
 *   stretchHalfStepThread<Calc,D>      one thread per walker, D <= 8: row, partner row and proposal live in
 *                                      registers; 16-byte vector loads/stores of the walker-major rows.
 *   stretchHalfStepWarp<Calc,NCHUNK>   one warp per walker, D <= 64*NCHUNK: lane l owns elements 64c+2l, 64c+2l+1
 *                                      (one double2 per chunk), the row pair is streamed once HBM -> registers,
 *                                      the log-posterior of a SEPARABLE Calculator (term()/finish()) is reduced
 *                                      in the canonical 64-way order of targets.h (bit-identical to its host form);
 *                                      any other Calculator gets its proposal staged in shared memory and is
 *                                      evaluated by lane 0 in the reference's sequential order.
 * All arithmetic on the parity path is written with explicit round-to-nearest intrinsics (no FMA contraction);
 * compile user Calculators with -fmad=false to keep them contraction-free too.
 */
#ifndef MCMCPP_B200_STRETCH_KERNELS_CUH
#define MCMCPP_B200_STRETCH_KERNELS_CUH

#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include "mcmcpp_b200/capi.h"
#include "mcmcpp_b200/device_rng.cuh"

namespace MCMC
{
namespace Device
{

enum { ThreadKernelMaxDim = 8, ThreadKernelThreads = 128, WarpKernelThreads = 256, WarpKernelMaxChunks = 16 };

/* ---- Calculator traits --------------------------------------------------------------------- */
template <class Calc>
struct IsSeparable
{
    template <class C> static char test(decltype(&C::term), decltype(&C::finish));
    template <class C> static long test(...);
    static const bool value = sizeof(test<Calc>(0, 0)) == sizeof(char);
};

/* ---- small helpers -------------------------------------------------------------------------- */
__device__ __forceinline__ double stretchPoint(double sel, double cur, double z)
{
    return __dadd_rn(sel, __dmul_rn(z, __dsub_rn(cur, sel)));
}

template <int D>
__device__ __forceinline__ void loadRow(const double* __restrict__ row, double (&out)[D])
{
    if ((D & 1) == 0) {
        const double2* r2 = reinterpret_cast<const double2*>(row);
#pragma unroll
        for (int d = 0; d < D / 2; ++d) { double2 v = r2[d]; out[2 * d] = v.x; out[2 * d + 1] = v.y; }
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) out[d] = row[d];
    }
}

template <int D>
__device__ __forceinline__ void loadRowReadOnly(const double* __restrict__ row, double (&out)[D])
{
    if ((D & 1) == 0) {
        const double2* r2 = reinterpret_cast<const double2*>(row);
#pragma unroll
        for (int d = 0; d < D / 2; ++d) { double2 v = __ldg(r2 + d); out[2 * d] = v.x; out[2 * d + 1] = v.y; }
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) out[d] = __ldg(row + d);
    }
}

template <int D>
__device__ __forceinline__ void storeRow(double* __restrict__ row, const double (&in)[D])
{
    if ((D & 1) == 0) {
        double2* r2 = reinterpret_cast<double2*>(row);
#pragma unroll
        for (int d = 0; d < D / 2; ++d) r2[d] = make_double2(in[2 * d], in[2 * d + 1]);
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) row[d] = in[d];
    }
}

struct WalkerRng
{
    U128 state, inc;
};

__device__ __forceinline__ WalkerRng loadRng(const mcmcpp_halfstep_args& a, int w)
{
    WalkerRng r;
    ulonglong2 s = reinterpret_cast<const ulonglong2*>(a.rng_state)[w];
    r.state.lo = s.x;
    r.state.hi = s.y;
    /* SEQUENTIAL: the sampler's one stream 0 (EnsembleSampler.h:217); PER_WALKER / PER_WALKER32: stream = walker index
     * (pcg32 keeps its 64-bit state and increment in the low words) */
    r.inc.lo = (a.rng_mode == MCMCPP_RNG_SEQUENTIAL) ? 1ULL : ((((uint64_t)w) << 1) | 1ULL);
    r.inc.hi = 0;
    return r;
}

/* leave the state at this walker's next draw: SEQUENTIAL jumps over the other W-1 walkers' draws */
__device__ __forceinline__ void storeRng(const mcmcpp_halfstep_args& a, int w, WalkerRng r)
{
    if (a.rng_mode == MCMCPP_RNG_SEQUENTIAL) {
        Affine128 m;
        m.mult.lo = a.skip_mult[0]; m.mult.hi = a.skip_mult[1];
        m.plus.lo = a.skip_plus[0]; m.plus.hi = a.skip_plus[1];
        r.state = applyAffine(m, r.state);
    }
    reinterpret_cast<ulonglong2*>(a.rng_state)[w] = make_ulonglong2(r.state.lo, r.state.hi);
}

__device__ __forceinline__ GwConstants gwOf(const mcmcpp_halfstep_args& a)
{
    GwConstants g;
    g.term1 = a.gw_term1;
    g.invSqrtAlpha = a.gw_inv_sqrt_alpha;
    return g;
}

/* the three draws of an update from whichever engine the sampler runs (uniform branch on a launch argument) */
__device__ __forceinline__ StretchDraws drawStretchFor(const mcmcpp_halfstep_args& a, WalkerRng& rng)
{
    if (a.rng_mode == MCMCPP_RNG_PER_WALKER32) return drawStretch32(rng.state.lo, rng.inc.lo, (uint32_t)a.K, (uint32_t)a.bound_threshold, a.bound_recip, gwOf(a));
    return drawStretch(rng.state, rng.inc, (uint64_t)a.K, a.bound_threshold, a.bound_recip, gwOf(a));
}

/* cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute: one sampler per GPU in one process (cfg.device)
 * must opt in on each of them.  One instance per kernel instantiation (a function-local static at the launch site). */
struct DynamicSmemOptIn
{
    std::atomic<size_t> bytes[64];
    template <class Kernel>
    bool ensure(Kernel kernel, size_t need)
    {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return false;
        if (dev < 0 || dev >= 64) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need) == cudaSuccess;
        if (bytes[dev].load(std::memory_order_relaxed) >= need) return true;
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need) != cudaSuccess) return false;
        bytes[dev].store(need, std::memory_order_relaxed);
        return true;
    }
};

__device__ __forceinline__ const double* partnerBase(const mcmcpp_halfstep_args& a, int j)
{
    /* PEER sharding: the partner row lives in the HBM of the rank that owns slot j (NVLink load) */
    return a.peer_pos ? a.peer_pos[j / a.slots_per_rank] : a.pos;
}

/* ---- kernel 1: one thread per walker ---------------------------------------------------------- */
/* D <= 2: 14 CTAs of 128 threads per SM (<= 36 registers) -- 2^19 walkers of a colour are then 1.98 waves on 148 SMs
 * instead of 2.31 at the 12 CTAs the natural 38 registers allow (a third, mostly empty, wave).
 * Load order: rng state, own row and log-prob first (independent of the draws); then the partner index alone, the
 * partner-row load, and only then z and log(1-u) -- ~250 instructions that cover the gather's L2 latency. */
template <class Calc, int D, bool TRACE, bool RNG32>
__global__ void __launch_bounds__(ThreadKernelThreads, (D <= 2 ? 14 : 1)) stretchHalfStepThread(const mcmcpp_halfstep_args a, const Calc calcIn)
{
    const int i = blockIdx.x * ThreadKernelThreads + threadIdx.x;
    int accepted = 0;
    if (i < a.slot_count) {
        Calc calc = calcIn;
        const int slot = a.slot_begin + i;
        const int w = a.half * a.K + slot;
        const int other0 = (1 - a.half) * a.K;

        WalkerRng rng = loadRng(a, w);
        double cur[D], prop[D];
        double* ownRow = a.pos + (size_t)w * D;
        loadRow<D>(ownRow, cur);
        const double oldp = a.logp[w];
        StretchDraws dr;
        if (RNG32) { dr.redraws = 0; dr.j = drawPartner32(rng.state.lo, rng.inc.lo, (uint32_t)a.K, (uint32_t)a.bound_threshold, a.bound_recip); }
        else dr.j = drawPartner(rng.state, rng.inc, (uint64_t)a.K, a.bound_threshold, a.bound_recip, dr.redraws);
        /* Programmatic dependent launch: everything above reads what THIS colour's previous launch (two launches back,
         * complete) wrote; only the partner rows come from the launch right before.  Wait for it here, then let the next
         * launch start its own preamble on the SM slots this one's last, partly filled wave leaves idle (it reads what
         * the previous launch wrote, which is complete once this wait has returned). */
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        loadRowReadOnly<D>(partnerBase(a, dr.j) + (size_t)(other0 + dr.j) * D, prop);
        if (RNG32) drawScaleAndTest32(rng.state.lo, rng.inc.lo, gwOf(a), dr.z, dr.logOneMinusU);
        else drawScaleAndTest(rng.state, rng.inc, gwOf(a), dr.z, dr.logOneMinusU);
        storeRng(a, w, rng);
        if (dr.redraws) atomicAdd(a.flags, dr.redraws);
#pragma unroll
        for (int d = 0; d < D; ++d) prop[d] = stretchPoint(prop[d], cur[d], dr.z);

        const double scaling = __dmul_rn(RNG32 ? logOfNormal(dr.z) : log(dr.z), (double)(D - 1));   /* the same bits (device_rng.cuh) */
        const double newp = calc.calcLogPostProb(prop);
        const double diff = __dsub_rn(__dadd_rn(scaling, newp), oldp);
        accepted = dr.logOneMinusU < diff ? 1 : 0;
        if (accepted) {
            storeRow<D>(ownRow, prop);
            a.logp[w] = newp;
        }
        if (a.chain_row) {
            double* chainRow = a.chain_row + (size_t)w * D;
            if (accepted) storeRow<D>(chainRow, prop); else storeRow<D>(chainRow, cur);
        }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = dr.j; t.accept = accepted; t.z = dr.z; t.e = dr.logOneMinusU; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
    /* acceptance counter: warp popcount, one fire-and-forget RED per warp into the CTA's slot (a CTA-wide count would
     * park every warp at a barrier until the CTA's slowest gather has come back, then expose a load for the +=) */
    const unsigned ballot = __ballot_sync(0xffffffffu, accepted != 0);
    if ((threadIdx.x & 31) == 0 && ballot) atomicAdd(a.accept_slots + blockIdx.x, (unsigned long long)__popc(ballot));
}

/* ---- kernel 2: one warp per walker ------------------------------------------------------------ */
template <class Calc, bool SEPARABLE>
struct WarpLogp;

/* separable Calculator: canonical 64-way strided partial sums + xor butterfly (targets.h canonical64Finish) */
template <class Calc>
struct WarpLogp<Calc, true>
{
    template <int NCHUNK>
    static __device__ __forceinline__ double eval(Calc& calc, const double2 (&prop)[NCHUNK], int D, int lane, double*)
    {
        double p0 = 0.0, p1 = 0.0;
#pragma unroll
        for (int c = 0; c < NCHUNK; ++c) {
            const int d = 64 * c + 2 * lane;
            if (d < D) p0 = __dadd_rn(p0, calc.term(prop[c].x, d));
            if (d + 1 < D) p1 = __dadd_rn(p1, calc.term(prop[c].y, d + 1));
        }
        double s = __dadd_rn(p0, p1);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) s = __dadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
        return calc.finish(s);
    }
};

/* any other Calculator: proposal staged in shared memory, lane 0 evaluates in the reference's order */
template <class Calc>
struct WarpLogp<Calc, false>
{
    template <int NCHUNK>
    static __device__ __forceinline__ double eval(Calc& calc, const double2 (&prop)[NCHUNK], int D, int lane, double* stage)
    {
#pragma unroll
        for (int c = 0; c < NCHUNK; ++c) {
            const int d = 64 * c + 2 * lane;
            if (d < D) stage[d] = prop[c].x;
            if (d + 1 < D) stage[d + 1] = prop[c].y;
        }
        __syncwarp();
        double v = 0.0;
        if (lane == 0) v = calc.calcLogPostProb(stage);
        v = __shfl_sync(0xffffffffu, v, 0);
        __syncwarp();
        return v;
    }
};

template <class Calc, int NCHUNK, bool VEC, bool TRACE>
__global__ void __launch_bounds__(WarpKernelThreads) stretchHalfStepWarp(const mcmcpp_halfstep_args a, const Calc calcIn)
{
    extern __shared__ double stageAll[];
    const int lane = threadIdx.x & 31;
    const int warpInBlock = threadIdx.x >> 5;
    const int i = blockIdx.x * (WarpKernelThreads / 32) + warpInBlock;
    if (i >= a.slot_count) return;          /* warp-uniform */
    Calc calc = calcIn;
    const int D = a.D;
    const int slot = a.slot_begin + i;
    const int w = a.half * a.K + slot;
    const int other0 = (1 - a.half) * a.K;

    /* every lane replays the same three draws (cheaper than a broadcast of six registers) */
    WalkerRng rng = loadRng(a, w);
    StretchDraws dr = drawStretchFor(a, rng);
    /* programmatic dependent launch (see stretchHalfStepThread): the draws read only this colour's rng state */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    double* ownRow = a.pos + (size_t)w * D;
    const double* selRow = partnerBase(a, dr.j) + (size_t)(other0 + dr.j) * D;
    double2 cur[NCHUNK], prop[NCHUNK];
#pragma unroll
    for (int c = 0; c < NCHUNK; ++c) {
        const int d = 64 * c + 2 * lane;
        cur[c] = make_double2(0.0, 0.0);
        prop[c] = make_double2(0.0, 0.0);
        if (VEC) {
            if (d < D) {
                cur[c] = *reinterpret_cast<const double2*>(ownRow + d);
                prop[c] = __ldg(reinterpret_cast<const double2*>(selRow + d));
            }
        } else {
            if (d < D) { cur[c].x = ownRow[d]; prop[c].x = __ldg(selRow + d); }
            if (d + 1 < D) { cur[c].y = ownRow[d + 1]; prop[c].y = __ldg(selRow + d + 1); }
        }
    }
#pragma unroll
    for (int c = 0; c < NCHUNK; ++c) {
        prop[c].x = stretchPoint(prop[c].x, cur[c].x, dr.z);
        prop[c].y = stretchPoint(prop[c].y, cur[c].y, dr.z);
    }

    const double scaling = __dmul_rn(log(dr.z), (double)(D - 1));
    const double newp = WarpLogp<Calc, IsSeparable<Calc>::value>::template eval<NCHUNK>(calc, prop, D, lane, stageAll + (size_t)warpInBlock * D);
    const double diff = __dsub_rn(__dadd_rn(scaling, newp), a.logp[w]);
    const bool accepted = dr.logOneMinusU < diff;

    if (accepted) {
#pragma unroll
        for (int c = 0; c < NCHUNK; ++c) {
            const int d = 64 * c + 2 * lane;
            if (VEC) {
                if (d < D) *reinterpret_cast<double2*>(ownRow + d) = prop[c];
            } else {
                if (d < D) ownRow[d] = prop[c].x;
                if (d + 1 < D) ownRow[d + 1] = prop[c].y;
            }
        }
    }
    if (a.chain_row) {
        double* chainRow = a.chain_row + (size_t)w * D;
#pragma unroll
        for (int c = 0; c < NCHUNK; ++c) {
            const int d = 64 * c + 2 * lane;
            const double2 v = accepted ? prop[c] : cur[c];
            if (VEC) {
                if (d < D) *reinterpret_cast<double2*>(chainRow + d) = v;
            } else {
                if (d < D) chainRow[d] = v.x;
                if (d + 1 < D) chainRow[d + 1] = v.y;
            }
        }
    }
    if (lane == 0) {
        storeRng(a, w, rng);
        if (dr.redraws) atomicAdd(a.flags, dr.redraws);
        if (accepted) {
            a.logp[w] = newp;
            atomicAdd(a.accept_slots + blockIdx.x, 1ULL);
        }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = dr.j; t.accept = accepted ? 1 : 0; t.z = dr.z; t.e = dr.logOneMinusU; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
}

/* ---- kernel 3: one warp per walker, rows staged HBM -> shared memory by TMA bulk copies ---------------- */
/*
 * For wide rows (D > 128) the register kernel above holds both rows in registers (187 regs/thread at D=1000,
 * 8 warps/SM) and cannot keep enough bytes in flight to saturate HBM (ncu r01: 56% of DRAM peak).  Here lane 0
 * issues two cp.async.bulk copies (own row, partner row; 8*D bytes each) that complete on an mbarrier; no
 * register is tied up while the data is in flight and up to 12 walkers (192 KB) per SM are outstanding.
 * The proposal is built in place of the partner row in shared memory, so a generic Calculator can be evaluated
 * on it directly; a separable one is reduced in the canonical 64-way order (bit-identical to its host form).
 * Needs 16-byte aligned rows: D even.
 */
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void bulkLoad(uint32_t dstSmem, const void* srcGlobal, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dstSmem), "l"(srcGlobal), "r"(bytes), "r"(bar) : "memory");
}

__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t phase)
{
    asm volatile("{\n"
                 ".reg .pred P1;\n"
                 "LAB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
                 "@P1 bra DONE;\n"
                 "bra LAB_WAIT;\n"
                 "DONE:\n"
                 "}" ::"r"(bar), "r"(phase) : "memory");
}

enum { BulkKernelMinDim = 130 };

/* The draws of a colour, one THREAD per owned walker, for the TMA kernels.  ~600 dependent integer/FP64 instructions
 * per walker (three 128-bit LCG steps, the stream skip, two logs): made inside a copy kernel they sit in front of its
 * bulk copies and leave its shared memory idle for ~2 us per CTA; here they run at full occupancy and the copy
 * kernel only reads (j, z, log(1-u)) from the draw scratch. */
static __global__ void __launch_bounds__(256) stretchDrawKernel(const mcmcpp_halfstep_args a)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= a.slot_count) return;
    const int w = a.half * a.K + a.slot_begin + i;
    WalkerRng rng = loadRng(a, w);
    StretchDraws dr = drawStretchFor(a, rng);
    storeRng(a, w, rng);
    if (dr.redraws) atomicAdd(a.flags, dr.redraws);
    a.dim_z[i] = dr.z; a.dim_e[i] = dr.logOneMinusU; a.dim_j[i] = dr.j;
}

template <class Calc, int WARPS, bool TRACE, bool PRE, bool ODD = false>
__global__ void __launch_bounds__(WARPS * 32) stretchHalfStepBulk(const mcmcpp_halfstep_args a, const Calc calcIn, const uint32_t rowStride)
{
    extern __shared__ __align__(128) unsigned char bulkSmem[];
    const int lane = threadIdx.x & 31;
    const int warpInBlock = threadIdx.x >> 5;
    const int i = blockIdx.x * WARPS + warpInBlock;
    if (i >= a.slot_count) return;          /* warp-uniform; warps of a CTA never synchronise with each other */
    Calc calc = calcIn;
    const int D = a.D;
    const uint32_t rowBytes = (uint32_t)D * 8u;
    double* ownS = reinterpret_cast<double*>(bulkSmem + (size_t)warpInBlock * 2u * rowStride);
    double* selS = reinterpret_cast<double*>(bulkSmem + (size_t)warpInBlock * 2u * rowStride + rowStride);
    const uint32_t bar = smemAddr(bulkSmem + (size_t)WARPS * 2u * rowStride + 8u * warpInBlock);
    const int slot = a.slot_begin + i;
    const int w = a.half * a.K + slot;
    const int other0 = (1 - a.half) * a.K;

    WalkerRng rng;
    StretchDraws dr;
    if (PRE) {                               /* stretchDrawKernel made them */
        dr.j = a.dim_j[i]; dr.z = a.dim_z[i]; dr.logOneMinusU = a.dim_e[i]; dr.redraws = 0;
    } else {
        rng = loadRng(a, w);
        dr = drawStretchFor(a, rng);
    }
    double* ownRow = a.pos + (size_t)w * D;
    const double* selRow = partnerBase(a, dr.j) + (size_t)(other0 + dr.j) * D;
    if (ODD) {
        /* odd numParams: every second row starts 8 bytes off the 16-byte grid a bulk copy needs -- plain coalesced loads stage
         * the rows instead (rows past 1024 odd parameters, which the register-resident warp kernel cannot hold) */
        for (int d = lane; d < D; d += 32) { ownS[d] = ownRow[d]; selS[d] = __ldg(selRow + d); }
        if (lane == 0) { ownS[D] = 0.0; selS[D] = 0.0; }      /* the pair partner of the last element (the row stride leaves room) */
    } else if (lane == 0) {
        mbarInit(bar, 1);
        mbarExpectTx(bar, 2u * rowBytes);
        bulkLoad(smemAddr(ownS), ownRow, rowBytes, bar);
        bulkLoad(smemAddr(selS), selRow, rowBytes, bar);
    }
    __syncwarp();
    /* overlapped with the copies: the scalar part of the update */
    const double scaling = __dmul_rn(log(dr.z), (double)(D - 1));
    const double oldp = a.logp[w];
    if (!ODD) mbarWait(bar, 0);

    double p0 = 0.0, p1 = 0.0;
    for (int d = 2 * lane; d < D; d += 64) {
        const double2 c = *reinterpret_cast<const double2*>(ownS + d);
        const double2 s = *reinterpret_cast<const double2*>(selS + d);
        double2 p;
        p.x = stretchPoint(s.x, c.x, dr.z);
        p.y = stretchPoint(s.y, c.y, dr.z);
        *reinterpret_cast<double2*>(selS + d) = p;       /* the proposal replaces the partner row */
        if constexpr (IsSeparable<Calc>::value) {
            p0 = __dadd_rn(p0, calc.term(p.x, d));
            if (!ODD || d + 1 < D) p1 = __dadd_rn(p1, calc.term(p.y, d + 1));
        }
    }
    double newp;
    if constexpr (IsSeparable<Calc>::value) {
        double s = __dadd_rn(p0, p1);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) s = __dadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
        newp = calc.finish(s);
    } else {
        __syncwarp();
        double v = 0.0;
        if (lane == 0) v = calc.calcLogPostProb(selS);
        newp = __shfl_sync(0xffffffffu, v, 0);
    }
    const double diff = __dsub_rn(__dadd_rn(scaling, newp), oldp);
    const bool accepted = dr.logOneMinusU < diff;

    if (ODD) {
        __syncwarp();
        if (accepted) for (int d = lane; d < D; d += 32) ownRow[d] = selS[d];
        if (a.chain_row) {
            double* chainRow = a.chain_row + (size_t)w * D;
            const double* src = accepted ? selS : ownS;
            for (int d = lane; d < D; d += 32) chainRow[d] = src[d];
        }
    } else {
        if (accepted) {
            for (int d = 2 * lane; d < D; d += 64) *reinterpret_cast<double2*>(ownRow + d) = *reinterpret_cast<const double2*>(selS + d);
        }
        if (a.chain_row) {
            double* chainRow = a.chain_row + (size_t)w * D;
            const double* src = accepted ? selS : ownS;
            for (int d = 2 * lane; d < D; d += 64) *reinterpret_cast<double2*>(chainRow + d) = *reinterpret_cast<const double2*>(src + d);
        }
    }
    if (lane == 0) {
        if (!PRE) {
            storeRng(a, w, rng);
            if (dr.redraws) atomicAdd(a.flags, dr.redraws);
        }
        if (accepted) {
            a.logp[w] = newp;
            atomicAdd(a.accept_slots + blockIdx.x, 1ULL);
        }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = dr.j; t.accept = accepted ? 1 : 0; t.z = dr.z; t.e = dr.logOneMinusU; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
}

/* ---- launchers ---------------------------------------------------------------------------------- */
inline int launchStatus()
{
    return cudaGetLastError() == cudaSuccess ? MCMCPP_OK : MCMCPP_ERR_CUDA;
}

template <class Calc, int D>
int launchThreadKernel(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
{
    const int grid = (a.slot_count + ThreadKernelThreads - 1) / ThreadKernelThreads;
    if (grid > a.max_grid) return MCMCPP_ERR_ARG;
    /* launched as a programmatic dependent of the kernel before it in the stream (see griddepcontrol.wait in the kernel);
     * MCMCPP_B200_PDL=0 falls back to plain stream order */
    static const bool pdl = [] { const char* e = getenv("MCMCPP_B200_PDL"); return !(e && e[0] == '0'); }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(ThreadKernelThreads); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    const bool rng32 = a.rng_mode == MCMCPP_RNG_PER_WALKER32;
    cudaError_t rc = a.trace ? (rng32 ? cudaLaunchKernelEx(&cfg, stretchHalfStepThread<Calc, D, true, true>, a, calc)
                                      : cudaLaunchKernelEx(&cfg, stretchHalfStepThread<Calc, D, true, false>, a, calc))
                             : (rng32 ? cudaLaunchKernelEx(&cfg, stretchHalfStepThread<Calc, D, false, true>, a, calc)
                                      : cudaLaunchKernelEx(&cfg, stretchHalfStepThread<Calc, D, false, false>, a, calc));
    return rc == cudaSuccess ? launchStatus() : MCMCPP_ERR_CUDA;
}

template <class Calc, int NCHUNK>
int launchWarpKernel(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
{
    const int warpsPerBlock = WarpKernelThreads / 32;
    const int grid = (a.slot_count + warpsPerBlock - 1) / warpsPerBlock;
    if (grid > a.max_grid) return MCMCPP_ERR_ARG;
    const size_t smem = IsSeparable<Calc>::value ? 0 : sizeof(double) * (size_t)a.D * warpsPerBlock;
    const bool vec = (a.D % 2) == 0;
    static const bool pdl = [] { const char* e = getenv("MCMCPP_B200_PDL"); return !(e && e[0] == '0'); }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(WarpKernelThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
#define MCMCPP_LAUNCH_WARP(V, T)                                                                             \
    do {                                                                                                     \
        static DynamicSmemOptIn optIn;                                                                       \
        if (smem > 48 * 1024 && !optIn.ensure(stretchHalfStepWarp<Calc, NCHUNK, V, T>, smem)) return MCMCPP_ERR_CUDA; \
        if (cudaLaunchKernelEx(&cfg, stretchHalfStepWarp<Calc, NCHUNK, V, T>, a, calc) != cudaSuccess) return MCMCPP_ERR_CUDA; \
    } while (0)
    if (vec) { if (a.trace) MCMCPP_LAUNCH_WARP(true, true); else MCMCPP_LAUNCH_WARP(true, false); }
    else     { if (a.trace) MCMCPP_LAUNCH_WARP(false, true); else MCMCPP_LAUNCH_WARP(false, false); }
#undef MCMCPP_LAUNCH_WARP
    return launchStatus();
}

template <class Calc, int WARPS, bool TRACE, bool PRE, bool ODD = false>
int launchBulkKernelWTP(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride, int grid, size_t smem)
{
    static DynamicSmemOptIn optIn;
    if (!optIn.ensure(stretchHalfStepBulk<Calc, WARPS, TRACE, PRE, ODD>, 227 * 1024)) return MCMCPP_ERR_CUDA;
    stretchHalfStepBulk<Calc, WARPS, TRACE, PRE, ODD><<<grid, WARPS * 32, smem, stream>>>(a, calc, rowStride);
    return launchStatus();
}

template <class Calc, int WARPS>
int launchBulkKernelW(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
{
    const int grid = (a.slot_count + WARPS - 1) / WARPS;
    if (grid > a.max_grid) return MCMCPP_ERR_ARG;
    const size_t smem = (size_t)WARPS * 2u * rowStride + 8u * WARPS;
    if (a.D & 1)                             /* odd numParams: rows staged by plain loads, draws in the kernel */
        return a.trace ? launchBulkKernelWTP<Calc, WARPS, true, false, true>(a, calc, stream, rowStride, grid, smem)
                       : launchBulkKernelWTP<Calc, WARPS, false, false, true>(a, calc, stream, rowStride, grid, smem);
    if (a.dim_z && a.dim_e && a.dim_j) {     /* draw scratch available: draws in their own kernel */
        stretchDrawKernel<<<(a.slot_count + 255) / 256, 256, 0, stream>>>(a);
        if (a.launch_count) *a.launch_count += 1;
        return a.trace ? launchBulkKernelWTP<Calc, WARPS, true, true>(a, calc, stream, rowStride, grid, smem)
                       : launchBulkKernelWTP<Calc, WARPS, false, true>(a, calc, stream, rowStride, grid, smem);
    }
    return a.trace ? launchBulkKernelWTP<Calc, WARPS, true, false>(a, calc, stream, rowStride, grid, smem)
                   : launchBulkKernelWTP<Calc, WARPS, false, false>(a, calc, stream, rowStride, grid, smem);
}

/* warps per CTA chosen so that >= 3 CTAs (12 walkers, 2 rows each) fit the 227 KB of one SM when rows allow */
template <class Calc>
int launchBulkKernel(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
{
    const uint32_t rowStride = (((uint32_t)a.D + ((uint32_t)a.D & 1u)) * 8u + 127u) & ~127u;      /* an odd row's last element gets a pair partner */
    const size_t perWarp = 2u * (size_t)rowStride + 8u;
    if (4 * perWarp <= 74 * 1024) return launchBulkKernelW<Calc, 4>(a, calc, stream, rowStride);
    if (2 * perWarp <= 74 * 1024) return launchBulkKernelW<Calc, 2>(a, calc, stream, rowStride);
    if (perWarp <= 226 * 1024) return launchBulkKernelW<Calc, 1>(a, calc, stream, rowStride);
    return MCMCPP_ERR_UNSUPPORTED;
}

inline bool bulkKernelDisabled()
{
    static const bool off = [] { const char* e = getenv("MCMCPP_B200_NO_BULK"); return e && e[0] == '1'; }();
    return off;
}

/* ---- small ensembles: the whole runMCMC loop in one launch (mcmcpp_fused_args) ------------------------------------
 * One CTA; thread t owns red walker t and black walker K + t (their rng states in registers for the whole call); the
 * ensemble lives in shared memory; __syncthreads() is the colour boundary.  Per update exactly the arithmetic of
 * stretchHalfStepThread, so chains, counters and rng states are the ones the launch-per-colour path produces. */
enum { FusedMaxThreads = 1024 };

template <class Calc, int D>
__global__ void __launch_bounds__(FusedMaxThreads, 1) stretchFusedSmall(const mcmcpp_fused_args f, const Calc calcIn)
{
    extern __shared__ double fusedSmem[];
    const mcmcpp_halfstep_args& a = f.base;
    const int K = a.K, W = 2 * K, t = threadIdx.x;
    double* sPos = fusedSmem;                                 /* [W][D] */
    double* sLogp = fusedSmem + (size_t)W * D;                /* [W] */
    for (int idx = t; idx < W * D; idx += blockDim.x) sPos[idx] = a.pos[idx];
    for (int idx = t; idx < W; idx += blockDim.x) sLogp[idx] = a.logp[idx];
    Calc calc = calcIn;
    const bool live = t < K;
    WalkerRng rng[2];
    if (live) { rng[0] = loadRng(a, t); rng[1] = loadRng(a, K + t); }
    Affine128 skip;
    skip.mult.lo = a.skip_mult[0]; skip.mult.hi = a.skip_mult[1];
    skip.plus.lo = a.skip_plus[0]; skip.plus.hi = a.skip_plus[1];
    const bool sequential = a.rng_mode == MCMCPP_RNG_SEQUENTIAL;
    unsigned long long accepted = 0;
    unsigned redraws = 0;
    __syncthreads();
    const long long total = (long long)f.stored_steps * f.interval;
    int sinceStore = 0;
    long long row = 0;
    for (long long step = 0; step < total; ++step) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            if (live) {
                const int w = half * K + t, other0 = (1 - half) * K;
                StretchDraws dr = drawStretchFor(a, rng[half]);
                if (sequential) rng[half].state = applyAffine(skip, rng[half].state);   /* storeRng: over the other walkers' draws */
                redraws += (unsigned)dr.redraws;
                double cur[D], prop[D];
#pragma unroll
                for (int d = 0; d < D; ++d) { cur[d] = sPos[(size_t)w * D + d]; prop[d] = sPos[(size_t)(other0 + dr.j) * D + d]; }
#pragma unroll
                for (int d = 0; d < D; ++d) prop[d] = stretchPoint(prop[d], cur[d], dr.z);
                const double scaling = __dmul_rn(log(dr.z), (double)(D - 1));
                const double newp = calc.calcLogPostProb(prop);
                const double diff = __dsub_rn(__dadd_rn(scaling, newp), sLogp[w]);
                if (dr.logOneMinusU < diff) {
#pragma unroll
                    for (int d = 0; d < D; ++d) sPos[(size_t)w * D + d] = prop[d];
                    sLogp[w] = newp;
                    ++accepted;
                }
            }
            __syncthreads();
        }
        if (++sinceStore == f.interval) {
            sinceStore = 0;
            if (f.chain_rows && live) {
                double* out = f.chain_rows + (size_t)row * W * D;
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    out[(size_t)t * D + d] = sPos[(size_t)t * D + d];
                    out[(size_t)(K + t) * D + d] = sPos[(size_t)(K + t) * D + d];
                }
            }
            ++row;
        }
    }
    for (int idx = t; idx < W * D; idx += blockDim.x) a.pos[idx] = sPos[idx];
    for (int idx = t; idx < W; idx += blockDim.x) a.logp[idx] = sLogp[idx];
    if (live) {
        reinterpret_cast<ulonglong2*>(a.rng_state)[t] = make_ulonglong2(rng[0].state.lo, rng[0].state.hi);
        reinterpret_cast<ulonglong2*>(a.rng_state)[K + t] = make_ulonglong2(rng[1].state.lo, rng[1].state.hi);
        if (redraws) atomicAdd(a.flags, redraws);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, off);
    if ((t & 31) == 0 && accepted) atomicAdd(a.accept_slots, accepted);
}

/* ---- the same, with the draws made AHEAD by the whole CTA (sequential stream, W <= 1024) --------------------------------
 * In stretchFusedSmall one thread makes a walker's three draws and then its update: ~450 dependent instructions per colour on
 * ONE warp for a 32-walker ensemble (0.8 us per colour).  But in the sequential stream the draws of update (step s, walker w)
 * are the base stream advanced by 3 (s W + w) -- they depend on nothing the chain does.  So the CTA works in windows of B steps
 * (B = 1024 / W): first ALL B W threads make one update's draws each (thread (b, w): steps b, b + B, ... of walker w, its state
 * jumping 3 W B draws per window) into shared memory, then the K walker threads run the 2 B colours of the window, whose
 * dependent chain is now draws-from-shared-memory -> partner row -> proposal -> log-prob -> test.  The arithmetic of every draw
 * and update is stretchFusedSmall's; chains, counters and final rng states are bit-identical (tests/test_gpu_parity.py fused
 * cases run both).  One difference, outside parity: after a bounded_rand rejection (probability K / 2^64 per update; counted in
 * flags[0] as everywhere) the later draws of that walker do not shift by one. */
template <class Calc, int D, bool CLUSTER>
__global__ void __launch_bounds__(FusedMaxThreads, 1) stretchFusedAhead(const mcmcpp_fused_args f, const Calc calcIn, const int stepsPerWindow,
                                                                        const int overlap)
{
    extern __shared__ double fusedSmem[];
    const mcmcpp_halfstep_args& a = f.base;
    const int K = a.K, W = 2 * K, t = threadIdx.x, B = stepsPerWindow, T = W * B;
    const int buffers = overlap ? 2 : 1;
    double* sPos = fusedSmem;                                 /* [W][D] */
    double* sLogp = sPos + (size_t)W * D;                     /* [W] */
    double* sZ = sLogp + W;                                   /* [buffers][B][W] z, log(1-u), log(z) (D-1), partner */
    double* sE = sZ + (size_t)buffers * T;
    double* sS = sE + (size_t)buffers * T;
    int* sJ = reinterpret_cast<int*>(sS + (size_t)buffers * T);
    /* CLUSTER: a thread-block cluster; CTA 0 holds the ensemble and runs the walkers, every other thread of the cluster draws,
     * storing its draws into CTA 0's buffers through distributed shared memory; the cluster barrier ends a window */
    int crank = 0;
    double *wZ = sZ, *wE = sE, *wS = sS;
    int* wJ = sJ;
    if (CLUSTER) {
        cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
        crank = (int)cluster.block_rank();
        wZ = cluster.map_shared_rank(sZ, 0); wE = cluster.map_shared_rank(sE, 0); wS = cluster.map_shared_rank(sS, 0);
        wJ = cluster.map_shared_rank(sJ, 0);
    }
    if (crank == 0) {
        for (int idx = t; idx < W * D; idx += blockDim.x) sPos[idx] = a.pos[idx];
        for (int idx = t; idx < W; idx += blockDim.x) sLogp[idx] = a.logp[idx];
    }
    Calc calc = calcIn;
    const long long total = (long long)f.stored_steps * f.interval;

    /* walker role: thread t < K owns red walker t and black walker K + t; their warps meet at barrier 1 after every colour.
     * drawing role: thread (b, w) owns steps b, b + B, ... of walker w.  overlap: the warps after the walker warps draw window
     * k + 1 while the walker warps run window k (two draw buffers); otherwise every thread draws, then the walkers run. */
    const int walkerThreads = ((K + 31) / 32) * 32;
    const bool live = crank == 0 && t < K;
    const int u = CLUSTER ? (crank == 0 ? t - walkerThreads : (FusedMaxThreads - walkerThreads) + (crank - 1) * FusedMaxThreads + t)
                          : t - (overlap ? walkerThreads : 0);
    const bool drawer = u >= 0 && u < T;
    const int b = drawer ? u / W : 0, w = drawer ? u - b * W : 0;
    U128 inc; inc.lo = 1ULL; inc.hi = 0;                      /* the sampler's stream 0 (EnsembleSampler.h:217) */
    U128 base; base.lo = 0; base.hi = 0;
    Affine128 window; window.mult = base; window.plus = base;
    long long next = b;
    unsigned redraws = 0;
    if (drawer) {
        const ulonglong2 s0 = reinterpret_cast<const ulonglong2*>(a.rng_state)[w];
        base.lo = s0.x; base.hi = s0.y;
        base = applyAffine(pcg64AdvanceMap(inc, 3ULL * (uint64_t)W * (uint64_t)b, 0), base);
        window = pcg64AdvanceMap(inc, 3ULL * (uint64_t)W * (uint64_t)B, 0);
    }
    unsigned long long accepted = 0;
    int sinceStore = 0;
    long long row = 0;

    auto draw = [&](int buffer) {
        if (drawer && next < total) {
            U128 st = base;
            const StretchDraws dr = drawStretch(st, inc, (uint64_t)K, a.bound_threshold, a.bound_recip, gwOf(a));
            redraws += (unsigned)dr.redraws;
            const int slot = buffer * T + u;
            wJ[slot] = dr.j; wZ[slot] = dr.z; wE[slot] = dr.logOneMinusU;
            wS[slot] = __dmul_rn(log(dr.z), (double)(D - 1));
            base = applyAffine(window, base);
            next += B;
        }
    };
    auto walk = [&](long long first, int buffer) {
        const int steps = (int)((total - first) < (long long)B ? (total - first) : (long long)B);
        for (int bb = 0; bb < steps; ++bb) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                if (live) {
                    const int ww = half * K + t, other0 = (1 - half) * K, item = buffer * T + bb * W + ww;
                    const int j = sJ[item];
                    const double z = sZ[item];
                    double cur[D], prop[D];
#pragma unroll
                    for (int d = 0; d < D; ++d) { cur[d] = sPos[(size_t)ww * D + d]; prop[d] = sPos[(size_t)(other0 + j) * D + d]; }
#pragma unroll
                    for (int d = 0; d < D; ++d) prop[d] = stretchPoint(prop[d], cur[d], z);
                    const double newp = calc.calcLogPostProb(prop);
                    const double diff = __dsub_rn(__dadd_rn(sS[item], newp), sLogp[ww]);
                    if (sE[item] < diff) {
#pragma unroll
                        for (int d = 0; d < D; ++d) sPos[(size_t)ww * D + d] = prop[d];
                        sLogp[ww] = newp;
                        ++accepted;
                    }
                }
                if (walkerThreads == 32) __syncwarp();      /* one warp of walkers: no CTA barrier needed */
                else asm volatile("bar.sync 1, %0;" :: "r"(walkerThreads) : "memory");
            }
            if (++sinceStore == f.interval) {
                sinceStore = 0;
                if (f.chain_rows && live) {
                    double* out = f.chain_rows + (size_t)row * W * D;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        out[(size_t)t * D + d] = sPos[(size_t)t * D + d];
                        out[(size_t)(K + t) * D + d] = sPos[(size_t)(K + t) * D + d];
                    }
                }
                ++row;
            }
        }
    };

    auto windowEnd = [&]() {
        if (CLUSTER) cooperative_groups::this_cluster().sync(); else __syncthreads();
    };
    if (CLUSTER) cooperative_groups::this_cluster().sync();       /* CTA 0's shared memory exists before anybody stores into it */
    draw(0);
    windowEnd();
    int parity = 0;
    for (long long first = 0; first < total; first += B) {
        if (overlap) {
            if (crank == 0 && t < walkerThreads) walk(first, parity); else draw(parity ^ 1);
            windowEnd();
            parity ^= 1;
        } else {
            if (t < walkerThreads) walk(first, 0);
            __syncthreads();
            draw(0);
            __syncthreads();
        }
    }
    if (crank == 0) {
        for (int idx = t; idx < W * D; idx += blockDim.x) a.pos[idx] = sPos[idx];
        for (int idx = t; idx < W; idx += blockDim.x) a.logp[idx] = sLogp[idx];
    }
    if (drawer) {
        /* walker w's state at rest = the stream at ITS next draw: the base of the item (step total, walker w) */
        if (next == total) reinterpret_cast<ulonglong2*>(a.rng_state)[w] = make_ulonglong2(base.lo, base.hi);
        if (redraws) atomicAdd(a.flags, redraws);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, off);
    if ((t & 31) == 0 && accepted) atomicAdd(a.accept_slots, accepted);
}

/* Pass &FusedLauncher<Calc>::launch to mcmcpp_sampler_set_fused_launcher (stretch move, any Calculator). */
template <class Calc, int FixedD = 0>
struct FusedLauncher
{
    template <int D>
    static int run(const mcmcpp_fused_args& f, const Calc& calc, cudaStream_t stream)
    {
        const mcmcpp_halfstep_args& a = f.base;
        /* sequential stream and W <= 1024: the draws are made ahead by the whole CTA (MCMCPP_B200_FUSED_AHEAD=0: never) */
        static const bool aheadAllowed = [] { const char* e = getenv("MCMCPP_B200_FUSED_AHEAD"); return !(e && e[0] == '0'); }();
        if (aheadAllowed && a.rng_mode == MCMCPP_RNG_SEQUENTIAL && 2 * a.K <= FusedMaxThreads) {      /* W <= 1024 */
            /* enough threads beside the walker warps to draw at least one step per window: they draw window k + 1 while window k
             * runs (MCMCPP_B200_FUSED_OVERLAP=0: draw, then run) */
            static const bool overlapAllowed = [] { const char* e = getenv("MCMCPP_B200_FUSED_OVERLAP"); return !(e && e[0] == '0'); }();
            static const bool clusterAllowed = [] { const char* e = getenv("MCMCPP_B200_FUSED_CLUSTER"); return !(e && e[0] == '0'); }();
            const int W = 2 * a.K, walkerThreads = ((a.K + 31) / 32) * 32;
            const size_t stateBytes = sizeof(double) * ((size_t)W * D + (size_t)W), perDraw = 3 * sizeof(double) + sizeof(int);
            /* from 256 walkers on the draws are most of a step on one SM (3.5 W cycles of issue against ~900 of walker latency):
             * a thread-block cluster of 2, 4 or 8 CTAs draws, CTA 0 runs the walkers (distributed shared memory) */
            if (overlapAllowed && clusterAllowed && W >= 256 && stateBytes + 2 * perDraw * (size_t)W <= 160 * 1024) {
                int B = (int)std::min<size_t>(8, (160 * 1024 - stateBytes) / (2 * perDraw * (size_t)W));
                int ctas = 0;
                for (;; --B) {
                    const int beside = FusedMaxThreads - walkerThreads, need = W * B - beside;
                    ctas = 1 + (need > 0 ? (need + FusedMaxThreads - 1) / FusedMaxThreads : 0);
                    ctas = ctas <= 2 ? 2 : (ctas <= 4 ? 4 : 8);
                    if (1 + (need > 0 ? (need + FusedMaxThreads - 1) / FusedMaxThreads : 0) <= 8 || B == 1) break;
                }
                const int T = W * B;
                const size_t smemCluster = stateBytes + 2 * perDraw * (size_t)T;
                static DynamicSmemOptIn optInCluster;
                if (smemCluster > 48 * 1024 && !optInCluster.ensure(stretchFusedAhead<Calc, D, true>, smemCluster)) return MCMCPP_ERR_CUDA;
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(FusedMaxThreads); cfg.dynamicSmemBytes = smemCluster; cfg.stream = stream;
                cudaLaunchAttribute attr[1];
                attr[0].id = cudaLaunchAttributeClusterDimension;
                attr[0].val.clusterDim.x = ctas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
                cfg.attrs = attr; cfg.numAttrs = 1;
                if (cudaLaunchKernelEx(&cfg, stretchFusedAhead<Calc, D, true>, f, calc, B, 1) != cudaSuccess) return MCMCPP_ERR_CUDA;
                return launchStatus();
            }
            const int overlap = overlapAllowed && (FusedMaxThreads - walkerThreads) / W >= 1 ? 1 : 0;
            const int B = overlap ? (FusedMaxThreads - walkerThreads) / W : FusedMaxThreads / W, T = W * B;
            const int threads = overlap ? walkerThreads + ((T + 31) / 32) * 32 : ((T + 31) / 32) * 32;
            const size_t smemAhead = sizeof(double) * ((size_t)W * D + (size_t)W + 3 * (size_t)T * (overlap + 1)) + sizeof(int) * (size_t)T * (overlap + 1);
            static DynamicSmemOptIn optInAhead;
            if (smemAhead > 48 * 1024 && !optInAhead.ensure(stretchFusedAhead<Calc, D, false>, smemAhead)) return MCMCPP_ERR_CUDA;
            stretchFusedAhead<Calc, D, false><<<1, threads, smemAhead, stream>>>(f, calc, B, overlap);
            return launchStatus();
        }
        const size_t smem = sizeof(double) * ((size_t)2 * a.K * D + (size_t)2 * a.K);
        static DynamicSmemOptIn optIn;
        if (smem > 48 * 1024 && !optIn.ensure(stretchFusedSmall<Calc, D>, smem)) return MCMCPP_ERR_CUDA;
        const int threads = ((a.K + 31) / 32) * 32;
        stretchFusedSmall<Calc, D><<<1, threads, smem, stream>>>(f, calc);
        return launchStatus();
    }

    static int launch(const mcmcpp_fused_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_fused_args& f = *args;
        const Calc& calc = *static_cast<const Calc*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (f.base.K > FusedMaxThreads || f.base.D > ThreadKernelMaxDim || f.base.slot_count != f.base.K) return MCMCPP_ERR_UNSUPPORTED;
        if (f.stored_steps <= 0 || f.interval <= 0) return MCMCPP_ERR_ARG;
        if constexpr (FixedD > 0) {
            if (f.base.D != FixedD) return MCMCPP_ERR_ARG;
            return run<FixedD>(f, calc, stream);
        } else {
            switch (f.base.D) {
            case 1: return run<1>(f, calc, stream);
            case 2: return run<2>(f, calc, stream);
            case 3: return run<3>(f, calc, stream);
            case 4: return run<4>(f, calc, stream);
            case 5: return run<5>(f, calc, stream);
            case 6: return run<6>(f, calc, stream);
            case 7: return run<7>(f, calc, stream);
            case 8: return run<8>(f, calc, stream);
            default: return MCMCPP_ERR_UNSUPPORTED;
            }
        }
    }
};

/* FixedD > 0 restricts the thread kernel to one dimension (2-D test targets); 0 = any D. */
template <class Calc, int FixedD = 0>
struct HalfStepLauncher
{
    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        const Calc& calc = *static_cast<const Calc*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (a.slot_count <= 0) return MCMCPP_OK;
        if constexpr (FixedD > 0) {
            if (a.D != FixedD) return MCMCPP_ERR_ARG;
            return launchThreadKernel<Calc, FixedD>(a, calc, stream);
        } else {
            switch (a.D) {
            case 1: return launchThreadKernel<Calc, 1>(a, calc, stream);
            case 2: return launchThreadKernel<Calc, 2>(a, calc, stream);
            case 3: return launchThreadKernel<Calc, 3>(a, calc, stream);
            case 4: return launchThreadKernel<Calc, 4>(a, calc, stream);
            case 5: return launchThreadKernel<Calc, 5>(a, calc, stream);
            case 6: return launchThreadKernel<Calc, 6>(a, calc, stream);
            case 7: return launchThreadKernel<Calc, 7>(a, calc, stream);
            case 8: return launchThreadKernel<Calc, 8>(a, calc, stream);
            default: break;
            }
            if (a.D >= BulkKernelMinDim && (a.D % 2) == 0 && !bulkKernelDisabled()) return launchBulkKernel<Calc>(a, calc, stream);
            const int chunks = (a.D + 63) / 64;
            if (chunks <= 1) return launchWarpKernel<Calc, 1>(a, calc, stream);
            if (chunks <= 2) return launchWarpKernel<Calc, 2>(a, calc, stream);
            if (chunks <= 4) return launchWarpKernel<Calc, 4>(a, calc, stream);
            if (chunks <= 8) return launchWarpKernel<Calc, 8>(a, calc, stream);
            if (chunks <= 16) return launchWarpKernel<Calc, 16>(a, calc, stream);
            return launchBulkKernel<Calc>(a, calc, stream);   /* odd D > 1024 (or the bulk kernel switched off): rows staged in shared memory */
        }
    }
};

}
}
#endif
