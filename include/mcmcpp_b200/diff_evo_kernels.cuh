/*
 * include/mcmcpp_b200/diff_evo_kernels.cuh -- MCMC::Mover::DifferentialEvolution (MCMCpp/Movers/DifferentialEvolution.h:
 * 113-142) behind the same half-step skeleton as the stretch move.
 *
 * Per update: ind1 = bounded(K); ind2 = bounded(K) until != ind1; proposal_d = cur_d + gamma (w1_d - w2_d) +
 * (lowEdge + u_d width) with one uniform per parameter; accept iff (newp - logp) > log(1 - u).  The number of draws per
 * update is not fixed (the ind2 loop), so the sequential sampler's single stream cannot be indexed by jump-ahead:
 * the device offers this mover with one pcg64 stream per walker only (MCMCPP_RNG_PER_WALKER; the reference's
 * per-thread discipline, RedBlkUpdater.h:86).  Arithmetic order and rounding are the reference's.
 *   D <= 8   diffEvoThread: one thread per walker
 *   else     diffEvoWarp:   one warp per walker; the D uniforms of an update are consecutive draws of the walker's
 *            stream, so lane l positions a copy of the engine at draw 64c + 2l by an affine jump
 *            (state -> M^n state + inc G_n; the tables of M^n and G_n = 1 + M + ... + M^(n-1) do not depend on the
 *            stream) and every lane makes its own two draws per 64-element chunk.
 */
#ifndef MCMCPP_B200_DIFF_EVO_KERNELS_CUH
#define MCMCPP_B200_DIFF_EVO_KERNELS_CUH

#include <cmath>
#include "mcmcpp_b200/stretch_kernels.cuh"

namespace MCMC
{
namespace Device
{

struct DiffEvoParams
{
    double gamma, lowEdge, width;       /* DifferentialEvolution.h:58,150-151 */
};

template <class Calc>
struct DiffEvoBlob                       /* the by-value "calculator" of mcmcpp_sampler_create for this mover */
{
    Calc calc;
    DiffEvoParams mv;
};

/* jump n draws ahead: state -> mult * state + inc * geom.  [l] n = 2l (l < 32); [32] n = 62; [33] n = D */
struct DiffEvoJumps
{
    uint64_t mult[34][2], geom[34][2];
};

inline DiffEvoJumps makeDiffEvoJumps(int D)
{
    DiffEvoJumps t;
    U128 one; one.lo = 1; one.hi = 0;
    for (int k = 0; k < 34; ++k) {
        const uint64_t n = k < 32 ? 2ULL * (uint64_t)k : (k == 32 ? 62ULL : (uint64_t)D);
        Affine128 m = pcg64AdvanceMap(one, n, 0);          /* inc = 1: plus = G_n */
        t.mult[k][0] = m.mult.lo; t.mult[k][1] = m.mult.hi;
        t.geom[k][0] = m.plus.lo; t.geom[k][1] = m.plus.hi;
    }
    return t;
}

__device__ __forceinline__ U128 deJump(const DiffEvoJumps& t, int k, U128 state, U128 inc)
{
    U128 m, g;
    m.lo = t.mult[k][0]; m.hi = t.mult[k][1];
    g.lo = t.geom[k][0]; g.hi = t.geom[k][1];
    return add128(mul128(m, state), mul128(inc, g));
}

__device__ __forceinline__ int deBounded(U128& state, U128 inc, const mcmcpp_halfstep_args& a)
{
    uint64_t r = pcg64Next(state, inc);
    while (r < a.bound_threshold) r = pcg64Next(state, inc);          /* pcg_extras.hpp:514-525 */
    return (int)modByRecip(r, (uint64_t)a.K, a.bound_recip);
}

__device__ __forceinline__ double deProposal(double cur, double w1, double w2, double u, const DiffEvoParams& mv)
{
    /* currState[i] + gamma*(w1State[i] - w2State[i]) + (lowEdge + u*width), left to right, no contraction */
    return __dadd_rn(__dadd_rn(cur, __dmul_rn(mv.gamma, __dsub_rn(w1, w2))), __dadd_rn(mv.lowEdge, __dmul_rn(u, mv.width)));
}

template <class Calc, int D, bool TRACE>
__global__ void __launch_bounds__(ThreadKernelThreads, (D <= 2 ? 12 : 1)) diffEvoThread(const mcmcpp_halfstep_args a, const DiffEvoBlob<Calc> blob)
{
    const int i = blockIdx.x * ThreadKernelThreads + threadIdx.x;
    int accepted = 0;
    if (i < a.slot_count) {
        Calc calc = blob.calc;
        const int slot = a.slot_begin + i;
        const int w = a.half * a.K + slot;
        const int other0 = (1 - a.half) * a.K;
        WalkerRng rng = loadRng(a, w);
        double cur[D], w1[D], w2[D], prop[D];
        double* ownRow = a.pos + (size_t)w * D;
        loadRow<D>(ownRow, cur);                              /* own row, log-prob: independent of the draws, issued first */
        const double oldp = a.logp[w];
        const int ind1 = deBounded(rng.state, rng.inc, a);
        int ind2;
        do { ind2 = deBounded(rng.state, rng.inc, a); } while (ind2 == ind1);
        /* programmatic dependent launch (see stretchHalfStepThread): only the partner rows come from the launch before */
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        loadRowReadOnly<D>(partnerBase(a, ind1) + (size_t)(other0 + ind1) * D, w1);
        loadRowReadOnly<D>(partnerBase(a, ind2) + (size_t)(other0 + ind2) * D, w2);
#pragma unroll
        for (int d = 0; d < D; ++d) prop[d] = deProposal(cur[d], w1[d], w2[d], unitFromU64(pcg64Next(rng.state, rng.inc)), blob.mv);
        const double newp = calc.calcLogPostProb(prop);
        const double diff = __dsub_rn(newp, oldp);
        const double e = log(1.0 - unitFromU64(pcg64Next(rng.state, rng.inc)));
        reinterpret_cast<ulonglong2*>(a.rng_state)[w] = make_ulonglong2(rng.state.lo, rng.state.hi);
        accepted = diff > e ? 1 : 0;
        if (accepted) { storeRow<D>(ownRow, prop); a.logp[w] = newp; }
        if (a.chain_row) {
            double* chainRow = a.chain_row + (size_t)w * D;
            if (accepted) storeRow<D>(chainRow, prop); else storeRow<D>(chainRow, cur);
        }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = ind1; t.accept = accepted; t.z = (double)ind2; t.e = e; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, accepted != 0);
    if ((threadIdx.x & 31) == 0 && ballot) atomicAdd(a.accept_slots + blockIdx.x, (unsigned long long)__popc(ballot));
}

/* ---- small ensembles: the whole runMCMC loop in one launch (mcmcpp_fused_args), as stretchFusedSmall ------------------------
 * One CTA; thread t owns red walker t and black walker K + t (both engines in registers for the whole call); the ensemble
 * lives in shared memory; __syncthreads() is the colour boundary.  Per update exactly the arithmetic of diffEvoThread, so
 * chains, counters and rng states are the ones the launch-per-colour path produces (tests/test_gpu_movers.py runs both). */
template <class Calc, int D>
__global__ void __launch_bounds__(FusedMaxThreads, 1) diffEvoFusedSmall(const mcmcpp_fused_args f, const DiffEvoBlob<Calc> blob)
{
    extern __shared__ double fusedSmem[];
    const mcmcpp_halfstep_args& a = f.base;
    const int K = a.K, W = 2 * K, t = threadIdx.x;
    double* sPos = fusedSmem;                                 /* [W][D] */
    double* sLogp = fusedSmem + (size_t)W * D;                /* [W] */
    for (int idx = t; idx < W * D; idx += blockDim.x) sPos[idx] = a.pos[idx];
    for (int idx = t; idx < W; idx += blockDim.x) sLogp[idx] = a.logp[idx];
    Calc calc = blob.calc;
    const bool live = t < K;
    WalkerRng rng[2];
    if (live) { rng[0] = loadRng(a, t); rng[1] = loadRng(a, K + t); }
    unsigned long long accepted = 0;
    __syncthreads();
    const long long total = (long long)f.stored_steps * f.interval;
    int sinceStore = 0;
    long long row = 0;
    for (long long step = 0; step < total; ++step) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            if (live) {
                const int w = half * K + t, other0 = (1 - half) * K;
                const int ind1 = deBounded(rng[half].state, rng[half].inc, a);
                int ind2;
                do { ind2 = deBounded(rng[half].state, rng[half].inc, a); } while (ind2 == ind1);
                double prop[D];
#pragma unroll
                for (int d = 0; d < D; ++d)
                    prop[d] = deProposal(sPos[(size_t)w * D + d], sPos[(size_t)(other0 + ind1) * D + d], sPos[(size_t)(other0 + ind2) * D + d],
                                         unitFromU64(pcg64Next(rng[half].state, rng[half].inc)), blob.mv);
                const double newp = calc.calcLogPostProb(prop);
                const double diff = __dsub_rn(newp, sLogp[w]);
                const double e = log(1.0 - unitFromU64(pcg64Next(rng[half].state, rng[half].inc)));
                if (diff > e) {
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
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, off);
    if ((t & 31) == 0 && accepted) atomicAdd(a.accept_slots, accepted);
}

/* Pass &DiffEvoFusedLauncher<Calc>::launch to mcmcpp_sampler_set_fused_launcher (the sampler's calculator is a DiffEvoBlob<Calc>). */
template <class Calc, int FixedD = 0>
struct DiffEvoFusedLauncher
{
    template <int D>
    static int run(const mcmcpp_fused_args& f, const DiffEvoBlob<Calc>& blob, cudaStream_t stream)
    {
        const mcmcpp_halfstep_args& a = f.base;
        const size_t smem = sizeof(double) * ((size_t)2 * a.K * D + (size_t)2 * a.K);
        static DynamicSmemOptIn optIn;
        if (smem > 48 * 1024 && !optIn.ensure(diffEvoFusedSmall<Calc, D>, smem)) return MCMCPP_ERR_CUDA;
        diffEvoFusedSmall<Calc, D><<<1, ((a.K + 31) / 32) * 32, smem, stream>>>(f, blob);
        return launchStatus();
    }

    static int launch(const mcmcpp_fused_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_fused_args& f = *args;
        const DiffEvoBlob<Calc>& blob = *static_cast<const DiffEvoBlob<Calc>*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (f.base.rng_mode != MCMCPP_RNG_PER_WALKER) return MCMCPP_ERR_UNSUPPORTED;
        if (f.base.K > FusedMaxThreads || f.base.K < 2 || f.base.D > ThreadKernelMaxDim || f.base.slot_count != f.base.K) return MCMCPP_ERR_UNSUPPORTED;
        if (f.stored_steps <= 0 || f.interval <= 0) return MCMCPP_ERR_ARG;
        if constexpr (FixedD > 0) {
            if (f.base.D != FixedD) return MCMCPP_ERR_ARG;
            return run<FixedD>(f, blob, stream);
        } else {
            switch (f.base.D) {
            case 1: return run<1>(f, blob, stream);
            case 2: return run<2>(f, blob, stream);
            case 3: return run<3>(f, blob, stream);
            case 4: return run<4>(f, blob, stream);
            case 5: return run<5>(f, blob, stream);
            case 6: return run<6>(f, blob, stream);
            case 7: return run<7>(f, blob, stream);
            case 8: return run<8>(f, blob, stream);
            default: return MCMCPP_ERR_UNSUPPORTED;
            }
        }
    }
};

enum { DiffEvoWarps = 4 };

template <class Calc, bool TRACE>
__global__ void __launch_bounds__(DiffEvoWarps * 32) diffEvoWarp(const mcmcpp_halfstep_args a, const DiffEvoBlob<Calc> blob, const DiffEvoJumps jumps)
{
    extern __shared__ double deStage[];                       /* [warps][D] proposals */
    const int lane = threadIdx.x & 31, warpInBlock = threadIdx.x >> 5;
    const int i = blockIdx.x * DiffEvoWarps + warpInBlock;
    if (i >= a.slot_count) return;                            /* warp-uniform; warps never synchronise with each other */
    Calc calc = blob.calc;
    const int D = a.D;
    double* stage = deStage + (size_t)warpInBlock * D;
    const int slot = a.slot_begin + i;
    const int w = a.half * a.K + slot;
    const int other0 = (1 - a.half) * a.K;
    WalkerRng rng = loadRng(a, w);                            /* every lane replays the two index draws */
    const int ind1 = deBounded(rng.state, rng.inc, a);
    int ind2;
    do { ind2 = deBounded(rng.state, rng.inc, a); } while (ind2 == ind1);
    double* ownRow = a.pos + (size_t)w * D;
    const double* w1Row = partnerBase(a, ind1) + (size_t)(other0 + ind1) * D;
    const double* w2Row = partnerBase(a, ind2) + (size_t)(other0 + ind2) * D;
    /* lane l: draws 64c + 2l and 64c + 2l + 1 of the D uniforms */
    U128 ls = deJump(jumps, lane, rng.state, rng.inc);
    for (int c = 0; c * 64 < D; ++c) {
        const int d = 64 * c + 2 * lane;
        const double u0 = unitFromU64(pcg64Next(ls, rng.inc));
        const double u1 = unitFromU64(pcg64Next(ls, rng.inc));
        if (d < D) stage[d] = deProposal(ownRow[d], __ldg(w1Row + d), __ldg(w2Row + d), u0, blob.mv);
        if (d + 1 < D) stage[d + 1] = deProposal(ownRow[d + 1], __ldg(w1Row + d + 1), __ldg(w2Row + d + 1), u1, blob.mv);
        ls = deJump(jumps, 32, ls, rng.inc);                  /* 62 more: next chunk */
    }
    __syncwarp();
    double newp;
    if constexpr (IsSeparable<Calc>::value) {
        double p0 = 0.0, p1 = 0.0;
        for (int d = 2 * lane; d < D; d += 64) {
            p0 = __dadd_rn(p0, calc.term(stage[d], d));
            if (d + 1 < D) p1 = __dadd_rn(p1, calc.term(stage[d + 1], d + 1));
        }
        double s = __dadd_rn(p0, p1);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) s = __dadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
        newp = calc.finish(s);
    } else {
        double v = 0.0;
        if (lane == 0) v = calc.calcLogPostProb(stage);
        newp = __shfl_sync(0xffffffffu, v, 0);
    }
    /* the accept draw follows the D uniforms */
    U128 fs = deJump(jumps, 33, rng.state, rng.inc);
    const double e = log(1.0 - unitFromU64(pcg64Next(fs, rng.inc)));
    const double diff = __dsub_rn(newp, a.logp[w]);
    const bool accepted = diff > e;
    if (a.chain_row) {
        double* chainRow = a.chain_row + (size_t)w * D;
        for (int d = lane; d < D; d += 32) chainRow[d] = accepted ? stage[d] : ownRow[d];
    }
    if (accepted) for (int d = lane; d < D; d += 32) ownRow[d] = stage[d];
    if (lane == 0) {
        reinterpret_cast<ulonglong2*>(a.rng_state)[w] = make_ulonglong2(fs.lo, fs.hi);
        if (accepted) { a.logp[w] = newp; atomicAdd(a.accept_slots + blockIdx.x, 1ULL); }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = ind1; t.accept = accepted ? 1 : 0; t.z = (double)ind2; t.e = e; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
}

/* Launcher: pass &DiffEvoLauncher<Calc>::launch and a DiffEvoBlob<Calc> to mcmcpp_sampler_create (rng_mode PER_WALKER). */
template <class Calc>
struct DiffEvoLauncher
{
    template <int D>
    static int thread(const mcmcpp_halfstep_args& a, const DiffEvoBlob<Calc>& blob, cudaStream_t stream)
    {
        const int grid = (a.slot_count + ThreadKernelThreads - 1) / ThreadKernelThreads;
        static const bool pdl = [] { const char* e = getenv("MCMCPP_B200_PDL"); return !(e && e[0] == '0'); }();
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(ThreadKernelThreads); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
        cudaError_t rc = a.trace ? cudaLaunchKernelEx(&cfg, diffEvoThread<Calc, D, true>, a, blob)
                                 : cudaLaunchKernelEx(&cfg, diffEvoThread<Calc, D, false>, a, blob);
        return rc == cudaSuccess ? launchStatus() : MCMCPP_ERR_CUDA;
    }

    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        const DiffEvoBlob<Calc>& blob = *static_cast<const DiffEvoBlob<Calc>*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (a.rng_mode != MCMCPP_RNG_PER_WALKER) return MCMCPP_ERR_UNSUPPORTED;   /* variable draws per update */
        if (a.slot_count <= 0) return MCMCPP_OK;
        switch (a.D) {
        case 1: return thread<1>(a, blob, stream);
        case 2: return thread<2>(a, blob, stream);
        case 3: return thread<3>(a, blob, stream);
        case 4: return thread<4>(a, blob, stream);
        case 5: return thread<5>(a, blob, stream);
        case 6: return thread<6>(a, blob, stream);
        case 7: return thread<7>(a, blob, stream);
        case 8: return thread<8>(a, blob, stream);
        default: break;
        }
        if (a.D > 4096) return MCMCPP_ERR_UNSUPPORTED;
        static thread_local int cachedD = -1;          /* per host thread: samplers of different D may launch from different threads */
        static thread_local DiffEvoJumps jumps;
        if (cachedD != a.D) { jumps = makeDiffEvoJumps(a.D); cachedD = a.D; }
        const int grid = (a.slot_count + DiffEvoWarps - 1) / DiffEvoWarps;
        if (grid > a.max_grid) return MCMCPP_ERR_ARG;
        const size_t smem = sizeof(double) * (size_t)DiffEvoWarps * a.D;
        if (smem > 48 * 1024) {
            cudaFuncSetAttribute(diffEvoWarp<Calc, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaFuncSetAttribute(diffEvoWarp<Calc, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        }
        if (a.trace) diffEvoWarp<Calc, true><<<grid, DiffEvoWarps * 32, smem, stream>>>(a, blob, jumps);
        else diffEvoWarp<Calc, false><<<grid, DiffEvoWarps * 32, smem, stream>>>(a, blob, jumps);
        return launchStatus();
    }
};

}
}
#endif
