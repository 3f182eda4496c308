/*
 * include/mcmcpp_b200/walk_move_kernels.cuh -- MCMC::Mover::WalkMove (MCMCpp/Movers/WalkMove.h:95-186) behind the same
 * half-step skeleton.
 *
 * Per update: S distinct partner indices by selection sampling (one uniform per examined walker, :117-136); S standard
 * normals r_j (libstdc++ std::normal_distribution: Marsaglia polar, the second value of a pair is kept for the next
 * call -- that saved value is per-engine state, args.mover_state); proposal_i = cur_i + (a_i - (b (c_i / S))) with
 * a_i = sum_j r_j v_ji, c_i = sum_j v_ji, b = sum_j r_j, all in order j (:139-183); accept iff log(1-u) < newp - logp.
 * Draw counts vary, so: one pcg64 stream per walker only (MCMCPP_RNG_PER_WALKER).  One warp per walker: lane 0 owns the
 * engine (selection, normals, the accept draw), the lanes share the parameters.  The normals go through the device's
 * log(), which can differ from glibc's in the last bit, so positions carry the tolerance tier (<= 1e-12 relative).
 */
#ifndef MCMCPP_B200_WALK_MOVE_KERNELS_CUH
#define MCMCPP_B200_WALK_MOVE_KERNELS_CUH

#include "mcmcpp_b200/stretch_kernels.cuh"

namespace MCMC
{
namespace Device
{

enum { WalkMoveWarps = 4, WalkMoveMaxPoints = 64 };

template <class Calc>
struct WalkMoveBlob                      /* the by-value "calculator" of mcmcpp_sampler_create for this mover */
{
    Calc calc;
    int numPoints;                       /* WalkMove(numParams, prngInit, calc, numSamples), WalkMove.h:58 */
};

/* std::normal_distribution<double>::operator() of libstdc++ (bits/random.tcc), mean 0, stddev 1 */
__device__ __forceinline__ double walkStdNormal(U128& state, U128 inc, double& saved, double& available)
{
    double ret;
    if (available != 0.0) { available = 0.0; ret = saved; }
    else {
        double x, y, r2;
        do {
            x = __dsub_rn(__dmul_rn(2.0, unitFromU64(pcg64Next(state, inc))), 1.0);
            y = __dsub_rn(__dmul_rn(2.0, unitFromU64(pcg64Next(state, inc))), 1.0);
            r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        } while (r2 > 1.0 || r2 == 0.0);
        const double mult = sqrt(__ddiv_rn(__dmul_rn(-2.0, log(r2)), r2));
        saved = __dmul_rn(x, mult); available = 1.0;
        ret = __dmul_rn(y, mult);
    }
    return __dadd_rn(__dmul_rn(ret, 1.0), 0.0);
}

template <class Calc, bool TRACE>
__global__ void __launch_bounds__(WalkMoveWarps * 32) walkMoveWarp(const mcmcpp_halfstep_args a, const WalkMoveBlob<Calc> blob)
{
    extern __shared__ double wmSmem[];                        /* [warps][D] proposals, then [warps][MaxPoints] normals + indices */
    const int lane = threadIdx.x & 31, warpInBlock = threadIdx.x >> 5;
    const int i = blockIdx.x * WalkMoveWarps + warpInBlock;
    if (i >= a.slot_count) return;                            /* warp-uniform; warps never synchronise with each other */
    Calc calc = blob.calc;
    const int D = a.D, S = blob.numPoints;
    double* stage = wmSmem + (size_t)warpInBlock * D;
    double* rnd = wmSmem + (size_t)WalkMoveWarps * D + (size_t)warpInBlock * WalkMoveMaxPoints;
    int* idx = reinterpret_cast<int*>(wmSmem + (size_t)WalkMoveWarps * D + (size_t)WalkMoveWarps * WalkMoveMaxPoints) + warpInBlock * WalkMoveMaxPoints;
    const int slot = a.slot_begin + i;
    const int w = a.half * a.K + slot;
    const int other0 = (1 - a.half) * a.K;
    const double ptCount = (double)S;
    WalkerRng rng;
    double saved = 0.0, available = 0.0, sumR = 0.0;
    if (lane == 0) {
        rng = loadRng(a, w);
        saved = a.mover_state[2 * (size_t)w]; available = a.mover_state[2 * (size_t)w + 1];
        int numSelected = 0, examined = 0;                    /* selectWalkers */
        while (numSelected < S) {
            const double u = unitFromU64(pcg64Next(rng.state, rng.inc));
            if (__dmul_rn((double)(a.K - examined), u) >= __dsub_rn(ptCount, (double)numSelected)) ++examined;
            else { idx[numSelected] = examined; ++examined; ++numSelected; }
        }
        for (int j = 0; j < S; ++j) {
            const double r = walkStdNormal(rng.state, rng.inc, saved, available);
            rnd[j] = r;
            sumR = __dadd_rn(sumR, r);
        }
    }
    sumR = __shfl_sync(0xffffffffu, sumR, 0);
    __syncwarp();
    double* ownRow = a.pos + (size_t)w * D;
    for (int d = lane; d < D; d += 32) {
        double i0 = 0.0, i2 = 0.0;
        for (int j = 0; j < S; ++j) {
            const double value = __ldg(partnerBase(a, idx[j]) + (size_t)(other0 + idx[j]) * D + d);
            i0 = __dadd_rn(i0, __dmul_rn(rnd[j], value));
            i2 = __dadd_rn(i2, value);
        }
        stage[d] = __dadd_rn(ownRow[d], __dsub_rn(i0, __dmul_rn(sumR, __ddiv_rn(i2, ptCount))));
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
    const double diff = __dsub_rn(newp, a.logp[w]);
    double e = 0.0;
    if (lane == 0) e = log(1.0 - unitFromU64(pcg64Next(rng.state, rng.inc)));
    e = __shfl_sync(0xffffffffu, e, 0);
    const bool accepted = e < diff;
    if (a.chain_row) {
        double* chainRow = a.chain_row + (size_t)w * D;
        for (int d = lane; d < D; d += 32) chainRow[d] = accepted ? stage[d] : ownRow[d];
    }
    if (accepted) for (int d = lane; d < D; d += 32) ownRow[d] = stage[d];
    if (lane == 0) {
        reinterpret_cast<ulonglong2*>(a.rng_state)[w] = make_ulonglong2(rng.state.lo, rng.state.hi);
        a.mover_state[2 * (size_t)w] = saved; a.mover_state[2 * (size_t)w + 1] = available;
        if (accepted) { a.logp[w] = newp; atomicAdd(a.accept_slots + blockIdx.x, 1ULL); }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace t;
                t.j = idx[0]; t.accept = accepted ? 1 : 0; t.z = rnd[0]; t.e = e; t.newp = newp; t.diff = diff;
                a.trace[slot] = t;
            }
        }
    }
}

/* ---- small ensembles: the whole runMCMC loop in one launch (mcmcpp_fused_args), as stretchFusedSmall ------------------------
 * One CTA; thread t owns red walker t and black walker K + t (engines and the normal distribution's saved value in registers
 * for the whole call); the ensemble lives in shared memory; __syncthreads() is the colour boundary.  Per update the draws and
 * the arithmetic of walkMoveWarp in its order (selection, S normals and their sum in order j; per parameter the two sums over
 * j; the Calculator's own evaluation order), so the results are the launch-per-colour path's bit for bit. */
template <class Calc, int D>
__global__ void __launch_bounds__(FusedMaxThreads, 1) walkMoveFusedSmall(const mcmcpp_fused_args f, const WalkMoveBlob<Calc> blob)
{
    extern __shared__ double fusedSmem[];
    const mcmcpp_halfstep_args& a = f.base;
    const int K = a.K, W = 2 * K, t = threadIdx.x, S = blob.numPoints;
    double* sPos = fusedSmem;                                 /* [W][D] */
    double* sLogp = fusedSmem + (size_t)W * D;                /* [W] */
    for (int idx = t; idx < W * D; idx += blockDim.x) sPos[idx] = a.pos[idx];
    for (int idx = t; idx < W; idx += blockDim.x) sLogp[idx] = a.logp[idx];
    Calc calc = blob.calc;
    const bool live = t < K;
    const double ptCount = (double)S;
    WalkerRng rng[2];
    double saved[2] = {0.0, 0.0}, available[2] = {0.0, 0.0};
    if (live) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            rng[h] = loadRng(a, h * K + t);
            saved[h] = a.mover_state[2 * (size_t)(h * K + t)]; available[h] = a.mover_state[2 * (size_t)(h * K + t) + 1];
        }
    }
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
                int sel[WalkMoveMaxPoints];
                int numSelected = 0, examined = 0;                    /* selectWalkers */
                while (numSelected < S) {
                    const double u = unitFromU64(pcg64Next(rng[half].state, rng[half].inc));
                    if (__dmul_rn((double)(K - examined), u) >= __dsub_rn(ptCount, (double)numSelected)) ++examined;
                    else { sel[numSelected] = examined; ++examined; ++numSelected; }
                }
                double i0[D], i2[D], sumR = 0.0;
#pragma unroll
                for (int d = 0; d < D; ++d) { i0[d] = 0.0; i2[d] = 0.0; }
                for (int j = 0; j < S; ++j) {
                    const double r = walkStdNormal(rng[half].state, rng[half].inc, saved[half], available[half]);
                    sumR = __dadd_rn(sumR, r);
                    const double* partner = sPos + (size_t)(other0 + sel[j]) * D;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        i0[d] = __dadd_rn(i0[d], __dmul_rn(r, partner[d]));
                        i2[d] = __dadd_rn(i2[d], partner[d]);
                    }
                }
                double prop[D];
#pragma unroll
                for (int d = 0; d < D; ++d) prop[d] = __dadd_rn(sPos[(size_t)w * D + d], __dsub_rn(i0[d], __dmul_rn(sumR, __ddiv_rn(i2[d], ptCount))));
                const double newp = calc.calcLogPostProb(prop);
                const double diff = __dsub_rn(newp, sLogp[w]);
                const double e = log(1.0 - unitFromU64(pcg64Next(rng[half].state, rng[half].inc)));
                if (e < diff) {
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
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            reinterpret_cast<ulonglong2*>(a.rng_state)[h * K + t] = make_ulonglong2(rng[h].state.lo, rng[h].state.hi);
            a.mover_state[2 * (size_t)(h * K + t)] = saved[h]; a.mover_state[2 * (size_t)(h * K + t) + 1] = available[h];
        }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, off);
    if ((t & 31) == 0 && accepted) atomicAdd(a.accept_slots, accepted);
}

/* Pass &WalkMoveFusedLauncher<Calc>::launch to mcmcpp_sampler_set_fused_launcher (the sampler's calculator is a WalkMoveBlob<Calc>). */
template <class Calc, int FixedD = 0>
struct WalkMoveFusedLauncher
{
    template <int D>
    static int run(const mcmcpp_fused_args& f, const WalkMoveBlob<Calc>& blob, cudaStream_t stream)
    {
        const mcmcpp_halfstep_args& a = f.base;
        const size_t smem = sizeof(double) * ((size_t)2 * a.K * D + (size_t)2 * a.K);
        static DynamicSmemOptIn optIn;
        if (smem > 48 * 1024 && !optIn.ensure(walkMoveFusedSmall<Calc, D>, smem)) return MCMCPP_ERR_CUDA;
        walkMoveFusedSmall<Calc, D><<<1, ((a.K + 31) / 32) * 32, smem, stream>>>(f, blob);
        return launchStatus();
    }

    static int launch(const mcmcpp_fused_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_fused_args& f = *args;
        const WalkMoveBlob<Calc>& blob = *static_cast<const WalkMoveBlob<Calc>*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (f.base.rng_mode != MCMCPP_RNG_PER_WALKER || !f.base.mover_state) return MCMCPP_ERR_UNSUPPORTED;
        if (blob.numPoints < 1 || blob.numPoints > WalkMoveMaxPoints || blob.numPoints > f.base.K) return MCMCPP_ERR_ARG;
        if (f.base.K > FusedMaxThreads || f.base.D > ThreadKernelMaxDim || f.base.slot_count != f.base.K) return MCMCPP_ERR_UNSUPPORTED;
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

/* Launcher: pass &WalkMoveLauncher<Calc>::launch and a WalkMoveBlob<Calc> to mcmcpp_sampler_create (rng_mode PER_WALKER;
 * the runtime must have been asked for mover state: mcmcpp_config.mover_state_doubles = 2). */
template <class Calc>
struct WalkMoveLauncher
{
    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        const WalkMoveBlob<Calc>& blob = *static_cast<const WalkMoveBlob<Calc>*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        if (a.rng_mode != MCMCPP_RNG_PER_WALKER || !a.mover_state) return MCMCPP_ERR_UNSUPPORTED;
        if (blob.numPoints < 1 || blob.numPoints > WalkMoveMaxPoints || blob.numPoints > a.K) return MCMCPP_ERR_ARG;   /* WalkMove.h:119 asserts */
        if (a.slot_count <= 0) return MCMCPP_OK;
        if (a.D > 4096) return MCMCPP_ERR_UNSUPPORTED;
        const int grid = (a.slot_count + WalkMoveWarps - 1) / WalkMoveWarps;
        if (grid > a.max_grid) return MCMCPP_ERR_ARG;
        const size_t smem = sizeof(double) * ((size_t)WalkMoveWarps * a.D + (size_t)WalkMoveWarps * WalkMoveMaxPoints) + sizeof(int) * WalkMoveWarps * WalkMoveMaxPoints;
        if (smem > 48 * 1024) {
            cudaFuncSetAttribute(walkMoveWarp<Calc, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaFuncSetAttribute(walkMoveWarp<Calc, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        }
        if (a.trace) walkMoveWarp<Calc, true><<<grid, WalkMoveWarps * 32, smem, stream>>>(a, blob);
        else walkMoveWarp<Calc, false><<<grid, WalkMoveWarps * 32, smem, stream>>>(a, blob);
        return launchStatus();
    }
};

}
}
#endif
