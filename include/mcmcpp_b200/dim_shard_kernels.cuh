/*
 * include/mcmcpp_b200/dim_shard_kernels.cuh -- MCMCPP_SHARD_DIM: the stretch-move half-step with the PARAMETER
 * dimension sharded over the GPUs of one box (SURVEY.md s.8(e) alternative 2).
 *
 * Every rank holds all W walkers but only its columns [dim_begin, dim_begin + D) of each row, makes the SAME draws
 * (j, z, log(1-u)) from the sampler's pcg64 stream, builds its slice of every proposal
 * (MCMCpp/Movers/StretchMove.h:103-109) and the partial sum of a separable Calculator (term()/finish(), targets.h)
 * over its columns.  One colour = three launches:
 *
 *   dimShardDraw    one thread per walker: the three draws of the update (j, z, log(1-u)) and log(z)(D-1).
 *   dimShardSum     one-shot CTAs, one tile of NW walkers each: lane q of a warp issues walker q's two TMA bulk
 *                   copies (own slice, partner slice) onto the walker's mbarrier; the warp reduces the proposals to canonical
 *                   64-way partial sums; the CTA stores them, one line of NW partials per rank, straight into EVERY rank's
 *                   table over NVLink (peer-mapped stores).  It never waits for anything but its own copies.
 *   dimShardFinish  one warp per group of 32 walkers: each lane takes the `world` partial sums of its walker from this rank's
 *                   table -- waiting for the other ranks' if they are not there yet --, adds them in rank order (the same bits
 *                   on every rank, so every rank takes the same Metropolis decision, StretchMove.h:112-122), and the warp
 *                   rebuilds the accepted walkers' slices (8-13 % of them at D = 1000).  Decision and commit in one kernel,
 *                   no work list, no atomics (round 1: decide kernel + list + commit kernel, behind a rendezvous).
 * Who waits for whom: finishing kernels wait for summing kernels (of every rank); summing kernels wait for nobody.  So there
 * is no cycle, on any number of GPUs -- and none either when all the ranks live in one process on one GPU
 * (mcmcpp_sampler_run_group, tests/test_gpu_rank_group.py): the summing kernels of all ranks are launched before any
 * finishing kernel.
 *
 * Exchange protocol: a table slot holds either the all-ones NaN (`DimEmpty`, no arithmetic produces it) or a partial
 * sum.  Writers store 8-byte values (single, indivisible stores); the one reader of a slot spins until it is not
 * DimEmpty, takes the value and puts DimEmpty back.  No flag, no fence, no rendezvous kernel and no NCCL call.
 * The table has a half per colour; a rank cannot write a slot of colour c again before its reader has emptied it: it
 * would have to be two half-steps ahead, and it cannot finish the half-step in between without this rank's partials of
 * that half-step, which this rank only sends after the kernel that emptied the slot has completed (stream order).
 * A poll that sees nothing for `timeout_ns` sets flags[1] and gives up (mcmcpp_sampler_run then returns
 * MCMCPP_ERR_STATE): a dead peer cannot wedge the other GPUs.
 *
 * The log-posterior is summed slice by slice (canonical 64-way order inside a slice, rank order across), so for
 * world > 1 it carries the tolerance tier (<= 1e-12 relative); with world == 1 it is the single-GPU order, bit-exact.
 */
#ifndef MCMCPP_B200_DIM_SHARD_KERNELS_CUH
#define MCMCPP_B200_DIM_SHARD_KERNELS_CUH

#include <cstring>
#include "mcmcpp_b200/stretch_kernels.cuh"

namespace MCMC
{
namespace Device
{

enum { DimFusedWarps = 4, DimFusedThreads = DimFusedWarps * 32 };
static const unsigned long long DimEmpty = ~0ULL;            /* all-ones NaN: "no partial here yet" */

/* ---- phase 0: the draws of a colour, one THREAD per walker --------------------------------------------------------
 * ~600 dependent integer/FP64 instructions per walker (three 128-bit LCG steps, the stream skip, two logs).  Made
 * inside the copy kernel they would sit in front of its bulk copies; here they run at full occupancy and the copy
 * kernel only reads (j, z, log(1-u)). */
static __global__ void __launch_bounds__(256) dimShardDraw(const mcmcpp_halfstep_args a)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= a.K) return;
    const int w = a.half * a.K + i;
    WalkerRng rng = loadRng(a, w);
    StretchDraws dr = drawStretchFor(a, rng);
    storeRng(a, w, rng);
    if (dr.redraws) atomicAdd(a.flags, dr.redraws);
    a.dim_z[i] = dr.z; a.dim_e[i] = dr.logOneMinusU; a.dim_j[i] = dr.j;
    a.dim_s[i] = __dmul_rn(log(dr.z), (double)(a.D_total - 1));          /* log(z) (D - 1)  (StretchMove.h:110) */
}

__device__ __forceinline__ unsigned long long globalTimerNs()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

/* wait until the slot holds a partial sum, take it, leave the slot empty for the colour after next.  `v` is what an
 * earlier look at the slot saw.  Backs off between looks (the waiting warps must not load the L2 the copies go through). */
__device__ __forceinline__ double dimTakePartial(double* slotPtr, unsigned long long v, unsigned long long timeoutNs, int32_t* flags)
{
    volatile unsigned long long* slot = reinterpret_cast<volatile unsigned long long*>(slotPtr);
    if (v == DimEmpty) {
        const unsigned long long t0 = globalTimerNs();
        unsigned spins = 0, nap = 64;
        while ((v = *slot) == DimEmpty) {
            __nanosleep(nap);
            if (nap < 1024) nap <<= 1;
            if ((++spins & 255u) == 0 && globalTimerNs() - t0 > timeoutNs) {
                atomicExch(flags + 1, 1);                    /* surfaced by mcmcpp_sampler_run */
                return 0.0;
            }
        }
    }
    *slot = DimEmpty;
    return __longlong_as_double((long long)v);
}

/* ---- the summing kernel: one tile of NW = 4 WPW walkers per CTA ------------------------------------------------------
 * TMA = rows are 16-byte multiples (even slice): staged with cp.async.bulk; otherwise the warps read their walkers' rows
 * straight from global memory (odd slices: tests and the last rank of an odd row).
 * Shared memory: [NW walkers][own row | partner row] (rowStride bytes each), NW mbarriers, NW partials. */
template <class Calc, int WPW, bool TMA>
__global__ void __launch_bounds__(DimFusedThreads) dimShardSum(const mcmcpp_halfstep_args a, const Calc calcIn, const uint32_t rowStride, const int waitAll)
{
    extern __shared__ __align__(128) unsigned char dimSmem[];
    constexpr int NW = DimFusedWarps * WPW;                  /* walkers per tile */
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D, K = a.K;
    const int t = (int)blockIdx.x;
    const int own0 = a.half * K, other0 = (1 - a.half) * K;
    Calc calc = calcIn;
    const uint32_t rowBytes = (uint32_t)D * 8u;
    unsigned char* const rows = dimSmem + (size_t)(warp * WPW) * 2u * rowStride;
    unsigned char* const barBase = dimSmem + (TMA ? (size_t)NW * 2u * rowStride : 0);     /* no staged rows without TMA */
    double* const ctaPartial = reinterpret_cast<double*>(barBase + 8u * NW);
    if (TMA) {
        if (threadIdx.x < NW) mbarInit(smemAddr(barBase + 8u * threadIdx.x), 1);
        __syncthreads();
    }
    /* lane q < WPW starts walker q's two copies */
    const int iA = t * NW + warp * WPW + lane;               /* this lane's walker of tile t */
    int jA = 0;
    double zA = 1.0;
    if (lane < WPW && iA < K) {
        jA = __ldg(a.dim_j + iA);
        zA = __ldg(a.dim_z + iA);
        if (TMA) {
            const uint32_t bar = smemAddr(barBase + 8u * (warp * WPW + lane));
            const uint32_t dst = smemAddr(rows + (size_t)lane * 2u * rowStride);
            mbarExpectTx(bar, 2u * rowBytes);
            bulkLoad(dst, a.pos + (size_t)(own0 + iA) * D, rowBytes, bar);
            bulkLoad(dst + rowStride, a.pos + (size_t)(other0 + jA) * D, rowBytes, bar);
        }
    }
    /* where this thread's share of the tile's line goes (loaded while the copies fly): thread u < world * NW stores partial
     * u % NW into rank u / NW's table */
    const int sendK = (int)threadIdx.x % NW;
    double* const sendSlot = ((int)threadIdx.x < a.world * NW && t * NW + sendK < K)
                                 ? a.dim_partial[(int)threadIdx.x / NW] + ((size_t)a.half * a.world + a.rank) * K + (size_t)t * NW + sendK : nullptr;
    /* partial sums of the warp's walkers: all copies awaited first, then WPW independent dependency chains side by side */
    if (TMA && (waitAll & 1)) {
#pragma unroll
        for (int q = 0; q < WPW; ++q)
            if (t * NW + warp * WPW + q < K) mbarWait(smemAddr(barBase + 8u * (warp * WPW + q)), 0);
    }
    double s[WPW];
#pragma unroll
    for (int q = 0; q < WPW; ++q) {
        const int i = t * NW + warp * WPW + q;
        const double zq = __shfl_sync(0xffffffffu, zA, q);
        double p0 = 0.0, p1 = 0.0;
        if (i < K) {                                                  /* warp-uniform */
            if (TMA) {
                if (!(waitAll & 1)) mbarWait(smemAddr(barBase + 8u * (warp * WPW + q)), 0);
                const double* ownS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride);
                const double* selS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride + rowStride);
                for (int d = 2 * lane; d < D; d += 64) {
                    const double2 c = *reinterpret_cast<const double2*>(ownS + d);
                    const double2 sl = *reinterpret_cast<const double2*>(selS + d);
                    p0 = __dadd_rn(p0, calc.term(stretchPoint(sl.x, c.x, zq), d));
                    p1 = __dadd_rn(p1, calc.term(stretchPoint(sl.y, c.y, zq), d + 1));
                }
            } else {
                const int jq = __shfl_sync(0xffffffffu, jA, q);
                const double* ownRow = a.pos + (size_t)(own0 + i) * D;
                const double* selRow = a.pos + (size_t)(other0 + jq) * D;
                for (int d = 2 * lane; d < D; d += 64) {
                    p0 = __dadd_rn(p0, calc.term(stretchPoint(selRow[d], ownRow[d], zq), d));
                    if (d + 1 < D) p1 = __dadd_rn(p1, calc.term(stretchPoint(selRow[d + 1], ownRow[d + 1], zq), d + 1));
                }
            }
        }
        s[q] = __dadd_rn(p0, p1);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
        for (int q = 0; q < WPW; ++q) s[q] = __dadd_rn(s[q], __shfl_xor_sync(0xffffffffu, s[q], off));
    }
#pragma unroll
    for (int q = 0; q < WPW; ++q)
        if (lane == 0) ctaPartial[warp * WPW + q] = s[q];
    __syncthreads();
    if (sendSlot) {
        if (waitAll & 2) *sendSlot = ctaPartial[sendK];                                    /* experiment: plain store */
        else *reinterpret_cast<volatile double*>(sendSlot) = ctaPartial[sendK];
    }
    for (int u = threadIdx.x + DimFusedThreads; u < a.world * NW; u += DimFusedThreads) {            /* world * NW > 128: the rest of the line */
        const int r = u / NW, k = u - r * NW;
        if (t * NW + k < K)
            *reinterpret_cast<volatile double*>(a.dim_partial[r] + ((size_t)a.half * a.world + a.rank) * K + (size_t)t * NW + k) = ctaPartial[k];
    }
}

/* ---- the finishing kernel: persistent warps, groups of 32 walkers in index order (warp g of the grid takes groups g, g + G, ...) */
enum { DimFinishWarps = 4, DimPollUnroll = 8 };   /* 4 warps x 93 registers: fits beside 7 summing CTAs of 48 registers */

template <class Calc, bool VEC, bool TRACE, int WALK, int CH, int MINB>
__global__ void __launch_bounds__(DimFinishWarps * 32, MINB) dimShardFinish(const mcmcpp_halfstep_args a, const Calc calcIn, const unsigned long long timeoutNs)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D, K = a.K;
    const int own0 = a.half * K, other0 = (1 - a.half) * K;
    Calc calc = calcIn;
    const int totalWarps = (int)gridDim.x * DimFinishWarps;
    const int numGroups = (K + 31) / 32;
    unsigned long long myAccepted = 0;
    /* consecutive groups go to different SMs: the groups that become ready at the same time are spread over the machine */
    for (int g = warp * (int)gridDim.x + (int)blockIdx.x; g < numGroups; g += totalWarps) {
        const int iF = g * 32 + lane;
        const bool mine = iF < K;
        bool accepted = false;
        double z = 1.0;
        int j = 0;
        if (mine) {
            double* const slot0 = a.dim_partial[a.rank] + (size_t)a.half * a.world * K + iF;        /* + r * K */
            unsigned long long seen[DimPollUnroll];
#pragma unroll
            for (int r = 0; r < DimPollUnroll; ++r)              /* all the ranks' slots at once (independent loads) */
                seen[r] = r < a.world ? *reinterpret_cast<volatile unsigned long long*>(slot0 + (size_t)r * K) : 0ULL;
            const int w = own0 + iF;
            z = __ldg(a.dim_z + iF);
            j = __ldg(a.dim_j + iF);
            const double e = __ldg(a.dim_e + iF), scaling = __ldg(a.dim_s + iF), oldp = a.logp[w];
            double sum = 0.0;
#pragma unroll
            for (int r = 0; r < DimPollUnroll; ++r) {
                if (r < a.world) {
                    const double p = dimTakePartial(slot0 + (size_t)r * K, seen[r], timeoutNs, a.flags);
                    sum = r == 0 ? p : __dadd_rn(sum, p);            /* rank order */
                }
            }
            for (int r = DimPollUnroll; r < a.world; ++r) sum = __dadd_rn(sum, dimTakePartial(slot0 + (size_t)r * K, DimEmpty, timeoutNs, a.flags));
            const double newp = calc.finish(sum);
            const double diff = __dsub_rn(__dadd_rn(scaling, newp), oldp);
            accepted = e < diff;
            if (accepted) a.logp[w] = newp;
            if (TRACE) {
                if (iF < a.trace_rows) {
                    mcmcpp_trace tr;
                    tr.j = j; tr.accept = accepted ? 1 : 0; tr.z = z; tr.e = e; tr.newp = newp; tr.diff = diff;
                    a.trace[iF] = tr;
                }
            }
        }
        const unsigned accMask = __ballot_sync(0xffffffffu, accepted);
        unsigned work = a.chain_row ? __ballot_sync(0xffffffffu, mine) : accMask;         /* stored steps: every walker's chain row */
        myAccepted += (unsigned long long)__popc(accMask);
        /* rebuild the slices of the walkers in `work`: the whole warp on WALK walkers at a time, CH 64-element chunks of each
         * in flight per lane (short slices: 4 walkers x 2 chunks, so that one memory latency covers four walkers) */
        while (work) {
            int qs[WALK];
#pragma unroll
            for (int k = 0; k < WALK; ++k) {
                qs[k] = work ? __ffs(work) - 1 : -1;
                if (work) work &= work - 1;
            }
            if (VEC) {
                double* ownRow[WALK];
                const double* selRow[WALK];
                double zq[WALK];
                bool acc[WALK];
#pragma unroll
                for (int k = 0; k < WALK; ++k) {
                    const int q = qs[k] < 0 ? 0 : qs[k];
                    acc[k] = qs[k] >= 0 && ((accMask >> q) & 1u);
                    zq[k] = __shfl_sync(0xffffffffu, z, q);
                    const int jq = __shfl_sync(0xffffffffu, j, q);
                    ownRow[k] = a.pos + (size_t)(own0 + g * 32 + q) * D;
                    selRow[k] = a.pos + (size_t)(other0 + jq) * D;
                }
                for (int d0 = 0; d0 < D; d0 += 64 * CH) {
                    double2 c[WALK][CH], sl[WALK][CH];
#pragma unroll
                    for (int k = 0; k < WALK; ++k) {
#pragma unroll
                        for (int u = 0; u < CH; ++u) {
                            const int d = d0 + 64 * u + 2 * lane;
                            if (qs[k] >= 0 && d < D) {
                                c[k][u] = *reinterpret_cast<const double2*>(ownRow[k] + d);
                                if (acc[k]) sl[k][u] = *reinterpret_cast<const double2*>(selRow[k] + d);
                            }
                        }
                    }
#pragma unroll
                    for (int k = 0; k < WALK; ++k) {
#pragma unroll
                        for (int u = 0; u < CH; ++u) {
                            const int d = d0 + 64 * u + 2 * lane;
                            if (qs[k] >= 0 && d < D) {
                                double2 v2 = c[k][u];
                                if (acc[k]) {
                                    v2.x = stretchPoint(sl[k][u].x, c[k][u].x, zq[k]);
                                    v2.y = stretchPoint(sl[k][u].y, c[k][u].y, zq[k]);
                                    *reinterpret_cast<double2*>(ownRow[k] + d) = v2;
                                }
                                if (a.chain_row) *reinterpret_cast<double2*>(a.chain_row + (ownRow[k] - a.pos) + d) = v2;
                            }
                        }
                    }
                }
            } else {
#pragma unroll
                for (int k = 0; k < WALK; ++k) {
                    if (qs[k] < 0) continue;
                    const int q = qs[k];
                    const bool acc = (accMask >> q) & 1u;
                    const double zq = __shfl_sync(0xffffffffu, z, q);
                    const int jq = __shfl_sync(0xffffffffu, j, q);
                    double* ownRow = a.pos + (size_t)(own0 + g * 32 + q) * D;
                    const double* selRow = a.pos + (size_t)(other0 + jq) * D;
                    double* chainRow = a.chain_row ? a.chain_row + (size_t)(own0 + g * 32 + q) * D : nullptr;
                    for (int d = lane; d < D; d += 32) {
                        double v1 = ownRow[d];
                        if (acc) { v1 = stretchPoint(selRow[d], v1, zq); ownRow[d] = v1; }
                        if (chainRow) chainRow[d] = v1;
                    }
                }
            }
        }
    }
    if (lane == 0 && myAccepted) atomicAdd(a.accept_slots + (((int)blockIdx.x * DimFinishWarps + warp) % a.max_grid), myAccepted);
}

/* Launcher of a DIM-sharded colour for a separable Calculator (a VIEW of this rank's columns: term(x, d) takes the
 * local column index).  Pass &DimShardLauncher<Calc>::launch to mcmcpp_sampler_create with shard_mode DIM. */
template <class Calc>
struct DimShardLauncher
{
    static_assert(IsSeparable<Calc>::value, "DIM sharding needs a separable Calculator: term(x, d) and finish(sum)");

    static int envInt(const char* name, int dflt)
    {
        const char* e = getenv(name);
        return e ? atoi(e) : dflt;
    }

    template <int WPW, bool TMA>
    static int sumW(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        constexpr int NW = DimFusedWarps * WPW;
        const size_t smem = TMA ? (size_t)NW * 2u * rowStride + 16u * NW : 16u * NW;      /* rows, mbarriers, partials */
        if (smem > 227u * 1024u) return MCMCPP_ERR_UNSUPPORTED;
        static DynamicSmemOptIn optIn;
        if (!optIn.ensure(dimShardSum<Calc, WPW, TMA>, smem)) return MCMCPP_ERR_CUDA;
        static const int waitAll = envInt("MCMCPP_B200_DIM_WAIT_ALL", 0);
        dimShardSum<Calc, WPW, TMA><<<(a.K + NW - 1) / NW, DimFusedThreads, smem, stream>>>(a, calc, rowStride, waitAll);
        return launchStatus();
    }

    /* finishing kernel of one rank on `stream`: one warp per group of 32 walkers */
    static int finish(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
    {
        static const unsigned long long timeoutNs = (unsigned long long)envInt("MCMCPP_B200_RDV_TIMEOUT_MS", 20000) * 1000000ULL;
        const int numGroups = (a.K + 31) / 32;
        const int grid = (numGroups + DimFinishWarps - 1) / DimFinishWarps;
        const bool vec = (a.D % 2) == 0;
#define MCMCPP_DIM_FINISH(V, T, WALK, CH, MINB) dimShardFinish<Calc, V, T, WALK, CH, MINB><<<grid, DimFinishWarps * 32, 0, stream>>>(a, calc, timeoutNs)
        static const int variant = envInt("MCMCPP_B200_DIM_FINISH_VARIANT", 0);
        if (a.trace) {
            if (vec) MCMCPP_DIM_FINISH(true, true, 1, 4, 4); else MCMCPP_DIM_FINISH(false, true, 1, 1, 4);
        } else if (!vec) {
            MCMCPP_DIM_FINISH(false, false, 1, 1, 4);
        } else if (variant == 1) {
            MCMCPP_DIM_FINISH(true, false, 1, 4, 8);
        } else if (variant == 2) {
            MCMCPP_DIM_FINISH(true, false, 2, 2, 6);
        } else if (a.D <= 128) {
            MCMCPP_DIM_FINISH(true, false, 4, 2, 4);
        } else if (a.D <= 256) {
            MCMCPP_DIM_FINISH(true, false, 2, 4, 4);
        } else {
            MCMCPP_DIM_FINISH(true, false, 1, 4, 4);
        }
#undef MCMCPP_DIM_FINISH
        return launchStatus();
    }

    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        if (!a.dim_partial || !a.dim_z || !a.dim_e || !a.dim_j || !a.dim_s) return MCMCPP_ERR_STATE;
        if (a.world > 32) return MCMCPP_ERR_UNSUPPORTED;
        const Calc& calc = *static_cast<const Calc*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        dimShardDraw<<<(a.K + 255) / 256, 256, 0, stream>>>(a);
        int rc = sumOnly(a, calc, stream);
        if (rc != MCMCPP_OK) return rc;
        if ((rc = finish(a, calc, stream)) != MCMCPP_OK) return rc;
        if (a.launch_count) *a.launch_count += 2;
        return MCMCPP_OK;
    }

    /* mcmcpp_group_fn: one colour of `count` ranks that live in this process, on one stream: every rank's draws and summing
     * kernel first (they wait for nobody), then every rank's finishing kernel */
    static int launchGroup(const mcmcpp_halfstep_args* args, const void* const* calculators, int count, void* cudaStream)
    {
        if (count < 1 || count > 8) return MCMCPP_ERR_UNSUPPORTED;
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        for (int r = 0; r < count; ++r) {
            const mcmcpp_halfstep_args& a = args[r];
            if (!a.dim_partial || !a.dim_z || !a.dim_e || !a.dim_j || !a.dim_s || a.world != count || a.rank != r || a.K != args[0].K) return MCMCPP_ERR_STATE;
            dimShardDraw<<<(a.K + 255) / 256, 256, 0, stream>>>(a);
            int rc = sumOnly(a, *static_cast<const Calc*>(calculators[r]), stream);
            if (rc != MCMCPP_OK) return rc;
        }
        for (int r = 0; r < count; ++r) {
            int rc = finish(args[r], *static_cast<const Calc*>(calculators[r]), stream);
            if (rc != MCMCPP_OK) return rc;
        }
        return MCMCPP_OK;
    }

    /* the summing kernel alone (the draws are already on the stream) */
    static int sumOnly(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
    {
        const uint32_t rowStride = ((uint32_t)a.D * 8u + 127u) & ~127u;
        static const size_t stage = (size_t)envInt("MCMCPP_B200_DIM_STAGE_KB", 36) * 1024;
        const size_t perWalker = 2u * (size_t)rowStride;
        const bool tma = (a.D % 2) == 0 && !bulkKernelDisabled();
#define MCMCPP_DIM_SUM(WPW) (tma ? sumW<WPW, true>(a, calc, stream, rowStride) : sumW<WPW, false>(a, calc, stream, rowStride))
        if (DimFusedWarps * 8 * perWalker <= stage) return MCMCPP_DIM_SUM(8);
        if (DimFusedWarps * 4 * perWalker <= stage) return MCMCPP_DIM_SUM(4);
        if (DimFusedWarps * 2 * perWalker <= stage) return MCMCPP_DIM_SUM(2);
        return MCMCPP_DIM_SUM(1);
#undef MCMCPP_DIM_SUM
    }
};

}
}
#endif
