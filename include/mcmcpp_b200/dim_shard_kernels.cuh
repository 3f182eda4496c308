/*
 * include/mcmcpp_b200/dim_shard_kernels.cuh -- MCMCPP_SHARD_DIM: the stretch-move half-step with the PARAMETER
 * dimension sharded over the GPUs of one box (SURVEY.md s.8(e) alternative 2).
 *
 * Every rank holds all W walkers but only its columns [dim_begin, dim_begin + D) of each row, makes the SAME draws
 * (j, z, log(1-u)) from the sampler's pcg64 stream, builds its slice of every proposal
 * (MCMCpp/Movers/StretchMove.h:103-109) and the partial sum of a separable Calculator (term()/finish(), targets.h)
 * over its columns.  One colour = two launches:
 *
 *   dimShardDraw    one thread per walker: the three draws of the update (j, z, log(1-u)) and log(z)(D-1) into the draw scratch.
 *   dimShardFused   one-shot CTAs, one TILE of NW walkers each; CTA t (= blockIdx.x: the hardware hands CTAs out in index order,
 *                   so a lower tile's CTA has started earlier)
 *                   has two kinds of warps that never wait for each other:
 *                     SUMMING (4 warps)  issue the TMA bulk copies of tile t (own slice + partner slice of every walker, one
 *                        mbarrier per walker), reduce the proposals to canonical 64-way partial sums and store them, one line of
 *                        NW partials per rank, straight into EVERY rank's table over NVLink (peer-mapped stores);
 *                     FINISHING (1-2 warps)  tile t - LAG: take the `world` partial sums of each of its walkers from this rank's
 *                        table, add them in rank order -- the same bits on every rank, so every rank takes the same Metropolis
 *                        decision (StretchMove.h:112-122) -- and rebuild the accepted walkers' slices (8-13 % of them at
 *                        D = 1000; their rows were streamed LAG tiles ago and are still in L2).
 *                   A CTA's sums never wait for anything but its own copies, and what a finishing warp waits for comes from CTAs
 *                   of lower index (started earlier) or from ranks that cannot be waiting for more than this rank has sent.
 *                   (A global ticket counter would make "started earlier" independent of the dispatch order, at the price of
 *                   32,768 atomics on one address and a barrier per CTA; every wait is bounded by the timeout either way.)
 *                   The hardware's CTA scheduler does the pipelining: 200+ KB of rows in flight per SM.
 *
 * Exchange protocol: a table slot holds either the all-ones NaN (`DimEmpty`, no arithmetic produces it) or a partial
 * sum.  Writers store 8-byte values (single, indivisible stores); the one reader of a slot spins until it is not
 * DimEmpty, takes the value and puts DimEmpty back.  No flag, no fence, no rendezvous kernel and no NCCL call.
 * The table has a half per colour; a rank cannot write a slot of colour c again before its reader has emptied it: it
 * would have to be two half-steps ahead, and it cannot finish the half-step in between without this rank's partials of
 * that half-step, which this rank only sends after the launch that emptied the slot has completed (stream order).
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
 * earlier look at the slot saw. */
__device__ __forceinline__ double dimTakePartial(double* slotPtr, unsigned long long v, unsigned long long timeoutNs, int32_t* flags)
{
    volatile unsigned long long* slot = reinterpret_cast<volatile unsigned long long*>(slotPtr);
    if (v == DimEmpty) {
        const unsigned long long t0 = globalTimerNs();
        unsigned spins = 0;
        while ((v = *slot) == DimEmpty) {
            __nanosleep(32);
            if ((++spins & 1023u) == 0 && globalTimerNs() - t0 > timeoutNs) {
                atomicExch(flags + 1, 1);                    /* surfaced by mcmcpp_sampler_run */
                return 0.0;
            }
        }
    }
    *slot = DimEmpty;
    return __longlong_as_double((long long)v);
}

/* ---- the fused kernel (see the header comment): one tile per CTA ----------------------------------------------------
 * TMA = rows are 16-byte multiples (even slice): staged with cp.async.bulk; otherwise the warps read their walkers' rows
 * straight from global memory (odd slices: tests and the last rank of an odd row).
 * Shared memory: [NW walkers][own row | partner row] (rowStride bytes each), NW mbarriers, NW partials.
 * `t` is the CTA's tile (its block index; in the rank-group kernel block index / ranks). */
enum { DimPollUnroll = 8 };

__device__ __forceinline__ void dimSumWarpsSync()
{
    asm volatile("bar.sync 1, 128;" ::: "memory");            /* the four summing warps only (DimFusedThreads) */
}

template <class Calc, int WPW, bool TMA, bool TRACE, int FW>
__device__ __forceinline__ void dimShardFusedBody(const mcmcpp_halfstep_args& a, const Calc& calcIn, const uint32_t rowStride,
                                                  const int lagTiles, const int t, const unsigned long long timeoutNs)
{
    extern __shared__ __align__(128) unsigned char dimSmem[];
    constexpr int NW = DimFusedWarps * WPW;                  /* walkers per tile */
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D, K = a.K;
    const int numTiles = (K + NW - 1) / NW;
    const int own0 = a.half * K, other0 = (1 - a.half) * K;
    Calc calc = calcIn;

    if (warp >= DimFusedWarps) {
        /* ================= FINISHING warps: tile t - LAG, independent of the summing warps of this CTA ================= */
        constexpr int PER = (NW + FW - 1) / FW;              /* walkers of the tile per finishing warp, one per lane */
        static_assert(PER <= 32, "a finishing warp decides at most 32 walkers");
        const int f = t - lagTiles;
        if (f < 0 || f >= numTiles) return;
        const int k0 = (warp - DimFusedWarps) * PER;
        const int iF = f * NW + k0 + lane;
        const bool mine = lane < PER && k0 + lane < NW && iF < K;
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
        if (lane == 0 && accMask) atomicAdd(a.accept_slots + (t % a.max_grid), (unsigned long long)__popc(accMask));
        /* rebuild the slices of the walkers in `work`, the whole warp on one walker at a time (rows come from L2) */
        while (work) {
            const int q = __ffs(work) - 1;
            work &= work - 1;
            const bool acc = (accMask >> q) & 1u;
            const int iq = f * NW + k0 + q;
            const double zq = __shfl_sync(0xffffffffu, z, q);
            const int jq = __shfl_sync(0xffffffffu, j, q);
            double* ownRow = a.pos + (size_t)(own0 + iq) * D;
            const double* selRow = a.pos + (size_t)(other0 + jq) * D;
            double* chainRow = a.chain_row ? a.chain_row + (size_t)(own0 + iq) * D : nullptr;
            if (TMA) {
                for (int d0 = 0; d0 < D; d0 += 256) {                      /* four vector loads per row in flight per lane */
                    double2 c[4], sl[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int d = d0 + 64 * u + 2 * lane;
                        if (d < D) {
                            c[u] = *reinterpret_cast<const double2*>(ownRow + d);
                            if (acc) sl[u] = *reinterpret_cast<const double2*>(selRow + d);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int d = d0 + 64 * u + 2 * lane;
                        if (d < D) {
                            double2 v2 = c[u];
                            if (acc) {
                                v2.x = stretchPoint(sl[u].x, c[u].x, zq);
                                v2.y = stretchPoint(sl[u].y, c[u].y, zq);
                                *reinterpret_cast<double2*>(ownRow + d) = v2;
                            }
                            if (chainRow) *reinterpret_cast<double2*>(chainRow + d) = v2;
                        }
                    }
                }
            } else {
                for (int d = lane; d < D; d += 32) {
                    double v1 = ownRow[d];
                    if (acc) { v1 = stretchPoint(selRow[d], v1, zq); ownRow[d] = v1; }
                    if (chainRow) chainRow[d] = v1;
                }
            }
        }
        return;
    }

    /* ================= SUMMING warps: tile t ================= */
    if (t >= numTiles) return;                                /* CTA-uniform: the finish-only CTAs behind the last tile */
    const uint32_t rowBytes = (uint32_t)D * 8u;
    unsigned char* const rows = dimSmem + (size_t)(warp * WPW) * 2u * rowStride;
    unsigned char* const barBase = dimSmem + (TMA ? (size_t)NW * 2u * rowStride : 0);     /* no staged rows without TMA */
    double* const ctaPartial = reinterpret_cast<double*>(barBase + 8u * NW);
    if (TMA) {
        if (threadIdx.x < NW) mbarInit(smemAddr(barBase + 8u * threadIdx.x), 1);
        dimSumWarpsSync();
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
    /* partial sums of the warp's walkers: all copies awaited first, then WPW independent dependency chains side by side */
    if (TMA) {
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
    dimSumWarpsSync();
    /* one line of NW partials to every rank */
    for (int u = threadIdx.x; u < a.world * NW; u += DimFusedThreads) {
        const int r = u / NW, k = u - r * NW;
        if (t * NW + k < K)
            *reinterpret_cast<volatile double*>(a.dim_partial[r] + ((size_t)a.half * a.world + a.rank) * K + (size_t)t * NW + k) = ctaPartial[k];
    }
}

template <class Calc, int WPW, bool TMA, bool TRACE, int FW, int MINB>
__global__ void __launch_bounds__(DimFusedThreads + 32 * FW, MINB) dimShardFused(const mcmcpp_halfstep_args a, const Calc calcIn, const uint32_t rowStride,
                                                                            const int lagTiles, const unsigned long long timeoutNs)
{
    dimShardFusedBody<Calc, WPW, TMA, TRACE, FW>(a, calcIn, rowStride, lagTiles, (int)blockIdx.x, timeoutNs);
}

/* ---- all ranks of a DIM-sharded ensemble in ONE launch of one process (mcmcpp_sampler_run_group) ------------------------
 * The ranks' CTAs wait on one another (each polls partial sums the others store), so as separate launches on one GPU
 * nothing would guarantee that they run at the same time.  Here CTA b works for rank b % R, takes its ticket from that rank's
 * counter and runs exactly the code of the one-rank kernel on that rank's own buffers, tables and column slice: the world > 1
 * protocol -- slot hand-over, rank-ordered sums, replicated decisions, per-rank slices (even ones through TMA, odd ones through
 * plain loads) -- runs on a one-GPU box.  Progress: a CTA sends before it ever blocks, and only blocks on LOWER tiles of any rank
 * (see lagFor below), whose CTAs have lower block indices: they have started or will start as resident ones retire. */
enum { DimGroupMax = 4 };
template <class Calc>
struct DimGroupArgs
{
    int count;
    uint32_t rowStride[DimGroupMax];
    mcmcpp_halfstep_args args[DimGroupMax];
    alignas(16) unsigned char calc[DimGroupMax][(sizeof(Calc) + 15) / 16 * 16];      /* Calculators by value (they need not be default-constructible) */
};

template <class Calc, int WPW, bool TRACE, int FW>
__global__ void __launch_bounds__(DimFusedThreads + 32 * FW) dimShardFusedGroup(const DimGroupArgs<Calc> g, const int lagTiles, const unsigned long long timeoutNs)
{
    const int r = (int)blockIdx.x % g.count, t = (int)blockIdx.x / g.count;
    const Calc& calc = *reinterpret_cast<const Calc*>(g.calc[r]);
    if ((g.args[r].D % 2) == 0) dimShardFusedBody<Calc, WPW, true, TRACE, FW>(g.args[r], calc, g.rowStride[r], lagTiles, t, timeoutNs);
    else dimShardFusedBody<Calc, WPW, false, TRACE, FW>(g.args[r], calc, g.rowStride[r], lagTiles, t, timeoutNs);
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

    /* LAG in tiles.  Any value >= 0 is safe: a CTA sends its own sums before it ever blocks, and what it blocks on comes from CTAs
     * of lower index -- handed out earlier by the hardware -- or from the other ranks, which cannot be blocked on more than this
     * rank has already sent.  Small: the finished walkers' rows are still in L2 and the finish
     * runs under this tile's TMA latency; too small: the other ranks' sums have not crossed NVLink yet and the finish moves behind
     * the tile's own work.  Default: a quarter of the CTAs resident at once. */
    static int lagFor(int residentCtas)
    {
        static const int lag = envInt("MCMCPP_B200_DIM_LAG", -1);
        return lag >= 0 ? lag : (residentCtas + 3) / 4;
    }

    template <int WPW, bool TMA, bool TRACE, int FW, int MINB>
    static int fusedFM(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride, size_t smem)
    {
        constexpr int NW = DimFusedWarps * WPW;
        static const unsigned long long timeoutNs = (unsigned long long)envInt("MCMCPP_B200_RDV_TIMEOUT_MS", 20000) * 1000000ULL;
        static DynamicSmemOptIn optIn;
        if (!optIn.ensure(dimShardFused<Calc, WPW, TMA, TRACE, FW, MINB>, smem)) return MCMCPP_ERR_CUDA;
        int perSm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dimShardFused<Calc, WPW, TMA, TRACE, FW, MINB>, DimFusedThreads + 32 * FW, smem) != cudaSuccess || perSm < 1)
            return MCMCPP_ERR_CUDA;
        const int lag = lagFor((a.sm_count > 0 ? a.sm_count : 148) * perSm);
        const int numTiles = (a.K + NW - 1) / NW;
        dimShardFused<Calc, WPW, TMA, TRACE, FW, MINB><<<numTiles + lag, DimFusedThreads + 32 * FW, smem, stream>>>(a, calc, rowStride, lag, timeoutNs);
        return launchStatus();
    }

    template <int WPW, bool TMA, bool TRACE>
    static int fused(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        constexpr int NW = DimFusedWarps * WPW;
        const size_t smem = TMA ? (size_t)NW * 2u * rowStride + 16u * NW : 16u * NW;      /* rows, mbarriers, partials */
        if (smem > 227u * 1024u) return MCMCPP_ERR_UNSUPPORTED;
        /* finishing warps beside the four summing warps; CTAs per SM the register allocation is asked to allow (short slices need many
         * CTAs in flight: the per-CTA chain load index -> copy -> sum -> send is latency, not bandwidth) */
        static const int fwEnv = envInt("MCMCPP_B200_DIM_FW", 0), minbEnv = envInt("MCMCPP_B200_DIM_MINB", 0);
        const bool two = fwEnv ? fwEnv == 2 : NW >= 16;
        const bool dense = (minbEnv ? minbEnv > 1 : true) && TMA && !TRACE && smem <= 33u * 1024u;
        if (NW >= 8 && two) {
            if (dense) return fusedFM<WPW, TMA, TRACE, 2, 6>(a, calc, stream, rowStride, smem);
            return fusedFM<WPW, TMA, TRACE, 2, 1>(a, calc, stream, rowStride, smem);
        }
        if (dense) return fusedFM<WPW, TMA, TRACE, 1, 7>(a, calc, stream, rowStride, smem);
        return fusedFM<WPW, TMA, TRACE, 1, 1>(a, calc, stream, rowStride, smem);
    }

    template <int WPW, bool TMA>
    static int fusedT(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        return a.trace ? fused<WPW, TMA, true>(a, calc, stream, rowStride) : fused<WPW, TMA, false>(a, calc, stream, rowStride);
    }

    template <int WPW, bool TRACE>
    static int fusedGroup(const DimGroupArgs<Calc>& g, cudaStream_t stream, uint32_t maxStride)
    {
        constexpr int NW = DimFusedWarps * WPW;
        static const unsigned long long timeoutNs = (unsigned long long)envInt("MCMCPP_B200_RDV_TIMEOUT_MS", 20000) * 1000000ULL;
        const size_t smem = (size_t)NW * 2u * maxStride + 16u * NW;
        if (smem > 227u * 1024u) return MCMCPP_ERR_UNSUPPORTED;
        constexpr int FW = NW >= 16 ? 2 : 1;
        static DynamicSmemOptIn optIn;
        if (!optIn.ensure(dimShardFusedGroup<Calc, WPW, TRACE, FW>, smem)) return MCMCPP_ERR_CUDA;
        int perSm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dimShardFusedGroup<Calc, WPW, TRACE, FW>, DimFusedThreads + 32 * FW, smem) != cudaSuccess || perSm < 1)
            return MCMCPP_ERR_CUDA;
        const mcmcpp_halfstep_args& a0 = g.args[0];
        const int lag = lagFor((a0.sm_count > 0 ? a0.sm_count : 148) * perSm);       /* resident CTAs of ALL ranks together */
        const int numTiles = (a0.K + NW - 1) / NW;
        dimShardFusedGroup<Calc, WPW, TRACE, FW><<<(numTiles + lag) * g.count, DimFusedThreads + 32 * FW, smem, stream>>>(g, lag, timeoutNs);
        return launchStatus();
    }

    /* mcmcpp_group_fn: one colour of `count` ranks that live in this process, on one stream */
    static int launchGroup(const mcmcpp_halfstep_args* args, const void* const* calculators, int count, void* cudaStream)
    {
        if (count < 1 || count > DimGroupMax) return MCMCPP_ERR_UNSUPPORTED;
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        DimGroupArgs<Calc> g;
        g.count = count;
        uint32_t maxStride = 0;
        bool trace = false;
        for (int r = 0; r < count; ++r) {
            const mcmcpp_halfstep_args& a = args[r];
            if (!a.dim_partial || !a.dim_z || !a.dim_e || !a.dim_j || !a.dim_s || a.world != count || a.rank != r || a.K != args[0].K) return MCMCPP_ERR_STATE;
            g.args[r] = a;
            memcpy(g.calc[r], calculators[r], sizeof(Calc));
            g.rowStride[r] = ((uint32_t)a.D * 8u + 127u) & ~127u;
            if (g.rowStride[r] > maxStride) maxStride = g.rowStride[r];
            trace = trace || a.trace != nullptr;
            dimShardDraw<<<(a.K + 255) / 256, 256, 0, stream>>>(a);                 /* every rank makes the same draws, into its own scratch */
        }
        for (int r = count; r < DimGroupMax; ++r) { g.args[r] = args[0]; memcpy(g.calc[r], g.calc[0], sizeof(Calc)); g.rowStride[r] = g.rowStride[0]; }
        const size_t perWalker = 2u * (size_t)maxStride, stage = 36u * 1024u;
#define MCMCPP_DIM_GROUP(WPW) (trace ? fusedGroup<WPW, true>(g, stream, maxStride) : fusedGroup<WPW, false>(g, stream, maxStride))
        if (DimFusedWarps * 8 * perWalker <= stage) return MCMCPP_DIM_GROUP(8);
        if (DimFusedWarps * 4 * perWalker <= stage) return MCMCPP_DIM_GROUP(4);
        if (DimFusedWarps * 2 * perWalker <= stage) return MCMCPP_DIM_GROUP(2);
        return MCMCPP_DIM_GROUP(1);
#undef MCMCPP_DIM_GROUP
    }

    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        if (!a.dim_partial || !a.dim_z || !a.dim_e || !a.dim_j || !a.dim_s) return MCMCPP_ERR_STATE;
        if (a.world > 32) return MCMCPP_ERR_UNSUPPORTED;
        const Calc& calc = *static_cast<const Calc*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        dimShardDraw<<<(a.K + 255) / 256, 256, 0, stream>>>(a);
        if (a.launch_count) *a.launch_count += 1;
        const uint32_t rowStride = ((uint32_t)a.D * 8u + 127u) & ~127u;
        /* a tile stages ~36 KB (4 warps x WPW walkers x 2 rows) whatever the slice length is: 6 CTAs per SM at short slices */
        static const size_t stage = (size_t)envInt("MCMCPP_B200_DIM_STAGE_KB", 36) * 1024;
        const size_t perWalker = 2u * (size_t)rowStride;
        const bool tma = (a.D % 2) == 0 && !bulkKernelDisabled();
#define MCMCPP_DIM_FUSED(WPW) (tma ? fusedT<WPW, true>(a, calc, stream, rowStride) : fusedT<WPW, false>(a, calc, stream, rowStride))
        if (DimFusedWarps * 8 * perWalker <= stage) return MCMCPP_DIM_FUSED(8);
        if (DimFusedWarps * 4 * perWalker <= stage) return MCMCPP_DIM_FUSED(4);
        if (DimFusedWarps * 2 * perWalker <= stage) return MCMCPP_DIM_FUSED(2);
        return MCMCPP_DIM_FUSED(1);
#undef MCMCPP_DIM_FUSED
    }
};

}
}
#endif
