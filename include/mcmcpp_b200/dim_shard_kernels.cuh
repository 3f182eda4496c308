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
 *   dimShardFused   persistent CTAs, each a software pipeline over its tiles of NW walkers with a ring of S
 *                   shared-memory stages:
 *                     issue(n+P)   lane q of a warp issues walker q's two cp.async.bulk copies (own slice, partner
 *                                  slice) onto the walker's mbarrier -- P tiles are always in flight;
 *                     partial(n)   canonical 64-way partial sums of the tile's proposals from shared memory; the NW
 *                                  partials go, as one line per rank, straight into EVERY rank's partial table over
 *                                  NVLink (peer-mapped stores);
 *                     finish(n-L)  for the tile staged L iterations ago: poll the `world` partials of each walker in
 *                                  this rank's table, add them in rank order -- the same bits on every rank, so every
 *                                  rank takes the same Metropolis decision (StretchMove.h:112-122) -- and commit the
 *                                  accepted walkers' slices FROM SHARED MEMORY (the rows are never read twice).
 *
 * Exchange protocol: a table slot holds either the all-ones NaN (`DimEmpty`, no arithmetic produces it) or a partial
 * sum.  Writers store 8-byte values (single, indivisible stores); the one reader of a slot spins until it is not
 * DimEmpty, takes the value and puts DimEmpty back.  No flag, no fence, no rendezvous kernel and no NCCL call: the
 * L iterations between partial(n) and finish(n) hide the NVLink latency.  The table has a half per colour; a rank
 * cannot write a slot of colour c again before its reader has emptied it: it would have to be two half-steps ahead,
 * and it cannot finish the half-step in between without this rank's partials of that half-step, which this rank only
 * sends after the launch that emptied the slot has completed (stream order).  A poll that sees nothing for
 * `timeout_ns` sets flags[1] and gives up (mcmcpp_sampler_run then returns MCMCPP_ERR_STATE): a dead peer cannot wedge
 * the other GPUs.
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
 * earlier (prefetched) look at the slot saw. */
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

/* ---- phase 1 + 2, one persistent kernel (see the header comment) ------------------------------------------------
 * TMA = rows are 16-byte multiples (even slice): staged with cp.async.bulk; otherwise the warp copies its walkers'
 * rows into the stage itself (odd slices: tests and the last rank of an odd row).
 * Shared memory: S stages x [NW walkers][own row | partner row] (rowStride bytes each), S x NW mbarriers, S x NW x 4 doubles of
 * per-walker scalars {z, log(1-u), log(z)(D-1), current log-prob}, 2 x NW partials.  S stages = P in flight + L waiting for
 * the other ranks + 1 being worked on; S and P are compile-time so that the ring indices cost no division.
 * Nothing on the per-tile critical path waits for global memory except the staged rows themselves: the scalars of tile
 * n+P+1 are loaded (lane q <-> walker q of the warp) while tile n is worked on, and the first look at the partial slots
 * of tile n-L is taken before the partial sums of tile n are computed.  At a 1 KB slice the kernel is bound by instruction
 * issue, not by HBM: all index arithmetic is carried as running pointers. */
template <class Calc, int WPW, bool TMA, bool TRACE, int S, int P>
__device__ __forceinline__ void dimShardFusedBody(const mcmcpp_halfstep_args& a, const Calc& calcIn, const uint32_t rowStride,
                                                  const unsigned long long timeoutNs, const int bid, const int nblocks)
{
    extern __shared__ __align__(128) unsigned char dimSmem[];
    constexpr int NW = DimFusedWarps * WPW;                  /* walkers per tile */
    constexpr int L = S - P - 1;
    static_assert(P >= 1 && L >= 1, "ring: P tiles in flight, L waiting, one in work");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D, K = a.K;
    const uint32_t rowBytes = (uint32_t)D * 8u;
    const uint32_t stageBytes = (uint32_t)NW * 2u * rowStride;
    unsigned char* const barBase = dimSmem + (size_t)S * stageBytes;
    double* const meta = reinterpret_cast<double*>(barBase + 8u * S * NW);              /* [S][NW][4] */
    double* const ctaPartial = meta + S * NW * 4;                                        /* [2][NW] */
    const int numTiles = (K + NW - 1) / NW;
    const int myTiles = bid < numTiles ? (numTiles - bid + nblocks - 1) / nblocks : 0;
    const int own0 = a.half * K, other0 = (1 - a.half) * K;
    double* const table = a.dim_partial[a.rank] + (size_t)a.half * a.world * K;          /* this rank's slots of the colour */
    Calc calc = calcIn;
    unsigned long long myAccepted = 0;

    /* this warp's part of a stage, as running offsets */
    unsigned char* const warpRows0 = dimSmem + (size_t)(warp * WPW) * 2u * rowStride;   /* + st * stageBytes */
    double* const warpMeta0 = meta + warp * WPW * 4;                                     /* + st * NW * 4 */
    const uint32_t warpBar0 = smemAddr(barBase) + 8u * (uint32_t)(warp * WPW);           /* + st * 8 * NW */
    const int walkerStep = nblocks * NW;                      /* walker index distance between consecutive tiles of this CTA */
    const int firstWalker = bid * NW + warp * WPW;            /* walker of slot (warp, 0) of local tile 0 */

    /* who polls what: one pass of the warp takes the partials of qpp walkers, lane (ql * world + r) reads rank r's slot of walker ql */
    const int qpp = 32 / a.world;
    const int pollQ = lane / a.world, pollR = lane - pollQ * a.world;
    const bool pollLane = pollQ < qpp;
    double* const pollSlot0 = table + (size_t)pollR * K;      /* + walker index */
    /* who sends what: thread t < world * NW stores partial k = t % NW of the tile into rank t / NW's table */
    const int sendK = (int)threadIdx.x % NW;
    const bool sender = (int)threadIdx.x < a.world * NW;
    double* const sendSlot0 = sender ? a.dim_partial[(int)threadIdx.x / NW] + ((size_t)a.half * a.world + a.rank) * K + sendK : nullptr;   /* + tile * NW */
    const bool manyRanks = a.world * NW > DimFusedThreads;    /* more than one store pass (world > 8 with wide tiles): generic loop */

    if (TMA) {
        for (int b = threadIdx.x; b < S * NW; b += DimFusedThreads) mbarInit(smemAddr(barBase + 8u * b), 1);
        __syncthreads();
    }

    /* scalars of the tile this lane will issue next (lane q < WPW <-> walker q of the warp), loaded one iteration ahead */
    int jN = 0; double zN = 1.0, eN = 0.0, sN = 0.0, pN = 0.0;
    int wPre = firstWalker + lane;                            /* walker the next prefetch is for (this lane's) */
    auto prefetch = [&]() {
        if (lane < WPW && wPre < K) {
            jN = __ldg(a.dim_j + wPre); zN = __ldg(a.dim_z + wPre); eN = __ldg(a.dim_e + wPre); sN = __ldg(a.dim_s + wPre);
            pN = a.logp[own0 + wPre];
        }
        wPre += walkerStep;
    };
    /* issue: walker slot (warp, q) of the next tile goes to stage stI; uses the scalars the last prefetch loaded */
    int stI = 0, wIss = firstWalker;                          /* stage and first walker (of this warp) of the tile to issue */
    auto issue = [&]() {
        unsigned char* rows = warpRows0 + (size_t)stI * stageBytes;
        const int i = wIss + lane;
        if (lane < WPW && i < K) {
            double2* mrow = reinterpret_cast<double2*>(warpMeta0 + stI * (NW * 4) + lane * 4);
            mrow[0] = make_double2(zN, eN); mrow[1] = make_double2(sN, pN);
            if (TMA) {
                /* the stage's last readers were this warp's own lanes, lanes whose loads have returned (their values were used):
                 * program order is enough, and a proxy fence here would wait for the warp's copies already in flight */
                const uint32_t bar = warpBar0 + 8u * (uint32_t)(stI * NW + lane);
                const uint32_t dst = smemAddr(rows) + (uint32_t)lane * 2u * rowStride;
                mbarExpectTx(bar, 2u * rowBytes);
                bulkLoad(dst, a.pos + (size_t)(own0 + i) * D, rowBytes, bar);
                bulkLoad(dst + rowStride, a.pos + (size_t)(other0 + jN) * D, rowBytes, bar);
            }
        }
        if (!TMA) {
#pragma unroll
            for (int q = 0; q < WPW; ++q) {
                const int iq = wIss + q;
                if (iq >= K) break;
                const int jq = __shfl_sync(0xffffffffu, jN, q);
                const double* ownRow = a.pos + (size_t)(own0 + iq) * D;
                const double* selRow = a.pos + (size_t)(other0 + jq) * D;
                double* ownS = reinterpret_cast<double*>(rows + (size_t)q * 2u * rowStride);
                double* selS = reinterpret_cast<double*>(rows + (size_t)q * 2u * rowStride + rowStride);
                for (int d = lane; d < D; d += 32) { ownS[d] = ownRow[d]; selS[d] = selRow[d]; }
            }
        }
        __syncwarp();
        stI = stI + 1 == S ? 0 : stI + 1;
        wIss += walkerStep;
    };

    prefetch();
    int issued = 0;
    for (; issued < P && issued < myTiles; ++issued) { issue(); prefetch(); }

    int stP = 0, wPart = firstWalker;                         /* stage / first walker of the tile whose partial sums are next */
    uint32_t parP = 0;                                        /* mbarrier phase parity of that stage's current use */
    int stF = 0, wFin = firstWalker;                          /* ... of the tile to finish next */
    int sendOff = bid * NW;                                   /* first walker of the tile being summed (send side) */
    for (int n = 0; n < myTiles + L; ++n) {
        if (issued < myTiles) { issue(); prefetch(); ++issued; }

        /* first look at the partial slots of tile n - L (pass 0 of the polls), taken before the arithmetic below */
        const bool finishing = n >= L;
        unsigned long long seen = DimEmpty;
        if (finishing && pollLane && pollQ < WPW && wFin + pollQ < K) seen = *reinterpret_cast<volatile unsigned long long*>(pollSlot0 + wFin + pollQ);

        /* ---- partial(n): proposal slices and their partial sums, one line of NW partials to every rank ---- */
        if (n < myTiles) {
            const unsigned char* rows = warpRows0 + (size_t)stP * stageBytes;
            const double* mrow = warpMeta0 + stP * (NW * 4);
            double s[WPW];
#pragma unroll
            for (int q = 0; q < WPW; ++q) {
                double p0 = 0.0, p1 = 0.0;
                if (wPart + q < K) {                                      /* warp-uniform */
                    const double zq = mrow[4 * q];
                    if (TMA) mbarWait(warpBar0 + 8u * (uint32_t)(stP * NW + q), parP);
                    const double* ownS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride);
                    const double* selS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride + rowStride);
                    for (int d = 2 * lane; d < D; d += 64) {
                        if (TMA || d + 1 < D) {
                            const double2 c = *reinterpret_cast<const double2*>(ownS + d);
                            const double2 t = *reinterpret_cast<const double2*>(selS + d);
                            p0 = __dadd_rn(p0, calc.term(stretchPoint(t.x, c.x, zq), d));
                            p1 = __dadd_rn(p1, calc.term(stretchPoint(t.y, c.y, zq), d + 1));
                        } else {
                            p0 = __dadd_rn(p0, calc.term(stretchPoint(selS[d], ownS[d], zq), d));
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
            double* mine = ctaPartial + (n & 1) * NW;
#pragma unroll
            for (int q = 0; q < WPW; ++q)
                if (lane == 0) mine[warp * WPW + q] = s[q];
            __syncthreads();
            if (!manyRanks) {
                if (sender && sendOff + sendK < K) *reinterpret_cast<volatile double*>(sendSlot0 + sendOff) = mine[sendK];
            } else {
                for (int t = threadIdx.x; t < a.world * NW; t += DimFusedThreads) {
                    const int r = t / NW, k = t - r * NW;
                    if (sendOff + k < K)
                        *reinterpret_cast<volatile double*>(a.dim_partial[r] + ((size_t)a.half * a.world + a.rank) * K + sendOff + k) = mine[k];
                }
            }
            sendOff += walkerStep;
            wPart += walkerStep;
            if (stP + 1 == S) { stP = 0; parP ^= 1u; } else ++stP;
        }

        /* ---- finish(n - L): rank-ordered sum, Metropolis test, commit from shared memory ---- */
        if (finishing) {
            const unsigned char* rows = warpRows0 + (size_t)stF * stageBytes;
            const double2* mrow = reinterpret_cast<const double2*>(warpMeta0 + stF * (NW * 4));
            for (int q0 = 0; q0 < WPW; q0 += qpp) {
                double part = 0.0;
                if (pollLane && q0 + pollQ < WPW && wFin + q0 + pollQ < K)
                    part = dimTakePartial(pollSlot0 + wFin + q0 + pollQ, q0 == 0 ? seen : DimEmpty, timeoutNs, a.flags);
                for (int q = q0; q < q0 + qpp && q < WPW; ++q) {
                    const int i = wFin + q;
                    if (i >= K) break;                                    /* warp-uniform */
                    const int w = own0 + i;
                    const int src = (q - q0) * a.world;
                    double sum = __shfl_sync(0xffffffffu, part, src);     /* rank order, every lane alike */
                    for (int r = 1; r < a.world; ++r) sum = __dadd_rn(sum, __shfl_sync(0xffffffffu, part, src + r));
                    const double2 ze = mrow[2 * q], sp = mrow[2 * q + 1];
                    const double z = ze.x, e = ze.y;
                    const double newp = calc.finish(sum);
                    const double diff = __dsub_rn(__dadd_rn(sp.x, newp), sp.y);
                    const bool accepted = e < diff;
                    if (accepted || a.chain_row) {
                        const double* ownS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride);
                        const double* selS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride + rowStride);
                        double* ownRow = a.pos + (size_t)w * D;
                        double* chainRow = a.chain_row ? a.chain_row + (size_t)w * D : nullptr;
                        for (int d = 2 * lane; d < D; d += 64) {
                            if (TMA || d + 1 < D) {
                                double2 v = *reinterpret_cast<const double2*>(ownS + d);
                                if (accepted) {
                                    const double2 t = *reinterpret_cast<const double2*>(selS + d);
                                    v.x = stretchPoint(t.x, v.x, z);
                                    v.y = stretchPoint(t.y, v.y, z);
                                }
                                if (TMA) {
                                    if (accepted) *reinterpret_cast<double2*>(ownRow + d) = v;
                                    if (chainRow) *reinterpret_cast<double2*>(chainRow + d) = v;
                                } else {
                                    if (accepted) { ownRow[d] = v.x; ownRow[d + 1] = v.y; }
                                    if (chainRow) { chainRow[d] = v.x; chainRow[d + 1] = v.y; }
                                }
                            } else {
                                double v = ownS[d];
                                if (accepted) { v = stretchPoint(selS[d], v, z); ownRow[d] = v; }
                                if (chainRow) chainRow[d] = v;
                            }
                        }
                    }
                    if (lane == 0) {
                        if (accepted) { a.logp[w] = newp; ++myAccepted; }
                        if (TRACE) {
                            if (i < a.trace_rows) {
                                mcmcpp_trace t;
                                t.j = a.dim_j[i]; t.accept = accepted ? 1 : 0; t.z = z; t.e = e; t.newp = newp; t.diff = diff;
                                a.trace[i] = t;
                            }
                        }
                    }
                }
            }
            __syncwarp();
            wFin += walkerStep;
            stF = stF + 1 == S ? 0 : stF + 1;
        }
    }
    if (lane == 0 && myAccepted) atomicAdd(a.accept_slots + bid, myAccepted);
}

template <class Calc, int WPW, bool TMA, bool TRACE, int S, int P>
__global__ void __launch_bounds__(DimFusedThreads) dimShardFused(const mcmcpp_halfstep_args a, const Calc calcIn, const uint32_t rowStride,
                                                                  const unsigned long long timeoutNs)
{
    dimShardFusedBody<Calc, WPW, TMA, TRACE, S, P>(a, calcIn, rowStride, timeoutNs, (int)blockIdx.x, (int)gridDim.x);
}

/* ---- all ranks of a DIM-sharded ensemble in ONE cooperative launch of one process (mcmcpp_sampler_run_group) ------------
 * The ranks' kernels wait on one another (each polls partial sums the others store), so as separate launches on one GPU
 * nothing would guarantee that they are resident together.  Here CTA b works as block b / R of rank b % R, on that rank's own
 * buffers, tables and column slice, with exactly the code of the one-rank kernel: the world > 1 protocol -- slot hand-over,
 * rank-ordered sums, replicated decisions, per-rank slices (even ones through the TMA ring, odd ones through plain loads) --
 * runs on a one-GPU box.  A cooperative launch either has every CTA resident or fails. */
enum { DimGroupMax = 4 };
template <class Calc>
struct DimGroupArgs
{
    int count;
    uint32_t rowStride[DimGroupMax];
    mcmcpp_halfstep_args args[DimGroupMax];
    alignas(16) unsigned char calc[DimGroupMax][(sizeof(Calc) + 15) / 16 * 16];      /* Calculators by value (they need not be default-constructible) */
};

enum { DimGroupStages = 4, DimGroupPrefetch = 2 };          /* the ring of the emulation kernel: small, so that all ranks fit */
template <class Calc, int WPW, bool TRACE>
__global__ void __launch_bounds__(DimFusedThreads) dimShardFusedGroup(const DimGroupArgs<Calc> g, const unsigned long long timeoutNs)
{
    const int r = (int)blockIdx.x % g.count, bid = (int)blockIdx.x / g.count, nblocks = (int)gridDim.x / g.count;
    const Calc& calc = *reinterpret_cast<const Calc*>(g.calc[r]);
    if ((g.args[r].D % 2) == 0)
        dimShardFusedBody<Calc, WPW, true, TRACE, DimGroupStages, DimGroupPrefetch>(g.args[r], calc, g.rowStride[r], timeoutNs, bid, nblocks);
    else
        dimShardFusedBody<Calc, WPW, false, TRACE, DimGroupStages, DimGroupPrefetch>(g.args[r], calc, g.rowStride[r], timeoutNs, bid, nblocks);
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

    template <int WPW, bool TMA, bool TRACE, int S, int P>
    static int fusedSP(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        constexpr int NW = DimFusedWarps * WPW;
        static const unsigned long long timeoutNs = (unsigned long long)envInt("MCMCPP_B200_RDV_TIMEOUT_MS", 20000) * 1000000ULL;
        const size_t stageBytes = (size_t)NW * 2u * rowStride;
        const size_t smem = (size_t)S * (stageBytes + 40u * NW) + 16u * NW;      /* rows + mbarrier + 4 scalars per walker slot */
        if (smem > 227u * 1024u) return MCMCPP_ERR_UNSUPPORTED;
        static DynamicSmemOptIn optIn;
        if (!optIn.ensure(dimShardFused<Calc, WPW, TMA, TRACE, S, P>, smem)) return MCMCPP_ERR_CUDA;
        int perSm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dimShardFused<Calc, WPW, TMA, TRACE, S, P>, DimFusedThreads, smem) != cudaSuccess || perSm < 1)
            perSm = 1;
        const int numTiles = (a.K + NW - 1) / NW;
        int grid = (a.sm_count > 0 ? a.sm_count : 148) * perSm;           /* persistent: everything resident at once */
        if (grid > numTiles) grid = numTiles;
        if (grid > a.max_grid) grid = a.max_grid;
        dimShardFused<Calc, WPW, TMA, TRACE, S, P><<<grid, DimFusedThreads, smem, stream>>>(a, calc, rowStride, timeoutNs);
        return launchStatus();
    }

    /* ring shapes: S stages = P tiles in flight + L waiting for the other ranks' partials + one being worked on */
    template <int WPW, bool TMA, bool TRACE>
    static int fused(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        constexpr int NW = DimFusedWarps * WPW;
        static const int ring = envInt("MCMCPP_B200_DIM_RING", 52);        /* two digits: stages, in flight */
        const size_t perStage = (size_t)NW * 2u * rowStride + 40u * NW;
        if (TMA && !TRACE) {
            if (ring == 42 && 4 * perStage <= 200u * 1024u) return fusedSP<WPW, TMA, TRACE, 4, 2>(a, calc, stream, rowStride);
            if (ring == 63 && 6 * perStage <= 200u * 1024u) return fusedSP<WPW, TMA, TRACE, 6, 3>(a, calc, stream, rowStride);
            if (ring == 74 && 7 * perStage <= 200u * 1024u) return fusedSP<WPW, TMA, TRACE, 7, 4>(a, calc, stream, rowStride);
        }
        if (5 * perStage <= 200u * 1024u) return fusedSP<WPW, TMA, TRACE, 5, 2>(a, calc, stream, rowStride);
        return fusedSP<WPW, TMA, TRACE, 3, 1>(a, calc, stream, rowStride);      /* wide slices: fewer stages */
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
        const size_t stageBytes = (size_t)NW * 2u * maxStride;
        const size_t smem = (size_t)DimGroupStages * (stageBytes + 40u * NW) + 16u * NW;
        if (smem > 227u * 1024u) return MCMCPP_ERR_UNSUPPORTED;
        static DynamicSmemOptIn optIn;
        if (!optIn.ensure(dimShardFusedGroup<Calc, WPW, TRACE>, smem)) return MCMCPP_ERR_CUDA;
        int perSm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dimShardFusedGroup<Calc, WPW, TRACE>, DimFusedThreads, smem) != cudaSuccess || perSm < 1)
            return MCMCPP_ERR_CUDA;
        const mcmcpp_halfstep_args& a0 = g.args[0];
        const int numTiles = (a0.K + NW - 1) / NW;
        int nblocks = ((a0.sm_count > 0 ? a0.sm_count : 148) * perSm) / g.count;      /* per rank; everything resident at once */
        if (nblocks > numTiles) nblocks = numTiles;
        if (nblocks > a0.max_grid) nblocks = a0.max_grid;
        if (nblocks < 1) return MCMCPP_ERR_UNSUPPORTED;
        unsigned long long t_ = timeoutNs;
        void* params[] = {(void*)&g, &t_};
        if (cudaLaunchCooperativeKernel((const void*)dimShardFusedGroup<Calc, WPW, TRACE>, dim3(nblocks * g.count), dim3(DimFusedThreads), params, smem,
                                        stream) != cudaSuccess)
            return MCMCPP_ERR_CUDA;
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
        const size_t perWalker = 2u * (size_t)maxStride, stage = 16u * 1024u;
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
        /* a tile stages ~16 KB (4 warps x WPW walkers x 2 rows) whatever the slice length is */
        static const size_t stage = (size_t)envInt("MCMCPP_B200_DIM_STAGE_KB", 16) * 1024;
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
