/*
 * include/mcmcpp_b200/dim_shard_kernels.cuh -- MCMCPP_SHARD_DIM: the stretch-move half-step with the PARAMETER
 * dimension sharded over the GPUs of one box (SURVEY.md s.8(e) alternative 2).
 *
 * Every rank holds all W walkers but only its columns [dim_begin, dim_begin + D) of each row, makes the SAME draws
 * (j, z, log(1-u)) from the sampler's pcg64 stream, builds its slice of every proposal
 * (MCMCpp/Movers/StretchMove.h:103-109) and the partial sum of a separable Calculator (term()/finish(), targets.h)
 * over its columns.  The K partial sums of a colour are exchanged INSIDE the update kernel: 16 walkers per CTA, so one
 * 128-byte line per peer is stored straight into every rank's partial table over NVLink (peer-mapped memory); a
 * stream-ordered rendezvous follows; the decision kernel adds the `world` partials in rank order -- the same
 * bits on every rank, so every rank takes the same Metropolis decision (StretchMove.h:112-122) -- and lists the
 * accepted walkers; the commit kernel rebuilds the accepted proposals' slices from the two rows (17 % of the
 * walkers at D = 1000; cheaper than keeping every proposal).  Traffic per rank is 1/world of the single-GPU kernel's, the exchange is 8 K bytes
 * per peer per colour whatever D is.
 *
 * The log-posterior is summed slice by slice (canonical 64-way order inside a slice, rank order across), so for
 * world > 1 it carries the tolerance tier (<= 1e-12 relative); with world == 1 it is the single-GPU order, bit-exact.
 */
#ifndef MCMCPP_B200_DIM_SHARD_KERNELS_CUH
#define MCMCPP_B200_DIM_SHARD_KERNELS_CUH

#include "mcmcpp_b200/stretch_kernels.cuh"

namespace MCMC
{
namespace Device
{

enum { DimWarps = 16, DimThreads = DimWarps * 32, DimBatch = 4 };

/* lane l owns elements 64c + 2l, 64c + 2l + 1 of a row slice (the canonical order of targets.h) */
template <bool VEC>
__device__ __forceinline__ double2 dimLoad2(const double* __restrict__ row, int d, int D)
{
    double2 v = make_double2(0.0, 0.0);
    if (VEC) {
        if (d < D) v = *reinterpret_cast<const double2*>(row + d);
    } else {
        if (d < D) v.x = row[d];
        if (d + 1 < D) v.y = row[d + 1];
    }
    return v;
}

template <bool VEC>
__device__ __forceinline__ void dimStore2(double* __restrict__ row, int d, int D, double2 v)
{
    if (VEC) {
        if (d < D) *reinterpret_cast<double2*>(row + d) = v;
    } else {
        if (d < D) row[d] = v.x;
        if (d + 1 < D) row[d + 1] = v.y;
    }
}

/* ---- in-kernel rendezvous (args.dim_epoch > 0) --------------------------------------------------------------------
 * Behind every rank's partial table sit 2 x world flag words.  The decide kernel runs after this rank's partial kernel in
 * stream order, so when it starts this rank's partials have landed everywhere: its first CTA stores the epoch into flag
 * [colour][this rank] of EVERY rank (NVLink stores, like the partials themselves), then every CTA waits until all `world`
 * flags of the colour have reached the epoch.  A rank can be at most one colour ahead of another (it needs the other's
 * flag to get further), and consecutive colours use different halves of the table and of the flags, so nothing is
 * overwritten while it is still being read.  (A fence + arrival counter at the end of the partial kernel instead costs
 * a system-scope fence per CTA, which waits for the CTA's NVLink stores: 2.5x slower at 2 GPUs.) */
__device__ __forceinline__ void dimRendezvous(const mcmcpp_halfstep_args& a)
{
    if (a.dim_epoch == 0 || a.world == 1) return;
    const size_t tableDoubles = 2 * (size_t)a.world * (size_t)a.K;
    if ((int)threadIdx.x < a.world) {
        if (blockIdx.x == 0) {
            volatile unsigned long long* theirs = reinterpret_cast<volatile unsigned long long*>(a.dim_partial[threadIdx.x] + tableDoubles);
            theirs[(size_t)a.half * a.world + a.rank] = a.dim_epoch;
        }
        const volatile unsigned long long* flag = reinterpret_cast<const volatile unsigned long long*>(a.dim_partial[a.rank] + tableDoubles)
                                                  + (size_t)a.half * a.world + threadIdx.x;
        while (*flag < a.dim_epoch) __nanosleep(64);
        __threadfence_system();
    }
    __syncthreads();
}

/* ---- phase 1: draws, proposal slice, partial sum, partial -> every rank's table ---------------------------- */
template <class Calc, bool VEC>
__global__ void __launch_bounds__(DimThreads) dimShardPartial(const mcmcpp_halfstep_args a, const Calc calcIn)
{
    __shared__ double ctaPartial[DimWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i = blockIdx.x * DimWarps + warp;           /* slot of this colour; DIM sharding owns every slot */
    const int D = a.D;
    double partial = 0.0;
    if (i < a.K) {
        Calc calc = calcIn;
        const int w = a.half * a.K + i;
        const int other0 = (1 - a.half) * a.K;
        WalkerRng rng = loadRng(a, w);
        StretchDraws dr = drawStretch(rng.state, rng.inc, (uint64_t)a.K, a.bound_threshold, a.bound_recip, gwOf(a));
        const double* ownRow = a.pos + (size_t)w * D;
        const double* selRow = a.pos + (size_t)(other0 + dr.j) * D;
        double p0 = 0.0, p1 = 0.0;
        for (int c0 = 0; c0 * 64 < D; c0 += DimBatch) {
            double2 cur[DimBatch], sel[DimBatch];
#pragma unroll
            for (int b = 0; b < DimBatch; ++b) {          /* all loads of the batch in flight before the first use */
                const int d = 64 * (c0 + b) + 2 * lane;
                cur[b] = dimLoad2<VEC>(ownRow, d, D);
                sel[b] = dimLoad2<VEC>(selRow, d, D);
            }
#pragma unroll
            for (int b = 0; b < DimBatch; ++b) {
                const int d = 64 * (c0 + b) + 2 * lane;
                if (d < D) p0 = __dadd_rn(p0, calc.term(stretchPoint(sel[b].x, cur[b].x, dr.z), d));
                if (d + 1 < D) p1 = __dadd_rn(p1, calc.term(stretchPoint(sel[b].y, cur[b].y, dr.z), d + 1));
            }
        }
        double s = __dadd_rn(p0, p1);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) s = __dadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
        partial = s;
        if (lane == 0) {
            storeRng(a, w, rng);
            if (dr.redraws) atomicAdd(a.flags, dr.redraws);
            a.dim_z[i] = dr.z; a.dim_e[i] = dr.logOneMinusU; a.dim_j[i] = dr.j;
        }
    }
    if (lane == 0) ctaPartial[warp] = partial;
    if (blockIdx.x == 0 && threadIdx.x == 0) *a.dim_count = 0;     /* the previous colour's commit is done (stream order) */
    __syncthreads();
    /* one 128-byte line per rank: thread t stores walker (t % 16)'s partial into rank (t / 16)'s table */
    const int slot0 = blockIdx.x * DimWarps;
    for (int t = threadIdx.x; t < a.world * DimWarps; t += DimThreads) {
        const int q = t / DimWarps, k = t - q * DimWarps;
        if (slot0 + k < a.K)
            a.dim_partial[q][((size_t)a.half * a.world + a.rank) * a.K + slot0 + k] = ctaPartial[k];
    }
}


/* ---- phase 0 (TMA form only): the draws of a colour, one THREAD per walker --------------------------------------
 * ~600 dependent integer/FP64 instructions per walker (three 128-bit LCG steps, the stream skip, two logs).  Made
 * inside the copy kernel they sit in front of the bulk copies and leave its shared memory idle for ~2 us per CTA;
 * here they run at full occupancy and the copy kernel only reads (j, z). */
static __global__ void __launch_bounds__(256) dimShardDraw(const mcmcpp_halfstep_args a)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= a.K) return;
    const int w = a.half * a.K + i;
    WalkerRng rng = loadRng(a, w);
    StretchDraws dr = drawStretch(rng.state, rng.inc, (uint64_t)a.K, a.bound_threshold, a.bound_recip, gwOf(a));
    storeRng(a, w, rng);
    if (dr.redraws) atomicAdd(a.flags, dr.redraws);
    a.dim_z[i] = dr.z; a.dim_e[i] = dr.logOneMinusU; a.dim_j[i] = dr.j;
    if (i == 0) *a.dim_count = 0;                              /* the previous colour's commit is done (stream order) */
}

/* ---- phase 1, TMA form (even slices): 4 warps per CTA, WPW walkers per warp all in flight at once ------------
 * Lane q < WPW reads walker q's draws (dimShardDraw) and issues that
 * walker's two cp.async.bulk copies (own slice, partner slice) onto the walker's own mbarrier; the warp then reduces
 * the walkers from shared memory.  WPW is chosen so that a CTA stages ~32 KB whatever the slice length is
 * (6-7 CTAs per SM: while some reduce, the others have their copies in flight). */
enum { DimBulkWarps = 4 };

template <class Calc, int WPW>
__global__ void __launch_bounds__(DimBulkWarps * 32) dimShardPartialBulk(const mcmcpp_halfstep_args a, const Calc calcIn, const uint32_t rowStride)
{
    extern __shared__ __align__(128) unsigned char dimSmem[];
    constexpr int NW = DimBulkWarps * WPW;                 /* walkers per CTA */
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D;
    const uint32_t rowBytes = (uint32_t)D * 8u;
    unsigned char* rows = dimSmem + (size_t)warp * WPW * 2u * rowStride;
    unsigned char* barBase = dimSmem + (size_t)NW * 2u * rowStride;
    double* ctaPartial = reinterpret_cast<double*>(barBase + 8u * NW);
    const int slot0 = blockIdx.x * NW;
    const int group0 = slot0 + warp * WPW;
    Calc calc = calcIn;

    double z = 1.0;
    if (lane < WPW && group0 + lane < a.K) {
        const int i = group0 + lane;
        const int w = a.half * a.K + i;
        const int other0 = (1 - a.half) * a.K;
        const int j = a.dim_j[i];
        const uint32_t bar = smemAddr(barBase + 8u * (warp * WPW + lane));
        mbarInit(bar, 1);
        mbarExpectTx(bar, 2u * rowBytes);
        bulkLoad(smemAddr(rows + (size_t)lane * 2u * rowStride), a.pos + (size_t)w * D, rowBytes, bar);
        bulkLoad(smemAddr(rows + (size_t)lane * 2u * rowStride + rowStride), a.pos + (size_t)(other0 + j) * D, rowBytes, bar);
        z = a.dim_z[i];
    }
    __syncwarp();
    /* partial sums of the warp's WPW walkers; the WPW butterflies run interleaved (independent dependency chains) */
    double s[WPW];
#pragma unroll
    for (int q = 0; q < WPW; ++q) {
        const double zq = __shfl_sync(0xffffffffu, z, q);
        double p0 = 0.0, p1 = 0.0;
        if (group0 + q < a.K) {                                  /* warp-uniform */
            mbarWait(smemAddr(barBase + 8u * (warp * WPW + q)), 0);
            const double* ownS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride);
            const double* selS = reinterpret_cast<const double*>(rows + (size_t)q * 2u * rowStride + rowStride);
            for (int d = 2 * lane; d < D; d += 64) {
                const double2 c = *reinterpret_cast<const double2*>(ownS + d);
                const double2 t = *reinterpret_cast<const double2*>(selS + d);
                p0 = __dadd_rn(p0, calc.term(stretchPoint(t.x, c.x, zq), d));
                p1 = __dadd_rn(p1, calc.term(stretchPoint(t.y, c.y, zq), d + 1));
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
    /* NW consecutive partials (64..256 bytes) to every rank's table */
    for (int t = threadIdx.x; t < a.world * NW; t += DimBulkWarps * 32) {
        const int r = t / NW, k = t - r * NW;
        if (slot0 + k < a.K)
            a.dim_partial[r][((size_t)a.half * a.world + a.rank) * a.K + slot0 + k] = ctaPartial[k];
    }
}

/* ---- phase 2a: rank-ordered sum and Metropolis test, one THREAD per walker; work list for the commit ---------
 * Accepted walkers are a minority (17 % at D = 1000), so the commit runs over a compacted list (warp-aggregated
 * append; the order of the list does not matter) instead of leaving 5 warps in 6 idle.  On stored steps every
 * walker is listed (its chain row must be written either way); bit 31 of an entry = accepted. */
template <class Calc, bool TRACE>
__global__ void __launch_bounds__(256) dimShardDecide(const mcmcpp_halfstep_args a, const Calc calcIn)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    bool accepted = false, listed = false;
    dimRendezvous(a);                                         /* every rank's partials of this colour have landed */
    if (i < a.K) {
        Calc calc = calcIn;
        const int w = a.half * a.K + i;
        const double* table = a.dim_partial[a.rank] + (size_t)a.half * a.world * a.K + i;
        double sum = table[0];
        for (int r = 1; r < a.world; ++r) sum = __dadd_rn(sum, table[(size_t)r * a.K]);
        const double z = a.dim_z[i], e = a.dim_e[i];
        const double scaling = __dmul_rn(log(z), (double)(a.D_total - 1));
        const double newp = calc.finish(sum);
        const double diff = __dsub_rn(__dadd_rn(scaling, newp), a.logp[w]);
        accepted = e < diff;
        listed = accepted || a.chain_row != nullptr;
        if (accepted) a.logp[w] = newp;
        if (TRACE) {
            if (i < a.trace_rows) {
                mcmcpp_trace t;
                t.j = a.dim_j[i]; t.accept = accepted ? 1 : 0; t.z = z; t.e = e; t.newp = newp; t.diff = diff;
                a.trace[i] = t;
            }
        }
    }
    /* CTA-aggregated append: ONE atomic per CTA on the list counter (a warp-aggregated append is 16,384 atomics on one
     * address per colour at K = 2^19, ~9 us of serialised L2 atomics in a 17 us kernel) */
    __shared__ int warpListed[8], warpBase[8], ctaBase;
    const unsigned ballot = __ballot_sync(0xffffffffu, listed);
    const int warp = threadIdx.x >> 5;
    if (lane == 0) warpListed[warp] = __popc(ballot);
    const int blockAccepted = __syncthreads_count(accepted);
    if (threadIdx.x == 0) {
        int sum = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) { warpBase[q] = sum; sum += warpListed[q]; }
        ctaBase = sum ? atomicAdd(a.dim_count, sum) : 0;
        if (blockAccepted) a.accept_slots[blockIdx.x] += (unsigned long long)blockAccepted;
    }
    __syncthreads();
    if (listed) a.dim_list[ctaBase + warpBase[warp] + __popc(ballot & ((1u << lane) - 1u))] = i | (accepted ? (int)0x80000000 : 0);
}

/* ---- phase 2b: commit the listed walkers' slices (persistent CTAs; vector loads, all of a warp's loads first) -- */
template <bool VEC>
__global__ void __launch_bounds__(DimThreads) dimShardCommit(const mcmcpp_halfstep_args a)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D;
    const int count = *a.dim_count;
    const int other0 = (1 - a.half) * a.K;
    for (int idx = blockIdx.x * DimWarps + warp; idx < count; idx += gridDim.x * DimWarps) {
        const int entry = a.dim_list[idx];
        const bool accepted = entry < 0;
        const int i = entry & 0x7fffffff;
        const int w = a.half * a.K + i;
        const double z = a.dim_z[i];
        double* ownRow = a.pos + (size_t)w * D;
        double* chainRow = a.chain_row ? a.chain_row + (size_t)w * D : nullptr;
        const double* selRow = a.pos + (size_t)(other0 + a.dim_j[i]) * D;
        for (int c0 = 0; c0 * 64 < D; c0 += 2 * DimBatch) {
            double2 cur[2 * DimBatch], sel[2 * DimBatch];
#pragma unroll
            for (int b = 0; b < 2 * DimBatch; ++b) {
                const int d = 64 * (c0 + b) + 2 * lane;
                cur[b] = dimLoad2<VEC>(ownRow, d, D);
                sel[b] = accepted ? dimLoad2<VEC>(selRow, d, D) : cur[b];
            }
#pragma unroll
            for (int b = 0; b < 2 * DimBatch; ++b) {
                const int d = 64 * (c0 + b) + 2 * lane;
                double2 v = cur[b];
                if (accepted) {
                    v.x = stretchPoint(sel[b].x, cur[b].x, z);
                    v.y = stretchPoint(sel[b].y, cur[b].y, z);
                    dimStore2<VEC>(ownRow, d, D, v);
                }
                if (chainRow) dimStore2<VEC>(chainRow, d, D, v);
            }
        }
    }
}

/* Short slices (<= 256 columns, even): a warp takes R = 8 / CPR listed walkers per pass so that 8 vector loads per
 * row array are in flight per lane whatever the slice length (one walker at a time leaves only 2 KB per warp in flight
 * at 126 columns). */
template <int CPR>
__global__ void __launch_bounds__(DimThreads) dimShardCommitShort(const mcmcpp_halfstep_args a)
{
    constexpr int R = 8 / CPR, SLOTS = R * CPR;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = a.D;
    const int count = *a.dim_count;
    const int other0 = (1 - a.half) * a.K;
    for (int idx0 = (blockIdx.x * DimWarps + warp) * R; idx0 < count; idx0 += gridDim.x * DimWarps * R) {
        int wRow[R];
        bool acc[R];
        double z[R];
        const double* selRow[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool live = idx0 + r < count;
            const int entry = live ? a.dim_list[idx0 + r] : 0;
            const int i = entry & 0x7fffffff;
            acc[r] = live && entry < 0;
            wRow[r] = live ? a.half * a.K + i : -1;
            z[r] = a.dim_z[i];
            selRow[r] = a.pos + (size_t)(other0 + a.dim_j[i]) * D;
        }
        double2 cur[SLOTS], sel[SLOTS];
#pragma unroll
        for (int b = 0; b < SLOTS; ++b) {
            const int r = b / CPR, d = 64 * (b % CPR) + 2 * lane;
            const bool ok = wRow[r] >= 0 && d < D;
            cur[b] = ok ? *reinterpret_cast<const double2*>(a.pos + (size_t)wRow[r] * D + d) : make_double2(0.0, 0.0);
            sel[b] = (ok && acc[r]) ? *reinterpret_cast<const double2*>(selRow[r] + d) : cur[b];
        }
#pragma unroll
        for (int b = 0; b < SLOTS; ++b) {
            const int r = b / CPR, d = 64 * (b % CPR) + 2 * lane;
            if (wRow[r] < 0 || d >= D) continue;
            double2 v = cur[b];
            if (acc[r]) {
                v.x = stretchPoint(sel[b].x, cur[b].x, z[r]);
                v.y = stretchPoint(sel[b].y, cur[b].y, z[r]);
                *reinterpret_cast<double2*>(a.pos + (size_t)wRow[r] * D + d) = v;
            }
            if (a.chain_row) *reinterpret_cast<double2*>(a.chain_row + (size_t)wRow[r] * D + d) = v;
        }
    }
}

/* Launcher of a DIM-sharded colour for a separable Calculator (a VIEW of this rank's columns: term(x, d) takes the
 * local column index).  Pass &DimShardLauncher<Calc>::launch to mcmcpp_sampler_create with shard_mode DIM. */
template <class Calc>
struct DimShardLauncher
{
    static_assert(IsSeparable<Calc>::value, "DIM sharding needs a separable Calculator: term(x, d) and finish(sum)");

    template <int WPW>
    static void partialBulk(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream, uint32_t rowStride)
    {
        constexpr int NW = DimBulkWarps * WPW;
        static bool attr = false;
        if (!attr) { cudaFuncSetAttribute(dimShardPartialBulk<Calc, WPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); attr = true; }
        const size_t smem = (size_t)NW * 2u * rowStride + 16u * NW;
        dimShardPartialBulk<Calc, WPW><<<(a.K + NW - 1) / NW, DimBulkWarps * 32, smem, stream>>>(a, calc, rowStride);
    }

    template <bool VEC>
    static int run(const mcmcpp_halfstep_args& a, const Calc& calc, cudaStream_t stream)
    {
        const int grid = (a.K + DimWarps - 1) / DimWarps;
        const uint32_t rowStride = ((uint32_t)a.D * 8u + 127u) & ~127u;
        const size_t perWalker = 2u * (size_t)rowStride;
        if (VEC && !bulkKernelDisabled() && DimBulkWarps * perWalker <= 226 * 1024) {
            dimShardDraw<<<(a.K + 255) / 256, 256, 0, stream>>>(a);
            if (a.launch_count) *a.launch_count += 1;
            static const size_t stage = [] { const char* e = getenv("MCMCPP_B200_DIM_STAGE_KB"); return (size_t)(e ? atoi(e) : 36) * 1024; }();
            if (DimBulkWarps * 8 * perWalker <= stage) partialBulk<8>(a, calc, stream, rowStride);
            else if (DimBulkWarps * 4 * perWalker <= stage) partialBulk<4>(a, calc, stream, rowStride);
            else if (DimBulkWarps * 2 * perWalker <= stage) partialBulk<2>(a, calc, stream, rowStride);
            else partialBulk<1>(a, calc, stream, rowStride);
        } else {
            dimShardPartial<Calc, VEC><<<grid, DimThreads, 0, stream>>>(a, calc);
        }
        if (a.world > 1 && a.dim_epoch == 0) {
            /* stream-ordered rendezvous (the in-kernel flags are off): every rank's partials have landed in every table */
            int rc = a.allreduce_sum ? a.allreduce_sum(a.allreduce_ctx, a.sync_word, 1, stream) : MCMCPP_ERR_STATE;
            if (rc != MCMCPP_OK) return rc;
        }
        const int dgrid = (a.K + 255) / 256;
        if (a.trace) dimShardDecide<Calc, true><<<dgrid, 256, 0, stream>>>(a, calc);
        else dimShardDecide<Calc, false><<<dgrid, 256, 0, stream>>>(a, calc);
        static int perSm = 0;                                       /* persistent grid: what is resident at once */
        if (perSm == 0) {
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dimShardCommit<VEC>, DimThreads, 0) != cudaSuccess || perSm < 1) perSm = 1;
        }
        int cgrid = (a.sm_count > 0 ? a.sm_count : 148) * perSm;
        if (cgrid > grid) cgrid = grid;
        const int cpr = (a.D + 63) / 64;
        if (VEC && cpr <= 4) {
            const int sgrid = cgrid > (a.sm_count > 0 ? a.sm_count : 148) ? (a.sm_count > 0 ? a.sm_count : 148) : cgrid;
            if (cpr == 1) dimShardCommitShort<1><<<sgrid, DimThreads, 0, stream>>>(a);
            else if (cpr == 2) dimShardCommitShort<2><<<sgrid, DimThreads, 0, stream>>>(a);
            else if (cpr == 3) dimShardCommitShort<3><<<sgrid, DimThreads, 0, stream>>>(a);
            else dimShardCommitShort<4><<<sgrid, DimThreads, 0, stream>>>(a);
        } else {
            dimShardCommit<VEC><<<cgrid, DimThreads, 0, stream>>>(a);
        }
        if (a.launch_count) *a.launch_count += 2;
        return launchStatus();
    }

    static int launch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
    {
        const mcmcpp_halfstep_args& a = *args;
        if (!a.dim_partial || !a.dim_z || !a.dim_list || !a.dim_count) return MCMCPP_ERR_STATE;
        const Calc& calc = *static_cast<const Calc*>(calculator);
        cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
        return (a.D % 2 == 0) ? run<true>(a, calc, stream) : run<false>(a, calc, stream);
    }
};

}
}
#endif
