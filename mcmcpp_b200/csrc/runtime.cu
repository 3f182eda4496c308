/*
 * mcmcpp_b200/csrc/runtime.cu -- the non-template runtime behind include/mcmcpp_b200/capi.h.
 *
 * Owns what the reference's EnsembleSampler owns (MCMCpp/EnsembleSampler.h:178-195: walkers, chain,
 * counters, sampling mode) as device buffers, and drives EnsembleSampler::runMCMC / performStep
 * (EnsembleSampler.h:284-310,341-360) as two kernel launches per ensemble step on one CUDA stream:
 *   walker state   pos[W][D] walker-major rows + logp[W]               (Walker/Walker.h:142-149)
 *   rng            one 128-bit pcg64 state per walker, pre-positioned  (Utility/MultiSampler.h:120)
 *   chain          2-slot device ring of step rows, each streamed to a pinned host array
 *                  [stored][W][D] on a side stream                      (Chain/Chain.h:93,204-236)
 *   counters       per-CTA accepted partial sums, reduced on query     (Walker/Walker.h:111-122)
 * The update kernels themselves are launched through the `launcher` function pointer supplied at
 * creation: built-in Calculators (builtin_targets.cu) or a user instantiation of
 * MCMC::Device::HalfStepLauncher<Calc> (include/mcmcpp_b200/stretch_kernels.cuh).
 */
#include <cuda_runtime.h>
#include <nccl.h>      /* types only: the library is bound at run time (see NcclApi) */
#include <dlfcn.h>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "mcmcpp_b200/capi.h"
#include "mcmcpp_b200/device_rng.cuh"
#include "builtin_targets.h"
#include "analysis.h"

using MCMC::Device::U128;

namespace
{

thread_local std::string g_lastError;

int fail(int code, const std::string& what)
{
    g_lastError = what;
    return code;
}

#define CUDA_TRY(expr)                                                                                \
    do {                                                                                              \
        cudaError_t err__ = (expr);                                                                   \
        if (err__ != cudaSuccess)                                                                     \
            return fail(MCMCPP_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(err__));      \
    } while (0)

/*
 * NCCL is bound with dlopen/dlsym the first time a communicator is needed, preferring a libnccl.so.2 that
 * is already in the process (a Python host that imported torch has torch's own NCCL loaded; linking a second
 * copy at load time would clash with it).  Single-GPU use never touches NCCL.
 */
struct NcclApi
{
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi& nccl()
{
    static NcclApi api;
    if (api.ok) return api;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return api;
    *(void**)&api.GetUniqueId = dlsym(h, "ncclGetUniqueId");
    *(void**)&api.CommInitRank = dlsym(h, "ncclCommInitRank");
    *(void**)&api.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void**)&api.AllGather = dlsym(h, "ncclAllGather");
    *(void**)&api.AllReduce = dlsym(h, "ncclAllReduce");
    *(void**)&api.GetErrorString = dlsym(h, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.GetErrorString;
    return api;
}

#define NCCL_TRY(expr)                                                                                \
    do {                                                                                              \
        if (!nccl().ok) return fail(MCMCPP_ERR_NCCL, "libnccl.so.2 could not be loaded");             \
        ncclResult_t err__ = (expr);                                                                  \
        if (err__ != ncclSuccess)                                                                     \
            return fail(MCMCPP_ERR_NCCL, std::string(#expr) + ": " + nccl().GetErrorString(err__));   \
    } while (0)

enum { ChainSlots = 2 };

/* every walker's pcg64 state at its first draw */
__global__ void initRngKernel(uint64_t* rngState, int W, int K, uint64_t seed, int rngMode)
{
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= W) return;
    MCMC::Device::Pcg64 g;
    if (rngMode == MCMCPP_RNG_PER_WALKER32) {
        uint64_t inc;                                         /* pcg32(seed, w): 64-bit state in the low word */
        g.state.lo = MCMC::Device::pcg32Seed(seed, (uint64_t)w, inc);
        g.state.hi = 0;
    } else if (rngMode == MCMCPP_RNG_PER_WALKER) {
        g = MCMC::Device::pcg64Seed(seed, (uint64_t)w);
    } else {
        /* sequential order inside step 0: red slots 0..K-1 then black slots 0..K-1, 3 draws each;
         * walker index w = h*K + i is exactly that order (EnsembleSampler.h:345-354) */
        g = MCMC::Device::pcg64Seed(seed, 0);
        MCMC::Device::Affine128 m = MCMC::Device::pcg64AdvanceMap(g.inc, 3ULL * (uint64_t)w, 0);
        g.state = MCMC::Device::applyAffine(m, g.state);
    }
    reinterpret_cast<ulonglong2*>(rngState)[w] = make_ulonglong2(g.state.lo, g.state.hi);
}

__global__ void sumSlotsKernel(const unsigned long long* slots, int n, unsigned long long* out)
{
    __shared__ unsigned long long partial[256];
    unsigned long long s = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += slots[i];
    partial[threadIdx.x] = s;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) partial[threadIdx.x] += partial[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0 && partial[0]) atomicAdd(out, partial[0]);     /* integer sum: order does not matter */
}

/* PEER sharding: device-side barrier over the ranks.  Behind every rank's position array sit `world` epoch words; thread
 * r stores this barrier's number into word [rank] of rank r (an NVLink store) and waits until word [r] of its own array has
 * reached it.  Runs after the update kernel in stream order, so this rank's rows are complete when its number goes out.
 * A wait that sees nothing for timeoutNs sets flags[1] (reported by mcmcpp_sampler_run) instead of spinning for ever. */
__global__ void peerBarrierKernel(const double* const* peerPos, size_t flagOffsetDoubles, int rank, int world, unsigned long long epoch,
                                  int32_t* flags, unsigned long long timeoutNs)
{
    const int r = threadIdx.x;
    if (r >= world) return;
    __threadfence_system();
    volatile unsigned long long* theirs = reinterpret_cast<volatile unsigned long long*>(const_cast<double*>(peerPos[r]) + flagOffsetDoubles);
    theirs[rank] = epoch;
    const volatile unsigned long long* mine = reinterpret_cast<const volatile unsigned long long*>(peerPos[rank] + flagOffsetDoubles) + r;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned spins = 0;
    while (*mine < epoch) {
        __nanosleep(64);
        if ((++spins & 1023u) == 0) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeoutNs) { atomicExch(flags + 1, 1); break; }
        }
    }
    __threadfence_system();
}

unsigned long long rendezvousTimeoutNs()
{
    static const unsigned long long ns = [] {
        const char* e = getenv("MCMCPP_B200_RDV_TIMEOUT_MS");
        return (unsigned long long)(e ? atoi(e) : 20000) * 1000000ULL;
    }();
    return ns;
}

}  // namespace

struct mcmcpp_sampler
{
    mcmcpp_config cfg;
    int W = 0, D = 0, K = 0;
    int Dl = 0, dimBegin = 0;              /* DIM sharding: local slice of the row (else Dl == D, dimBegin == 0) */
    int slotBegin = 0, slotCount = 0;      /* owned slots of each colour */
    mcmcpp_halfstep_fn launcher = nullptr;
    std::vector<char> calculator;
    double* dParams = nullptr;             /* built-in target parameters */
    void* builtinAux = nullptr;            /* built-in target scratch (dense kernels) */

    cudaStream_t stream = nullptr, copyStream = nullptr, sideStream = nullptr;     /* sideStream + sideEvent: handed to launchers (capi.h side_stream); the built-in ones do not use them */
    cudaEvent_t sideEvent[2] = {nullptr, nullptr};
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    cudaEvent_t slotReady[ChainSlots] = {nullptr, nullptr}, slotFree[ChainSlots] = {nullptr, nullptr};
    bool slotBusy[ChainSlots] = {false, false};

    double* dPos = nullptr;
    double* dLogp = nullptr;
    uint64_t* dRng = nullptr;
    unsigned long long* dAcceptSlots = nullptr;
    unsigned long long* dAcceptSum = nullptr;
    int32_t* dFlags = nullptr;
    mcmcpp_trace* dTrace = nullptr;
    double* dChain[ChainSlots] = {nullptr, nullptr};
    int maxGrid = 0;

    double* dStore = nullptr;              /* device-resident chain [devCapRows][W][Dl] (cfg.device_chain_bytes > 0) */
    long long devCapRows = 0, devRows = 0; /* rows it can hold / holds: the LAST devRows rows of the chain live here */
    double* hChain = nullptr;              /* pinned [capacityRows][W][D] */
    long long hChainRows = 0;              /* allocated rows */
    long long capRows = 0;                 /* byte cap in rows (0 = chain off) */
    long long stored = 0;                  /* storedSteps (EnsembleSampler.h:190) */

    bool subSampling = false;
    int interval = 1;
    bool initialised = false;
    unsigned long long acceptedBase = 0;   /* initial placement counts as accepted (Walker.h:162-170) */
    unsigned long long totalSteps = 0;     /* per-walker totals summed over owned walkers */
    long long ensembleSteps = 0;           /* ensemble steps since set_initial (trace indexing) */
    unsigned long long launches = 0;

    mcmcpp_halfstep_args base;             /* constant part of the launch arguments */

    ncclComm_t comm = nullptr;
    int32_t* dBarrier = nullptr;
    std::vector<void*> peerOpened;
    const double** dPeerPos = nullptr;
    double* dMoverState = nullptr;         /* [W][moverDoubles] */
    int moverDoubles = 0;
    /* PostStepAction with a device-resident step view (EnsembleSampler.h:356-359) */
    mcmcpp_step_fn stepFn = nullptr;
    unsigned long long peerEpoch = 0;      /* PEER: barriers issued so far (every rank makes the same calls) */
    bool peerNeedsBarrier = false;         /* PEER: rows changed by the host since the last barrier (set_initial / restore) */
    mcmcpp_fused_fn fusedFn = nullptr;     /* small ensembles: the whole runMCMC loop in one launch */
    mcmcpp_group_fn groupFn = nullptr;     /* DIM: all ranks of the ensemble in one cooperative launch (mcmcpp_sampler_run_group) */
    int groupDepth = 0;                    /* > 0 while mcmcpp_sampler_run_group drives this sampler (no cross-process rendezvous then) */
    bool tablesWired = false;              /* DIM: every rank's partial table is known (peer_import or peer_import_local) */
    double* dFusedChain = nullptr;         /* [fusedRows][W][D] rows filled by the fused kernel */
    long long fusedRows = 0;
    void* stepCtx = nullptr;
    /* on-the-fly CovarianceMatrix over the stored steps (analysis.cu) */
    mcmcpp_cov_accum* cov = nullptr;
    int covInterval = 1;
    long long covIndex = 0, covSteps = 0;
    /* DIM sharding */
    double* dDimPartial = nullptr;         /* [2][world][K] */
    double** dDimTable = nullptr;          /* [world] device table of every rank's partial table */
    double* dDimZ = nullptr;
    double* dDimE = nullptr;
    double* dDimS = nullptr;
    int32_t* dDimJ = nullptr;
};

namespace
{

size_t rowDoubles(const mcmcpp_sampler* s) { return (size_t)s->W * (size_t)s->D; }       /* host-facing step row */
size_t devRowDoubles(const mcmcpp_sampler* s) { return (size_t)s->W * (size_t)s->Dl; }   /* what this rank holds of it */

void dimSlice(int D, int rank, int world, int* begin, int* count)
{
    /* deal out PAIRS of columns: every slice but possibly the last is even (16-byte aligned rows) */
    const int pairs = D / 2, base = pairs / world, rem = pairs % world;
    const int b = 2 * (rank * base + std::min(rank, rem));
    int c = 2 * (base + (rank < rem ? 1 : 0));
    if (rank == world - 1) c = D - b;
    *begin = b; *count = c;
}

/* device slice <-> host-facing full-width rows (plain copy when the rank holds whole rows) */
cudaError_t copyToHostRows(mcmcpp_sampler* s, double* hostRow, const double* devRow, cudaStream_t st, cudaMemcpyKind kind = cudaMemcpyDeviceToHost)
{
    if (s->Dl == s->D) return cudaMemcpyAsync(hostRow, devRow, sizeof(double) * rowDoubles(s), kind, st);
    return cudaMemcpy2DAsync(hostRow + s->dimBegin, sizeof(double) * s->D, devRow, sizeof(double) * s->Dl, sizeof(double) * s->Dl,
                             (size_t)s->W, kind, st);
}

cudaError_t copyFromFullRows(mcmcpp_sampler* s, double* devRow, const double* fullRow, cudaStream_t st, cudaMemcpyKind kind)
{
    if (s->Dl == s->D) return cudaMemcpyAsync(devRow, fullRow, sizeof(double) * rowDoubles(s), kind, st);
    return cudaMemcpy2DAsync(devRow, sizeof(double) * s->Dl, fullRow + s->dimBegin, sizeof(double) * s->D, sizeof(double) * s->Dl,
                             (size_t)s->W, kind, st);
}

int ensureHostChain(mcmcpp_sampler* s, long long rowsNeeded)
{
    if (s->capRows == 0) return MCMCPP_OK;
    rowsNeeded = std::min(rowsNeeded, s->capRows);
    if (rowsNeeded <= s->hChainRows) return MCMCPP_OK;
    long long newRows = std::min(s->capRows, std::max(rowsNeeded, s->hChainRows * 2));
    double* fresh = nullptr;
    CUDA_TRY(cudaMallocHost(&fresh, sizeof(double) * rowDoubles(s) * (size_t)newRows));
    if (s->Dl != s->D) std::memset(fresh, 0, sizeof(double) * rowDoubles(s) * (size_t)newRows);   /* columns of other ranks read 0 */
    if (s->hChain) {
        std::memcpy(fresh, s->hChain, sizeof(double) * rowDoubles(s) * (size_t)std::min(s->stored, s->hChainRows));
        cudaFreeHost(s->hChain);
    }
    s->hChain = fresh;
    s->hChainRows = newRows;
    return MCMCPP_OK;
}

int ensureDeviceChain(mcmcpp_sampler* s)
{
    if (s->capRows == 0 || s->dChain[0]) return MCMCPP_OK;
    for (int i = 0; i < ChainSlots; ++i) {
        CUDA_TRY(cudaMalloc(&s->dChain[i], sizeof(double) * devRowDoubles(s)));
        /* rows of walkers this rank does not own stay zero so the host layer can sum shards */
        CUDA_TRY(cudaMemsetAsync(s->dChain[i], 0, sizeof(double) * devRowDoubles(s), s->stream));
    }
    return MCMCPP_OK;
}

long long keptRows(const mcmcpp_sampler* s) { return s->capRows > 0 ? std::min(s->stored, s->capRows) : 0; }

/* device-resident chain: move every row held in HBM to its place in the pinned host chain */
int spillDeviceRows(mcmcpp_sampler* s)
{
    if (!s->dStore || s->devRows == 0) return MCMCPP_OK;
    const long long hostRows = keptRows(s) - s->devRows;
    int rc = ensureHostChain(s, hostRows + s->devRows);
    if (rc != MCMCPP_OK) return rc;
    for (long long r = 0; r < s->devRows; ++r)
        CUDA_TRY(copyToHostRows(s, s->hChain + rowDoubles(s) * (size_t)(hostRows + r), s->dStore + devRowDoubles(s) * (size_t)r, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->devRows = 0;
    return MCMCPP_OK;
}

/* device-resident chain: the slot the next stored row goes to (spills first when the block is full) */
int nextDeviceSlot(mcmcpp_sampler* s, double** slot)
{
    if (s->devRows == s->devCapRows) {
        int rc = spillDeviceRows(s);
        if (rc != MCMCPP_OK) return rc;
    }
    *slot = s->dStore + devRowDoubles(s) * (size_t)s->devRows;
    return MCMCPP_OK;
}

int peerBarrier(mcmcpp_sampler* s)
{
    if (s->cfg.shard_mode == MCMCPP_SHARD_DIM)      /* the epoch words sit behind the partial table */
        peerBarrierKernel<<<1, 32, 0, s->stream>>>(s->dDimTable, 2 * (size_t)s->cfg.world * (size_t)s->K, s->cfg.rank, s->cfg.world, ++s->peerEpoch,
                                                   s->dFlags, rendezvousTimeoutNs());
    else
    peerBarrierKernel<<<1, 32, 0, s->stream>>>(s->dPeerPos, devRowDoubles(s), s->cfg.rank, s->cfg.world, ++s->peerEpoch, s->dFlags,
                                               rendezvousTimeoutNs());
    CUDA_TRY(cudaGetLastError());
    ++s->launches;
    return MCMCPP_OK;
}

/* one colour: kernel launch (+ the exchange that makes the updated half visible to every rank) */
int halfStep(mcmcpp_sampler* s, int half, double* chainRow)
{
    mcmcpp_halfstep_args a = s->base;
    a.half = half;
    a.chain_row = chainRow;
    a.trace = nullptr;
    a.trace_rows = 0;
    if (s->dTrace) {
        long long first = s->ensembleSteps * (long long)s->W + (long long)half * s->K;
        long long rows = (long long)s->cfg.trace_updates - first;
        if (rows > 0) {
            a.trace = s->dTrace + first;
            a.trace_rows = (int32_t)std::min<long long>(rows, s->K);
        }
    }
    int rc = s->launcher(&a, s->calculator.data(), s->stream);
    if (rc != MCMCPP_OK) return fail(rc, std::string("half-step launch failed: ") + cudaGetErrorString(cudaGetLastError()));
    ++s->launches;
    if (s->cfg.world > 1) {
        if (s->cfg.shard_mode == MCMCPP_SHARD_ALLGATHER) {
            /* in-place all-gather of the colour just updated: contiguous rows [half*K, half*K+K) */
            double* halfBase = s->dPos + (size_t)half * s->K * s->D;
            NCCL_TRY(nccl().AllGather(halfBase + (size_t)s->slotBegin * s->D, halfBase, (size_t)s->slotCount * s->D,
                                   ncclDouble, s->comm, s->stream));
        } else if (s->cfg.shard_mode == MCMCPP_SHARD_PEER) {
            /* nobody starts the other colour before everyone has finished this one */
            int rc2 = peerBarrier(s);
            if (rc2 != MCMCPP_OK) return rc2;
        }
    }
    return MCMCPP_OK;
}

/* mcmcpp_allreduce_fn of DATA sharding: in-place ncclAllReduce(sum) on the launcher's stream */
int allreduceSum(void* ctx, double* devBuf, size_t count, void* cudaStream)
{
    mcmcpp_sampler* s = static_cast<mcmcpp_sampler*>(ctx);
    NCCL_TRY(nccl().AllReduce(devBuf, devBuf, count, ncclDouble, ncclSum, s->comm, static_cast<cudaStream_t>(cudaStream)));
    return MCMCPP_OK;
}

int appendRowFromDeviceSlot(mcmcpp_sampler* s, int slot)
{
    /* side stream: wait for the black half-step, copy the row to its place in the pinned host chain */
    CUDA_TRY(cudaEventRecord(s->slotReady[slot], s->stream));
    CUDA_TRY(cudaStreamWaitEvent(s->copyStream, s->slotReady[slot], 0));
    CUDA_TRY(copyToHostRows(s, s->hChain + rowDoubles(s) * (size_t)s->stored, s->dChain[slot], s->copyStream));
    CUDA_TRY(cudaEventRecord(s->slotFree[slot], s->copyStream));
    s->slotBusy[slot] = true;
    return MCMCPP_OK;
}

int createCommon(const mcmcpp_config* cfg, mcmcpp_sampler** out)
{
    if (!cfg || !out) return fail(MCMCPP_ERR_ARG, "null argument");
    const int W = cfg->num_walkers, D = cfg->num_params;
    /* the reference's ctor asserts (EnsembleSampler.h:207-208) */
    if (W <= 0 || D <= 0 || (W % 2) != 0 || W <= 2 * D) return fail(MCMCPP_ERR_ARG, "numWalkers must be even and exceed 2*numParams");
    if (cfg->alpha_num <= 0 || cfg->alpha_denom <= 0) return fail(MCMCPP_ERR_ARG, "alpha must be positive");
    if (cfg->rng_mode < MCMCPP_RNG_SEQUENTIAL || cfg->rng_mode > MCMCPP_RNG_PER_WALKER32) return fail(MCMCPP_ERR_ARG, "unknown rng mode");
    if (cfg->chain_block_steps < 0) return fail(MCMCPP_ERR_ARG, "chain_block_steps must be >= 0");
    if (cfg->world < 1 || cfg->rank < 0 || cfg->rank >= cfg->world) return fail(MCMCPP_ERR_ARG, "bad rank/world");
    if (cfg->shard_mode == MCMCPP_SHARD_DIM && D / 2 < cfg->world)
        return fail(MCMCPP_ERR_ARG, "DIM sharding needs at least two columns per GPU");
    if (cfg->shard_mode != MCMCPP_SHARD_DATA && cfg->shard_mode != MCMCPP_SHARD_DIM && (W / 2) % cfg->world != 0)
        return fail(MCMCPP_ERR_ARG, "walkers per colour must divide evenly over the GPUs");
    if (cfg->world > 1 && cfg->shard_mode == MCMCPP_SHARD_NONE) return fail(MCMCPP_ERR_ARG, "world > 1 needs a shard mode");

    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(MCMCPP_ERR_CUDA, "no CUDA device: mcmcpp_b200 has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= count) return fail(MCMCPP_ERR_ARG, "bad device ordinal");
    CUDA_TRY(cudaSetDevice(cfg->device));

    mcmcpp_sampler* s = new mcmcpp_sampler();
    s->cfg = *cfg;
    s->W = W; s->D = D; s->K = W / 2;
    s->Dl = D; s->dimBegin = 0;
    if (cfg->shard_mode == MCMCPP_SHARD_DIM) dimSlice(D, cfg->rank, cfg->world, &s->dimBegin, &s->Dl);
    if (cfg->shard_mode == MCMCPP_SHARD_DATA || cfg->shard_mode == MCMCPP_SHARD_DIM) {   /* replicated walkers: every rank updates all */
        s->slotCount = s->K;
        s->slotBegin = 0;
    } else {
        s->slotCount = s->K / cfg->world;
        s->slotBegin = cfg->rank * s->slotCount;
    }
    s->capRows = (long long)(cfg->max_chain_bytes / (sizeof(double) * (unsigned long long)W * (unsigned long long)D));
    if (cfg->chain_block_steps > 0 && cfg->max_chain_bytes > 0) {
        /* the reference's granularity: maxBlocks = maxSize / (sizeof * numParams * BlockSize * numWalkers) whole blocks, and the
         * first block is allocated whatever maxBlocks says (Chain/Chain.h:178-186,218-233) */
        const unsigned long long blockBytes = sizeof(double) * (unsigned long long)W * (unsigned long long)D * (unsigned long long)cfg->chain_block_steps;
        const unsigned long long blocks = std::max<unsigned long long>(1ULL, cfg->max_chain_bytes / blockBytes);
        s->capRows = (long long)(blocks * (unsigned long long)cfg->chain_block_steps);
    }
    if (cfg->device_chain_bytes > 0) {
        const unsigned long long devRowBytes = sizeof(double) * (unsigned long long)W * (unsigned long long)s->Dl;
        s->devCapRows = (long long)std::max<unsigned long long>(1ULL, cfg->device_chain_bytes / devRowBytes);
        if (s->capRows == 0) s->capRows = s->devCapRows;            /* no host cap given: the device block is the chain */
        s->devCapRows = std::min(s->devCapRows, s->capRows);
    }
    *out = s;   /* destroy() cleans up on later failure */

    CUDA_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&s->copyStream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreate(&s->evStart));
    CUDA_TRY(cudaEventCreate(&s->evStop));
    for (int i = 0; i < ChainSlots; ++i) {
        CUDA_TRY(cudaEventCreateWithFlags(&s->slotReady[i], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&s->slotFree[i], cudaEventDisableTiming));
    }
    s->maxGrid = s->K;   /* upper bound for any launcher: one CTA per walker */
    /* PEER: `world` barrier epoch words sit behind the rows, inside the one allocation the IPC handle exports */
    const size_t posTail = cfg->shard_mode == MCMCPP_SHARD_PEER ? (size_t)cfg->world + 2 : 0;
    CUDA_TRY(cudaMalloc(&s->dPos, sizeof(double) * (devRowDoubles(s) + posTail)));
    if (posTail) CUDA_TRY(cudaMemset(s->dPos + devRowDoubles(s), 0, sizeof(double) * posTail));
    CUDA_TRY(cudaMalloc(&s->dLogp, sizeof(double) * (size_t)W));
    CUDA_TRY(cudaMalloc(&s->dRng, sizeof(uint64_t) * 2 * (size_t)W));
    CUDA_TRY(cudaMalloc(&s->dAcceptSlots, sizeof(unsigned long long) * (size_t)s->maxGrid));
    CUDA_TRY(cudaMalloc(&s->dAcceptSum, sizeof(unsigned long long)));
    CUDA_TRY(cudaMalloc(&s->dFlags, sizeof(int32_t) * 4));
    CUDA_TRY(cudaMalloc(&s->dBarrier, sizeof(int32_t)));
    CUDA_TRY(cudaMemset(s->dAcceptSlots, 0, sizeof(unsigned long long) * (size_t)s->maxGrid));
    CUDA_TRY(cudaMemset(s->dFlags, 0, sizeof(int32_t) * 4));
    CUDA_TRY(cudaMemset(s->dBarrier, 0, sizeof(int32_t)));
    CUDA_TRY(cudaMemset(s->dPos, 0, sizeof(double) * devRowDoubles(s)));
    if (s->devCapRows > 0) {
        CUDA_TRY(cudaMalloc(&s->dStore, sizeof(double) * devRowDoubles(s) * (size_t)s->devCapRows));
        CUDA_TRY(cudaMemset(s->dStore, 0, sizeof(double) * devRowDoubles(s) * (size_t)s->devCapRows));   /* unowned walkers' rows read 0 */
    }
    s->moverDoubles = cfg->mover_state_doubles > 0 ? cfg->mover_state_doubles : 0;
    if (s->moverDoubles) {
        CUDA_TRY(cudaMalloc(&s->dMoverState, sizeof(double) * (size_t)W * (size_t)s->moverDoubles));
        CUDA_TRY(cudaMemset(s->dMoverState, 0, sizeof(double) * (size_t)W * (size_t)s->moverDoubles));
    }
    /* draw scratch (z, log(1-u), j of the colour in flight): kernels whose draws run in their own launch */
    CUDA_TRY(cudaMalloc(&s->dDimZ, sizeof(double) * 2 * (size_t)s->K));      /* 2K each: room to double-buffer by colour */
    CUDA_TRY(cudaMalloc(&s->dDimE, sizeof(double) * 2 * (size_t)s->K));
    CUDA_TRY(cudaMalloc(&s->dDimJ, sizeof(int32_t) * 2 * (size_t)s->K));
    CUDA_TRY(cudaMalloc(&s->dDimS, sizeof(double) * 2 * (size_t)s->K));
    if (cfg->shard_mode == MCMCPP_SHARD_DIM) {
        /* partial table [2 colours][world][K]; every slot starts EMPTY (all-ones, dim_shard_kernels.cuh) */
        const size_t tableDoubles = 2 * (size_t)cfg->world * (size_t)s->K, tail = (size_t)cfg->world + 2;   /* + epoch words of the start-of-run rendezvous */
        CUDA_TRY(cudaMalloc(&s->dDimPartial, sizeof(double) * (tableDoubles + tail)));
        CUDA_TRY(cudaMemset(s->dDimPartial, 0xFF, sizeof(double) * tableDoubles));
        CUDA_TRY(cudaMemset(s->dDimPartial + tableDoubles, 0, sizeof(double) * tail));
        CUDA_TRY(cudaMalloc(&s->dDimTable, sizeof(double*) * (size_t)cfg->world));
        if (cfg->world == 1) CUDA_TRY(cudaMemcpy(s->dDimTable, &s->dDimPartial, sizeof(double*), cudaMemcpyHostToDevice));
        CUDA_TRY(cudaStreamCreateWithFlags(&s->sideStream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) CUDA_TRY(cudaEventCreateWithFlags(&s->sideEvent[i], cudaEventDisableTiming));
    }
    if (cfg->trace_updates > 0) {
        CUDA_TRY(cudaMalloc(&s->dTrace, sizeof(mcmcpp_trace) * (size_t)cfg->trace_updates));
        CUDA_TRY(cudaMemset(s->dTrace, 0, sizeof(mcmcpp_trace) * (size_t)cfg->trace_updates));
    }

    /* constant launch arguments */
    mcmcpp_halfstep_args& a = s->base;
    std::memset(&a, 0, sizeof a);
    a.pos = s->dPos; a.logp = s->dLogp; a.rng_state = s->dRng;
    a.accept_slots = s->dAcceptSlots; a.flags = s->dFlags;
    a.W = W; a.D = s->Dl; a.K = s->K;
    a.D_total = D; a.dim_begin = s->dimBegin; a.rank = cfg->rank; a.world = cfg->world;
    a.dim_partial = s->dDimTable; a.dim_z = s->dDimZ; a.dim_e = s->dDimE; a.dim_j = s->dDimJ; a.dim_s = s->dDimS;
    a.mover_state = s->dMoverState;
    a.side_stream = s->sideStream; a.side_event[0] = s->sideEvent[0]; a.side_event[1] = s->sideEvent[1];
    { int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device); a.sm_count = sms; }
    a.slot_begin = s->slotBegin; a.slot_count = s->slotCount; a.slots_per_rank = s->slotCount;
    a.rng_mode = cfg->rng_mode;
    const uint64_t K64 = (uint64_t)s->K;
    a.bound_threshold = (0ULL - K64) % K64;                                    /* pcg_extras.hpp:517 */
    if (cfg->rng_mode == MCMCPP_RNG_PER_WALKER32) a.bound_threshold = (uint64_t)((0u - (uint32_t)K64) % (uint32_t)K64);   /* the same in 32 bits */
    a.bound_recip = (K64 == 1) ? ~0ULL : (uint64_t)((((unsigned __int128)1) << 64) / K64);
    U128 inc; inc.lo = 1; inc.hi = 0;                                          /* stream 0 */
    MCMC::Device::Affine128 skip = MCMC::Device::pcg64AdvanceMap(inc, 3ULL * (uint64_t)W - 3ULL, 0);
    a.skip_mult[0] = skip.mult.lo; a.skip_mult[1] = skip.mult.hi;
    a.skip_plus[0] = skip.plus.lo; a.skip_plus[1] = skip.plus.hi;
    MCMC::Device::GwConstants gw = MCMC::Device::makeGwConstants(cfg->alpha_num, cfg->alpha_denom);
    a.gw_term1 = gw.term1; a.gw_inv_sqrt_alpha = gw.invSqrtAlpha;
    a.max_grid = s->maxGrid;
    a.launch_count = &s->launches;
    /* The engine is seeded ONCE, here, as the reference seeds its mover in the sampler's constructor (EnsembleSampler.h:217);
     * setInitialWalkerPos never touches it (EnsembleSampler.h:220-230), so a second placement continues the stream. */
    initRngKernel<<<(s->W + 255) / 256, 256, 0, s->stream>>>(s->dRng, s->W, s->K, (uint64_t)cfg->seed, cfg->rng_mode);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int afterInitialState(mcmcpp_sampler* s, const double* hostRowOrNull)
{
    /* counters as after Walker::setFirstPoint (Walker.h:76,162-170): one more accepted step per walker, nothing reset;
     * the engine and the mover's state keep going (the reference's setInitialWalkerPos touches neither) */
    s->peerNeedsBarrier = true;
    const unsigned long long owned = 2ULL * (unsigned long long)s->slotCount;
    s->acceptedBase += owned;
    s->totalSteps += owned;
    s->ensembleSteps = 0;
    /* chain row for the initial placement + incrementChainStep (EnsembleSampler.h:228-229) */
    if (s->capRows > 0 && s->stored < s->capRows) {
        if (s->dStore) {
            double* slot = nullptr;
            int rc = nextDeviceSlot(s, &slot);
            if (rc != MCMCPP_OK) return rc;
            CUDA_TRY(cudaMemcpyAsync(slot, s->dPos, sizeof(double) * devRowDoubles(s), cudaMemcpyDeviceToDevice, s->stream));
            ++s->devRows;
        } else {
            int rc = ensureHostChain(s, s->stored + 1);
            if (rc != MCMCPP_OK) return rc;
            double* dst = s->hChain + rowDoubles(s) * (size_t)s->stored;
            if (hostRowOrNull && s->Dl == s->D) std::memcpy(dst, hostRowOrNull, sizeof(double) * rowDoubles(s));
            else CUDA_TRY(copyToHostRows(s, dst, s->dPos, s->stream));
        }
    }
    ++s->stored;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->initialised = true;
    return MCMCPP_OK;
}

}  // namespace

/* both logs on the same inputs (mcmcpp_device_log_check) */
static __global__ void logCheckKernel(const double* x, size_t n, double* kernelLog, double* cudaLog)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    kernelLog[i] = MCMC::Device::logOfNormal(x[i]);
    cudaLog[i] = log(x[i]);
}

/* internal (analysis.h): lets the other translation units of the library leave a message for mcmcpp_last_error */
int mcmcpp_internal_fail(int code, const char* what) { return fail(code, what ? what : ""); }

extern "C" {

const char* mcmcpp_last_error(void) { return g_lastError.c_str(); }

int mcmcpp_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    return count;
}

int mcmcpp_device_log_check(int device, const double* x, size_t n, double* kernel_log, double* cuda_log)
{
    if (!x || !kernel_log || !cuda_log || n == 0) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(device));
    double* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, sizeof(double) * 3 * n));
    int rc = MCMCPP_OK;
    if (cudaMemcpy(d, x, sizeof(double) * n, cudaMemcpyHostToDevice) != cudaSuccess) rc = MCMCPP_ERR_CUDA;
    if (rc == MCMCPP_OK) {
        logCheckKernel<<<(unsigned)((n + 255) / 256), 256>>>(d, n, d + n, d + 2 * n);
        if (cudaMemcpy(kernel_log, d + n, sizeof(double) * n, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(cuda_log, d + 2 * n, sizeof(double) * n, cudaMemcpyDeviceToHost) != cudaSuccess) rc = MCMCPP_ERR_CUDA;
    }
    cudaFree(d);
    return rc == MCMCPP_OK ? rc : fail(rc, std::string("log check failed: ") + cudaGetErrorString(cudaGetLastError()));
}

int mcmcpp_sampler_create(const mcmcpp_config* cfg, mcmcpp_halfstep_fn launcher, const void* calculator, size_t calc_bytes,
                          mcmcpp_sampler** out)
{
    if (!launcher || (!calculator && calc_bytes)) return fail(MCMCPP_ERR_ARG, "null launcher/calculator");
    int rc = createCommon(cfg, out);
    if (rc != MCMCPP_OK) { if (out && *out) { mcmcpp_sampler_destroy(*out); *out = nullptr; } return rc; }
    (*out)->launcher = launcher;
    (*out)->calculator.assign((const char*)calculator, (const char*)calculator + calc_bytes);
    if ((*out)->calculator.empty()) (*out)->calculator.resize(1);
    return MCMCPP_OK;
}

int mcmcpp_sampler_create_builtin(const mcmcpp_config* cfg, int target_kind, const double* params, size_t num_params_values,
                                  mcmcpp_sampler** out)
{
    int rc = createCommon(cfg, out);
    if (rc != MCMCPP_OK) { if (out && *out) { mcmcpp_sampler_destroy(*out); *out = nullptr; } return rc; }
    mcmcpp_sampler* s = *out;
    if (num_params_values) {
        if (!params) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(MCMCPP_ERR_ARG, "null params"); }
        cudaError_t e = cudaMalloc(&s->dParams, sizeof(double) * num_params_values);
        if (e == cudaSuccess) e = cudaMemcpy(s->dParams, params, sizeof(double) * num_params_values, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(MCMCPP_ERR_CUDA, cudaGetErrorString(e)); }
    }
    rc = mcmcpp_builtin_bind(target_kind, s->W, s->D, params, num_params_values, s->dParams, &s->launcher, &s->calculator,
                             &s->builtinAux, cfg->shard_mode, s->dimBegin, s->Dl);
    if (rc != MCMCPP_OK) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(rc, "unknown built-in target or bad parameter count"); }
    s->fusedFn = mcmcpp_builtin_fused(target_kind, s->D, cfg->shard_mode);
    s->groupFn = mcmcpp_builtin_group(target_kind, cfg->shard_mode);
    return MCMCPP_OK;
}

int mcmcpp_sampler_create_builtin_mover(const mcmcpp_config* cfg, int mover_kind, const double* mover_params, size_t num_mover_params,
                                        int target_kind, const double* params, size_t num_params_values, mcmcpp_sampler** out)
{
    if (mover_kind == MCMCPP_MOVER_STRETCH) return mcmcpp_sampler_create_builtin(cfg, target_kind, params, num_params_values, out);
    if (mover_kind != MCMCPP_MOVER_DIFF_EVOLUTION && mover_kind != MCMCPP_MOVER_WALK) return fail(MCMCPP_ERR_UNSUPPORTED, "unknown mover");
    if (!cfg) return fail(MCMCPP_ERR_ARG, "null argument");
    if (cfg->rng_mode != MCMCPP_RNG_PER_WALKER)
        return fail(MCMCPP_ERR_UNSUPPORTED, "this mover draws a variable number of values per update: per-walker streams only");
    if (cfg->shard_mode == MCMCPP_SHARD_DIM || cfg->shard_mode == MCMCPP_SHARD_DATA)
        return fail(MCMCPP_ERR_UNSUPPORTED, "this mover: walker sharding (allgather / peer) only");
    mcmcpp_config withState = *cfg;
    if (mover_kind == MCMCPP_MOVER_WALK) withState.mover_state_doubles = 2;       /* the normal distribution's saved value + flag */
    cfg = &withState;
    int rc = createCommon(cfg, out);
    if (rc != MCMCPP_OK) { if (out && *out) { mcmcpp_sampler_destroy(*out); *out = nullptr; } return rc; }
    mcmcpp_sampler* s = *out;
    if (num_params_values) {
        if (!params) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(MCMCPP_ERR_ARG, "null params"); }
        cudaError_t e = cudaMalloc(&s->dParams, sizeof(double) * num_params_values);
        if (e == cudaSuccess) e = cudaMemcpy(s->dParams, params, sizeof(double) * num_params_values, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(MCMCPP_ERR_CUDA, cudaGetErrorString(e)); }
    }
    rc = mcmcpp_builtin_bind_mover(mover_kind, target_kind, s->W, s->D, mover_params, num_mover_params, params, num_params_values, s->dParams,
                                   &s->launcher, &s->calculator, &s->fusedFn);
    if (rc != MCMCPP_OK) { mcmcpp_sampler_destroy(s); *out = nullptr; return fail(rc, "mover: unsupported target or bad parameters"); }
    return MCMCPP_OK;
}

int mcmcpp_sampler_destroy(mcmcpp_sampler* s)
{
    if (!s) return MCMCPP_OK;
    cudaSetDevice(s->cfg.device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->copyStream) cudaStreamSynchronize(s->copyStream);
    if (s->sideStream) { cudaStreamSynchronize(s->sideStream); cudaStreamDestroy(s->sideStream); }
    for (int i = 0; i < 2; ++i) if (s->sideEvent[i]) cudaEventDestroy(s->sideEvent[i]);
    if (s->comm && nccl().ok) nccl().CommDestroy(s->comm);
    for (void* p : s->peerOpened) cudaIpcCloseMemHandle(p);
    cudaFree(s->dPeerPos);
    cudaFree(s->dMoverState);
    cudaFree(s->dFusedChain);
    cudaFree(s->dStore);
    cudaFree(s->dDimPartial); cudaFree(s->dDimTable); cudaFree(s->dDimZ); cudaFree(s->dDimE); cudaFree(s->dDimS); cudaFree(s->dDimJ);
    cudaFree(s->dPos); cudaFree(s->dLogp); cudaFree(s->dRng); cudaFree(s->dAcceptSlots); cudaFree(s->dAcceptSum);
    cudaFree(s->dFlags); cudaFree(s->dTrace); cudaFree(s->dParams); cudaFree(s->dBarrier);
    mcmcpp_builtin_release(s->builtinAux);
    mcmcpp_cov_destroy(s->cov);
    for (int i = 0; i < ChainSlots; ++i) {
        cudaFree(s->dChain[i]);
        if (s->slotReady[i]) cudaEventDestroy(s->slotReady[i]);
        if (s->slotFree[i]) cudaEventDestroy(s->slotFree[i]);
    }
    if (s->hChain) cudaFreeHost(s->hChain);
    if (s->evStart) cudaEventDestroy(s->evStart);
    if (s->evStop) cudaEventDestroy(s->evStop);
    if (s->stream) cudaStreamDestroy(s->stream);
    if (s->copyStream) cudaStreamDestroy(s->copyStream);
    delete s;
    return MCMCPP_OK;
}

int mcmcpp_sampler_set_initial(mcmcpp_sampler* s, const double* positions_host, const double* aux_host)
{
    if (!s || !positions_host || !aux_host) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(copyFromFullRows(s, s->dPos, positions_host, s->stream, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpyAsync(s->dLogp, aux_host, sizeof(double) * (size_t)s->W, cudaMemcpyHostToDevice, s->stream));
    return afterInitialState(s, positions_host);
}

int mcmcpp_sampler_set_initial_device(mcmcpp_sampler* s, const double* positions_dev, const double* aux_dev)
{
    if (!s || !positions_dev || !aux_dev) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(copyFromFullRows(s, s->dPos, positions_dev, s->stream, cudaMemcpyDeviceToDevice));
    CUDA_TRY(cudaMemcpyAsync(s->dLogp, aux_dev, sizeof(double) * (size_t)s->W, cudaMemcpyDeviceToDevice, s->stream));
    return afterInitialState(s, nullptr);
}

/* the rank's own slice of the rows, contiguous [W][D_local] (DIM sharding; otherwise the same as the full-row calls) */
int mcmcpp_sampler_set_initial_local(mcmcpp_sampler* s, const double* positions_local_host, const double* aux_host)
{
    if (!s || !positions_local_host || !aux_host) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(cudaMemcpyAsync(s->dPos, positions_local_host, sizeof(double) * devRowDoubles(s), cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->dLogp, aux_host, sizeof(double) * (size_t)s->W, cudaMemcpyHostToDevice, s->stream));
    return afterInitialState(s, s->Dl == s->D ? positions_local_host : nullptr);
}

int mcmcpp_sampler_get_state_local(mcmcpp_sampler* s, double* positions_local_host, double* aux_host)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    if (positions_local_host)
        CUDA_TRY(cudaMemcpyAsync(positions_local_host, s->dPos, sizeof(double) * devRowDoubles(s), cudaMemcpyDeviceToHost, s->stream));
    if (aux_host) CUDA_TRY(cudaMemcpyAsync(aux_host, s->dLogp, sizeof(double) * (size_t)s->W, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int mcmcpp_sampler_set_slicing(mcmcpp_sampler* s, int use_slicing, int interval)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (interval <= 0) return fail(MCMCPP_ERR_ARG, "slicing interval must be positive");   /* EnsembleSampler.h:327 */
    s->subSampling = use_slicing != 0;
    s->interval = interval;
    return MCMCPP_OK;
}

int mcmcpp_sampler_run(mcmcpp_sampler* s, int num_steps, float* device_ms)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (!s->initialised) return fail(MCMCPP_ERR_STATE, "runMCMC before setInitialWalkerPos");
    if (num_steps <= 0) return fail(MCMCPP_ERR_ARG, "numSteps must be positive");           /* RedBlkCtrlerSpinLock.h:188 */
    if (s->cfg.world > 1 && !s->comm && (s->cfg.shard_mode == MCMCPP_SHARD_ALLGATHER || s->cfg.shard_mode == MCMCPP_SHARD_DATA))
        return fail(MCMCPP_ERR_STATE, "world > 1 but no communicator (mcmcpp_sampler_comm_init)");
    if (s->cfg.world > 1 && s->cfg.shard_mode == MCMCPP_SHARD_PEER && !s->dPeerPos)
        return fail(MCMCPP_ERR_STATE, "PEER sharding needs mcmcpp_sampler_peer_import");
    if (s->cfg.world > 1 && s->cfg.shard_mode == MCMCPP_SHARD_DIM && !s->tablesWired)
        return fail(MCMCPP_ERR_STATE, "DIM sharding needs mcmcpp_sampler_peer_import");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    const bool chainOn = s->capRows > 0;
    int rc = MCMCPP_OK;
    if (chainOn && !s->dStore) {
        if ((rc = ensureHostChain(s, s->stored + num_steps)) != MCMCPP_OK) return rc;
        if ((rc = ensureDeviceChain(s)) != MCMCPP_OK) return rc;
    }
    const int perStored = s->subSampling ? s->interval : 1;
    int status = MCMCPP_OK;
    if (s->cfg.world > 1 && s->cfg.shard_mode == MCMCPP_SHARD_PEER && s->peerNeedsBarrier) {
        /* the first colour reads partner rows from the other ranks: their setInitialWalkerPos / restore copies must have landed */
        if ((rc = peerBarrier(s)) != MCMCPP_OK) return rc;
        s->peerNeedsBarrier = false;
    }
    if (s->cfg.world > 1 && s->cfg.shard_mode == MCMCPP_SHARD_DIM && s->groupDepth == 0) {
        /* every rank has entered runMCMC before the clock of this call starts: the ranks' host threads arrive at different times, and
         * a rank that is early would otherwise spend that time inside its first finishing kernel (device-side barrier, bounded) */
        if ((rc = peerBarrier(s)) != MCMCPP_OK) return rc;
    }
    CUDA_TRY(cudaEventRecord(s->evStart, s->stream));
    static const bool fusedAllowed = [] { const char* e = getenv("MCMCPP_B200_FUSED"); return !(e && e[0] == '0'); }();
    const bool fused = fusedAllowed && s->fusedFn && s->cfg.world == 1 && s->cfg.shard_mode == MCMCPP_SHARD_NONE && !s->stepFn && !s->cov
                       && !s->dTrace && s->K <= 1024 && s->D <= 8;
    if (fused) {
        /* same bookkeeping as the loop below, one launch per chunk of stored steps */
        const size_t rowBytes = sizeof(double) * rowDoubles(s);
        if (chainOn && !s->dStore && !s->dFusedChain) {
            s->fusedRows = std::max<long long>(1, std::min<long long>(16384, (long long)((32u << 20) / rowBytes)));
            CUDA_TRY(cudaMalloc(&s->dFusedChain, rowBytes * (size_t)s->fusedRows));
        }
        long long remaining = num_steps;
        while (remaining > 0) {
            long long chunk = std::min<long long>(remaining, (1 << 18) / perStored + 1);
            const bool keep = chainOn && s->stored < s->capRows;
            double* devSlot = nullptr;
            if (keep && s->dStore) {                           /* device-resident chain: the kernel writes the rows where they stay */
                if ((rc = nextDeviceSlot(s, &devSlot)) != MCMCPP_OK) return rc;
                chunk = std::min(chunk, std::min(s->devCapRows - s->devRows, s->capRows - s->stored));
            } else if (chainOn) chunk = keep ? std::min(chunk, std::min(s->fusedRows, s->capRows - s->stored)) : 1;
            mcmcpp_fused_args f;
            f.base = s->base;
            f.base.half = 0; f.base.chain_row = nullptr; f.base.trace = nullptr; f.base.trace_rows = 0;
            f.stored_steps = (int32_t)chunk;
            f.interval = perStored;
            f.chain_rows = keep ? (devSlot ? devSlot : s->dFusedChain) : nullptr;
            rc = s->fusedFn(&f, s->calculator.data(), s->stream);
            if (rc != MCMCPP_OK) return fail(rc, std::string("fused launch failed: ") + cudaGetErrorString(cudaGetLastError()));
            ++s->launches;
            if (keep && !devSlot) CUDA_TRY(cudaMemcpyAsync(s->hChain + rowDoubles(s) * (size_t)s->stored, s->dFusedChain, rowBytes * (size_t)chunk,
                                                           cudaMemcpyDeviceToHost, s->stream));
            if (devSlot) s->devRows += chunk;
            s->stored += chunk;
            s->ensembleSteps += chunk * perStored;
            s->totalSteps += 2ULL * (unsigned long long)s->slotCount * (unsigned long long)perStored * (unsigned long long)chunk;
            remaining -= chunk;
            if (chainOn && s->stored >= s->capRows) { status = MCMCPP_CHAIN_FULL; break; }
            if (keep && !devSlot) CUDA_TRY(cudaStreamSynchronize(s->stream));      /* the staging rows are reused by the next chunk */
        }
    }
    for (int i = 0; !fused && i < num_steps; ++i) {
        /* the interval-1 unsaved steps come first (EnsembleSampler.h:300-303) */
        for (int j = 1; j < perStored; ++j) {
            if ((rc = halfStep(s, 0, nullptr)) != MCMCPP_OK) return rc;
            if ((rc = halfStep(s, 1, nullptr)) != MCMCPP_OK) return rc;
            ++s->ensembleSteps;
        }
        const bool keepRow = chainOn && s->stored < s->capRows;
        const int slot = (int)(s->stored % ChainSlots);
        double* row = nullptr;
        if (keepRow && s->dStore) {
            if ((rc = nextDeviceSlot(s, &row)) != MCMCPP_OK) return rc;
        } else if (keepRow) {
            if (s->slotBusy[slot]) {
                CUDA_TRY(cudaStreamWaitEvent(s->stream, s->slotFree[slot], 0));
                s->slotBusy[slot] = false;
            }
            row = s->dChain[slot];
        }
        if ((rc = halfStep(s, 0, row)) != MCMCPP_OK) return rc;
        if ((rc = halfStep(s, 1, row)) != MCMCPP_OK) return rc;
        ++s->ensembleSteps;
        if (s->stepFn) s->stepFn(s->stepCtx, s->dPos, s->stored, (void*)s->stream);   /* stream-ordered view of the step just made */
        if (s->cov) {                                   /* the stored row IS the state after the black half-step */
            if ((s->covIndex % s->covInterval) == 0) {
                if ((rc = mcmcpp_cov_accumulate(s->cov, s->dPos, s->W, s->stream)) != MCMCPP_OK) return fail(rc, "covariance accumulation failed");
                ++s->covSteps;
            }
            ++s->covIndex;
        }
        if (keepRow && s->dStore) ++s->devRows;
        else if (keepRow && (rc = appendRowFromDeviceSlot(s, slot)) != MCMCPP_OK) return rc;
        ++s->stored;
        s->totalSteps += 2ULL * (unsigned long long)s->slotCount * (unsigned long long)perStored;
        if (chainOn && s->stored >= s->capRows) {   /* incrementChainStep() == EndOfChain (Chain.h:229-233) */
            status = MCMCPP_CHAIN_FULL;
            break;
        }
    }
    CUDA_TRY(cudaEventRecord(s->evStop, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->copyStream));
    if (device_ms) CUDA_TRY(cudaEventElapsedTime(device_ms, s->evStart, s->evStop));
    if (s->cfg.world > 1) {
        int32_t timedOut = 0;
        CUDA_TRY(cudaMemcpy(&timedOut, s->dFlags + 1, sizeof timedOut, cudaMemcpyDeviceToHost));
        if (timedOut) return fail(MCMCPP_ERR_STATE, "a cross-GPU wait timed out (a peer rank died, or never made the matching runMCMC call); "
                                                    "walker state is undefined from here (MCMCPP_B200_RDV_TIMEOUT_MS sets the limit)");
    }
    return status;
}

int mcmcpp_sampler_set_fused_launcher(mcmcpp_sampler* s, mcmcpp_fused_fn launcher)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    s->fusedFn = launcher;
    return MCMCPP_OK;
}

int mcmcpp_sampler_get_counts(mcmcpp_sampler* s, unsigned long long* accepted, unsigned long long* total)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(cudaMemsetAsync(s->dAcceptSum, 0, sizeof(unsigned long long), s->stream));
    sumSlotsKernel<<<std::max(1, std::min(592, (s->maxGrid + 255) / 256)), 256, 0, s->stream>>>(s->dAcceptSlots, s->maxGrid, s->dAcceptSum);
    CUDA_TRY(cudaGetLastError());
    unsigned long long sum = 0;
    CUDA_TRY(cudaMemcpyAsync(&sum, s->dAcceptSum, sizeof sum, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (accepted) *accepted = s->acceptedBase + sum;
    if (total) *total = s->totalSteps;
    return MCMCPP_OK;
}

int mcmcpp_sampler_get_stored_steps(mcmcpp_sampler* s) { return s ? (int)s->stored : 0; }

int mcmcpp_sampler_reset(mcmcpp_sampler* s)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    /* storedSteps = 0, chain reset, walker counters zeroed; positions kept (EnsembleSampler.h:312-322) */
    s->stored = 0;
    s->devRows = 0;
    s->acceptedBase = 0;
    s->totalSteps = 0;
    CUDA_TRY(cudaMemsetAsync(s->dAcceptSlots, 0, sizeof(unsigned long long) * (size_t)s->maxGrid, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int mcmcpp_sampler_store_current(mcmcpp_sampler* s)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (!s->initialised) return fail(MCMCPP_ERR_STATE, "no walker positions yet");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    if (s->capRows > 0 && s->stored < s->capRows) {
        if (s->dStore) {
            double* slot = nullptr;
            int rc = nextDeviceSlot(s, &slot);
            if (rc != MCMCPP_OK) return rc;
            CUDA_TRY(cudaMemcpyAsync(slot, s->dPos, sizeof(double) * devRowDoubles(s), cudaMemcpyDeviceToDevice, s->stream));
            ++s->devRows;
        } else {
            int rc = ensureHostChain(s, s->stored + 1);
            if (rc != MCMCPP_OK) return rc;
            CUDA_TRY(copyToHostRows(s, s->hChain + rowDoubles(s) * (size_t)s->stored, s->dPos, s->stream));
        }
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    ++s->stored;    /* EnsembleSampler.h:240-241 */
    return MCMCPP_OK;
}

int mcmcpp_sampler_slice_and_burn(mcmcpp_sampler* s, int interval, int burn_in)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (interval <= 0 || burn_in < 0) return fail(MCMCPP_ERR_ARG, "interval must be > 0 and burnIn >= 0");  /* EnsembleSampler.h:335-336 */
    /* Chain::resetChainForSubSampling (Chain/Chain.h:268-305): rows burn_in, burn_in + interval, ... move to the front, in place.
     * The chain is the host rows followed by the rows still held in HBM; each part is compacted where it lives (host: memmove,
     * device: row copies on the sampler's stream, every destination at or before its source). */
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    const long long steps = s->capRows > 0 ? keptRows(s) : s->stored;
    if (burn_in == 0 && interval == 1) return MCMCPP_OK;
    if ((steps - burn_in) < interval || steps <= burn_in) { s->stored = 0; s->devRows = 0; return MCMCPP_OK; }
    const long long hostRows = steps - s->devRows;
    long long kept = 0, keptHost = 0;
    const size_t row = rowDoubles(s), devRow = devRowDoubles(s);
    for (long long r = burn_in; r < steps; r += interval, ++kept) {
        if (r < hostRows) {
            if (s->hChain && r != kept) std::memmove(s->hChain + row * (size_t)kept, s->hChain + row * (size_t)r, sizeof(double) * row);
            keptHost = kept + 1;
        } else {
            const long long src = r - hostRows, dst = kept - keptHost;
            if (src != dst)
                CUDA_TRY(cudaMemcpyAsync(s->dStore + devRow * (size_t)dst, s->dStore + devRow * (size_t)src, sizeof(double) * devRow,
                                         cudaMemcpyDeviceToDevice, s->stream));
        }
    }
    if (s->dStore) CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->devRows = kept - keptHost;
    s->stored = kept;
    return MCMCPP_OK;
}

int mcmcpp_sampler_chain(mcmcpp_sampler* s, const double** rows_host, long long* stored_steps)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (s->dStore && s->devRows > 0) {                     /* device-resident rows cross PCIe now, because the host asked */
        CUDA_TRY(cudaSetDevice(s->cfg.device));
        int rc = spillDeviceRows(s);
        if (rc != MCMCPP_OK) return rc;
    }
    if (rows_host) *rows_host = s->hChain;
    if (stored_steps) *stored_steps = s->capRows > 0 ? std::min(s->stored, s->capRows) : 0;
    return MCMCPP_OK;
}

int mcmcpp_sampler_chain_consume(mcmcpp_sampler* s, long long rows)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (s->dStore && s->devRows > 0) {
        CUDA_TRY(cudaSetDevice(s->cfg.device));
        int rc = spillDeviceRows(s);
        if (rc != MCMCPP_OK) return rc;
    }
    const long long have = s->capRows > 0 ? std::min(s->stored, s->capRows) : s->stored;
    if (rows < 0 || rows > have) return fail(MCMCPP_ERR_ARG, "cannot consume more rows than are stored");
    if (rows == 0) return MCMCPP_OK;
    const size_t row = rowDoubles(s);
    if (s->hChain && have > rows) std::memmove(s->hChain, s->hChain + row * (size_t)rows, sizeof(double) * row * (size_t)(have - rows));
    s->stored -= rows;
    for (int i = 0; i < ChainSlots; ++i) s->slotBusy[i] = false;      /* run() left the side stream idle */
    return MCMCPP_OK;
}

int mcmcpp_sampler_chain_device(mcmcpp_sampler* s, const double** rows_dev, long long* first_step, long long* rows)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (!s->dStore) return fail(MCMCPP_ERR_STATE, "no device-resident chain (cfg.device_chain_bytes == 0)");
    if (rows_dev) *rows_dev = s->dStore;
    if (first_step) *first_step = keptRows(s) - s->devRows;
    if (rows) *rows = s->devRows;
    return MCMCPP_OK;
}

int mcmcpp_sampler_chain_covariance(mcmcpp_sampler* s, int slice_interval, double* mean, double* cov, double* corr, long long* points)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (slice_interval <= 0) return fail(MCMCPP_ERR_ARG, "slice interval must be positive");
    if (!s->dStore) return fail(MCMCPP_ERR_STATE, "no device-resident chain (cfg.device_chain_bytes == 0)");
    if (s->Dl != s->D || (s->cfg.world > 1 && s->cfg.shard_mode != MCMCPP_SHARD_DATA))
        return fail(MCMCPP_ERR_UNSUPPORTED, "chain analysis on the device needs whole rows of every walker on this rank");
    if (s->devRows == 0 || s->devRows != keptRows(s)) return fail(MCMCPP_ERR_STATE, "part of the chain has spilled to the host (or it is empty)");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    mcmcpp_cov_accum* c = mcmcpp_cov_create(s->D);
    if (!c) return fail(MCMCPP_ERR_UNSUPPORTED, "covariance accumulator: numParams must be <= 128");
    int rc = mcmcpp_cov_zero(c, s->stream);
    long long used = 0;
    for (long long r = 0; rc == MCMCPP_OK && r < s->devRows; r += slice_interval, ++used)       /* calculateCovar's slicing (CovarianceMatrix.h:126-131) */
        rc = mcmcpp_cov_accumulate(c, s->dStore + devRowDoubles(s) * (size_t)r, s->W, s->stream);
    if (rc == MCMCPP_OK) rc = mcmcpp_cov_finalize(c, used * (long long)s->W, mean, cov, corr, s->stream);
    if (points) *points = used * (long long)s->W;
    mcmcpp_cov_destroy(c);
    return rc == MCMCPP_OK ? rc : fail(rc, "covariance over the device chain failed");
}

int mcmcpp_sampler_chain_autocorr_times(mcmcpp_sampler* s, int window_scaling, int reference_buffer_reuse, double* iat_out)
{
    if (!s || !iat_out) return fail(MCMCPP_ERR_ARG, "null argument");
    if (!s->dStore) return fail(MCMCPP_ERR_STATE, "no device-resident chain (cfg.device_chain_bytes == 0)");
    if (s->Dl != s->D || (s->cfg.world > 1 && s->cfg.shard_mode != MCMCPP_SHARD_DATA))
        return fail(MCMCPP_ERR_UNSUPPORTED, "chain analysis on the device needs whole rows of every walker on this rank");
    if (s->devRows < 2 || s->devRows != keptRows(s)) return fail(MCMCPP_ERR_STATE, "part of the chain has spilled to the host (or it is too short)");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    int rc = mcmcpp_autocorr_device(s->dStore, s->devRows, s->W, s->D, window_scaling, reference_buffer_reuse, iat_out);
    return rc == MCMCPP_OK ? rc : fail(rc, "autocorrelation times over the device chain failed");
}

int mcmcpp_sampler_get_state(mcmcpp_sampler* s, double* positions_host, double* aux_host)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    if (positions_host) CUDA_TRY(copyToHostRows(s, positions_host, s->dPos, s->stream));   /* DIM: this rank's columns only */
    if (aux_host) CUDA_TRY(cudaMemcpyAsync(aux_host, s->dLogp, sizeof(double) * (size_t)s->W, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int mcmcpp_sampler_state_device(mcmcpp_sampler* s, double** positions_dev, double** aux_dev)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (positions_dev) *positions_dev = s->dPos;
    if (aux_dev) *aux_dev = s->dLogp;
    return MCMCPP_OK;
}

int mcmcpp_sampler_get_trace(mcmcpp_sampler* s, mcmcpp_trace* out, int max_rows, int* rows)
{
    if (!s || !out) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    long long have = std::min<long long>(s->cfg.trace_updates, s->ensembleSteps * (long long)s->W);
    int n = (int)std::min<long long>(have, max_rows);
    if (n > 0) {
        CUDA_TRY(cudaMemcpyAsync(out, s->dTrace, sizeof(mcmcpp_trace) * (size_t)n, cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    if (rows) *rows = n;
    return MCMCPP_OK;
}

int mcmcpp_sampler_rng_redraws(mcmcpp_sampler* s, int* redraws)
{
    if (!s || !redraws) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    CUDA_TRY(cudaMemcpyAsync(redraws, s->dFlags, sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int mcmcpp_sampler_launch_count(mcmcpp_sampler* s, unsigned long long* kernel_launches)
{
    if (!s || !kernel_launches) return fail(MCMCPP_ERR_ARG, "null argument");
    *kernel_launches = s->launches;
    return MCMCPP_OK;
}

void* mcmcpp_sampler_stream(mcmcpp_sampler* s) { return s ? (void*)s->stream : nullptr; }

/* ---- multi-GPU ---------------------------------------------------------------------------------- */
int mcmcpp_comm_unique_id(void* id128)
{
    if (!id128) return fail(MCMCPP_ERR_ARG, "null id");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NCCL_TRY(nccl().GetUniqueId(&id));
    std::memcpy(id128, &id, sizeof id);
    return MCMCPP_OK;
}

int mcmcpp_sampler_comm_init(mcmcpp_sampler* s, const void* id128)
{
    if (!s || !id128) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    NCCL_TRY(nccl().CommInitRank(&s->comm, s->cfg.world, id, s->cfg.rank));
    if (s->cfg.shard_mode == MCMCPP_SHARD_DATA) {
        s->base.allreduce_sum = &allreduceSum;
        s->base.allreduce_ctx = s;
    }
    return MCMCPP_OK;
}

int mcmcpp_sampler_peer_export(mcmcpp_sampler* s, void* handle64)
{
    if (!s || !handle64) return fail(MCMCPP_ERR_ARG, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, s->cfg.shard_mode == MCMCPP_SHARD_DIM ? (void*)s->dDimPartial : (void*)s->dPos));
    std::memcpy(handle64, &h, sizeof h);
    return MCMCPP_OK;
}

int mcmcpp_sampler_peer_import(mcmcpp_sampler* s, const void* handles64_all_ranks)
{
    if (!s || !handles64_all_ranks) return fail(MCMCPP_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    const bool dim = s->cfg.shard_mode == MCMCPP_SHARD_DIM;
    std::vector<const double*> ptrs((size_t)s->cfg.world, nullptr);
    for (int r = 0; r < s->cfg.world; ++r) {
        if (r == s->cfg.rank) { ptrs[(size_t)r] = dim ? s->dDimPartial : s->dPos; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, (const char*)handles64_all_ranks + 64 * (size_t)r, sizeof h);
        void* p = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        s->peerOpened.push_back(p);
        ptrs[(size_t)r] = (const double*)p;
    }
    if (dim) {
        CUDA_TRY(cudaMemcpy(s->dDimTable, ptrs.data(), sizeof(double*) * (size_t)s->cfg.world, cudaMemcpyHostToDevice));
        s->tablesWired = true;
        return MCMCPP_OK;
    }
    CUDA_TRY(cudaMalloc(&s->dPeerPos, sizeof(double*) * (size_t)s->cfg.world));
    CUDA_TRY(cudaMemcpy(s->dPeerPos, ptrs.data(), sizeof(double*) * (size_t)s->cfg.world, cudaMemcpyHostToDevice));
    s->base.peer_pos = s->dPeerPos;
    return MCMCPP_OK;
}

/* ---- the ranks of one DIM-sharded ensemble inside one process on one device (capi.h) ------------------------------------ */
int mcmcpp_sampler_set_group_launcher(mcmcpp_sampler* s, mcmcpp_group_fn launcher)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    s->groupFn = launcher;
    return MCMCPP_OK;
}

namespace
{
int checkGroup(mcmcpp_sampler* const* ss, int count)
{
    if (!ss || count < 2 || count > 8) return fail(MCMCPP_ERR_ARG, "a rank group has 2 to 8 samplers");
    for (int r = 0; r < count; ++r) {
        mcmcpp_sampler* s = ss[r];
        if (!s) return fail(MCMCPP_ERR_ARG, "null sampler in the group");
        if (s->cfg.shard_mode != MCMCPP_SHARD_DIM || s->cfg.world != count || s->cfg.rank != r)
            return fail(MCMCPP_ERR_ARG, "samplers[r] must be rank r of a DIM-sharded ensemble of `count` ranks");
        if (s->W != ss[0]->W || s->D != ss[0]->D || s->cfg.device != ss[0]->cfg.device || s->cfg.seed != ss[0]->cfg.seed ||
            s->cfg.rng_mode != ss[0]->cfg.rng_mode || s->capRows != ss[0]->capRows || s->dStore || s->cfg.trace_updates != ss[0]->cfg.trace_updates)
            return fail(MCMCPP_ERR_ARG, "the samplers of a group must agree on walkers, parameters, device, seed, rng mode and chain cap (no device-resident chain)");
    }
    return MCMCPP_OK;
}
}

int mcmcpp_sampler_peer_import_local(mcmcpp_sampler* const* samplers, int count)
{
    int rc = checkGroup(samplers, count);
    if (rc != MCMCPP_OK) return rc;
    CUDA_TRY(cudaSetDevice(samplers[0]->cfg.device));
    std::vector<double*> ptrs((size_t)count);
    for (int r = 0; r < count; ++r) ptrs[(size_t)r] = samplers[r]->dDimPartial;
    for (int r = 0; r < count; ++r) {
        CUDA_TRY(cudaMemcpy(samplers[r]->dDimTable, ptrs.data(), sizeof(double*) * (size_t)count, cudaMemcpyHostToDevice));
        samplers[r]->tablesWired = true;
    }
    return MCMCPP_OK;
}

int mcmcpp_sampler_run_group(mcmcpp_sampler* const* ss, int count, int num_steps)
{
    int rc = checkGroup(ss, count);
    if (rc != MCMCPP_OK) return rc;
    if (num_steps <= 0) return fail(MCMCPP_ERR_ARG, "numSteps must be positive");
    mcmcpp_sampler* lead = ss[0];
    if (!lead->groupFn) return fail(MCMCPP_ERR_UNSUPPORTED, "this Calculator has no rank-group launcher (mcmcpp_sampler_set_group_launcher)");
    for (int r = 0; r < count; ++r) {
        if (!ss[r]->initialised) return fail(MCMCPP_ERR_STATE, "runMCMC before setInitialWalkerPos");
        if (!ss[r]->tablesWired) return fail(MCMCPP_ERR_STATE, "the group needs mcmcpp_sampler_peer_import_local");
        if (ss[r]->stored != lead->stored || ss[r]->subSampling != lead->subSampling || ss[r]->interval != lead->interval)
            return fail(MCMCPP_ERR_STATE, "the samplers of a group must have made the same calls so far");
    }
    CUDA_TRY(cudaSetDevice(lead->cfg.device));
    /* every rank's work goes to the lead's stream for the duration of the call: one stream, one cooperative launch per colour */
    std::vector<cudaStream_t> own((size_t)count);
    for (int r = 0; r < count; ++r) { CUDA_TRY(cudaStreamSynchronize(ss[r]->stream)); own[(size_t)r] = ss[r]->stream; ss[r]->stream = lead->stream; }
    struct Restore { mcmcpp_sampler* const* ss; std::vector<cudaStream_t>& own; int count;
                     ~Restore() { for (int r = 0; r < count; ++r) ss[r]->stream = own[(size_t)r]; } } restore{ss, own, count};
    const bool chainOn = lead->capRows > 0;
    if (chainOn)
        for (int r = 0; r < count; ++r) {
            if ((rc = ensureHostChain(ss[r], ss[r]->stored + num_steps)) != MCMCPP_OK) return rc;
            if ((rc = ensureDeviceChain(ss[r])) != MCMCPP_OK) return rc;
        }
    const int perStored = lead->subSampling ? lead->interval : 1;
    std::vector<mcmcpp_halfstep_args> args((size_t)count);
    std::vector<const void*> calcs((size_t)count);
    for (int r = 0; r < count; ++r) calcs[(size_t)r] = ss[r]->calculator.data();
    auto colour = [&](int half, bool store) -> int {
        for (int r = 0; r < count; ++r) {
            mcmcpp_sampler* s = ss[r];
            mcmcpp_halfstep_args a = s->base;
            a.half = half;
            a.chain_row = store ? s->dChain[(int)(s->stored % ChainSlots)] : nullptr;
            a.trace = nullptr; a.trace_rows = 0;
            if (s->dTrace) {
                long long first = s->ensembleSteps * (long long)s->W + (long long)half * s->K;
                long long rows = (long long)s->cfg.trace_updates - first;
                if (rows > 0) { a.trace = s->dTrace + first; a.trace_rows = (int32_t)std::min<long long>(rows, s->K); }
            }
            args[(size_t)r] = a;
        }
        int rcl = lead->groupFn(args.data(), calcs.data(), count, lead->stream);
        if (rcl != MCMCPP_OK) return fail(rcl, std::string("rank-group launch failed: ") + cudaGetErrorString(cudaGetLastError()));
        for (int r = 0; r < count; ++r) ss[r]->launches += 3;
        return MCMCPP_OK;
    };
    int status = MCMCPP_OK;
    for (int i = 0; i < num_steps; ++i) {
        for (int j = 1; j < perStored; ++j) {
            if ((rc = colour(0, false)) != MCMCPP_OK) return rc;
            if ((rc = colour(1, false)) != MCMCPP_OK) return rc;
            for (int r = 0; r < count; ++r) ++ss[r]->ensembleSteps;
        }
        const bool keepRow = chainOn && lead->stored < lead->capRows;
        const int slot = (int)(lead->stored % ChainSlots);
        if (keepRow)
            for (int r = 0; r < count; ++r)
                if (ss[r]->slotBusy[slot]) { CUDA_TRY(cudaStreamWaitEvent(lead->stream, ss[r]->slotFree[slot], 0)); ss[r]->slotBusy[slot] = false; }
        if ((rc = colour(0, keepRow)) != MCMCPP_OK) return rc;
        if ((rc = colour(1, keepRow)) != MCMCPP_OK) return rc;
        for (int r = 0; r < count; ++r) {
            mcmcpp_sampler* s = ss[r];
            ++s->ensembleSteps;
            if (keepRow && (rc = appendRowFromDeviceSlot(s, slot)) != MCMCPP_OK) return rc;
            ++s->stored;
            s->totalSteps += 2ULL * (unsigned long long)s->slotCount * (unsigned long long)perStored;
        }
        if (chainOn && lead->stored >= lead->capRows) { status = MCMCPP_CHAIN_FULL; break; }
    }
    CUDA_TRY(cudaStreamSynchronize(lead->stream));
    for (int r = 0; r < count; ++r) {
        CUDA_TRY(cudaStreamSynchronize(ss[r]->copyStream));
        int32_t timedOut = 0;
        CUDA_TRY(cudaMemcpy(&timedOut, ss[r]->dFlags + 1, sizeof timedOut, cudaMemcpyDeviceToHost));
        if (timedOut) return fail(MCMCPP_ERR_STATE, "a wait on another rank's partial sums timed out inside the rank group");
    }
    return status;
}

/* ---- Analysis::CovarianceMatrix on the device (Analysis/CovarianceMatrix.h:120-215) -------------------------- */
int mcmcpp_sampler_cov_begin(mcmcpp_sampler* s, int slice_interval, int include_current)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (slice_interval <= 0) return fail(MCMCPP_ERR_ARG, "slice interval must be positive");
    if (s->Dl != s->D || (s->cfg.world > 1 && s->cfg.shard_mode != MCMCPP_SHARD_DATA))
        return fail(MCMCPP_ERR_UNSUPPORTED, "on-the-fly covariance needs whole rows of every walker on this rank");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    if (!s->cov && !(s->cov = mcmcpp_cov_create(s->D))) return fail(MCMCPP_ERR_UNSUPPORTED, "covariance accumulator: numParams must be <= 128");
    int rc = mcmcpp_cov_zero(s->cov, s->stream);
    if (rc != MCMCPP_OK) return fail(rc, "covariance accumulator reset failed");
    s->covInterval = slice_interval;
    s->covIndex = 0;
    s->covSteps = 0;
    if (include_current) {
        if (!s->initialised) return fail(MCMCPP_ERR_STATE, "no walker positions yet");
        if ((rc = mcmcpp_cov_accumulate(s->cov, s->dPos, s->W, s->stream)) != MCMCPP_OK) return fail(rc, "covariance accumulation failed");
        s->covIndex = 1;
        s->covSteps = 1;
    }
    return MCMCPP_OK;
}

int mcmcpp_sampler_cov_result(mcmcpp_sampler* s, double* mean, double* cov, double* corr, long long* points)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    if (!s->cov) return fail(MCMCPP_ERR_STATE, "mcmcpp_sampler_cov_begin was not called");
    if (s->covSteps == 0) return fail(MCMCPP_ERR_STATE, "no step accumulated yet");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    const long long n = s->covSteps * (long long)s->W;
    if (points) *points = n;
    int rc = mcmcpp_cov_finalize(s->cov, n, mean, cov, corr, s->stream);
    return rc == MCMCPP_OK ? rc : fail(rc, "covariance finalize failed");
}

int mcmcpp_sampler_cov_end(mcmcpp_sampler* s)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    mcmcpp_cov_destroy(s->cov);
    s->cov = nullptr;
    return MCMCPP_OK;
}

int mcmcpp_covariance(const double* chain_host, long long steps, int num_walkers, int num_params, int slice_interval, int device,
                      double* mean, double* cov, double* corr, long long* points)
{
    if (!chain_host || steps <= 0 || num_walkers <= 0 || num_params <= 0 || slice_interval <= 0) return fail(MCMCPP_ERR_ARG, "bad argument");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fail(MCMCPP_ERR_CUDA, "no CUDA device: mcmcpp_b200 has no CPU fallback");
    CUDA_TRY(cudaSetDevice(device));
    mcmcpp_cov_accum* c = mcmcpp_cov_create(num_params);
    if (!c) return fail(MCMCPP_ERR_UNSUPPORTED, "covariance accumulator: numParams must be <= 128");
    const size_t row = (size_t)num_walkers * (size_t)num_params;
    /* stage up to ~256 MB of used steps per upload */
    long long perUpload = std::max<long long>(1, (long long)((size_t)(256u << 20) / (row * sizeof(double))));
    double* dRows = nullptr;
    cudaStream_t st = nullptr;
    int rc = MCMCPP_OK;
    long long used = 0;
    if (cudaMalloc(&dRows, sizeof(double) * row * (size_t)perUpload) != cudaSuccess || cudaStreamCreate(&st) != cudaSuccess) rc = MCMCPP_ERR_CUDA;
    if (rc == MCMCPP_OK) rc = mcmcpp_cov_zero(c, st);
    long long staged = 0;
    for (long long index = 0; rc == MCMCPP_OK && index < steps; ++index) {
        if ((index % slice_interval) != 0) continue;              /* calculateCovar's slicing (CovarianceMatrix.h:126-131) */
        if (cudaMemcpyAsync(dRows + row * (size_t)staged, chain_host + row * (size_t)index, sizeof(double) * row, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = MCMCPP_ERR_CUDA;
        ++staged; ++used;
        if (staged == perUpload || index + slice_interval >= steps) {
            if (rc == MCMCPP_OK) rc = mcmcpp_cov_accumulate(c, dRows, staged * (long long)num_walkers, st);
            if (rc == MCMCPP_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = MCMCPP_ERR_CUDA;
            staged = 0;
        }
    }
    if (rc == MCMCPP_OK) rc = mcmcpp_cov_finalize(c, used * (long long)num_walkers, mean, cov, corr, st);
    if (points) *points = used * (long long)num_walkers;
    if (st) cudaStreamDestroy(st);
    cudaFree(dRows);
    mcmcpp_cov_destroy(c);
    return rc == MCMCPP_OK ? rc : fail(rc, "covariance computation failed");
}

int mcmcpp_sampler_set_step_action(mcmcpp_sampler* s, mcmcpp_step_fn fn, void* ctx)
{
    if (!s) return fail(MCMCPP_ERR_ARG, "null sampler");
    s->stepFn = fn;
    s->stepCtx = ctx;
    return MCMCPP_OK;
}

/* ---- checkpoint / resume (the reference has none) -------------------------------------
 * Everything the next step depends on: positions (this rank's slice), log-probs, every walker's pcg64 state, the
 * step index and the counters.  The host chain is the caller's to keep; a restored sampler starts with an empty one. */
namespace
{
struct CheckpointHeader
{
    uint64_t magic;                 /* "MCMCPB2\0" */
    int32_t W, D, Dl, dimBegin, rngMode, shardMode, rank, world, moverDoubles, pad;
    int64_t seed, ensembleSteps;
    uint64_t accepted, total;
};
const uint64_t CheckpointMagic = 0x00324250434d434dULL;

size_t checkpointBytes(const mcmcpp_sampler* s)
{
    return sizeof(CheckpointHeader) + sizeof(double) * devRowDoubles(s) + sizeof(double) * (size_t)s->W + sizeof(uint64_t) * 2 * (size_t)s->W +
           sizeof(double) * (size_t)s->W * (size_t)s->moverDoubles;
}
}

int mcmcpp_sampler_checkpoint_bytes(mcmcpp_sampler* s, size_t* bytes)
{
    if (!s || !bytes) return fail(MCMCPP_ERR_ARG, "null argument");
    *bytes = checkpointBytes(s);
    return MCMCPP_OK;
}

int mcmcpp_sampler_checkpoint_save(mcmcpp_sampler* s, void* buffer, size_t bytes)
{
    if (!s || !buffer) return fail(MCMCPP_ERR_ARG, "null argument");
    if (!s->initialised) return fail(MCMCPP_ERR_STATE, "no walker positions yet");
    if (bytes < checkpointBytes(s)) return fail(MCMCPP_ERR_ARG, "checkpoint buffer too small");
    CheckpointHeader h;
    std::memset(&h, 0, sizeof h);
    h.magic = CheckpointMagic;
    h.W = s->W; h.D = s->D; h.Dl = s->Dl; h.dimBegin = s->dimBegin; h.rngMode = s->cfg.rng_mode; h.shardMode = s->cfg.shard_mode;
    h.moverDoubles = s->moverDoubles;
    h.rank = s->cfg.rank; h.world = s->cfg.world; h.seed = s->cfg.seed; h.ensembleSteps = s->ensembleSteps;
    unsigned long long acc = 0, tot = 0;
    int rc = mcmcpp_sampler_get_counts(s, &acc, &tot);
    if (rc != MCMCPP_OK) return rc;
    h.accepted = acc; h.total = tot;
    char* out = static_cast<char*>(buffer);
    std::memcpy(out, &h, sizeof h); out += sizeof h;
    CUDA_TRY(cudaMemcpyAsync(out, s->dPos, sizeof(double) * devRowDoubles(s), cudaMemcpyDeviceToHost, s->stream)); out += sizeof(double) * devRowDoubles(s);
    CUDA_TRY(cudaMemcpyAsync(out, s->dLogp, sizeof(double) * (size_t)s->W, cudaMemcpyDeviceToHost, s->stream)); out += sizeof(double) * (size_t)s->W;
    CUDA_TRY(cudaMemcpyAsync(out, s->dRng, sizeof(uint64_t) * 2 * (size_t)s->W, cudaMemcpyDeviceToHost, s->stream)); out += sizeof(uint64_t) * 2 * (size_t)s->W;
    if (s->moverDoubles)
        CUDA_TRY(cudaMemcpyAsync(out, s->dMoverState, sizeof(double) * (size_t)s->W * (size_t)s->moverDoubles, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return MCMCPP_OK;
}

int mcmcpp_sampler_checkpoint_restore(mcmcpp_sampler* s, const void* buffer, size_t bytes)
{
    if (!s || !buffer) return fail(MCMCPP_ERR_ARG, "null argument");
    if (bytes < checkpointBytes(s)) return fail(MCMCPP_ERR_ARG, "checkpoint buffer too small for this sampler");
    CheckpointHeader h;
    std::memcpy(&h, buffer, sizeof h);
    if (h.magic != CheckpointMagic) return fail(MCMCPP_ERR_ARG, "not a mcmcpp_b200 checkpoint");
    if (h.W != s->W || h.D != s->D || h.Dl != s->Dl || h.dimBegin != s->dimBegin || h.rngMode != s->cfg.rng_mode || h.seed != s->cfg.seed ||
        h.moverDoubles != s->moverDoubles)
        return fail(MCMCPP_ERR_ARG, "checkpoint was taken from a different sampler configuration (walkers, parameters, slice, rng mode or seed)");
    CUDA_TRY(cudaSetDevice(s->cfg.device));
    const char* in = static_cast<const char*>(buffer) + sizeof h;
    CUDA_TRY(cudaMemcpyAsync(s->dPos, in, sizeof(double) * devRowDoubles(s), cudaMemcpyHostToDevice, s->stream)); in += sizeof(double) * devRowDoubles(s);
    CUDA_TRY(cudaMemcpyAsync(s->dLogp, in, sizeof(double) * (size_t)s->W, cudaMemcpyHostToDevice, s->stream)); in += sizeof(double) * (size_t)s->W;
    CUDA_TRY(cudaMemcpyAsync(s->dRng, in, sizeof(uint64_t) * 2 * (size_t)s->W, cudaMemcpyHostToDevice, s->stream)); in += sizeof(uint64_t) * 2 * (size_t)s->W;
    if (s->moverDoubles)
        CUDA_TRY(cudaMemcpyAsync(s->dMoverState, in, sizeof(double) * (size_t)s->W * (size_t)s->moverDoubles, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemsetAsync(s->dAcceptSlots, 0, sizeof(unsigned long long) * (size_t)s->maxGrid, s->stream));
    CUDA_TRY(cudaMemsetAsync(s->dFlags, 0, sizeof(int32_t) * 4, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->acceptedBase = h.accepted;
    s->totalSteps = h.total;
    s->ensembleSteps = h.ensembleSteps;
    s->stored = 0;
    s->devRows = 0;
    s->initialised = true;
    s->peerNeedsBarrier = true;
    return MCMCPP_OK;
}

int mcmcpp_dim_slice(int num_params, int rank, int world, int* begin, int* count)
{
    if (!begin || !count || world < 1 || rank < 0 || rank >= world || num_params / 2 < world) return fail(MCMCPP_ERR_ARG, "bad slice request");
    dimSlice(num_params, rank, world, begin, count);
    return MCMCPP_OK;
}

}  /* extern "C" */
