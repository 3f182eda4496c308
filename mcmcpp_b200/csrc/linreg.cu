/*
 * mcmcpp_b200/csrc/linreg.cu -- BASELINE config 5: stretch-move half-step for the Bayesian linear-regression
 * log-posterior  logp(th) = -1/2 ||y - X th||^2 - 1/2 ||th||^2 / priorVar  (include/mcmcpp_b200/targets.h
 * LinearRegression), with the likelihood of ALL proposals of a colour evaluated as one dense GEMM on the FP64
 * tensor cores.
 *
 * Same update as StretchMove::updateWalker (MCMCpp/Movers/StretchMove.h:100-123); calc.calcLogPostProb(proposal)
 * (StretchMove.h:111) -- 2 M P flops per walker in the reference -- becomes, per colour,
 *     R[m][k] = sum_p Xp[m][p] * Th[k][p]          Xp = [X | y | 0..]  (M x Dp),  Th = [proposal_k | -1 | 0..]  (K x Dp)
 *     sse[k]  = sum_m R[m][k]^2                     ( R = X th - y, so sse = ||y - X th||^2 )
 * i.e. an [M x Dp] x [Dp x K] contraction whose epilogue squares and column-sums in registers; R is never stored.
 * A half-step is four launches on the sampler's stream:
 *   1. linregPropose   thread per walker: draws (j, z, log(1-u)), proposal row -> Th, prior term; reference arithmetic
 *   2. linregSse       DMMA m8n8k4 GEMM; grid (K/256 proposal tiles, S data splits); each warp keeps its 32 proposals'
 *                      B fragments in registers for the whole launch and streams X tiles (64 rows) through a 4-stage
 *                      cp.async.bulk + mbarrier ring; column sums live in registers, one partial row per split
 *   3. linregReduce    fixed-order sum of the S partials (+ all-reduce over ranks when the DATA is sharded)
 *   4. linregAccept    thread per walker: Metropolis test and commit (StretchMove.h:112-122, Walker.h:162-179)
 * Draws, partner choice and proposals are bit-identical to the reference; sse is summed in tensor-core / split
 * order, so log-probs carry the tolerance tier (<= 1e-12 relative, SURVEY.md s.8(c) tier B).
 */
#include <cuda_runtime.h>
#include <cstring>
#include <new>
#include "builtin_targets.h"
#include "mcmcpp_b200/stretch_kernels.cuh"

namespace
{
using namespace MCMC::Device;

enum { LrThreads = 256, LrWarps = 8, LrNI = 4, LrTileN = LrWarps * LrNI * 8 /* 256 proposals per CTA */, LrTileM = 64, LrStages = 4 };

struct LinregState
{
    double* Xp = nullptr;       /* [Mpad][stride]  packed data rows (X | y | 0) */
    double* Th = nullptr;       /* [Kpad][stride]  proposal rows (theta' | -1 | 0) */
    double* z = nullptr;        /* [Kpad] */
    double* e = nullptr;        /* [Kpad] log(1-u) */
    double* tt = nullptr;       /* [Kpad] ||theta'||^2 */
    int* j = nullptr;           /* [Kpad] */
    double* partial = nullptr;  /* [S][Kpad] */
    double* sse = nullptr;      /* [Kpad] */
    long M = 0, Mpad = 0;
    int P = 0, KS = 0, stride = 0, S = 1, Kpad = 0, numTiles = 0, tilesPerSplit = 0;
    double priorVar = 1.0;
};

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

/* (X | y) row-major -> Xp with zero padding (columns > P and rows >= M) */
__global__ void linregPack(const double* __restrict__ X, const double* __restrict__ y, double* __restrict__ Xp, long M, long Mpad, int P,
                           int stride)
{
    const long total = Mpad * stride;
    for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long)gridDim.x * blockDim.x) {
        const long m = idx / stride;
        const int c = (int)(idx - m * stride);
        double v = 0.0;
        if (m < M) v = c < P ? X[m * P + c] : (c == P ? y[m] : 0.0);
        Xp[idx] = v;
    }
}

/* 1. draws + proposal rows */
__global__ void __launch_bounds__(LrThreads) linregPropose(const mcmcpp_halfstep_args a, const LinregState st)
{
    const int i = blockIdx.x * LrThreads + threadIdx.x;
    if (i >= st.Kpad) return;
    double* row = st.Th + (size_t)i * st.stride;
    if (i >= a.slot_count) {                       /* padding columns of the GEMM */
        for (int d = 0; d < st.stride; ++d) row[d] = 0.0;
        return;
    }
    const int D = a.D;
    const int w = a.half * a.K + a.slot_begin + i;
    const int other0 = (1 - a.half) * a.K;
    WalkerRng rng = loadRng(a, w);
    StretchDraws dr = drawStretchFor(a, rng);
    storeRng(a, w, rng);
    if (dr.redraws) atomicAdd(a.flags, dr.redraws);
    const double* own = a.pos + (size_t)w * D;
    const double* sel = partnerBase(a, dr.j) + (size_t)(other0 + dr.j) * D;
    double tt = 0.0;
    for (int d = 0; d < D; ++d) {
        const double x = stretchPoint(sel[d], own[d], dr.z);
        row[d] = x;
        tt = __dadd_rn(tt, __dmul_rn(x, x));       /* LinearRegression::calcLogPostProb prior loop, same order */
    }
    row[D] = -1.0;
    for (int d = D + 1; d < st.stride; ++d) row[d] = 0.0;
    st.z[i] = dr.z; st.e[i] = dr.logOneMinusU; st.tt[i] = tt; st.j[i] = dr.j;
}

/* 2. sse partials: the GEMM */
template <int KS>
__global__ void __launch_bounds__(LrThreads, 1) linregSse(const LinregState st)
{
    extern __shared__ __align__(128) double lrSmem[];
    __shared__ __align__(8) unsigned long long fullBar[LrStages];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int stride = st.stride;
    const uint32_t tileBytes = (uint32_t)(LrTileM * stride * sizeof(double));
    const int tile0 = blockIdx.y * st.tilesPerSplit;
    int nTiles = st.numTiles - tile0;
    if (nTiles > st.tilesPerSplit) nTiles = st.tilesPerSplit;

    if (tid == 0) {
        for (int s = 0; s < LrStages; ++s) mbarInit(smemAddr(&fullBar[s]), 1);
    }
    __syncthreads();
    if (tid == 0) {
        for (int s = 0; s < LrStages && s < nTiles; ++s) {
            mbarExpectTx(smemAddr(&fullBar[s]), tileBytes);
            bulkLoad(smemAddr(lrSmem + (size_t)s * LrTileM * stride), st.Xp + (size_t)(tile0 + s) * LrTileM * stride, tileBytes,
                     smemAddr(&fullBar[s]));
        }
    }
    /* B fragments of this warp's 32 proposals, resident for the whole launch: lane (g,t) holds Th[n0 + 8 ni + g][4 ks + t] */
    double B[LrNI][KS];
    {
        const double* th = st.Th + (size_t)(blockIdx.x * LrTileN + warp * LrNI * 8 + g) * stride + t;
#pragma unroll
        for (int ni = 0; ni < LrNI; ++ni)
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) B[ni][ks] = th[(size_t)ni * 8 * stride + ks * 4];
    }
    double sse[LrNI][2];
#pragma unroll
    for (int ni = 0; ni < LrNI; ++ni) { sse[ni][0] = 0.0; sse[ni][1] = 0.0; }

    for (int it = 0; it < nTiles; ++it) {
        const int s = it % LrStages;
        mbarWait(smemAddr(&fullBar[s]), (uint32_t)((it / LrStages) & 1));
        const double* xs = lrSmem + (size_t)s * LrTileM * stride + (size_t)g * stride + t;
#pragma unroll 1
        for (int mp = 0; mp < LrTileM / 16; ++mp) {
            double c0[LrNI][2], c1[LrNI][2];
#pragma unroll
            for (int ni = 0; ni < LrNI; ++ni) { c0[ni][0] = 0.0; c0[ni][1] = 0.0; c1[ni][0] = 0.0; c1[ni][1] = 0.0; }
            const double* x0 = xs + (size_t)(mp * 16) * stride;
            const double* x1 = x0 + (size_t)8 * stride;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                const double a0 = x0[ks * 4], a1 = x1[ks * 4];
#pragma unroll
                for (int ni = 0; ni < LrNI; ++ni) {
                    dmma884(c0[ni][0], c0[ni][1], a0, B[ni][ks]);
                    dmma884(c1[ni][0], c1[ni][1], a1, B[ni][ks]);
                }
            }
#pragma unroll
            for (int ni = 0; ni < LrNI; ++ni) {
                sse[ni][0] = fma(c0[ni][0], c0[ni][0], sse[ni][0]);
                sse[ni][1] = fma(c0[ni][1], c0[ni][1], sse[ni][1]);
                sse[ni][0] = fma(c1[ni][0], c1[ni][0], sse[ni][0]);
                sse[ni][1] = fma(c1[ni][1], c1[ni][1], sse[ni][1]);
            }
        }
        __syncthreads();                                       /* every warp is done with stage s */
        if (tid == 0 && it + LrStages < nTiles) {
            mbarExpectTx(smemAddr(&fullBar[s]), tileBytes);
            bulkLoad(smemAddr(lrSmem + (size_t)s * LrTileM * stride), st.Xp + (size_t)(tile0 + it + LrStages) * LrTileM * stride,
                     tileBytes, smemAddr(&fullBar[s]));
        }
    }
    /* column sums: this lane holds columns 8 ni + 2t, +1 for rows g (mod 8): fold the 8 row groups */
    double* out = st.partial + (size_t)blockIdx.y * st.Kpad + blockIdx.x * LrTileN + warp * LrNI * 8;
#pragma unroll
    for (int ni = 0; ni < LrNI; ++ni) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            double v = sse[ni][h];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (g == 0) out[ni * 8 + 2 * t + h] = v;
        }
    }
}

/* 3. fixed-order sum over the data splits */
__global__ void __launch_bounds__(LrThreads) linregReduce(const LinregState st)
{
    const int i = blockIdx.x * LrThreads + threadIdx.x;
    if (i >= st.Kpad) return;
    double s = 0.0;
    for (int sp = 0; sp < st.S; ++sp) s += st.partial[(size_t)sp * st.Kpad + i];
    st.sse[i] = s;
}

/* 4. Metropolis decision + commit */
template <bool TRACE>
__global__ void __launch_bounds__(LrThreads) linregAccept(const mcmcpp_halfstep_args a, const LinregState st)
{
    const int i = blockIdx.x * LrThreads + threadIdx.x;
    int accepted = 0;
    if (i < a.slot_count) {
        const int D = a.D;
        const int slot = a.slot_begin + i;
        const int w = a.half * a.K + slot;
        const double z = st.z[i];
        const double scaling = __dmul_rn(log(z), (double)(D - 1));
        /* -0.5*ss - 0.5*(tt/priorVar), as LinearRegression::calcLogPostProb finishes */
        const double newp = __dsub_rn(__dmul_rn(-0.5, st.sse[i]), __dmul_rn(0.5, __ddiv_rn(st.tt[i], st.priorVar)));
        const double diff = __dsub_rn(__dadd_rn(scaling, newp), a.logp[w]);
        accepted = st.e[i] < diff ? 1 : 0;
        const double* row = st.Th + (size_t)i * st.stride;
        double* own = a.pos + (size_t)w * D;
        if (a.chain_row) {
            double* chainRow = a.chain_row + (size_t)w * D;
            for (int d = 0; d < D; ++d) chainRow[d] = accepted ? row[d] : own[d];
        }
        if (accepted) {
            for (int d = 0; d < D; ++d) own[d] = row[d];
            a.logp[w] = newp;
        }
        if (TRACE) {
            if (slot < a.trace_rows) {
                mcmcpp_trace tr;
                tr.j = st.j[i]; tr.accept = accepted; tr.z = z; tr.e = st.e[i]; tr.newp = newp; tr.diff = diff;
                a.trace[slot] = tr;
            }
        }
    }
    const int blockAccepted = __syncthreads_count(accepted);
    if (threadIdx.x == 0 && blockAccepted) a.accept_slots[blockIdx.x] += (unsigned long long)blockAccepted;
}

template <int KS>
int launchSse(const LinregState& st, cudaStream_t stream)
{
    static DynamicSmemOptIn optIn;                            /* per device */
    const size_t smem = sizeof(double) * (size_t)LrStages * LrTileM * st.stride;
    if (!optIn.ensure(linregSse<KS>, 200 * 1024)) return MCMCPP_ERR_CUDA;
    dim3 grid((unsigned)(st.Kpad / LrTileN), (unsigned)st.S);
    linregSse<KS><<<grid, LrThreads, smem, stream>>>(st);
    return launchStatus();
}

int linregLaunch(const mcmcpp_halfstep_args* args, const void* calculator, void* cudaStream)
{
    const mcmcpp_halfstep_args& a = *args;
    const LinregState& st = **static_cast<LinregState* const*>(calculator);
    cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
    if (a.slot_count <= 0) return MCMCPP_OK;
    if (a.slot_count > st.Kpad) return MCMCPP_ERR_ARG;
    const int blocks = st.Kpad / LrThreads;
    linregPropose<<<blocks, LrThreads, 0, stream>>>(a, st);
    int rc;
    switch (st.KS) {
    case 4: rc = launchSse<4>(st, stream); break;
    case 8: rc = launchSse<8>(st, stream); break;
    case 13: rc = launchSse<13>(st, stream); break;
    case 16: rc = launchSse<16>(st, stream); break;
    default: return MCMCPP_ERR_UNSUPPORTED;
    }
    if (rc != MCMCPP_OK) return rc;
    linregReduce<<<blocks, LrThreads, 0, stream>>>(st);
    if (a.allreduce_sum) {
        rc = a.allreduce_sum(a.allreduce_ctx, st.sse, (size_t)st.Kpad, cudaStream);
        if (rc != MCMCPP_OK) return rc;
    }
    if (a.trace) linregAccept<true><<<blocks, LrThreads, 0, stream>>>(a, st);
    else linregAccept<false><<<blocks, LrThreads, 0, stream>>>(a, st);
    if (a.launch_count) *a.launch_count += 3;      /* the runtime counts one launch per half-step itself */
    return launchStatus();
}

}  // namespace

/* Tensor-core path for the linear-regression target (P + 1 <= 64 columns; MCMCPP_B200_DENSE_GENERIC=1 keeps the
 * reference-order generic kernel).  X, y: device views of the caller's parameter vector. */
bool mcmcpp_linreg_bind(int W, int D, long M, double priorVar, const double* devX, const double* devY, mcmcpp_halfstep_fn* launcher,
                        std::vector<char>* calculator, void** aux)
{
    const char* env = getenv("MCMCPP_B200_DENSE_GENERIC");
    if (D + 1 > 64 || (env && env[0] == '1')) return false;
    LinregState* st = new (std::nothrow) LinregState();
    if (!st) return false;
    const int need = (D + 1 + 3) / 4;
    st->KS = need <= 4 ? 4 : need <= 8 ? 8 : need <= 13 ? 13 : 16;
    st->stride = 4 * st->KS;
    if ((st->stride & 7) != 4) st->stride += 4;                 /* stride = 4 mod 8 doubles: conflict-free A-fragment loads */
    st->P = D; st->M = M; st->priorVar = priorVar;
    st->numTiles = (int)((M + LrTileM - 1) / LrTileM);
    st->Mpad = (long)st->numTiles * LrTileM;
    const int K = W / 2;
    st->Kpad = ((K + LrTileN - 1) / LrTileN) * LrTileN;
    /* data splits: fill whole waves of the SMs; at least 8 tiles per split */
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int nTilesN = st->Kpad / LrTileN;
    int S = 1;
    for (int waves = 1; waves <= 64; ++waves) {
        const int cand = (sms * waves) / nTilesN;
        if (cand < 1) continue;
        if ((st->numTiles + cand - 1) / cand < 8) break;
        S = cand;
        if (waves >= 8) break;
    }
    st->tilesPerSplit = (st->numTiles + S - 1) / S;
    st->S = (st->numTiles + st->tilesPerSplit - 1) / st->tilesPerSplit;
    bool ok = cudaMalloc(&st->Xp, sizeof(double) * (size_t)st->Mpad * st->stride) == cudaSuccess
        && cudaMalloc(&st->Th, sizeof(double) * (size_t)st->Kpad * st->stride) == cudaSuccess
        && cudaMalloc(&st->z, sizeof(double) * st->Kpad) == cudaSuccess && cudaMalloc(&st->e, sizeof(double) * st->Kpad) == cudaSuccess
        && cudaMalloc(&st->tt, sizeof(double) * st->Kpad) == cudaSuccess && cudaMalloc(&st->j, sizeof(int) * st->Kpad) == cudaSuccess
        && cudaMalloc(&st->partial, sizeof(double) * (size_t)st->S * st->Kpad) == cudaSuccess
        && cudaMalloc(&st->sse, sizeof(double) * st->Kpad) == cudaSuccess;
    if (ok) {
        linregPack<<<sms * 8, 256>>>(devX, devY, st->Xp, M, st->Mpad, D, st->stride);
        ok = cudaDeviceSynchronize() == cudaSuccess;
    }
    *aux = st;
    if (!ok) return false;
    calculator->resize(sizeof(LinregState*));
    std::memcpy(calculator->data(), &st, sizeof(LinregState*));
    *launcher = &linregLaunch;
    return true;
}

void mcmcpp_linreg_release(void* aux)
{
    LinregState* st = static_cast<LinregState*>(aux);
    if (!st) return;
    cudaFree(st->Xp); cudaFree(st->Th); cudaFree(st->z); cudaFree(st->e); cudaFree(st->tt); cudaFree(st->j);
    cudaFree(st->partial); cudaFree(st->sse);
    delete st;
}
