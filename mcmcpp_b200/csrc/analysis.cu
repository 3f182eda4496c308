/*
 * mcmcpp_b200/csrc/analysis.cu -- MCMC::Analysis::CovarianceMatrix (MCMCpp/Analysis/CovarianceMatrix.h:120-215) on the
 * device, so the covariance check of a 1M-walker run does not need the chain on the host.
 *
 * The reference adds every point to D Kahan-summed averages and D(D+1)/2 Kahan-summed products in chain order.  Here
 * the points are spread over a fixed grid; every CTA keeps its own Kahan (sum, compensation) pair per item in HBM
 * across calls, and finalize adds the CTAs' pairs in CTA order (Kahan again).  Same definitions (n = steps*W, cov =
 * E[xy] - E[x]E[y], correlation normalised by rows then columns); the summation ORDER differs, so results carry the
 * tolerance tier (<= 1e-12 relative on the sums; tests state it), not bit equality.
 *   D <= 4   covSmallKernel: one thread per point, items in registers, CTA tree, one Kahan step per CTA per call
 *   D <= 128 covItemKernel:  one thread per item (i, j<=i), 32-point chunks staged in shared memory
 */
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>
#include "analysis.h"
#include "mcmcpp_b200/capi.h"

namespace
{
enum { CovThreads = 512, CovChunk = 32, CovItemsPerThread = 17, CovSmallThreads = 256 };

__device__ __forceinline__ void kahanAdd(double& sum, double& comp, double val)
{
    const double y = __dsub_rn(val, comp);
    const double t = __dadd_rn(sum, y);
    comp = __dsub_rn(__dsub_rn(t, sum), y);
    sum = t;
}

/* item p: p < D -> average of x_p; else the (i, j <= i) product with q = p - D = i(i+1)/2 + j */
__device__ __forceinline__ void itemIndices(int p, int D, int& i, int& j)
{
    if (p < D) { i = p; j = -1; return; }
    const int q = p - D;
    int r = (int)((sqrt(8.0 * (double)q + 1.0) - 1.0) * 0.5);
    while ((r + 1) * (r + 2) / 2 <= q) ++r;
    while (r * (r + 1) / 2 > q) --r;
    i = r; j = q - r * (r + 1) / 2;
}

__global__ void __launch_bounds__(CovThreads) covItemKernel(const double* __restrict__ rows, long long nPoints, int D, int nItems,
                                                            double2* __restrict__ acc)
{
    extern __shared__ double chunk[];                       /* [CovChunk][D] */
    int ij[CovItemsPerThread];
    double sum[CovItemsPerThread], comp[CovItemsPerThread];
    double2* mine = acc + (size_t)blockIdx.x * nItems;
#pragma unroll
    for (int k = 0; k < CovItemsPerThread; ++k) {
        const int p = threadIdx.x + k * CovThreads;
        int i = 0, j = -1;
        sum[k] = 0.0; comp[k] = 0.0;
        if (p < nItems) {
            itemIndices(p, D, i, j);
            const double2 v = mine[p];
            sum[k] = v.x; comp[k] = v.y;
        }
        ij[k] = (i << 8) | (j & 0xff);
    }
    const long long nChunks = (nPoints + CovChunk - 1) / CovChunk;
    for (long long c = blockIdx.x; c < nChunks; c += gridDim.x) {
        const long long p0 = c * CovChunk;
        const int pts = (int)((nPoints - p0) < CovChunk ? (nPoints - p0) : CovChunk);
        __syncthreads();
        for (int idx = threadIdx.x; idx < pts * D; idx += CovThreads) chunk[idx] = rows[(size_t)p0 * D + idx];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < CovItemsPerThread; ++k) {
            if (threadIdx.x + k * CovThreads >= nItems) break;
            const int i = ij[k] >> 8, jj = ij[k] & 0xff;
            double s = 0.0;
            if (jj == 0xff) { for (int pt = 0; pt < pts; ++pt) s = __dadd_rn(s, chunk[pt * D + i]); }
            else { for (int pt = 0; pt < pts; ++pt) s = __dadd_rn(s, __dmul_rn(chunk[pt * D + i], chunk[pt * D + jj])); }
            kahanAdd(sum[k], comp[k], s);
        }
    }
#pragma unroll
    for (int k = 0; k < CovItemsPerThread; ++k) {
        const int p = threadIdx.x + k * CovThreads;
        if (p < nItems) mine[p] = make_double2(sum[k], comp[k]);
    }
}

template <int D>
__global__ void __launch_bounds__(CovSmallThreads) covSmallKernel(const double* __restrict__ rows, long long nPoints, double2* __restrict__ acc)
{
    constexpr int NI = D + D * (D + 1) / 2;
    __shared__ double red[CovSmallThreads / 32][NI];
    double sum[NI], comp[NI];
#pragma unroll
    for (int k = 0; k < NI; ++k) { sum[k] = 0.0; comp[k] = 0.0; }
    for (long long p = (long long)blockIdx.x * CovSmallThreads + threadIdx.x; p < nPoints; p += (long long)gridDim.x * CovSmallThreads) {
        double x[D];
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = rows[(size_t)p * D + d];
#pragma unroll
        for (int i = 0; i < D; ++i) {
            kahanAdd(sum[i], comp[i], x[i]);
#pragma unroll
            for (int j = 0; j <= i; ++j) kahanAdd(sum[D + i * (i + 1) / 2 + j], comp[D + i * (i + 1) / 2 + j], __dmul_rn(x[i], x[j]));
        }
    }
    /* CTA total of (sum - comp) per item: warp butterflies, then warp 0 over the warps */
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NI; ++k) {
        double v = __dsub_rn(sum[k], comp[k]);
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NI) {
        double v = 0.0;
        for (int w = 0; w < CovSmallThreads / 32; ++w) v = __dadd_rn(v, red[w][threadIdx.x]);
        double2 a = acc[(size_t)blockIdx.x * NI + threadIdx.x];
        kahanAdd(a.x, a.y, v);
        acc[(size_t)blockIdx.x * NI + threadIdx.x] = a;
    }
}

/* totals[p] = Kahan sum over CTAs of (sum, comp) pairs, CTA order */
__global__ void covTotalsKernel(const double2* __restrict__ acc, int nBlocks, int nItems, double* __restrict__ totals)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nItems) return;
    double s = 0.0, c = 0.0, carried = 0.0, cc = 0.0;
    for (int b = 0; b < nBlocks; ++b) {
        const double2 v = acc[(size_t)b * nItems + p];
        kahanAdd(s, c, v.x);
        kahanAdd(carried, cc, v.y);
    }
    totals[p] = __dsub_rn(__dsub_rn(s, c), __dsub_rn(carried, cc));
}
}  // namespace

struct mcmcpp_cov_accum
{
    int D = 0, nItems = 0, nBlocks = 0;
    double2* acc = nullptr;
    double* totals = nullptr;
};

mcmcpp_cov_accum* mcmcpp_cov_create(int D)
{
    if (D < 1 || D > 128) return nullptr;
    mcmcpp_cov_accum* c = new mcmcpp_cov_accum();
    c->D = D;
    c->nItems = D + D * (D + 1) / 2;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    c->nBlocks = D <= 4 ? sms * 8 : sms * 2;
    if (cudaMalloc(&c->acc, sizeof(double2) * (size_t)c->nBlocks * c->nItems) != cudaSuccess ||
        cudaMalloc(&c->totals, sizeof(double) * (size_t)c->nItems) != cudaSuccess) {
        mcmcpp_cov_destroy(c);
        return nullptr;
    }
    return c;
}

void mcmcpp_cov_destroy(mcmcpp_cov_accum* c)
{
    if (!c) return;
    cudaFree(c->acc); cudaFree(c->totals);
    delete c;
}

int mcmcpp_cov_zero(mcmcpp_cov_accum* c, cudaStream_t stream)
{
    return cudaMemsetAsync(c->acc, 0, sizeof(double2) * (size_t)c->nBlocks * c->nItems, stream) == cudaSuccess ? MCMCPP_OK : MCMCPP_ERR_CUDA;
}

int mcmcpp_cov_accumulate(mcmcpp_cov_accum* c, const double* devRows, long long nPoints, cudaStream_t stream)
{
    if (nPoints <= 0) return MCMCPP_OK;
    switch (c->D) {
    case 1: covSmallKernel<1><<<c->nBlocks, CovSmallThreads, 0, stream>>>(devRows, nPoints, c->acc); break;
    case 2: covSmallKernel<2><<<c->nBlocks, CovSmallThreads, 0, stream>>>(devRows, nPoints, c->acc); break;
    case 3: covSmallKernel<3><<<c->nBlocks, CovSmallThreads, 0, stream>>>(devRows, nPoints, c->acc); break;
    case 4: covSmallKernel<4><<<c->nBlocks, CovSmallThreads, 0, stream>>>(devRows, nPoints, c->acc); break;
    default:
        covItemKernel<<<c->nBlocks, CovThreads, sizeof(double) * CovChunk * c->D, stream>>>(devRows, nPoints, c->D, c->nItems, c->acc);
    }
    return cudaGetLastError() == cudaSuccess ? MCMCPP_OK : MCMCPP_ERR_CUDA;
}

int mcmcpp_cov_finalize(mcmcpp_cov_accum* c, long long nPoints, double* mean, double* cov, double* corr, cudaStream_t stream)
{
    const int D = c->D;
    covTotalsKernel<<<(c->nItems + 255) / 256, 256, 0, stream>>>(c->acc, c->nBlocks, c->nItems, c->totals);
    std::vector<double> t((size_t)c->nItems);
    if (cudaMemcpyAsync(t.data(), c->totals, sizeof(double) * t.size(), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess)
        return MCMCPP_ERR_CUDA;
    /* finalizeMatrix, CovarianceMatrix.h:143-185 */
    const double ptCount = (double)nPoints;
    std::vector<double> avg((size_t)D), cv((size_t)D * D), cr((size_t)D * D);
    for (int i = 0; i < D; ++i) avg[(size_t)i] = t[(size_t)i] / ptCount;
    for (int i = 0; i < D; ++i)
        for (int j = 0; j <= i; ++j) {
            const double temp = (t[(size_t)(D + i * (i + 1) / 2 + j)] / ptCount) - (avg[(size_t)i] * avg[(size_t)j]);
            cv[(size_t)i * D + j] = temp;
            cv[(size_t)j * D + i] = temp;
        }
    for (int i = 0; i < D; ++i) {
        const double sv = std::sqrt(cv[(size_t)i * D + i]);
        for (int j = 0; j < D; ++j) cr[(size_t)i * D + j] = cv[(size_t)i * D + j] / sv;
    }
    for (int j = 0; j < D; ++j) {
        const double sv = std::sqrt(cv[(size_t)j * D + j]);
        for (int i = 0; i < D; ++i) cr[(size_t)i * D + j] /= sv;
    }
    for (int i = 0; i < D; ++i) if (mean) mean[i] = avg[(size_t)i];
    for (size_t k = 0; k < (size_t)D * D; ++k) { if (cov) cov[k] = cv[k]; if (corr) corr[k] = cr[k]; }
    return MCMCPP_OK;
}

/* =====================================================================================================================
 * MCMC::Analysis::AutoCorrCalc / Detail::AutoCov (Analysis/AutoCorrCalc.h:150-262, Analysis/Detail/AutoCov.h:131-290)
 * on the device: integrated autocorrelation time of every parameter from the walker-averaged, normalised, circular
 * autocovariance (radix-2 FFT of length 2^ceil(log2 S), no zero padding, as the reference).
 *
 * Every arithmetic step is the reference's, in the reference's order, with its twiddles (host cos/sin of the rational
 * 2 pi, uploaded): per-(walker, parameter) Kahan means over the steps; bit-reversed load; butterflies (independent
 * inside a stage, so running them in parallel changes nothing); |X|^2; inverse butterflies; divide by element 0; the
 * header's compensated walker average; the windowed Kahan sum.  One CTA per parameter walks the walkers in order, so
 * the result is bit-identical to the CPU restatement the tests check it against.  quirk != 0 reproduces transferWalker's accumulation into the
 * un-cleared buffer (AutoCorrCalc.h:228-232), which chains every walker and parameter to the previous one: one CTA
 * does everything, in order.  Work buffers live in HBM/L2 (one CTA owns them; __syncthreads orders the stages).
 * ===================================================================================================================== */
namespace
{
enum { AcThreads = 1024 };

__device__ __forceinline__ unsigned acBitrev(unsigned v, int lg) { return __brev(v) >> (32 - lg); }

/* chain averages: thread per (walker, parameter), Kahan over the steps (AutoCorrCalc.h:234-262) */
__global__ void acMeansKernel(const double* __restrict__ chain, int S, int WD, double* __restrict__ mean)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= WD) return;
    double s = 0.0, c = 0.0;
    for (int t = 0; t < S; ++t) kahanAdd(s, c, chain[(size_t)t * WD + j]);
    mean[j] = __ddiv_rn(s, (double)S);
}

__device__ void acStages(double* re, double* im, const double* __restrict__ wr, const double* __restrict__ wi, double wiSign, int lg, int n)
{
    for (int s = 1; s <= lg; ++s) {
        const int m2 = 1 << (s - 1), stride = n >> s;
        for (int b = threadIdx.x; b < n / 2; b += AcThreads) {
            const int j = b & (m2 - 1), k = (b >> (s - 1)) << s;
            const double c = wr[j * stride], d = wiSign * wi[j * stride];
            const double xr = re[k + j + m2], xi = im[k + j + m2];
            const double t1r = __dsub_rn(__dmul_rn(c, xr), __dmul_rn(d, xi));
            const double t1i = __dadd_rn(__dmul_rn(c, xi), __dmul_rn(d, xr));
            const double t2r = re[k + j], t2i = im[k + j];
            re[k + j] = __dadd_rn(t2r, t1r); im[k + j] = __dadd_rn(t2i, t1i);
            re[k + j + m2] = __dsub_rn(t2r, t1r); im[k + j + m2] = __dsub_rn(t2i, t1i);
        }
        __syncthreads();
    }
}

/* grid (D or 1, G): CTA (p, g) walks walkers [g * walkersPerGroup, ...) of parameter p.  G == 1 (partOut == NULL): the whole
 * computation, bit-identical to the sequential reference order.  G > 1 (clean mode only): the group's compensated sum of the
 * normalised autocovariances goes to partOut[p][g][S]; acCombineKernel adds the groups in order and takes the windowed sum. */
__global__ void __launch_bounds__(AcThreads) acKernel(const double* __restrict__ chain, int S, int W, int D, int lg,
                                                      const double* __restrict__ twCos, const double* __restrict__ twSin,
                                                      const double* __restrict__ mean, int quirk, int windowScaling,
                                                      double* __restrict__ scratch, double* __restrict__ iat,
                                                      int walkersPerGroup, double* __restrict__ partOut)
{
    const int n = 1 << lg;
    double* base = scratch + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * (4 * (size_t)n + 3 * (size_t)S);
    double *re1 = base, *im1 = re1 + n, *re2 = im1 + n, *im2 = re2 + n, *acov = im2 + n, *avgA = acov + S, *avgC = avgA + S;
    const int WD = W * D;
    for (int i = threadIdx.x; i < S; i += AcThreads) acov[i] = 0.0;                   /* fresh buffer */
    const int pBegin = quirk ? 0 : blockIdx.x, pEnd = quirk ? D : blockIdx.x + 1;
    for (int p = pBegin; p < pEnd; ++p) {
        for (int i = threadIdx.x; i < S; i += AcThreads) { avgA[i] = 0.0; avgC[i] = 0.0; }
        __syncthreads();
        const int wBegin = blockIdx.y * walkersPerGroup, wEnd = min(W, wBegin + walkersPerGroup);
        for (int w = wBegin; w < wEnd; ++w) {
            const int off = w * D + p;
            const double mu = mean[off];
            /* transferWalker + copyCenterAndBitReverseChain */
            for (int i = threadIdx.x; i < n; i += AcThreads) {
                const unsigned r = acBitrev((unsigned)i, lg);
                double v = 0.0;
                if (i < S) {
                    double x = chain[(size_t)i * WD + off];
                    if (quirk) x = __dadd_rn(x, acov[i]);
                    v = __dsub_rn(x, mu);
                }
                re1[r] = v; im1[r] = 0.0;
            }
            __syncthreads();
            acStages(re1, im1, twCos, twSin, -1.0, lg, n);
            for (int i = threadIdx.x; i < n; i += AcThreads) {
                const unsigned r = acBitrev((unsigned)i, lg);
                re2[r] = __dadd_rn(__dmul_rn(re1[i], re1[i]), __dmul_rn(im1[i], im1[i]));
                im2[r] = 0.0;
            }
            __syncthreads();
            acStages(re2, im2, twCos, twSin, 1.0, lg, n);
            const double norm = re2[0];
            __syncthreads();
            for (int i = threadIdx.x; i < S; i += AcThreads) {
                const double a = __ddiv_rn(re2[i], norm);
                acov[i] = a;
                const double value = __dadd_rn(a, avgC[i]);                              /* averageAutocovarianceFunctions */
                const double temp = __dadd_rn(avgA[i], value);
                avgC[i] = __dsub_rn(__dsub_rn(temp, avgA[i]), value);
                avgA[i] = temp;
            }
            __syncthreads();
        }
        if (partOut) {                                                                  /* one group of many: hand the sums on */
            double* out = partOut + ((size_t)p * gridDim.y + blockIdx.y) * (size_t)S;
            for (int i = threadIdx.x; i < S; i += AcThreads) out[i] = avgA[i];
            return;
        }
        if (threadIdx.x == 0) {                                                         /* finalize + windowed sum */
            const double wcount = (double)W, factor = (double)windowScaling;
            double sum = -__ddiv_rn(avgA[0], wcount), comp = 0.0;
            bool converged = false;
            for (int i = 0; i < S; ++i) {
                const double value = __dsub_rn(__dmul_rn(2.0, __ddiv_rn(avgA[i], wcount)), comp);
                const double temp = __dadd_rn(sum, value);
                comp = __dsub_rn(__dsub_rn(temp, sum), value);
                sum = temp;
                if ((double)i > __dmul_rn(factor, sum)) { converged = true; break; }
            }
            iat[p] = converged ? sum : -sum;
        }
        __syncthreads();
    }
}
}  // namespace

namespace
{
/* grid D: lag i of parameter p = compensated sum over the walker groups, in group order; then the reference's windowed sum
 * (AutoCorrCalc.h:170-203) by thread 0 */
__global__ void __launch_bounds__(AcThreads) acCombineKernel(const double* __restrict__ part, int S, int W, int G, int windowScaling,
                                                             double* __restrict__ avg, double* __restrict__ iat)
{
    const int p = blockIdx.x;
    double* a = avg + (size_t)p * S;
    for (int i = threadIdx.x; i < S; i += AcThreads) {
        double sum = 0.0, comp = 0.0;
        for (int g = 0; g < G; ++g) kahanAdd(sum, comp, part[((size_t)p * G + g) * (size_t)S + i]);
        a[i] = sum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double wcount = (double)W, factor = (double)windowScaling;
        double sum = -__ddiv_rn(a[0], wcount), comp = 0.0;
        bool converged = false;
        for (int i = 0; i < S; ++i) {
            const double value = __dsub_rn(__dmul_rn(2.0, __ddiv_rn(a[i], wcount)), comp);
            const double temp = __dadd_rn(sum, value);
            comp = __dsub_rn(__dsub_rn(temp, sum), value);
            sum = temp;
            if ((double)i > __dmul_rn(factor, sum)) { converged = true; break; }
        }
        iat[p] = converged ? sum : -sum;
    }
}
}  // namespace

int mcmcpp_autocorr_device(const double* dChain, long long steps, int num_walkers, int num_params, int window_scaling,
                           int reference_buffer_reuse, double* iat_out)
{
    if (!dChain || !iat_out || steps < 2 || steps > (1LL << 24) || num_walkers <= 0 || num_params <= 0 || window_scaling <= 0) return MCMCPP_ERR_ARG;
    const int S = (int)steps, W = num_walkers, D = num_params;
    const int lg = (int)std::ceil(std::log2((double)S)), n = 1 << lg;          /* AutoCov::findNextPowerOfTwo, AutoCov.h:285-290 */
    /* AutoCov::calculateRootsOfUnity (AutoCov.h:270-281) on the host: the same libm values the reference uses */
    std::vector<double> cs((size_t)n / 2 + 1), sn((size_t)n / 2 + 1);
    const double twoPi = 2.0 * ((double)3141592653589793239ULL / (double)1000000000000000000ULL);
    for (int i = 0; i < n / 2; ++i) {
        const double angle = twoPi * (double)i / (double)n;
        cs[(size_t)i] = std::cos(angle);
        sn[(size_t)i] = std::sin(angle);
    }
    /* clean mode with many walkers: walker groups, so that the whole GPU works on it (2-3 CTAs per SM) */
    int G = 1;
    if (!reference_buffer_reuse && W >= 2048) {
        int sms = 148, dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        G = std::max(1, std::min(W / 64, (2 * sms + D - 1) / D));
    }
    const int perGroup = (W + G - 1) / G;
    G = (W + perGroup - 1) / perGroup;
    const dim3 grid(reference_buffer_reuse ? 1 : D, G);
    double *dCos = nullptr, *dSin = nullptr, *dMean = nullptr, *dScratch = nullptr, *dIat = nullptr, *dPart = nullptr;
    int rc = MCMCPP_OK;
    if (cudaMalloc(&dCos, sizeof(double) * cs.size()) != cudaSuccess ||
        cudaMalloc(&dSin, sizeof(double) * sn.size()) != cudaSuccess || cudaMalloc(&dMean, sizeof(double) * (size_t)W * D) != cudaSuccess ||
        cudaMalloc(&dScratch, sizeof(double) * (size_t)grid.x * grid.y * (4 * (size_t)n + 3 * (size_t)S)) != cudaSuccess ||
        cudaMalloc(&dIat, sizeof(double) * (size_t)D) != cudaSuccess ||
        (G > 1 && cudaMalloc(&dPart, sizeof(double) * ((size_t)D * G + D) * (size_t)S) != cudaSuccess))
        rc = MCMCPP_ERR_CUDA;
    if (rc == MCMCPP_OK) {
        cudaMemcpy(dCos, cs.data(), sizeof(double) * cs.size(), cudaMemcpyHostToDevice);
        cudaMemcpy(dSin, sn.data(), sizeof(double) * sn.size(), cudaMemcpyHostToDevice);
        acMeansKernel<<<(W * D + 255) / 256, 256>>>(dChain, S, W * D, dMean);
        acKernel<<<grid, AcThreads>>>(dChain, S, W, D, lg, dCos, dSin, dMean, reference_buffer_reuse ? 1 : 0, window_scaling, dScratch, dIat,
                                      perGroup, G > 1 ? dPart : nullptr);
        if (G > 1) acCombineKernel<<<D, AcThreads>>>(dPart, S, W, G, window_scaling, dPart + (size_t)D * G * S, dIat);
        if (cudaMemcpy(iat_out, dIat, sizeof(double) * (size_t)D, cudaMemcpyDeviceToHost) != cudaSuccess || cudaGetLastError() != cudaSuccess)
            rc = MCMCPP_ERR_CUDA;
    }
    cudaFree(dCos); cudaFree(dSin); cudaFree(dMean); cudaFree(dScratch); cudaFree(dIat); cudaFree(dPart);
    return rc;
}

extern "C" int mcmcpp_autocorr_times(const double* chain_host, long long steps, int num_walkers, int num_params, int window_scaling,
                                     int reference_buffer_reuse, int device, double* iat_out)
{
    if (!chain_host || !iat_out || steps < 2 || steps > (1LL << 24) || num_walkers <= 0 || num_params <= 0 || window_scaling <= 0) return MCMCPP_ERR_ARG;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || cudaSetDevice(device) != cudaSuccess) return MCMCPP_ERR_CUDA;
    const size_t chainBytes = sizeof(double) * (size_t)steps * num_walkers * num_params;
    double* dChain = nullptr;
    if (cudaMalloc(&dChain, chainBytes) != cudaSuccess) return MCMCPP_ERR_CUDA;
    int rc = cudaMemcpy(dChain, chain_host, chainBytes, cudaMemcpyHostToDevice) == cudaSuccess ? MCMCPP_OK : MCMCPP_ERR_CUDA;
    if (rc == MCMCPP_OK) rc = mcmcpp_autocorr_device(dChain, steps, num_walkers, num_params, window_scaling, reference_buffer_reuse, iat_out);
    cudaFree(dChain);
    return rc;
}

/* =====================================================================================================================
 * mcmcpp_covariance_device_noalloc (capi.h): the covariance kernels above on scratch that is NOT allocated.  No counterpart
 * in the reference; added for callers that cannot call cudaMalloc.
 *
 * The per-CTA Kahan pairs and the totals live in one file-scope device array.  Its name is only a stand-in in host code:
 * the address the kernels get is taken with cudaGetSymbolAddress, once per device, with that device current.  Every call
 * on a device uses the same array, which keeps the previous call's contents: it is cleared on the caller's stream first.
 * Calls that overlap on one GPU race on it.
 * ===================================================================================================================== */
namespace
{
enum {
    NoAllocMaxD = MCMCPP_NOALLOC_COV_MAX_PARAMS,
    NoAllocMaxItems = NoAllocMaxD + NoAllocMaxD * (NoAllocMaxD + 1) / 2,        /* 560 */
    NoAllocBlocksSmall = 148 * 8, NoAllocItemsSmall = 4 + 4 * 5 / 2,            /* D <= 4: 1184 CTAs x 14 items */
    NoAllocBlocks = 148 * 2,                                                    /* D > 4:   296 CTAs x <= 560 items */
    NoAllocPairs = NoAllocBlocks * NoAllocMaxItems,                             /* 165,760 >= 1184 x 14 */
    NoAllocDevices = 64
};
static_assert(NoAllocPairs >= NoAllocBlocksSmall * NoAllocItemsSmall, "scratch must hold the small-D layout too");

/* [NoAllocPairs] (sum, compensation) pairs, then [NoAllocMaxItems] totals: 2,656,640 bytes */
static __device__ __align__(256) double2 g_noAllocScratch[NoAllocPairs + (NoAllocMaxItems + 1) / 2];

/* device address of the array on each device (0 = not asked yet); the same value whichever thread stores it */
std::atomic<uintptr_t> g_noAllocAddress[NoAllocDevices];
}  // namespace

int mcmcpp_covariance_device_noalloc(const double* rows_dev, long long num_points, int num_params, int device,
                                     double* mean, double* cov, double* corr, void* cuda_stream)
{
    if (!rows_dev || num_points <= 0 || num_params <= 0) return mcmcpp_internal_fail(MCMCPP_ERR_ARG, "bad argument");
    if (num_params > NoAllocMaxD) {
        char msg[160];
        snprintf(msg, sizeof msg, "mcmcpp_covariance_device_noalloc: num_params = %d exceeds MCMCPP_NOALLOC_COV_MAX_PARAMS = %d "
                                  "(the static scratch array is sized for that)", num_params, (int)NoAllocMaxD);
        return mcmcpp_internal_fail(MCMCPP_ERR_UNSUPPORTED, msg);
    }
    if (device < 0 || device >= NoAllocDevices) return mcmcpp_internal_fail(MCMCPP_ERR_ARG, "device ordinal out of range");
    if (cudaSetDevice(device) != cudaSuccess) return mcmcpp_internal_fail(MCMCPP_ERR_CUDA, "cudaSetDevice failed");
    uintptr_t address = g_noAllocAddress[device].load(std::memory_order_acquire);
    if (!address) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noAllocScratch);       /* this device is current */
        if (e != cudaSuccess || !p) {
            char msg[160];
            snprintf(msg, sizeof msg, "cudaGetSymbolAddress of the static scratch failed on device %d: %s", device, cudaGetErrorString(e));
            return mcmcpp_internal_fail(MCMCPP_ERR_CUDA, msg);
        }
        address = (uintptr_t)p;
        g_noAllocAddress[device].store(address, std::memory_order_release);
    }
    double2* acc = reinterpret_cast<double2*>(address);
    double* totals = reinterpret_cast<double*>(acc + NoAllocPairs);
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const int D = num_params, nItems = D + D * (D + 1) / 2;
    const int nBlocks = D <= 4 ? NoAllocBlocksSmall : NoAllocBlocks;

    /* the array holds the last call's pairs: the ones this call reads must start at zero */
    if (cudaMemsetAsync(acc, 0, sizeof(double2) * (size_t)nBlocks * nItems, stream) != cudaSuccess)
        return mcmcpp_internal_fail(MCMCPP_ERR_CUDA, "clearing the static scratch failed");
    switch (D) {
    case 1: covSmallKernel<1><<<nBlocks, CovSmallThreads, 0, stream>>>(rows_dev, num_points, acc); break;
    case 2: covSmallKernel<2><<<nBlocks, CovSmallThreads, 0, stream>>>(rows_dev, num_points, acc); break;
    case 3: covSmallKernel<3><<<nBlocks, CovSmallThreads, 0, stream>>>(rows_dev, num_points, acc); break;
    case 4: covSmallKernel<4><<<nBlocks, CovSmallThreads, 0, stream>>>(rows_dev, num_points, acc); break;
    default: covItemKernel<<<nBlocks, CovThreads, sizeof(double) * CovChunk * D, stream>>>(rows_dev, num_points, D, nItems, acc);
    }
    covTotalsKernel<<<(nItems + 255) / 256, 256, 0, stream>>>(acc, nBlocks, nItems, totals);
    if (cudaGetLastError() != cudaSuccess) return mcmcpp_internal_fail(MCMCPP_ERR_CUDA, "covariance kernel launch failed");

    double t[NoAllocMaxItems];                                                  /* host stack: nothing is allocated */
    if (cudaMemcpyAsync(t, totals, sizeof(double) * (size_t)nItems, cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess) {
        char msg[160];
        snprintf(msg, sizeof msg, "covariance on the static scratch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return mcmcpp_internal_fail(MCMCPP_ERR_CUDA, msg);
    }
    /* finalizeMatrix, CovarianceMatrix.h:143-185 (as mcmcpp_cov_finalize, on fixed-size arrays) */
    const double ptCount = (double)num_points;
    double avg[NoAllocMaxD], cv[NoAllocMaxD * NoAllocMaxD], cr[NoAllocMaxD * NoAllocMaxD];
    for (int i = 0; i < D; ++i) avg[i] = t[i] / ptCount;
    for (int i = 0; i < D; ++i)
        for (int j = 0; j <= i; ++j) {
            const double temp = (t[D + i * (i + 1) / 2 + j] / ptCount) - (avg[i] * avg[j]);
            cv[i * D + j] = temp;
            cv[j * D + i] = temp;
        }
    for (int i = 0; i < D; ++i) {
        const double sv = std::sqrt(cv[i * D + i]);
        for (int j = 0; j < D; ++j) cr[i * D + j] = cv[i * D + j] / sv;
    }
    for (int j = 0; j < D; ++j) {
        const double sv = std::sqrt(cv[j * D + j]);
        for (int i = 0; i < D; ++i) cr[i * D + j] /= sv;
    }
    for (int i = 0; i < D; ++i) if (mean) mean[i] = avg[i];
    for (int k = 0; k < D * D; ++k) { if (cov) cov[k] = cv[k]; if (corr) corr[k] = cr[k]; }
    return MCMCPP_OK;
}
