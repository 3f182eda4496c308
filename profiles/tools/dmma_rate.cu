// DMMA issue-rate probe: warps per SM sub-partition x independent accumulator chains per warp
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int CH>
__global__ void k(double* out, int iters)
{
    double c[CH][2];
    for (int i = 0; i < CH; ++i) { c[i][0] = 0; c[i][1] = 0; }
    double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-4;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) dmma(c[i][0], c[i][1], a, b);
    }
    double s = 0;
    for (int i = 0; i < CH; ++i) s += c[i][0] + c[i][1];
    if (s == 1.2345) out[0] = s;
}
// smem-fed variant: per k-step 1 A LDS + CH B LDS, like the dense kernel
template <int CH>
__global__ void ks(double* out, int iters)
{
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < 108 * 108 + 32 * 108; i += blockDim.x) sm[i] = 1e-3 * i;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    double c[CH][2];
    for (int i = 0; i < CH; ++i) { c[i][0] = 0; c[i][1] = 0; }
    const double* arow = sm + 108 * 108 + ((warp & 3) * 8 + g) * 108 + t;
    const double* bcol = sm + t * 108 + g + ((warp >> 2) & 1) * 48;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 2
        for (int kk = 0; kk < 104; kk += 4) {
            const double av = arow[kk];
            const double* bk = bcol + kk * 108;
#pragma unroll
            for (int i = 0; i < CH; ++i) dmma(c[i][0], c[i][1], av, bk[i * 8]);
        }
    }
    double s = 0;
    for (int i = 0; i < CH; ++i) s += c[i][0] + c[i][1];
    if (s == 1.2345) out[0] = s;
}
template <class F> float timeit(F f)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main()
{
    double* out; cudaMalloc(&out, 8);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 20000;
    for (int warps : {4, 8, 16, 32}) {
        float m7 = timeit([&] { k<7><<<sms, warps * 32>>>(out, iters); });
        float m14 = timeit([&] { k<14><<<sms, warps * 32>>>(out, iters); });
        float m4 = timeit([&] { k<4><<<sms, warps * 32>>>(out, iters); });
        auto tf = [&](float ms, int ch) { return (double)sms * warps * iters * ch * 512.0 / (ms * 1e-3) / 1e12; };
        printf("reg-fed  warps/SM %2d: 4 chains %.1f TF, 7 chains %.1f TF, 14 chains %.1f TF\n", warps, tf(m4, 4), tf(m7, 7), tf(m14, 14));
    }
    const int smem = (108 * 108 + 32 * 108) * 8;
    cudaFuncSetAttribute(ks<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(ks<14>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int warps : {8, 16}) {
        float m7 = timeit([&] { ks<7><<<sms, warps * 32, smem>>>(out, 800); });
        float m14 = timeit([&] { ks<14><<<sms, warps * 32, smem>>>(out, 800); });
        auto tf = [&](float ms, int ch) { return (double)sms * warps * 800 * 26 * ch * 512.0 / (ms * 1e-3) / 1e12; };
        printf("smem-fed warps/SM %2d: 7 chains %.1f TF, 14 chains (2x LDS B) %.1f TF\n", warps, tf(m7, 7), tf(m14, 14));
    }
    return 0;
}
