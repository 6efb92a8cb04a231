// fma_probe.cu -- measures the FP32 issue ceiling the cascade kernels run against on this GPU:
// scalar FFMA vs packed FFMA2 (fma.rn.f32x2), at 1/2/4/8 warps per SM sub-partition, 8 independent
// dependency chains per thread.  Prints FMA lanes per clock per SM.  Development tool, not product.
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
template <int CH>
__global__ void k_scalar(float *out, float a, float b, int iters) {
    float v[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) v[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int i = 0; i < CH; i++) v[i] = __fmaf_rn(v[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += v[i];
    if (s == 12345.678f) out[0] = s;
}
template <int CH>
__global__ void k_packed(float *out, u64 a, u64 b, int iters) {
    u64 v[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) v[i] = a + threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int i = 0; i < CH; i++) v[i] = ffma2(v[i], a, b);
    }
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s ^= v[i];
    if (s == 12345ull) out[0] = 1.f;
}
// distinct-operand variants: every instruction reads three different (pairs of) registers, as the cascade does
template <int CH>
__global__ void k_packed3(float *out, const u64 *tab, int iters) {
    u64 v[CH], a[CH], b[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) { v[i] = tab[i] + threadIdx.x; a[i] = tab[CH + i] + (threadIdx.x & 1); b[i] = tab[2 * CH + i] + (threadIdx.x & 1); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int i = 0; i < CH; i++) v[i] = ffma2(v[i], a[(i + r) % CH], b[(i + 3 * r + 1) % CH]);
    }
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s ^= v[i];
    if (s == 12345ull) out[0] = 1.f;
}
template <int CH>
__global__ void k_scalar3(float *out, const float *tab, int iters) {
    float v[CH], a[CH], b[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) { v[i] = tab[i] + threadIdx.x; a[i] = tab[CH + i] + (threadIdx.x & 1); b[i] = tab[2 * CH + i] + (threadIdx.x & 1); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int i = 0; i < CH; i++) v[i] = __fmaf_rn(v[i], a[(i + r) % CH], b[(i + 3 * r + 1) % CH]);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += v[i];
    if (s == 12345.678f) out[0] = s;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float *d;
    cudaMalloc(&d, 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    const float one = 1.0000001f, eps = 1e-9f;
    unsigned ob, eb;
    memcpy(&ob, &one, 4);
    memcpy(&eb, &eps, 4);
    const u64 a2 = ((u64) ob << 32) | ob, b2 = ((u64) eb << 32) | eb;
    printf("%s  SMs %d  clock %d kHz (max; rates below assume it)\n", p.name, p.multiProcessorCount, khz);
    for (int wps = 1; wps <= 8; wps *= 2) {
        const int threads = wps * 4 * 32, blocks = p.multiProcessorCount;
        for (int packed = 0; packed < 2; packed++) {
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                if (packed) k_packed<8><<<blocks, threads>>>(d, a2, b2, iters);
                else k_scalar<8><<<blocks, threads>>>(d, one, eps, iters);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
            }
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            const double fma = (double) blocks * threads * iters * 64.0 * (packed ? 2 : 1);
            printf("warps/SMSP %d  %-6s  %.3f ms  %.3e FMA/s  %.1f FMA/clk/SM\n", wps, packed ? "FFMA2" : "FFMA", ms, fma / (ms * 1e-3),
                   fma / (ms * 1e-3) / (khz * 1e3) / blocks);
        }
    }
    {
        float hf[24];
        u64 hu[24];
        for (int i = 0; i < 24; i++) { hf[i] = i < 8 ? 0.5f + i : (i < 16 ? 1.0f - 1e-7f * i : 1e-9f * i); }
        for (int i = 0; i < 24; i++) { unsigned w; memcpy(&w, &hf[i], 4); hu[i] = ((u64) w << 32) | w; }
        float *df; u64 *du;
        cudaMalloc(&df, sizeof(hf)); cudaMalloc(&du, sizeof(hu));
        cudaMemcpy(df, hf, sizeof(hf), cudaMemcpyHostToDevice); cudaMemcpy(du, hu, sizeof(hu), cudaMemcpyHostToDevice);
        for (int wps = 1; wps <= 4; wps *= 2) {
            const int threads = wps * 4 * 32, blocks = p.multiProcessorCount;
            for (int packed = 0; packed < 2; packed++) {
                for (int rep = 0; rep < 2; rep++) {
                    cudaEventRecord(e0);
                    if (packed) k_packed3<8><<<blocks, threads>>>(d, du, iters);
                    else k_scalar3<8><<<blocks, threads>>>(d, df, iters);
                    cudaEventRecord(e1);
                    cudaEventSynchronize(e1);
                }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                const double fma = (double) blocks * threads * iters * 64.0 * (packed ? 2 : 1);
                printf("distinct operands: warps/SMSP %d  %-6s  %.3f ms  %.3e FMA/s  %.1f FMA/clk/SM\n", wps, packed ? "FFMA2" : "FFMA", ms,
                       fma / (ms * 1e-3), fma / (ms * 1e-3) / (khz * 1e3) / blocks);
            }
        }
    }
    return cudaGetLastError() != cudaSuccess;
}
