// Micro-benchmark: issue rate of the legacy warp-level tensor path (mma.sync m16n8k8 tf32) on sm_100a,
// in the access pattern a 3xTF32 compositor would use: 8 accumulator tiles per warp, 3 chained MMAs each.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mma_tf32_rate mma_tf32_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int CHAIN>
__global__ void __launch_bounds__(32) k_mma(float* out, int iters, long long* cyc)
{
    float acc[8][4];
    for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) acc[t][i] = 0.f;
    uint32_t a[4], b[2];
    for (int i = 0; i < 4; i++) a[i] = __float_as_uint(1.0f + threadIdx.x * 0.001f + i);
    for (int i = 0; i < 2; i++) b[i] = __float_as_uint(0.5f + threadIdx.x * 0.002f + i);
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int t = 0; t < 8; t++) {
#pragma unroll
            for (int c = 0; c < CHAIN; c++) mma_tf32(acc[t], a, b);
        }
    }
    long long t1 = clock64();
    float s = 0.f;
    for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) s += acc[t][i];
    out[blockIdx.x * 32 + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// FFMA2 stream of the same shape as the current blend, for comparison: 16 packed FMAs per "splat"
__global__ void __launch_bounds__(32) k_ffma2(float* out, int iters, long long* cyc)
{
    float2 C[16];
    for (int i = 0; i < 16; i++) C[i] = make_float2(0.f, 0.f);
    float2 w = make_float2(1.0f + threadIdx.x * 1e-3f, 1.0f), f = make_float2(0.5f, 0.25f);
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int s = 0; s < 8; s++) {
#pragma unroll
            for (int i = 0; i < 16; i++) C[i] = __ffma2_rn(w, f, C[i]);
        }
    }
    long long t1 = clock64();
    float s = 0.f;
    for (int i = 0; i < 16; i++) s += C[i].x + C[i].y;
    out[blockIdx.x * 32 + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main()
{
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 64 * 32 * sizeof(float));
    cudaMallocManaged(&cyc, 8);
    const int iters = 20000;
    for (int wps : {1, 4, 8, 12, 16, 24, 32}) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float ms;
        k_mma<3><<<148 * wps, 32>>>(out, 10, cyc); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_mma<3><<<148 * wps, 32>>>(out, iters, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
        cudaEventElapsedTime(&ms, e0, e1);
        double mmas = (double)148 * wps * iters * 24;
        printf("mma chain3 warps/SM=%2d: %.3f ms, %.2f cycles per mma per SM-subpartition (clock64 %.2f cyc/mma/warp), %.1f TFLOP/s\n",
               wps, ms, ms * 1e-3 * 1.965e9 / ((double)wps * iters * 24 / 4.0 > 0 ? ((double)wps * iters * 24 / 4.0) : 1), (double)*cyc / (iters * 24.0),
               mmas * 2048 / (ms * 1e-3) / 1e12);
        k_mma<1><<<148 * wps, 32>>>(out, 10, cyc); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_mma<1><<<148 * wps, 32>>>(out, iters, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
        cudaEventElapsedTime(&ms, e0, e1);
        mmas = (double)148 * wps * iters * 8;
        printf("mma chain1 warps/SM=%2d: %.3f ms, %.1f TFLOP/s (clock64 %.2f cyc/mma/warp)\n", wps, ms, mmas * 2048 / (ms * 1e-3) / 1e12,
               (double)*cyc / (iters * 8.0));
        k_ffma2<<<148 * wps, 32>>>(out, 10, cyc); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_ffma2<<<148 * wps, 32>>>(out, iters, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
        cudaEventElapsedTime(&ms, e0, e1);
        printf("ffma2      warps/SM=%2d: %.3f ms, %.2f cyc per FFMA2 per warp (clock64), %.1f TFLOP/s fp32\n", wps, ms,
               (double)*cyc / (iters * 128.0), (double)148 * wps * iters * 128 * 32 * 4 / (ms * 1e-3) / 1e12);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
