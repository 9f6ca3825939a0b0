// Micro-benchmark: do HMMA.1688.F32.TF32 (warp-level tensor path) and FFMA2 issued by DIFFERENT warps of one SM
// sub-partition overlap on sm_100a?  One 512-thread CTA per SM; warps 0-7 run the tensor loop, warps 8-15 the FFMA2
// loop (each sub-partition hosts two of each).  Times: tensor alone, FFMA2 alone, both.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mma_ffma_overlap mma_ffma_overlap.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// mode bit 0: tensor warps active, bit 1: FFMA2 warps active, bit 2: ONE warp alternates both (same-warp overlap)
__global__ void __launch_bounds__(512) k(float* out, int iters, int mode)
{
    const int warp = threadIdx.x >> 5;
    float s = 0.f;
    if (mode & 4) {
        float acc[8][4];
        for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) acc[t][i] = 0.f;
        uint32_t a[4], b[2];
        for (int i = 0; i < 4; i++) a[i] = __float_as_uint(1.0f + threadIdx.x * 0.001f + i);
        for (int i = 0; i < 2; i++) b[i] = __float_as_uint(0.5f + threadIdx.x * 0.002f + i);
        float2 C[16];
        for (int i = 0; i < 16; i++) C[i] = make_float2(0.f, 0.f);
        float2 w = make_float2(1.0f + threadIdx.x * 1e-3f, 1.0f), f = make_float2(0.5f, 0.25f);
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int t = 0; t < 8; t++) {
#pragma unroll
                for (int c = 0; c < 3; c++) mma_tf32(acc[t], a, b);
#pragma unroll
                for (int i = 0; i < 16; i++) C[i] = __ffma2_rn(w, f, C[i]);   // 16 FFMA2 per 3 HMMA
            }
        }
        for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) s += acc[t][i];
        for (int i = 0; i < 16; i++) s += C[i].x + C[i].y;
    } else if (warp < 8) {
        if (mode & 1) {
            float acc[8][4];
            for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) acc[t][i] = 0.f;
            uint32_t a[4], b[2];
            for (int i = 0; i < 4; i++) a[i] = __float_as_uint(1.0f + threadIdx.x * 0.001f + i);
            for (int i = 0; i < 2; i++) b[i] = __float_as_uint(0.5f + threadIdx.x * 0.002f + i);
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int t = 0; t < 8; t++) {
#pragma unroll
                    for (int c = 0; c < 3; c++) mma_tf32(acc[t], a, b);
                }
            }
            for (int t = 0; t < 8; t++) for (int i = 0; i < 4; i++) s += acc[t][i];
        }
    } else {
        if (mode & 2) {
            float2 C[16];
            for (int i = 0; i < 16; i++) C[i] = make_float2(0.f, 0.f);
            float2 w = make_float2(1.0f + threadIdx.x * 1e-3f, 1.0f), f = make_float2(0.5f, 0.25f);
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int q = 0; q < 4; q++) {
#pragma unroll
                    for (int i = 0; i < 16; i++) C[i] = __ffma2_rn(w, f, C[i]);
                }
            }
            for (int i = 0; i < 16; i++) s += C[i].x + C[i].y;
        }
    }
    out[blockIdx.x * 512 + threadIdx.x] = s;
}

int main()
{
    float* out;
    cudaMalloc(&out, 148 * 512 * sizeof(float));
    const int iters = 20000;
    const char* names[] = {"", "tensor warps alone (8 warps x 24 HMMA/iter)", "FFMA2 warps alone (8 warps x 64 FFMA2/iter)",
                           "both, different warps", "", "", "", ""};
    for (int mode : {1, 2, 3, 4}) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float ms;
        k<<<148, 512>>>(out, 10, mode); cudaDeviceSynchronize();
        cudaEventRecord(e0); k<<<148, 512>>>(out, iters, mode); cudaEventRecord(e1); cudaDeviceSynchronize();
        cudaEventElapsedTime(&ms, e0, e1);
        if (mode == 4)
            printf("16 warps, each interleaving 24 HMMA + 128 FFMA2 per iter: %.3f ms (%.1f cycles per iter per sub-partition of 4 warps)\n", ms,
                   ms * 1e-3 * 1.965e9 / iters);
        else
            printf("%s: %.3f ms (%.1f cycles per iter)\n", names[mode], ms, ms * 1e-3 * 1.965e9 / iters);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
