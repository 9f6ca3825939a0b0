// Micro-benchmark: the reference's sort call, verbatim, on this box.
//
// DGR32/cuda_rasterizer/rasterizer_impl.cu:306-311:
//     cub::DeviceRadixSort::SortPairs(list_sorting_space, sorting_size, point_list_keys_unsorted, point_list_keys,
//                                     point_list_unsorted, point_list, num_rendered, 0, 32 + bit)
// with the CUB of the toolkit this box builds the reference with (CUDA 12.9 -> CCCL 2.8.2, Policy1000 onesweep).
// Built by tools/sort_vs_cub.py with nvcc for sm_100a into tools/ubench/_build/; not part of the product library.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

extern "C" {

// Returns the average milliseconds of one SortPairs call over `iters` calls (CUDA events on `stream`), after
// `warm` untimed calls; < 0 on a CUDA error.  Buffers are device pointers; the sorted result of the last call is left
// in keys_out / vals_out.
float cub_sort_pairs_ms(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out, int n,
                        int end_bit, int warm, int iters, void* stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    size_t temp_bytes = 0;
    if (cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, stream) !=
        cudaSuccess)
        return -1.0f;
    void* temp = nullptr;
    if (cudaMalloc(&temp, temp_bytes ? temp_bytes : 1) != cudaSuccess) return -2.0f;
    for (int i = 0; i < warm; i++)
        cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, stream);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, stream);
    for (int i = 0; i < iters; i++)
        cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, stream);
    cudaEventRecord(b, stream);
    float ms = -3.0f;
    if (cudaEventSynchronize(b) == cudaSuccess) {
        cudaEventElapsedTime(&ms, a, b);
        ms /= (float)(iters > 0 ? iters : 1);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(temp);
    return ms;
}

// The rest of the reference's binning chain that CUB provides: InclusiveSum over tiles_touched
// (rasterizer_impl.cu:280), same timing recipe.
float cub_inclusive_sum_ms(const uint32_t* in, uint32_t* out, int n, int warm, int iters, void* stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    size_t temp_bytes = 0;
    if (cub::DeviceScan::InclusiveSum(nullptr, temp_bytes, in, out, n, stream) != cudaSuccess) return -1.0f;
    void* temp = nullptr;
    if (cudaMalloc(&temp, temp_bytes ? temp_bytes : 1) != cudaSuccess) return -2.0f;
    for (int i = 0; i < warm; i++) cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, n, stream);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, stream);
    for (int i = 0; i < iters; i++) cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, n, stream);
    cudaEventRecord(b, stream);
    float ms = -3.0f;
    if (cudaEventSynchronize(b) == cudaSuccess) {
        cudaEventElapsedTime(&ms, a, b);
        ms /= (float)(iters > 0 ? iters : 1);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(temp);
    return ms;
}

}  // extern "C"
