// K4: stable LSD radix sort of (u64 key, i32 value) pairs.
// Stands where gsplat's isect_tiles sorts its keys (reference call
// gslam/rasterization.py:261-272).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

GSB_API size_t gsb_sort_workspace_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DoubleBuffer<uint64_t> k((uint64_t*)nullptr, (uint64_t*)nullptr);
    cub::DoubleBuffer<int32_t> v((int32_t*)nullptr, (int32_t*)nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, n, 0, 64);
    return bytes + 256;
}

GSB_API int gsb_sort_pairs(int64_t* keys_in, int32_t* vals_in, int64_t* keys_out, int32_t* vals_out,
                           int64_t n, int end_bit, void* workspace, size_t workspace_bytes,
                           void* stream) {
    if (n == 0) return 0;
    GSB_CHECK_ARG(keys_in && vals_in && keys_out && vals_out && workspace);
    GSB_CHECK_ARG(end_bit > 0 && end_bit <= 64);
    cub::DoubleBuffer<uint64_t> k(reinterpret_cast<uint64_t*>(keys_in), reinterpret_cast<uint64_t*>(keys_out));
    cub::DoubleBuffer<int32_t> v(vals_in, vals_out);
    size_t need = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, k, v, n, 0, end_bit);
    if (need > workspace_bytes) { gsb_set_error("sort workspace too small"); return GSB_EWORKSPACE; }
    GSB_CUDA(cub::DeviceRadixSort::SortPairs(workspace, need, k, v, n, 0, end_bit, gsb_stream(stream)));
    // the result must end in the *_out buffers
    if (k.Current() != reinterpret_cast<uint64_t*>(keys_out))
        GSB_CUDA(cudaMemcpyAsync(keys_out, k.Current(), n * sizeof(uint64_t), cudaMemcpyDeviceToDevice,
                                 gsb_stream(stream)));
    if (v.Current() != vals_out)
        GSB_CUDA(cudaMemcpyAsync(vals_out, v.Current(), n * sizeof(int32_t), cudaMemcpyDeviceToDevice,
                                 gsb_stream(stream)));
    return 0;
}
