// K4/K5: tile-intersection keys, scan, offsets.
// Replaces gsplat isect_tiles / isect_offset_encode as called at reference
// gslam/rasterization.py:261-274.  (The counting half lives in project_fwd.)
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "common.cuh"

// ---------------------------------------------------------------------------
// inclusive scan int32 -> int64
// ---------------------------------------------------------------------------
struct PositiveFlag {
    __host__ __device__ __forceinline__ int64_t operator()(const int32_t& x) const { return x > 0 ? 1 : 0; }
};
struct Widen {
    __host__ __device__ __forceinline__ int64_t operator()(const int32_t& x) const { return (int64_t)x; }
};

GSB_API size_t gsb_scan_workspace_bytes(int64_t n) {
    size_t bytes = 0;
    cub::TransformInputIterator<int64_t, Widen, const int32_t*> it((const int32_t*)nullptr, Widen());
    cub::DeviceScan::InclusiveSum(nullptr, bytes, it, (int64_t*)nullptr, n);
    return bytes + 256;
}

GSB_API int gsb_scan_i32_i64(const int32_t* in, int64_t n, int flag_positive, int64_t* out,
                             void* workspace, size_t workspace_bytes, void* stream) {
    if (n == 0) return 0;
    GSB_CHECK_ARG(in && out && workspace);
    size_t need = 0;
    if (flag_positive) {
        cub::TransformInputIterator<int64_t, PositiveFlag, const int32_t*> it(in, PositiveFlag());
        cub::DeviceScan::InclusiveSum(nullptr, need, it, out, n);
        if (need > workspace_bytes) { gsb_set_error("scan workspace too small"); return GSB_EWORKSPACE; }
        GSB_CUDA(cub::DeviceScan::InclusiveSum(workspace, need, it, out, n, gsb_stream(stream)));
    } else {
        cub::TransformInputIterator<int64_t, Widen, const int32_t*> it(in, Widen());
        cub::DeviceScan::InclusiveSum(nullptr, need, it, out, n);
        if (need > workspace_bytes) { gsb_set_error("scan workspace too small"); return GSB_EWORKSPACE; }
        GSB_CUDA(cub::DeviceScan::InclusiveSum(workspace, need, it, out, n, gsb_stream(stream)));
    }
    return 0;
}

// ---------------------------------------------------------------------------
// key emission: one thread per (camera, Gaussian)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
isect_emit_kernel(const float* __restrict__ means2d, const int32_t* __restrict__ radii,
                  const float* __restrict__ rec, const int64_t* __restrict__ cum, int64_t CN, int N,
                  int tile_w, int tile_h, int tile_n_bits, int64_t* __restrict__ isect_ids,
                  int32_t* __restrict__ flatten_ids) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    if (r <= 0) return;
    float2 m = reinterpret_cast<const float2*>(means2d)[i];
    int x0, x1, y0, y1;
    tile_bbox(m.x, m.y, r, tile_w, tile_h, x0, x1, y0, y1);
    int cnt = (x1 - x0) * (y1 - y0);
    if (cnt == 0) return;
    int64_t o = cum[i] - cnt;
    int64_t cam = i / N;
    uint32_t dbits = __float_as_uint(rec[i * GSB_REC + 9]);
    int64_t hi = cam << tile_n_bits;
    for (int ty = y0; ty < y1; ++ty)
        for (int tx = x0; tx < x1; ++tx) {
            int64_t tile = (int64_t)ty * tile_w + tx;
            isect_ids[o] = ((hi | tile) << 32) | (int64_t)dbits;
            flatten_ids[o] = (int32_t)i;
            ++o;
        }
}

GSB_API int gsb_isect_emit(const float* means2d, const int32_t* radii, const float* rec,
                           const int64_t* cum, int C, int N, int tile_w, int tile_h, int tile_n_bits,
                           int64_t* isect_ids, int32_t* flatten_ids, void* stream) {
    int64_t CN = (int64_t)C * N;
    if (CN == 0) return 0;
    GSB_CHECK_ARG(means2d && radii && rec && cum);
    GSB_CHECK_ARG(CN < ((int64_t)1 << 31));
    int blocks = (int)((CN + 255) / 256);
    isect_emit_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(means2d, radii, rec, cum, CN, N, tile_w,
                                                              tile_h, tile_n_bits, isect_ids,
                                                              flatten_ids);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// offsets: one thread per sorted key
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
isect_offsets_kernel(const int64_t* __restrict__ isect_ids, int64_t n_isects, int n_tiles,
                     int64_t total_tiles, int tile_n_bits, int32_t* __restrict__ offsets) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_isects) return;
    int64_t k = isect_ids[i] >> 32;
    int64_t cur = (k >> tile_n_bits) * n_tiles + (k & (((int64_t)1 << tile_n_bits) - 1));
    if (i == 0) {
        for (int64_t t = 0; t <= cur; ++t) offsets[t] = 0;
    } else {
        int64_t kp = isect_ids[i - 1] >> 32;
        int64_t prev = (kp >> tile_n_bits) * n_tiles + (kp & (((int64_t)1 << tile_n_bits) - 1));
        for (int64_t t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)i;
    }
    if (i == n_isects - 1)
        for (int64_t t = cur + 1; t < total_tiles; ++t) offsets[t] = (int32_t)n_isects;
}

GSB_API int gsb_isect_offsets(const int64_t* isect_ids, int64_t n_isects, int C, int tile_w,
                              int tile_h, int tile_n_bits, int32_t* offsets, void* stream) {
    int64_t total = (int64_t)C * tile_w * tile_h;
    if (total == 0) return 0;
    GSB_CHECK_ARG(offsets);
    if (n_isects == 0) {
        GSB_CUDA(cudaMemsetAsync(offsets, 0, total * sizeof(int32_t), gsb_stream(stream)));
        return 0;
    }
    GSB_CHECK_ARG(isect_ids && n_isects < ((int64_t)1 << 31));
    int blocks = (int)((n_isects + 255) / 256);
    isect_offsets_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(isect_ids, n_isects, tile_w * tile_h,
                                                                 total, tile_n_bits, offsets);
    GSB_LAUNCH_CHECK();
    return 0;
}
