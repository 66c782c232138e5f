// K4/K5, B200-native form: counting sort by tile + per-tile depth sort in shared memory.
//
// gsplat sorts ALL (tile | depth) 64-bit keys with a global radix sort (7 passes
// over I keys) and then searches the sorted keys for the tile boundaries
// (reference call sites gslam/rasterization.py:261-274).  The keys are
// (camera, tile) major, so the same order is produced by
//   1. counting intersections per (camera, tile)          (atomics in project_fwd),
//   2. an exclusive scan over the C*T tile counters          -> isect_offsets, I,
//   3. scattering each (depth bits << 32 | flatten_id) into its tile's segment,
//   4. sorting every segment on that 64-bit word inside one CTA's shared memory.
// Ties in depth are broken by flatten_id, which is exactly the order a stable
// sort of the upstream emission order (c-major, n, tiles) gives, so the result
// is bit-identical to the global stable sort.  Nothing here needs the
// intersection count on the host: I lives in offsets[C*T].
#include "common.cuh"

// ---------------------------------------------------------------------------
// exclusive scan over the tile counters (one CTA; C*T is at most a few 10^4)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
tile_scan_kernel(const int32_t* __restrict__ counts, int n, int32_t* __restrict__ offsets,
                 int32_t* __restrict__ cursors, int64_t* __restrict__ n_isects_out) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        int i = base + tid;
        int v = (i < n) ? counts[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;   // inclusive over warps
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + (warp ? warp_tot[warp - 1] : 0) + x - v;
        if (i < n) { offsets[i] = excl; cursors[i] = excl; }
        __syncthreads();
        if (tid == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
    if (tid == 0) {
        offsets[n] = carry_s;
        if (n_isects_out) *n_isects_out = carry_s;
    }
}

GSB_API int gsb_tile_scan(const int32_t* tile_counts, int C, int tile_w, int tile_h, int32_t* offsets,
                          int32_t* cursors, int64_t* n_isects, void* stream) {
    int n = C * tile_w * tile_h;
    GSB_CHECK_ARG(n >= 0 && offsets && cursors);
    GSB_CHECK_ARG(n == 0 || tile_counts);
    tile_scan_kernel<<<1, 1024, 0, gsb_stream(stream)>>>(tile_counts, n, offsets, cursors, n_isects);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// scatter: one thread per (camera, Gaussian)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
isect_scatter_kernel(const float* __restrict__ means2d, const int32_t* __restrict__ radii,
                     const float* __restrict__ rec, int64_t CN, int N, int tile_w, int tile_h,
                     int32_t* __restrict__ cursors, int64_t capacity, uint64_t* __restrict__ dk,
                     int32_t* __restrict__ overflow) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    if (r <= 0) return;
    float2 m = reinterpret_cast<const float2*>(means2d)[i];
    int x0, x1, y0, y1;
    tile_bbox(m.x, m.y, r, tile_w, tile_h, x0, x1, y0, y1);
    int cam = (int)(i / N);
    uint64_t word = ((uint64_t)__float_as_uint(rec[i * GSB_REC + 9]) << 32) | (uint32_t)i;
    int32_t* cur = cursors + (size_t)cam * tile_w * tile_h;
    for (int ty = y0; ty < y1; ++ty)
        for (int tx = x0; tx < x1; ++tx) {
            int pos = atomicAdd(cur + ty * tile_w + tx, 1);
            if (pos < capacity) dk[pos] = word;
            else *overflow = 1;
        }
}

GSB_API int gsb_isect_scatter(const float* means2d, const int32_t* radii, const float* rec, int C, int N,
                              int tile_w, int tile_h, int32_t* cursors, int64_t capacity,
                              int64_t* depth_id, int32_t* overflow, void* stream) {
    int64_t CN = (int64_t)C * N;
    if (CN == 0) return 0;
    GSB_CHECK_ARG(means2d && radii && rec && cursors && overflow);
    GSB_CHECK_ARG(capacity == 0 || depth_id);
    GSB_CHECK_ARG(CN < ((int64_t)1 << 31) && capacity < ((int64_t)1 << 31));
    int blocks = (int)((CN + 255) / 256);
    isect_scatter_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(
        means2d, radii, rec, CN, N, tile_w, tile_h, cursors, capacity,
        reinterpret_cast<uint64_t*>(depth_id), overflow);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// per-tile segment sort (bitonic network with all compare-exchanges ascending,
// so a non-power-of-two length needs no padding: out-of-range partners are +inf)
// ---------------------------------------------------------------------------
#define SEG_THREADS 256
#define SEG_SMEM_ELEMS 8192   // 64 KB of shared memory; longer segments sort in global memory

template <typename Ptr>
__device__ __forceinline__ void bitonic_sort(Ptr a, int L) {
    int n2 = 1;
    while (n2 < L) n2 <<= 1;
    for (int k = 2; k <= n2; k <<= 1) {
        // first step of the merge: partner mirrors inside the block of size k
        for (int t = threadIdx.x; t < n2 / 2; t += SEG_THREADS) {
            int blk = t / (k / 2), off = t % (k / 2);
            int i = blk * k + off, p = blk * k + (k - 1 - off);
            if (p < L) {
                uint64_t x = a[i], y = a[p];
                if (x > y) { a[i] = y; a[p] = x; }
            }
        }
        __syncthreads();
        for (int j = k >> 2; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < n2 / 2; t += SEG_THREADS) {
                int i = ((t / j) * 2 * j) + (t % j), p = i + j;
                if (p < L) {
                    uint64_t x = a[i], y = a[p];
                    if (x > y) { a[i] = y; a[p] = x; }
                }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(SEG_THREADS)
segment_sort_kernel(uint64_t* __restrict__ dk, const int32_t* __restrict__ offsets, int n_tiles,
                    int tile_n_bits, int64_t capacity, int64_t* __restrict__ isect_ids,
                    int32_t* __restrict__ flatten_ids) {
    extern __shared__ uint64_t sm[];
    const int lin = blockIdx.x;                 // camera * n_tiles + tile
    int64_t s = offsets[lin], e = offsets[lin + 1];
    if (e > capacity) e = capacity;
    if (s >= e) return;
    const int L = (int)(e - s);
    uint64_t* seg = dk + s;
    const int cam = lin / n_tiles, tile = lin % n_tiles;
    const int64_t hi = (((int64_t)cam << tile_n_bits) | tile) << 32;
    if (L <= SEG_SMEM_ELEMS) {
        for (int t = threadIdx.x; t < L; t += SEG_THREADS) sm[t] = seg[t];
        __syncthreads();
        if (L > 1) bitonic_sort(sm, L);
        for (int t = threadIdx.x; t < L; t += SEG_THREADS) {
            uint64_t w = sm[t];
            isect_ids[s + t] = hi | (int64_t)(w >> 32);
            flatten_ids[s + t] = (int32_t)(uint32_t)w;
        }
    } else {
        bitonic_sort(seg, L);    // same network through L2; rare (a tile with > 8192 splats)
        for (int t = threadIdx.x; t < L; t += SEG_THREADS) {
            uint64_t w = seg[t];
            isect_ids[s + t] = hi | (int64_t)(w >> 32);
            flatten_ids[s + t] = (int32_t)(uint32_t)w;
        }
    }
}

GSB_API int gsb_segment_sort(int64_t* depth_id, const int32_t* offsets, int C, int tile_w, int tile_h,
                             int tile_n_bits, int64_t capacity, int64_t* isect_ids,
                             int32_t* flatten_ids, void* stream) {
    int n = C * tile_w * tile_h;
    if (n == 0 || capacity == 0) return 0;
    GSB_CHECK_ARG(depth_id && offsets && isect_ids && flatten_ids);
    static bool attr_set[64] = {false};
    size_t smem = SEG_SMEM_ELEMS * sizeof(uint64_t);
    int dev = 0;
    GSB_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev]) {   // the opt-in above 48 KB is a per-device attribute
        GSB_CUDA(cudaFuncSetAttribute(segment_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
        if (dev < 64) attr_set[dev] = true;
    }
    segment_sort_kernel<<<n, SEG_THREADS, smem, gsb_stream(stream)>>>(
        reinterpret_cast<uint64_t*>(depth_id), offsets, tile_w * tile_h, tile_n_bits, capacity, isect_ids,
        flatten_ids);
    GSB_LAUNCH_CHECK();
    return 0;
}
