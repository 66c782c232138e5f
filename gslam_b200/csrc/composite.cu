// K6/K7: per-tile front-to-back alpha compositing, forward and backward.
// Replaces gsplat rasterize_to_pixels (fork variant with n_touched /
// visibility_min_T) as called at reference gslam/rasterization.py:325-339.
//
// One CTA of 256 threads owns one 16x16 tile of one camera; a thread owns one
// pixel.  Sorted splat records (48 B) are gathered in batches of 256 into
// shared memory as three float4 planes so the inner loop reads them as
// conflict-free broadcasts.  The conic is pre-scaled by log2(e) at staging so
// that alpha = opacity * ex2(-sigma') is one FMUL + MUFU.EX2.
#include "composite.cuh"

template <bool COUNT, bool STATS>
__global__ void __launch_bounds__(256)
composite_fwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds, int C,
                     int width, int height, int tile_w, int tile_h, int64_t n_isects, float min_T,
                     float* __restrict__ render, float* __restrict__ alphas,
                     int32_t* __restrict__ last_ids, int32_t* __restrict__ n_touched,
                     unsigned long long* __restrict__ pair_count) {
    __shared__ float4 s0[BATCH];   // x, y, opacity, a'
    __shared__ float4 s1[BATCH];   // b', c', r, g
    __shared__ float4 s2[BATCH];   // b, depth, beta, -
    __shared__ int32_t s_id[BATCH];
    __shared__ int32_t s_cnt[BATCH];

    const int cam = blockIdx.z, tile = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31;
    const int px_i = blockIdx.x * GSB_TILE + (tid & 15), py_i = blockIdx.y * GSB_TILE + (tid >> 4);
    const float px = (float)px_i + 0.5f, py = (float)py_i + 0.5f;
    const bool inside = px_i < width && py_i < height;
    bool done = !inside;

    TileRange rg = tile_range(offsets, cam, tile, tile_w * tile_h, C, n_isects);
    float T = 1.0f;
    float acc[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    int cur = 0;
    int visited = 0;   // STATS: pairs this pixel evaluated before it stopped (the work unit P)

    for (int b0 = rg.start; b0 < rg.end; b0 += BATCH) {
        // barrier: everyone is past the previous batch; leave when every pixel is saturated
        if (__syncthreads_count(done) == 256) break;
        int bs = min(BATCH, rg.end - b0);
        if (tid < bs) {
            int id = flatten_ids[b0 + tid];
            const float4* rp = rec + (size_t)id * 3;
            float4 r0 = rp[0], r1 = rp[1], r2 = rp[2];
            r0.w *= 0.5f * LOG2E;
            r1.x *= LOG2E;
            r1.y *= 0.5f * LOG2E;
            s0[tid] = r0; s1[tid] = r1; s2[tid] = r2;
            if (COUNT) { s_id[tid] = id; s_cnt[tid] = 0; }
        }
        __syncthreads();
        for (int t = 0; t < bs; ++t) {
            if ((t & 7) == 0 && __all_sync(0xffffffffu, done)) break;
            float4 g0 = s0[t], g1 = s1[t];
            float dx = g0.x - px, dy = g0.y - py;
            float sigma = g0.w * dx * dx + g1.y * dy * dy + g1.x * dx * dy;
            float alpha = fminf(ALPHA_MAX, g0.z * fast_ex2(-sigma));
            bool hit = !done && sigma >= 0.f && alpha >= ALPHA_MIN;
            bool touched = false;
            if (STATS && !done) ++visited;
            if (hit) {
                float next_T = T * (1.0f - alpha);
                if (next_T <= T_STOP) {
                    done = true;
                } else {
                    float4 g2 = s2[t];
                    float vis = alpha * T;
                    acc[0] += g1.z * vis; acc[1] += g1.w * vis; acc[2] += g2.x * vis;
                    acc[3] += g2.y * vis; acc[4] += g2.z * vis;
                    touched = T > min_T;
                    cur = b0 + t;
                    T = next_T;
                }
            }
            if (COUNT) {
                unsigned m = __ballot_sync(0xffffffffu, touched);
                if (m && lane == 0) atomicAdd(&s_cnt[t], __popc(m));
            }
        }
        if (COUNT) {
            __syncthreads();
            if (tid < bs && s_cnt[tid] > 0) atomicAdd(n_touched + s_id[tid], s_cnt[tid]);
        }
    }
    if (inside) {
        size_t pix = ((size_t)cam * height + py_i) * width + px_i;
        float bg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
        if (backgrounds) {
#pragma unroll
            for (int k = 0; k < GSB_CH; ++k) bg[k] = backgrounds[cam * GSB_CH + k];
        }
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) render[pix * GSB_CH + k] = acc[k] + T * bg[k];
        alphas[pix] = 1.0f - T;
        last_ids[pix] = cur;
    }
    if (STATS) {
        int v = __reduce_add_sync(0xffffffffu, inside ? visited : 0);
        if (lane == 0 && v) atomicAdd(pair_count, (unsigned long long)v);
    }
}

GSB_API int gsb_composite_fwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                              const float* backgrounds, int C, int N, int width, int height, int tile_w,
                              int tile_h, int64_t n_isects, float visibility_min_T, float* render,
                              float* alphas, int32_t* last_ids, int32_t* n_touched,
                              int64_t* pair_count, void* stream) {
    (void)N;
    GSB_CHECK_ARG(C >= 0 && width > 0 && height > 0 && tile_w > 0 && tile_h > 0);
    if (C == 0) return 0;
    GSB_CHECK_ARG(offsets && render && alphas && last_ids);
    GSB_CHECK_ARG(n_isects <= 0 || (rec && flatten_ids));
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31) && C <= 65535);
    dim3 grid(tile_w, tile_h, C);
    const float4* r4 = reinterpret_cast<const float4*>(rec);
    unsigned long long* pc = reinterpret_cast<unsigned long long*>(pair_count);
#define GSB_FWD(COUNT, STATS)                                                                     \
    composite_fwd_kernel<COUNT, STATS><<<grid, 256, 0, gsb_stream(stream)>>>(                     \
        r4, flatten_ids, offsets, backgrounds, C, width, height, tile_w, tile_h, n_isects,        \
        visibility_min_T, render, alphas, last_ids, n_touched, pc)
    if (n_touched && pc) GSB_FWD(true, true);
    else if (n_touched) GSB_FWD(true, false);
    else if (pc) GSB_FWD(false, true);
    else GSB_FWD(false, false);
#undef GSB_FWD
    GSB_LAUNCH_CHECK();
    return 0;
}

