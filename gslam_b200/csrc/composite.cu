// K6/K7: per-tile front-to-back alpha compositing, forward and backward.
// Replaces gsplat rasterize_to_pixels (fork variant with n_touched /
// visibility_min_T) as called at reference gslam/rasterization.py:325-339.
//
// One CTA of 256 threads owns one 16x16 tile of one camera; a thread owns one
// pixel.  Sorted splat records (48 B) are gathered in batches of 256 into
// shared memory as three float4 planes so the inner loop reads them as
// conflict-free broadcasts.  The conic is pre-scaled by log2(e) at staging so
// that alpha = opacity * ex2(-sigma') is one FMUL + MUFU.EX2.
#include "common.cuh"

#define BATCH 256
#define ALPHA_MAX 0.999f
#define ALPHA_MIN (1.0f / 255.0f)
#define T_STOP 1e-4f
#define LOG2E 1.4426950408889634f

struct TileRange {
    int start, end;
};

__device__ __forceinline__ TileRange tile_range(const int32_t* offsets, int cam, int tile, int n_tiles,
                                                int C, int64_t n_isects) {
    TileRange r;
    int lin = cam * n_tiles + tile;
    r.start = offsets[lin];
    // n_isects < 0: offsets has C*T+1 entries (device-side count, gsb_tile_scan)
    r.end = (lin == C * n_tiles - 1 && n_isects >= 0) ? (int)n_isects : offsets[lin + 1];
    return r;
}

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

// ---------------------------------------------------------------------------
// Backward: back-to-front over the same sorted lists.  Per (pixel, Gaussian)
// gradients are summed over the warp with shuffles, combined across the 8
// warps of the tile in shared memory, and leave the CTA as ONE vector atomic
// per record row per (tile, Gaussian).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
composite_bwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds, int C,
                     int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     const float* __restrict__ render_alphas, const int32_t* __restrict__ last_ids,
                     const float* __restrict__ v_render, const float* __restrict__ v_alphas,
                     float* __restrict__ v_rec, float* __restrict__ v_backgrounds) {
    __shared__ float4 s0[BATCH];
    __shared__ float4 s1[BATCH];
    __shared__ float4 s2[BATCH];
    __shared__ int32_t s_id[BATCH];
    __shared__ float s_acc[BATCH][GSB_REC];
    __shared__ int s_max;

    const int cam = blockIdx.z, tile = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31;
    const int px_i = blockIdx.x * GSB_TILE + (tid & 15), py_i = blockIdx.y * GSB_TILE + (tid >> 4);
    const float px = (float)px_i + 0.5f, py = (float)py_i + 0.5f;
    const bool inside = px_i < width && py_i < height;

    TileRange rg = tile_range(offsets, cam, tile, tile_w * tile_h, C, n_isects);
    size_t pix = ((size_t)cam * height + min(py_i, height - 1)) * width + min(px_i, width - 1);

    float T_final = 1.f, vr[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f}, v_a = 0.f;
    int bin_final = -1;
    if (inside) {
        T_final = 1.0f - render_alphas[pix];
        bin_final = last_ids[pix];
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) vr[k] = v_render[pix * GSB_CH + k];
        v_a = v_alphas ? v_alphas[pix] : 0.f;
        // a pixel that accumulated nothing has last_ids == 0 but must not touch index 0
        if (T_final == 1.0f && bin_final == 0) {
            // ambiguous: either nothing hit, or only index 0 hit with alpha ~ 0; treat as nothing
            // unless index 0 belongs to this tile's range (then the recompute decides).
            if (rg.start != 0) bin_final = -1;
        }
    }
    if (v_backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) {
            float s = warp_sum(inside ? T_final * vr[k] : 0.f);
            if (lane == 0 && s != 0.f) atomicAdd(v_backgrounds + cam * GSB_CH + k, s);
        }
    }
    // fold the background and alpha terms that multiply T_final * ra into one scalar
    float bg_dot = 0.f;
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bg_dot += backgrounds[cam * GSB_CH + k] * vr[k];
    }
    const float tail = T_final * (v_a - bg_dot);

    if (tid == 0) s_max = -1;
    __syncthreads();
    if (bin_final >= 0) atomicMax(&s_max, bin_final);
    __syncthreads();
    const int block_last = s_max;
    if (block_last < rg.start || rg.end <= rg.start) return;
    const int warp_last = __reduce_max_sync(0xffffffffu, bin_final);

    float T = T_final;
    float buf[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    const int n_batches = (min(rg.end, block_last + 1) - rg.start + BATCH - 1) / BATCH;
    for (int b = n_batches - 1; b >= 0; --b) {
        const int b0 = rg.start + b * BATCH;
        const int bs = min(BATCH, min(rg.end, block_last + 1) - b0);
        __syncthreads();
        if (tid < bs) {
            int id = flatten_ids[b0 + tid];
            const float4* rp = rec + (size_t)id * 3;
            float4 r0 = rp[0], r1 = rp[1];
            // same pre-scaled conic as the forward pass, so both passes take identical hit decisions
            r0.w *= 0.5f * LOG2E;
            r1.x *= LOG2E;
            r1.y *= 0.5f * LOG2E;
            s0[tid] = r0; s1[tid] = r1; s2[tid] = rp[2];
            s_id[tid] = id;
        }
#pragma unroll
        for (int k = 0; k < GSB_REC; ++k) s_acc[tid][k] = 0.f;
        __syncthreads();
        for (int t = min(bs - 1, warp_last - b0); t >= 0; --t) {
            const int idx = b0 + t;
            float4 g0 = s0[t], g1 = s1[t];
            float dx = g0.x - px, dy = g0.y - py;
            float sigma = g0.w * dx * dx + g1.y * dy * dy + g1.x * dx * dy;   // sigma * log2(e)
            float vis = fast_ex2(-sigma);
            float alpha = fminf(ALPHA_MAX, g0.z * vis);
            bool valid = inside && idx <= bin_final && sigma >= 0.f && alpha >= ALPHA_MIN;
            if (!__any_sync(0xffffffffu, valid)) continue;
            float g[11];   // v_x v_y v_opac v_a v_b v_c v_r v_g v_b v_depth v_beta
#pragma unroll
            for (int k = 0; k < 11; ++k) g[k] = 0.f;
            if (valid) {
                float4 g2 = s2[t];
                float col[GSB_CH] = {g1.z, g1.w, g2.x, g2.y, g2.z};
                float ra = 1.0f / (1.0f - alpha);
                T *= ra;
                float fac = alpha * T;
                float v_alpha = 0.f;
#pragma unroll
                for (int k = 0; k < GSB_CH; ++k) {
                    g[6 + k] = fac * vr[k];
                    v_alpha += (col[k] * T - buf[k] * ra) * vr[k];
                    buf[k] += col[k] * fac;
                }
                v_alpha += tail * ra;
                if (g0.z * vis <= ALPHA_MAX) {
                    float v_sigma = -g0.z * vis * v_alpha;
                    g[3] = 0.5f * v_sigma * dx * dx;
                    g[4] = v_sigma * dx * dy;
                    g[5] = 0.5f * v_sigma * dy * dy;
                    float vsp = v_sigma * (1.0f / LOG2E);
                    g[0] = vsp * (2.f * g0.w * dx + g1.x * dy);
                    g[1] = vsp * (g1.x * dx + 2.f * g1.y * dy);
                    g[2] = vis * v_alpha;
                }
            }
#pragma unroll
            for (int k = 0; k < 11; ++k) g[k] = warp_sum(g[k]);
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < 11; ++k) atomicAdd(&s_acc[t][k], g[k]);
            }
        }
        __syncthreads();
        if (tid < bs) {
            float4 a0 = make_float4(s_acc[tid][0], s_acc[tid][1], s_acc[tid][2], s_acc[tid][3]);
            float4 a1 = make_float4(s_acc[tid][4], s_acc[tid][5], s_acc[tid][6], s_acc[tid][7]);
            float4 a2 = make_float4(s_acc[tid][8], s_acc[tid][9], s_acc[tid][10], 0.f);
            bool nz = a0.x != 0.f || a0.y != 0.f || a0.z != 0.f || a0.w != 0.f || a1.x != 0.f ||
                      a1.y != 0.f || a1.z != 0.f || a1.w != 0.f || a2.x != 0.f || a2.y != 0.f ||
                      a2.z != 0.f;
            if (nz) {
                float4* dst = reinterpret_cast<float4*>(v_rec) + (size_t)s_id[tid] * 3;
                atomicAdd(dst, a0);
                atomicAdd(dst + 1, a1);
                atomicAdd(dst + 2, a2);
            }
        }
    }
}

GSB_API int gsb_composite_bwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                              const float* backgrounds, int C, int N, int width, int height, int tile_w,
                              int tile_h, int64_t n_isects, const float* render_alphas,
                              const int32_t* last_ids, const float* v_render, const float* v_alphas,
                              float* v_rec, float* v_backgrounds, void* stream) {
    (void)N;
    GSB_CHECK_ARG(C >= 0 && width > 0 && height > 0 && tile_w > 0 && tile_h > 0);
    if (C == 0 || n_isects == 0) return 0;
    GSB_CHECK_ARG(rec && flatten_ids && offsets && render_alphas && last_ids && v_render && v_rec);
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31) && C <= 65535);
    dim3 grid(tile_w, tile_h, C);
    composite_bwd_kernel<<<grid, 256, 0, gsb_stream(stream)>>>(
        reinterpret_cast<const float4*>(rec), flatten_ids, offsets, backgrounds, C, width, height, tile_w,
        tile_h, n_isects, render_alphas, last_ids, v_render, v_alphas, v_rec, v_backgrounds);
    GSB_LAUNCH_CHECK();
    return 0;
}
