// K6: per-tile front-to-back alpha compositing, forward.
// Replaces gsplat rasterize_to_pixels (fork variant with n_touched /
// visibility_min_T) as called at reference gslam/rasterization.py:325-339.
//
// One CTA of 2 warps owns one 16x16 tile of one camera and every thread owns FOUR
// pixels (slot j of lane l = pixel (l & 7, l >> 3) of the j-th 8x4 block of the
// warp's 16x8 half-tile), the same layout as the backward kernel.  Four
// independent pixels per thread amortise the shared-memory reads and the loop
// overhead and give the scheduler ILP; the small CTA (64 threads, ~60 registers)
// lets every tile of a 640x480 view be resident in a single wave on 148 SMs.
// Sorted splat records (48 B) are gathered in batches into shared memory as three
// float4 planes and read as conflict-free broadcasts.  The conic is pre-scaled by
// log2(e) at staging so that alpha = opacity * ex2(-sigma') is FMUL + MUFU.EX2.
#include "composite.cuh"

#define FWD_THREADS 64
#define FWD_BATCH 128
#define PPT 4

template <bool COUNT, bool STATS>
__global__ void __launch_bounds__(FWD_THREADS)
composite_fwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds, int C,
                     int width, int height, int tile_w, int tile_h, int64_t n_isects, float min_T,
                     float* __restrict__ render, float* __restrict__ alphas,
                     int32_t* __restrict__ last_ids, int32_t* __restrict__ n_touched,
                     unsigned long long* __restrict__ pair_count) {
    __shared__ float4 s0[FWD_BATCH];   // x, y, opacity, a'
    __shared__ float4 s1[FWD_BATCH];   // b', c', r, g
    __shared__ float4 s2[FWD_BATCH];   // b, depth, beta, -
    __shared__ int32_t s_id[FWD_BATCH];
    __shared__ int32_t s_cnt[FWD_BATCH];

    const int cam = blockIdx.z, tile = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx = blockIdx.x * GSB_TILE + (lane & 7);
    const int by = blockIdx.y * GSB_TILE + warp * 8 + (lane >> 3);
    const float pxf = (float)bx + 0.5f, pyf = (float)by + 0.5f;

    TileRange rg = tile_range(offsets, cam, tile, tile_w * tile_h, C, n_isects);
    float T[PPT], acc[PPT][GSB_CH];
    int cur[PPT], visited = 0;
    unsigned live = 0;      // bit j: slot j still accumulating
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        T[j] = 1.0f; cur[j] = 0;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) acc[j][k] = 0.f;
        if (bx + (j & 1) * 8 < width && by + (j >> 1) * 4 < height) live |= 1u << j;
    }
    const unsigned inside = live;

    for (int b0 = rg.start; b0 < rg.end; b0 += FWD_BATCH) {
        // barrier: everyone is past the previous batch; leave when every pixel is saturated
        if (__syncthreads_count(live == 0) == FWD_THREADS) break;
        const int bs = min(FWD_BATCH, rg.end - b0);
        for (int t = tid; t < bs; t += FWD_THREADS) {
            const int id = flatten_ids[b0 + t];
            stage_splat(rec + (size_t)id * 3, s0[t], s1[t], s2[t]);
            if (COUNT) { s_id[t] = id; s_cnt[t] = 0; }
        }
        __syncthreads();
        for (int t = 0; t < bs; ++t) {
            if ((t & 7) == 0 && __all_sync(0xffffffffu, live == 0)) break;
            const float4 g0 = s0[t];        // x, y, a', b'
            const float4 g1 = s1[t];        // c', thr, opacity, r
            // Phase 1, branch-free: sigma and alpha of the four pixels (independent chains overlap)
            const float dx0 = g0.x - pxf, dx1 = dx0 - 8.0f;
            const float dy0 = g0.y - pyf, dy1 = dy0 - 4.0f;
            const float ax0 = g0.z * dx0 * dx0, ax1 = g0.z * dx1 * dx1;
            const float cy0 = g1.x * dy0 * dy0, cy1 = g1.x * dy1 * dy1;
            const float bx0 = g0.w * dx0, bx1 = g0.w * dx1;
            float sigma[PPT], alpha[PPT];
            sigma[0] = fmaf(bx0, dy0, ax0 + cy0);
            sigma[1] = fmaf(bx1, dy0, ax1 + cy0);
            sigma[2] = fmaf(bx0, dy1, ax0 + cy1);
            sigma[3] = fmaf(bx1, dy1, ax1 + cy1);
            bool hit[PPT];
            unsigned any = 0;
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                alpha[j] = fminf(ALPHA_MAX, g1.z * fast_ex2(-sigma[j]));
                const bool alive = (live >> j) & 1u;
                if (STATS && alive) ++visited;
                hit[j] = alive && sigma[j] >= 0.f && alpha[j] >= ALPHA_MIN;
                any |= hit[j] ? 1u : 0u;
            }
            // Phase 2: one vote; most (tile, splat) pairs touch no pixel of this warp's half-tile
            if (!__any_sync(0xffffffffu, any != 0)) continue;
            const float4 g2 = s2[t];        // g, b, depth, beta
            int n_touch = 0;
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                // predicated, no divergence: a miss composites alpha = 0
                const float al = hit[j] ? alpha[j] : 0.f;
                const float next_T = T[j] * (1.0f - al);
                const bool stop = hit[j] && next_T <= T_STOP;
                const bool take = hit[j] && !stop;
                const float vis = take ? al * T[j] : 0.f;
                acc[j][0] = fmaf(g1.w, vis, acc[j][0]);
                acc[j][1] = fmaf(g2.x, vis, acc[j][1]);
                acc[j][2] = fmaf(g2.y, vis, acc[j][2]);
                acc[j][3] = fmaf(g2.z, vis, acc[j][3]);
                acc[j][4] = fmaf(g2.w, vis, acc[j][4]);
                const bool touched = take && T[j] > min_T;
                cur[j] = take ? b0 + t : cur[j];
                T[j] = take ? next_T : T[j];
                if (stop) live &= ~(1u << j);
                if (COUNT) n_touch += __popc(__ballot_sync(0xffffffffu, touched));
            }
            if (COUNT && n_touch && lane == 0) atomicAdd(&s_cnt[t], n_touch);
        }
        if (COUNT) {
            __syncthreads();
            for (int t = tid; t < bs; t += FWD_THREADS)
                if (s_cnt[t] > 0) atomicAdd(n_touched + s_id[t], s_cnt[t]);
        }
    }
    float bg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bg[k] = backgrounds[cam * GSB_CH + k];
    }
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (!((inside >> j) & 1u)) continue;
        const int x = bx + (j & 1) * 8, y = by + (j >> 1) * 4;
        const size_t pix = ((size_t)cam * height + y) * width + x;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) render[pix * GSB_CH + k] = acc[j][k] + T[j] * bg[k];
        alphas[pix] = 1.0f - T[j];
        last_ids[pix] = cur[j];
    }
    if (STATS) {
        int v = __reduce_add_sync(0xffffffffu, visited);
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
    composite_fwd_kernel<COUNT, STATS><<<grid, FWD_THREADS, 0, gsb_stream(stream)>>>(             \
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

