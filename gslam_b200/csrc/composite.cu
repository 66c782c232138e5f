// K6: per-tile front-to-back alpha compositing, forward.
// Replaces gsplat rasterize_to_pixels (fork variant with n_touched /
// visibility_min_T) as called at reference gslam/rasterization.py:325-339.
//
// See composite.cuh for the decomposition: one warp per 16x8 half-tile, four
// pixels per thread, cp.async-staged rounds of 32 list entries, exact per-half-tile
// culling with ballot compaction.  The warp then walks only the kept entries:
//   phase 1 (branch-free): sigma', ex2 and alpha of its four pixels;
//   one ballot per 8x4 slot; slots without a hit are skipped warp-uniformly;
//   phase 2 (predicated): blend, transmittance update, stop / n_touched bookkeeping.
// The conic is pre-scaled by log2(e) at staging so alpha = opacity * ex2(-sigma')
// is one FMUL + MUFU.EX2.
#include "composite.cuh"

template <bool COUNT, bool STATS>
__global__ void __launch_bounds__(CMP_THREADS)
composite_fwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds,
                     int n_lin, int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     float min_T, float* __restrict__ render, float* __restrict__ alphas,
                     int32_t* __restrict__ last_ids, int32_t* __restrict__ n_touched,
                     unsigned long long* __restrict__ pair_count) {
    __shared__ WarpStage s_ws[CMP_WARPS];
    __shared__ int32_t s_ids[CMP_WARPS][NSTAGE][RND];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int half_id = blockIdx.x * CMP_WARPS + warp;
    const int lin = half_id >> 1, half = half_id & 1;     // lin = camera * T + tile
    if (lin >= n_lin) return;
    WarpStage& ws = s_ws[warp];
    const int n_tiles = tile_w * tile_h;
    const int cam = lin / n_tiles, tile = lin - cam * n_tiles;
    const int ty = tile / tile_w, tx = tile - ty * tile_w;
    const int bx = tx * GSB_TILE + (lane & 7);
    const int by = ty * GSB_TILE + half * 8 + (lane >> 3);
    const float pxf = (float)bx + 0.5f, pyf = (float)by + 0.5f;
    Rect rc;
    rc.x0 = (float)(tx * GSB_TILE) + 0.5f; rc.x1 = rc.x0 + 15.0f;
    rc.y0 = (float)(ty * GSB_TILE + half * 8) + 0.5f; rc.y1 = rc.y0 + 7.0f;

    const TileRange rg = tile_range(offsets, lin, n_lin, n_isects);
    float T[PPT], acc[PPT][GSB_CH];
    int cur[PPT], stop_at[PPT];
    unsigned live = 0;      // bit j: slot j still accumulating
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        T[j] = 1.0f; cur[j] = 0; stop_at[j] = -1;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) acc[j][k] = 0.f;
        if (bx + (j & 1) * 8 < width && by + (j >> 1) * 4 < height) live |= 1u << j;
    }
    const unsigned inside = live;

    const int n = rg.end - rg.start;
    const int n_rounds = (n + RND - 1) / RND;
    auto load_id = [&](int r) -> int {
        const int li = rg.start + r * RND + lane;
        return (r < n_rounds && li < rg.end) ? flatten_ids[li] : -1;
    };
    if (n_rounds > 0) {
        int id0 = load_id(0), id1 = load_id(1);
        int id_next = load_id(2);
        s_ids[warp][0][lane] = id0;
        issue_round(ws, 0, lane, rec, id0);
        s_ids[warp][1][lane] = id1;
        issue_round(ws, 1, lane, rec, id1);
        int stage = 0;          // ring slot of round r
        for (int r = 0; r < n_rounds; ++r) {
            if (__all_sync(full, live == 0)) break;
            int st2 = stage + 2; if (st2 >= NSTAGE) st2 -= NSTAGE;
            s_ids[warp][st2][lane] = id_next;
            issue_round(ws, st2, lane, rec, id_next);
            id_next = load_id(r + 3);
            cp_async_wait<2>();
            __syncwarp();
            // ---- cull + compact this round's 32 entries ------------------------------------
            const int li = rg.start + r * RND + lane;
            int n_keep;
            {
                const float4* raw = &ws.raw[stage][lane * 3];
                float4 o0, o1, o2;
                const bool keep = (li < rg.end) && stage_and_cull(raw[0], raw[1], raw[2], rc, li, o0, o1, o2);
                const unsigned m = __ballot_sync(full, keep);
                n_keep = __popc(m);
                if (keep) {
                    const int pos = __popc(m & ((1u << lane) - 1u));
                    ws.k0[pos] = o0; ws.k1[pos] = o1; ws.k2[pos] = o2;
                    if (COUNT) ws.gid[pos] = s_ids[warp][stage][lane];
                }
            }
            __syncwarp();
            // ---- composite the kept entries ----------------------------------------------------
            int cnt_reg = 0;    // lane k: pixels that counted kept entry k as "touched"
            for (int k = 0; k < n_keep; ++k) {
                if ((k & 7) == 7 && __all_sync(full, live == 0)) break;
                const float4 g0 = ws.k0[k];        // x, y, a', b'
                const float4 g1 = ws.k1[k];        // c', list index, opacity, r
                const float dx0 = g0.x - pxf, dx1 = dx0 - 8.0f;
                const float dy0 = g0.y - pyf, dy1 = dy0 - 4.0f;
                const float ax0 = g0.z * dx0, ax1 = g0.z * dx1;
                const float cy0 = g1.x * dy0 * dy0, cy1 = g1.x * dy1 * dy1;
                float sigma[PPT], alpha[PPT];
                sigma[0] = fmaf(dx0, fmaf(g0.w, dy0, ax0), cy0);
                sigma[1] = fmaf(dx1, fmaf(g0.w, dy0, ax1), cy0);
                sigma[2] = fmaf(dx0, fmaf(g0.w, dy1, ax0), cy1);
                sigma[3] = fmaf(dx1, fmaf(g0.w, dy1, ax1), cy1);
                bool hit[PPT];
                unsigned votes[PPT];
#pragma unroll
                for (int j = 0; j < PPT; ++j) {
                    alpha[j] = fminf(ALPHA_MAX, g1.z * fast_ex2(-sigma[j]));
                    hit[j] = ((live >> j) & 1u) && sigma[j] >= 0.f && alpha[j] >= ALPHA_MIN;
                }
#pragma unroll
                for (int j = 0; j < PPT; ++j) votes[j] = __ballot_sync(full, hit[j]);
                if ((votes[0] | votes[1] | votes[2] | votes[3]) == 0u) continue;
                const float4 g2 = ws.k2[k];        // g, b, depth, beta
                const int idx = __float_as_int(g1.y);
                int n_touch = 0;
#pragma unroll
                for (int j = 0; j < PPT; ++j) {
                    if (votes[j] == 0u) continue;              // warp-uniform
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
                    if (COUNT) n_touch += __popc(__ballot_sync(full, take && T[j] > min_T));
                    cur[j] = take ? idx : cur[j];
                    T[j] = take ? next_T : T[j];
                    if (stop) {
                        live &= ~(1u << j);
                        if (STATS) stop_at[j] = idx;
                    }
                }
                if (COUNT && lane == k) cnt_reg = n_touch;
            }
            if (COUNT && cnt_reg > 0) atomicAdd(n_touched + ws.gid[lane], cnt_reg);
            __syncwarp();
            if (++stage == NSTAGE) stage = 0;
        }
        cp_async_wait<0>();
    }
    float bg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bg[k] = backgrounds[cam * GSB_CH + k];
    }
    int visited = 0;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (!((inside >> j) & 1u)) continue;
        const int x = bx + (j & 1) * 8, y = by + (j >> 1) * 4;
        const size_t pix = ((size_t)cam * height + y) * width + x;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) render[pix * GSB_CH + k] = acc[j][k] + T[j] * bg[k];
        alphas[pix] = 1.0f - T[j];
        last_ids[pix] = cur[j];
        // P: entries walked while the pixel was still accumulating (the stopping one included)
        if (STATS) visited += stop_at[j] >= 0 ? stop_at[j] - rg.start + 1 : n;
    }
    if (STATS) {
        int v = __reduce_add_sync(full, visited);
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
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31));
    GSB_CHECK_ARG(tile_w == (width + GSB_TILE - 1) / GSB_TILE && tile_h == (height + GSB_TILE - 1) / GSB_TILE);
    const int64_t n_lin64 = (int64_t)C * tile_w * tile_h;
    GSB_CHECK_ARG(n_lin64 * 2 < ((int64_t)1 << 31));
    const int n_lin = (int)n_lin64;
    const int grid = (2 * n_lin + CMP_WARPS - 1) / CMP_WARPS;
    const float4* r4 = reinterpret_cast<const float4*>(rec);
    unsigned long long* pc = reinterpret_cast<unsigned long long*>(pair_count);
#define GSB_FWD(COUNT, STATS)                                                                     \
    composite_fwd_kernel<COUNT, STATS><<<grid, CMP_THREADS, 0, gsb_stream(stream)>>>(             \
        r4, flatten_ids, offsets, backgrounds, n_lin, width, height, tile_w, tile_h, n_isects,    \
        visibility_min_T, render, alphas, last_ids, n_touched, pc)
    if (n_touched && pc) GSB_FWD(true, true);
    else if (n_touched) GSB_FWD(true, false);
    else if (pc) GSB_FWD(false, true);
    else GSB_FWD(false, false);
#undef GSB_FWD
    GSB_LAUNCH_CHECK();
    return 0;
}
