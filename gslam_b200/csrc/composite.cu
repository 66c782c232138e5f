// K6: per-tile front-to-back alpha compositing, forward.
// Replaces gsplat rasterize_to_pixels (fork variant with n_touched /
// visibility_min_T) as called at reference gslam/rasterization.py:325-339.
//
// See composite.cuh for the decomposition: one warp per 8x8 quarter-tile, two
// pixels per thread, cp.async-staged rounds of 32 list entries, exact per-unit
// culling with ballot compaction.  The warp then walks only the kept entries:
//   phase 1 (branch-free): sigma', ex2 and alpha of its pixels;
//   one ballot per 8x4 slot; slots without a hit are skipped warp-uniformly;
//   phase 2 (predicated): blend, transmittance update, stop / n_touched bookkeeping.
// The conic is pre-scaled by log2(e) at staging so alpha = opacity * ex2(-sigma')
// is one FMUL + MUFU.EX2.
//
// Residency: 64 registers and 5 KB of shared memory per one-warp CTA put the SM's limit of 32
// CTAs in charge.  Measured alternatives (tools/trace_units.py): fewer resident
// warps are slower at every setting (12/16/20/24 per SM: 90/83/81/82 us against 78), and two-warp
// CTAs squeezed to 56 registers to make all 4800 units of a 640x480 view resident spill and lose
// more than the missing second wave gains.
#ifndef PPT
#define PPT 2
#endif
#define CMP_WARPS 1
#include "composite.cuh"
#ifdef GSB_TRACE
#include "trace.cuh"
GSB_TRACE_DEFINE(g_trace_fwd, gsb_debug_trace_fwd)
#endif

#ifndef FWD_MINB
#define FWD_MINB 32         // 32 one-warp CTAs per SM = 64 registers: without the bound ptxas takes 76 and only 25 fit
#endif                      // (measured 75.6 against 78.2 us; the backward kernel is the other way round: 103 registers
                            // unbounded beat 93 at BWD_MINB=20, 110.7 against 112.6 us)
template <bool COUNT, bool STATS>
__global__ void __launch_bounds__(CMP_THREADS, FWD_MINB)
composite_fwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds,
                     int n_lin, int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     float min_T, float* __restrict__ render, float* __restrict__ alphas,
                     int32_t* __restrict__ last_ids, unsigned long long* __restrict__ n_touched,
                     unsigned long long* __restrict__ pair_count) {
    __shared__ WarpStage s_ws[CMP_WARPS];
    __shared__ int32_t s_ids[CMP_WARPS][NSTAGE][RND];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int unit = blockIdx.x * CMP_WARPS + warp;
    if (unit >= n_lin * UNITS_PER_TILE) return;
    const Unit u = decode_unit(unit, lane, tile_w, tile_w * tile_h);
#ifdef GSB_TRACE
    const unsigned long long t_begin = gsb_gtimer();
    int n_kept_total = 0;
#endif
    WarpStage& ws = s_ws[warp];
    unsigned k_base = smem_addr(ws.raw);        // shared-window address of ring slot `stage`
    const unsigned k_end = k_base + NSTAGE * STAGE_BYTES;
    const float pxf = (float)u.bx + 0.5f, pyf = (float)u.by + 0.5f;

    const TileRange rg = tile_range(offsets, u.lin, n_lin, n_isects);
    // alive[j] doubles as the alpha threshold: 1/255 while the pixel accumulates, +inf once it has
    // stopped (or lies outside the image), so "alive && alpha >= 1/255" is ONE compare
    float T[PPT], acc[PPT][GSB_CH], alive[PPT];
    int cur[PPT], stop_at[PPT];
    unsigned inside = 0;
    const float DEAD = __int_as_float(0x7f800000);
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        T[j] = 1.0f; cur[j] = 0; stop_at[j] = -1; alive[j] = DEAD;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) acc[j][k] = 0.f;
        if (u.bx + SLOT_DX(j) < width && u.by + SLOT_DY(j) < height) { inside |= 1u << j; alive[j] = ALPHA_MIN; }
    }
    auto all_dead = [&]() -> bool {
        bool d = true;
#pragma unroll
        for (int j = 0; j < PPT; ++j) d = d && alive[j] == DEAD;
        return __all_sync(full, d);
    };

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
            if (all_dead()) break;
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
                float4* slot = ws.raw[stage];
                const float4 r0 = slot[lane * 3], r1 = slot[lane * 3 + 1], r2 = slot[lane * 3 + 2];
                __syncwarp();       // every lane holds its record: the slot may be overwritten
                float4 o0, o1, o2;
                const bool keep = stage_and_cull(r0, r1, r2, u.rc, li, o0, o1, o2) && (li < rg.end);
                const unsigned m = __ballot_sync(full, keep);
                n_keep = __popc(m);
                if (keep) {
                    const int pos = __popc(m & ((1u << lane) - 1u));
                    slot[pos] = o0; slot[RND + pos] = o1; slot[2 * RND + pos] = o2;
                    if (COUNT) ws.gid[pos] = s_ids[warp][stage][lane];
                }
            }
#ifdef GSB_TRACE
            n_kept_total += n_keep;
#endif
            __syncwarp();
            // ---- composite the kept entries ----------------------------------------------------
            int cnt_reg = 0;    // lane k: pixels that counted kept entry k as "touched"
            for (int k = 0; k < n_keep; ++k) {
                if ((k & 7) == 7 && all_dead()) break;
                const unsigned ka = k_base + (unsigned)k * 16u;
                const float4 g0 = lds128(ka);                        // x, y, a', b'
                const float4 g1 = lds128(ka + K_PLANE_BYTES);        // c', list index, opacity, r
                float dx[PPT], dy[PPT], sigma[PPT], alpha[PPT];
                eval_sigma(g0, g1.x, pxf, pyf, dx, dy, sigma);
                bool hit[PPT];
                unsigned votes[PPT], any = 0;
#pragma unroll
                for (int j = 0; j < PPT; ++j) {
                    alpha[j] = fminf(ALPHA_MAX, g1.z * fast_ex2(-sigma[j]));
                    hit[j] = alpha[j] >= alive[j] && sigma[j] >= 0.f;
                }
#pragma unroll
                for (int j = 0; j < PPT; ++j) { votes[j] = __ballot_sync(full, hit[j]); any |= votes[j]; }
                if (any == 0u) continue;
                const float4 g2 = lds128(ka + 2 * K_PLANE_BYTES);    // g, b, depth, beta
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
                    alive[j] = stop ? DEAD : alive[j];
                    if (STATS && stop) stop_at[j] = idx;
                }
                if (COUNT && lane == k) cnt_reg = n_touch;
            }
            if (COUNT && cnt_reg > 0) atomicAdd(n_touched + ws.gid[lane], (unsigned long long)cnt_reg);
            __syncwarp();
            k_base += STAGE_BYTES;
            if (++stage == NSTAGE) { stage = 0; k_base = k_end - NSTAGE * STAGE_BYTES; }
        }
        cp_async_wait<0>();
    }
    float bg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bg[k] = backgrounds[u.cam * GSB_CH + k];
    }
    int visited = 0;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (!((inside >> j) & 1u)) continue;
        const int x = u.bx + SLOT_DX(j), y = u.by + SLOT_DY(j);
        const size_t pix = ((size_t)u.cam * height + y) * width + x;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) render[pix * GSB_CH + k] = acc[j][k] + T[j] * bg[k];
        alphas[pix] = 1.0f - T[j];
        last_ids[pix] = cur[j];
        // P: entries walked while the pixel was still accumulating (the stopping one included)
        if (STATS) visited += stop_at[j] >= 0 ? stop_at[j] - rg.start + 1 : n;
    }
#ifdef GSB_TRACE
    gsb_trace_unit(g_trace_fwd, unit, lane, t_begin, n, n_kept_total);
#endif
    if (STATS) {
        int v = __reduce_add_sync(full, visited);
        if (lane == 0 && v) atomicAdd(pair_count, (unsigned long long)v);
    }
}

GSB_API int gsb_composite_fwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                              const float* backgrounds, int C, int N, int width, int height, int tile_w,
                              int tile_h, int64_t n_isects, float visibility_min_T, float* render,
                              float* alphas, int32_t* last_ids, int64_t* n_touched,
                              int64_t* pair_count, void* stream) {
    (void)N;
    GSB_CHECK_ARG(C >= 0 && width > 0 && height > 0 && tile_w > 0 && tile_h > 0);
    if (C == 0) return 0;
    GSB_CHECK_ARG(offsets && render && alphas && last_ids);
    GSB_CHECK_ARG(n_isects <= 0 || (rec && flatten_ids));
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31));
    GSB_CHECK_ARG(tile_w == (width + GSB_TILE - 1) / GSB_TILE && tile_h == (height + GSB_TILE - 1) / GSB_TILE);
    const int64_t n_lin64 = (int64_t)C * tile_w * tile_h;
    GSB_CHECK_ARG(n_lin64 * UNITS_PER_TILE < ((int64_t)1 << 31));
    const int n_lin = (int)n_lin64;
    const int grid = (UNITS_PER_TILE * n_lin + CMP_WARPS - 1) / CMP_WARPS;
    const float4* r4 = reinterpret_cast<const float4*>(rec);
    unsigned long long* pc = reinterpret_cast<unsigned long long*>(pair_count);
#define GSB_FWD(COUNT, STATS)                                                                     \
    composite_fwd_kernel<COUNT, STATS><<<grid, CMP_THREADS, 0, gsb_stream(stream)>>>(             \
        r4, flatten_ids, offsets, backgrounds, n_lin, width, height, tile_w, tile_h, n_isects,    \
        visibility_min_T, render, alphas, last_ids,                                              \
        reinterpret_cast<unsigned long long*>(n_touched), pc)
    if (n_touched && pc) GSB_FWD(true, true);
    else if (n_touched) GSB_FWD(true, false);
    else if (pc) GSB_FWD(false, true);
    else GSB_FWD(false, false);
#undef GSB_FWD
    GSB_LAUNCH_CHECK();
    return 0;
}
