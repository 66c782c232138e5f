// K7: compositing backward (gsplat rasterize_to_pixels bwd; autograd of the call at
// reference gslam/rasterization.py:325-339).
//
// Back-to-front over the tile's sorted list, like the forward pass reversed: the
// transmittance is rebuilt by dividing T_final by (1 - alpha), the colour behind
// each splat is a running suffix sum.
//
// Same decomposition as the forward kernel (composite.cuh): one independent warp per
// 8x8 quarter-tile, TWO pixels per thread, cp.async-staged rounds of 32 entries that are
// culled against the unit and compacted before the per-pixel loop.  Consequences:
//   * a thread sums the gradients of its pixels in registers before any
//     cross-lane traffic, so the warp reduction runs once per 64 pixels;
//   * a slot is a compact 8x4 block, so small splats leave whole slots without a
//     hit and the warp skips their gradient math after one vote;
//   * the reduction is a transposing butterfly: 16 values over 32 lanes in
//     8+4+2+1+1 = 16 shuffles (a plain butterfly needs 11 x 5 = 55), after which
//     16 lanes each hold one finished sum and park it in the warp's shared-memory row;
//   * per (unit, kept Gaussian) one lane issues three float4 atomics to v_rec,
//     32 Gaussians at a time.
#include "composite.cuh"

// Sum v[0..15] over the warp.  On return lane l holds in `out` the total of value
// index idx(l) = 8*b4 + 4*b3 + 2*b2 + b1 (b_k = bit k of l); lanes l and l^1 agree.
__device__ __forceinline__ float transpose_reduce16(const float (&v)[16], int lane, int& idx) {
    const unsigned full = 0xffffffffu;
    float a[8], b[4], c[2], d;
    bool up = lane & 16;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float keep = up ? v[k + 8] : v[k], send = up ? v[k] : v[k + 8];
        a[k] = keep + __shfl_xor_sync(full, send, 16);
    }
    up = lane & 8;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float keep = up ? a[k + 4] : a[k], send = up ? a[k] : a[k + 4];
        b[k] = keep + __shfl_xor_sync(full, send, 8);
    }
    up = lane & 4;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        float keep = up ? b[k + 2] : b[k], send = up ? b[k] : b[k + 2];
        c[k] = keep + __shfl_xor_sync(full, send, 4);
    }
    up = lane & 2;
    {
        float keep = up ? c[1] : c[0], send = up ? c[0] : c[1];
        d = keep + __shfl_xor_sync(full, send, 2);
    }
    d += __shfl_xor_sync(full, d, 1);
    idx = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    return d;
}

__global__ void __launch_bounds__(CMP_THREADS)
composite_bwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds,
                     int n_lin, int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     const float* __restrict__ render_alphas, const int32_t* __restrict__ last_ids,
                     const float* __restrict__ v_render, const float* __restrict__ v_alphas,
                     float* __restrict__ v_rec, float* __restrict__ v_backgrounds) {
    __shared__ WarpStage s_ws[CMP_WARPS];
    __shared__ int32_t s_ids[CMP_WARPS][NSTAGE][RND];
    __shared__ __align__(16) float s_acc[CMP_WARPS][RND][12];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int unit = blockIdx.x * CMP_WARPS + warp;
    if (unit >= n_lin * UNITS_PER_TILE) return;
    const Unit u = decode_unit(unit, lane, tile_w, tile_w * tile_h);
    WarpStage& ws = s_ws[warp];
    float (*acc_rows)[12] = s_acc[warp];
    const float pxf = (float)u.bx + 0.5f, pyf = (float)u.by + 0.5f;

    const TileRange rg = tile_range(offsets, u.lin, n_lin, n_isects);

    float T[PPT], tail[PPT], vr[PPT][GSB_CH], buf[PPT][GSB_CH];
    int bin_final[PPT];
    float bgv[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bgv[k] = backgrounds[u.cam * GSB_CH + k];
    }
    float vbg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    int my_last = -1;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const int x = u.bx + SLOT_DX(j), y = u.by + SLOT_DY(j);
        const bool inside = x < width && y < height;
        T[j] = 1.f; tail[j] = 0.f; bin_final[j] = -1;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) { vr[j][k] = 0.f; buf[j][k] = 0.f; }
        if (inside) {
            const size_t pix = ((size_t)u.cam * height + y) * width + x;
            const float T_final = 1.0f - render_alphas[pix];
            T[j] = T_final;
            bin_final[j] = last_ids[pix];
            // last_ids == 0 also encodes "nothing accumulated"; only the tile that owns index 0
            // may recompute it (the hit test then decides)
            if (T_final == 1.0f && bin_final[j] == 0 && rg.start != 0) bin_final[j] = -1;
            float bg_dot = 0.f;
#pragma unroll
            for (int k = 0; k < GSB_CH; ++k) {
                vr[j][k] = v_render[pix * GSB_CH + k];
                bg_dot += bgv[k] * vr[j][k];
                vbg[k] += T_final * vr[j][k];
            }
            const float va = v_alphas ? v_alphas[pix] : 0.f;
            tail[j] = T_final * (va - bg_dot);     // the terms of dL/dalpha that scale with T_final / (1 - alpha)
            my_last = max(my_last, bin_final[j]);
        }
    }
    if (v_backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) {
            float s = warp_sum(vbg[k]);
            if (lane == 0 && s != 0.f) atomicAdd(v_backgrounds + u.cam * GSB_CH + k, s);
        }
    }
    const int warp_last = __reduce_max_sync(full, my_last);
    if (warp_last < rg.start || rg.end <= rg.start) return;

    const int range_end = min(rg.end, warp_last + 1);
    const int n_rounds = (range_end - rg.start + RND - 1) / RND;
    auto load_id = [&](int r) -> int {          // rounds are walked from the back
        const int li = rg.start + r * RND + lane;
        return (r >= 0 && li < range_end) ? flatten_ids[li] : -1;
    };
    int id0 = load_id(n_rounds - 1), id1 = load_id(n_rounds - 2);
    int id_next = load_id(n_rounds - 3);
    s_ids[warp][0][lane] = id0;
    issue_round(ws, 0, lane, rec, id0);
    s_ids[warp][1][lane] = id1;
    issue_round(ws, 1, lane, rec, id1);
    int stage = 0;
    for (int r = n_rounds - 1; r >= 0; --r) {
        int st2 = stage + 2; if (st2 >= NSTAGE) st2 -= NSTAGE;
        s_ids[warp][st2][lane] = id_next;
        issue_round(ws, st2, lane, rec, id_next);
        id_next = load_id(r - 3);
        cp_async_wait<2>();
        __syncwarp();
        // ---- cull + compact ---------------------------------------------------------------
        const int li = rg.start + r * RND + lane;
        int n_keep;
        {
            const float4* raw = &ws.raw[stage][lane * 3];
            float4 o0, o1, o2;
            const bool keep = (li < range_end) && stage_and_cull(raw[0], raw[1], raw[2], u.rc, li, o0, o1, o2);
            const unsigned m = __ballot_sync(full, keep);
            n_keep = __popc(m);
            if (keep) {
                const int pos = __popc(m & ((1u << lane) - 1u));
                ws.k0[pos] = o0; ws.k1[pos] = o1; ws.k2[pos] = o2;
                ws.gid[pos] = s_ids[warp][stage][lane];
            }
        }
        __syncwarp();
        unsigned wrote = 0;      // bit k: kept entry k produced a gradient row
        for (int k = n_keep - 1; k >= 0; --k) {
            const float4 g0 = ws.k0[k];        // x, y, a', b'
            const float4 g1 = ws.k1[k];        // c', list index, opacity, r
            const int idx = __float_as_int(g1.y);
            // Phase 1, branch-free: sigma', exp and alpha of the lane's pixels (chains overlap)
            float dx[PPT], dy[PPT], sigma[PPT], vis[PPT], alpha[PPT];
            eval_sigma(g0, g1.x, pxf, pyf, dx, dy, sigma);
            bool valid[PPT];
            unsigned votes[PPT], any = 0;
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                vis[j] = fast_ex2(-sigma[j]);
                alpha[j] = fminf(ALPHA_MAX, g1.z * vis[j]);
                valid[j] = idx <= bin_final[j] && sigma[j] >= 0.f && alpha[j] >= ALPHA_MIN;
            }
#pragma unroll
            for (int j = 0; j < PPT; ++j) { votes[j] = __ballot_sync(full, valid[j]); any |= votes[j]; }
            if (any == 0u) continue;
            const float4 g2 = ws.k2[k];        // g, b, depth, beta
            const float col[GSB_CH] = {g1.w, g2.x, g2.y, g2.z, g2.w};
            float g[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) g[q] = 0.f;
            // Phase 2: per 8x4 block, a warp-uniform branch; inside, misses are predicated (alpha = 0)
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                if (votes[j] == 0u) continue;
                const float al = valid[j] ? alpha[j] : 0.f;
                const float ra = fast_rcp(1.0f - al);
                const float Tn = T[j] * ra;
                const float fac = al * Tn;
                float v_alpha = tail[j] * ra;
#pragma unroll
                for (int q = 0; q < GSB_CH; ++q) {
                    g[6 + q] = fmaf(fac, vr[j][q], g[6 + q]);
                    v_alpha = fmaf(fmaf(col[q], Tn, -buf[j][q] * ra), vr[j][q], v_alpha);
                    buf[j][q] = fmaf(col[q], fac, buf[j][q]);
                }
                T[j] = Tn;
                const float ov = g1.z * vis[j];
                const float v_sigma = (valid[j] && ov <= ALPHA_MAX) ? -ov * v_alpha : 0.f;
                const float v_op = (valid[j] && ov <= ALPHA_MAX) ? vis[j] * v_alpha : 0.f;
                const float vsp = v_sigma * (1.0f / LOG2E);
                // d sigma'/d(x,y) with sigma' = a'dx^2 + b'dxdy + c'dy^2
                g[0] = fmaf(vsp, 2.f * g0.z * dx[j] + g0.w * dy[j], g[0]);
                g[1] = fmaf(vsp, g0.w * dx[j] + 2.f * g1.x * dy[j], g[1]);
                g[2] += v_op;
                g[3] = fmaf(0.5f * v_sigma * dx[j], dx[j], g[3]);
                g[4] = fmaf(v_sigma * dx[j], dy[j], g[4]);
                g[5] = fmaf(0.5f * v_sigma * dy[j], dy[j], g[5]);
            }
            int slot;
            const float total = transpose_reduce16(g, lane, slot);
            if (!(lane & 1) && slot < 12) acc_rows[k][slot] = total;     // slot 11 is always 0
            wrote |= 1u << k;
        }
        __syncwarp();
        if ((wrote >> lane) & 1u) {
            const float4* a = reinterpret_cast<const float4*>(&acc_rows[lane][0]);
            const float4 a0 = a[0], a1 = a[1], a2 = a[2];
            float4* dst = reinterpret_cast<float4*>(v_rec) + (size_t)ws.gid[lane] * 3;
            atomicAdd(dst, a0);
            atomicAdd(dst + 1, a1);
            atomicAdd(dst + 2, a2);
        }
        __syncwarp();
        if (++stage == NSTAGE) stage = 0;
    }
    cp_async_wait<0>();
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
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31));
    GSB_CHECK_ARG(tile_w == (width + GSB_TILE - 1) / GSB_TILE && tile_h == (height + GSB_TILE - 1) / GSB_TILE);
    const int64_t n_lin64 = (int64_t)C * tile_w * tile_h;
    GSB_CHECK_ARG(n_lin64 * UNITS_PER_TILE < ((int64_t)1 << 31));
    const int n_lin = (int)n_lin64;
    const int grid = (UNITS_PER_TILE * n_lin + CMP_WARPS - 1) / CMP_WARPS;
    composite_bwd_kernel<<<grid, CMP_THREADS, 0, gsb_stream(stream)>>>(
        reinterpret_cast<const float4*>(rec), flatten_ids, offsets, backgrounds, n_lin, width, height,
        tile_w, tile_h, n_isects, render_alphas, last_ids, v_render, v_alphas, v_rec, v_backgrounds);
    GSB_LAUNCH_CHECK();
    return 0;
}
