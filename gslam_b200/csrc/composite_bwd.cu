// K7: compositing backward (gsplat rasterize_to_pixels bwd; autograd of the call at
// reference gslam/rasterization.py:325-339).
//
// Back-to-front over the tile's sorted list, like the forward pass reversed: the
// transmittance is rebuilt by dividing T_final by (1 - alpha), the colour behind
// each splat is a running suffix sum.
//
// Same decomposition as the forward kernel (composite.cuh): one independent warp per
// unit (16x8 half-tile, FOUR pixels per thread), cp.async-staged rounds of 32 entries that are
// culled against the unit and compacted before the per-pixel loop.  Consequences:
//   * a thread sums the gradients of its pixels in registers before any
//     cross-lane traffic, so the warp reduction runs once per 128 pixels;
//   * a slot is a compact 8x4 block, so small splats leave whole slots without a
//     hit and the warp skips their gradient math after one vote;
//   * the 11 sums of an entry cross the warp through shared memory: 11 conflict-free stores, four
//     16-byte loads and 15 adds on lanes 2c+h, one shuffle (a shuffle butterfly costs 55, a
//     transposing one 16 shuffles + 30 selects); the colour-behind state is ONE scalar per pixel
//     (behind . v_render) instead of a 5-vector;
//   * per (unit, kept Gaussian) one lane issues three float4 atomics to v_rec,
//     32 Gaussians at a time.
//
// Residency: 16x8 half-tile units (four pixels per thread) here, although the forward kernel uses
// 8x8 -- at 91 registers and 10.6 KB of shared memory 21 warps fit an SM, so all 2400 units of a
// 640x480 view are resident at once; with 8x8 units (4800 of them, 21 per SM) the second wave
// drains through a half-empty machine and costs more than the blunter culling of the larger unit
// (121 us against 113; fewer resident warps are slower still -- DESIGN.md 4.3).
#ifndef PPT
#define PPT 4
#endif
#define CMP_WARPS 1
#include "composite.cuh"
#ifdef GSB_TRACE
#include "trace.cuh"
GSB_TRACE_DEFINE(g_trace_bwd, gsb_debug_trace_bwd)
#endif

#define GRAD_N 11                       // gradient sums per (unit, Gaussian)
#define RED_STRIDE 36                   // floats per row of the reduction tile: 32 lanes + 4 pad, so the
                                        // 16-byte row reads of lanes 2c+h fall in distinct bank groups
#define RED_BYTES (GRAD_N * RED_STRIDE * 4)

__device__ __forceinline__ void sts32(unsigned addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// DET: instead of three float4 atomics, a finished row goes to det_rows[(q * det_cap + li) * 3] (q: the
// unit inside its tile, li: the entry's list index) and its bit is set in det_mask; the gather kernel
// below then sums the rows of every Gaussian in a fixed order (bit-reproducible gradients).
#ifndef BWD_MINB
#define BWD_MINB 1          // A/B knob (make variant EXTRA=-DBWD_MINB=..): resident CTAs per SM the registers must allow
#endif
template <bool DET>
__global__ void __launch_bounds__(CMP_THREADS, BWD_MINB)
composite_bwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds,
                     int n_lin, int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     const float* __restrict__ render_alphas, const int32_t* __restrict__ last_ids,
                     const float* __restrict__ v_render, const float* __restrict__ v_alphas,
                     float* __restrict__ v_rec, float* __restrict__ v_backgrounds,
                     float4* __restrict__ det_rows, unsigned* __restrict__ det_mask, int64_t det_cap) {
    __shared__ WarpStage s_ws[CMP_WARPS];
    __shared__ int32_t s_ids[CMP_WARPS][NSTAGE][RND];
    __shared__ __align__(16) float s_acc[CMP_WARPS][RND][12];
    __shared__ __align__(16) float s_red[CMP_WARPS][2][GRAD_N * RED_STRIDE];
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
    float (*acc_rows)[12] = s_acc[warp];
    unsigned k_base = smem_addr(ws.raw);        // shared-window address of ring slot `stage`
    const unsigned k_end = k_base + NSTAGE * STAGE_BYTES;
    const unsigned red_base = smem_addr(s_red[warp]);
    const float pxf = (float)u.bx + 0.5f, pyf = (float)u.by + 0.5f;

    const TileRange rg = tile_range(offsets, u.lin, n_lin, n_isects);

    float T[PPT], tail[PPT], vr[PPT][GSB_CH], behind[PPT];
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
        for (int k = 0; k < GSB_CH; ++k) vr[j][k] = 0.f;
        behind[j] = 0.f;
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
        float4* const slot = ws.raw[stage];
        {
            const float4 r0 = slot[lane * 3], r1 = slot[lane * 3 + 1], r2 = slot[lane * 3 + 2];
            __syncwarp();       // every lane holds its record: the slot may be overwritten
            float4 o0, o1, o2;
            const bool keep = stage_and_cull(r0, r1, r2, u.rc, li, o0, o1, o2) && (li < range_end);
            const unsigned m = __ballot_sync(full, keep);
            n_keep = __popc(m);
            if (keep) {
                const int pos = __popc(m & ((1u << lane) - 1u));
                slot[pos] = o0; slot[RND + pos] = o1; slot[2 * RND + pos] = o2;
                ws.gid[pos] = s_ids[warp][stage][lane];
            }
        }
#ifdef GSB_TRACE
        n_kept_total += n_keep;
#endif
        __syncwarp();
        unsigned wrote = 0;      // bit k: kept entry k produced a gradient row
        unsigned red_flip = 0;
        for (int k = n_keep - 1; k >= 0; --k) {
            const unsigned ka = k_base + (unsigned)k * 16u;
            const float4 g0 = lds128(ka);                        // x, y, a', b'
            const float4 g1 = lds128(ka + K_PLANE_BYTES);        // c', list index, opacity, r
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
            const float4 g2 = lds128(ka + 2 * K_PLANE_BYTES);    // g, b, depth, beta
            const float col[GSB_CH] = {g1.w, g2.x, g2.y, g2.z, g2.w};
            // raw sums of this lane: [0] sum w_x  [1] sum w_y  [2] v_opacity  [3] sum w_x dx
            // [4] sum w_x dy  [5] sum w_y dy  [6..10] v_colour, with w_x = v_sigma dx, w_y = v_sigma dy;
            // the conic-dependent finish (v_x, v_y, the halves) happens once per entry at the flush
            float g[GRAD_N];
#pragma unroll
            for (int q = 0; q < GRAD_N; ++q) g[q] = 0.f;
            // Phase 2: per 8x4 block, a warp-uniform branch; inside, misses are predicated (alpha = 0)
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
#ifndef BWD_NOSKIP
                if (votes[j] == 0u) continue;
#endif
                const float al = valid[j] ? alpha[j] : 0.f;
                const float ra = fast_rcp(1.0f - al);
                const float Tn = T[j] * ra;          // transmittance in front of this splat
                const float fac = al * Tn;
                float E = col[0] * vr[j][0];         // <colour, v_render>
#pragma unroll
                for (int q = 1; q < GSB_CH; ++q) E = fmaf(col[q], vr[j][q], E);
#pragma unroll
                for (int q = 0; q < GSB_CH; ++q) g[6 + q] = fmaf(fac, vr[j][q], g[6 + q]);
                // dL/dalpha = T E - (behind . v_render + T_final (bg . v_render - v_alpha)) / (1 - alpha)
                const float v_alpha = fmaf(Tn, E, ra * (tail[j] - behind[j]));
                behind[j] = fmaf(fac, E, behind[j]);
                T[j] = Tn;
                const float ov = g1.z * vis[j];
                const bool live = valid[j] && ov <= ALPHA_MAX;
                const float v_sigma = live ? -ov * v_alpha : 0.f;
                g[2] += live ? vis[j] * v_alpha : 0.f;
                const float wx = v_sigma * dx[j], wy = v_sigma * dy[j];
                g[0] += wx;
                g[1] += wy;
                g[3] = fmaf(wx, dx[j], g[3]);
                g[4] = fmaf(wx, dy[j], g[4]);
                g[5] = fmaf(wy, dy[j], g[5]);
            }
            // ---- warp reduction through shared memory: 11 conflict-free stores, then lane 2c+h sums
            // half a row with four 16-byte loads; one shuffle joins the halves
            const unsigned rb = red_base + red_flip;
            red_flip ^= RED_BYTES;      // double buffer: the next entry's stores cannot race these loads
#pragma unroll
            for (int q = 0; q < GRAD_N; ++q) sts32(rb + (unsigned)(q * RED_STRIDE * 4) + lane * 4u, g[q]);
            __syncwarp();
            float total = 0.f;
            if (lane < 2 * GRAD_N) {
                const unsigned ra4 = rb + (unsigned)((lane >> 1) * RED_STRIDE * 4 + (lane & 1) * 64);
                const float4 v0 = lds128(ra4), v1 = lds128(ra4 + 16), v2 = lds128(ra4 + 32), v3 = lds128(ra4 + 48);
                total = (((v0.x + v0.y) + (v0.z + v0.w)) + ((v1.x + v1.y) + (v1.z + v1.w))) +
                        (((v2.x + v2.y) + (v2.z + v2.w)) + ((v3.x + v3.y) + (v3.z + v3.w)));
            }
            total += __shfl_xor_sync(full, total, 1);
            if (lane < 2 * GRAD_N && !(lane & 1)) acc_rows[k][lane >> 1] = total;
            wrote |= 1u << k;
        }
        __syncwarp();
        if ((wrote >> lane) & 1u) {
            const float4* a = reinterpret_cast<const float4*>(&acc_rows[lane][0]);
            float4 a0 = a[0], a1 = a[1], a2 = a[2];
            // finish with this entry's conic: d sigma'/d(x,y) = (2a'dx + b'dy, b'dx + 2c'dy), sigma' = sigma log2(e)
            const float4 c0 = slot[lane];
            const float cc = slot[RND + lane].x;
            const float sx = a0.x * (1.0f / LOG2E), sy = a0.y * (1.0f / LOG2E);
            a0.x = 2.f * c0.z * sx + c0.w * sy;
            a0.y = c0.w * sx + 2.f * cc * sy;
            a0.w *= 0.5f;            // d sigma / d conic_a = dx^2 / 2
            a1.y *= 0.5f;            // d sigma / d conic_c = dy^2 / 2
            a2.w = 0.f;
            if (DET) {
                const int64_t row = (int64_t)(unit % UNITS_PER_TILE) * det_cap + __float_as_int(slot[RND + lane].y);
                float4* dst = det_rows + row * 3;
                dst[0] = a0; dst[1] = a1; dst[2] = a2;
                atomicOr(det_mask + (row >> 5), 1u << (row & 31));
            } else {
                float4* dst = reinterpret_cast<float4*>(v_rec) + (size_t)ws.gid[lane] * 3;
                atomicAdd(dst, a0);
                atomicAdd(dst + 1, a1);
                atomicAdd(dst + 2, a2);
            }
        }
        __syncwarp();
        k_base += STAGE_BYTES;
        if (++stage == NSTAGE) { stage = 0; k_base = k_end - NSTAGE * STAGE_BYTES; }
    }
    cp_async_wait<0>();
#ifdef GSB_TRACE
    gsb_trace_unit(g_trace_bwd, unit, lane, t_begin, range_end - rg.start, n_kept_total);
#endif
}

// ---------------------------------------------------------------------------
// Deterministic accumulation, second half: one thread per (camera, Gaussian) sums the rows the
// units wrote for it -- its tile rectangle in row-major order, the units of a tile in order -- so the
// gradient is a fixed-order sum, bit-identical from run to run.  The entry's position in a tile's
// sorted list is found by a binary search for its depth bits (ties: scan for the flatten id; the
// list is sorted by (depth bits, flatten id)).  v_rec is written for every (c, n), so the caller need
// not zero it.
// ---------------------------------------------------------------------------
struct Row12 {
    float4 a, b, c;
};

__device__ __forceinline__ void row_add(Row12& r, const float4 a, const float4 b, const float4 c) {
    r.a.x += a.x; r.a.y += a.y; r.a.z += a.z; r.a.w += a.w;
    r.b.x += b.x; r.b.y += b.y; r.b.z += b.z; r.b.w += b.w;
    r.c.x += c.x; r.c.y += c.y; r.c.z += c.z; r.c.w += c.w;
}

__device__ __forceinline__ void gather_tile(Row12& acc, int lin, uint32_t dbits, int id,
                                            const int32_t* __restrict__ offsets, int n_lin, int64_t n_isects,
                                            const int64_t* __restrict__ isect_ids,
                                            const int32_t* __restrict__ flatten_ids,
                                            const float4* __restrict__ det_rows,
                                            const unsigned* __restrict__ det_mask, int64_t det_cap) {
    const TileRange rg = tile_range(offsets, lin, n_lin, n_isects);
    int lo = rg.start, hi = rg.end;
    while (lo < hi) {       // first entry whose depth bits are >= dbits
        const int mid = (lo + hi) >> 1;
        if ((uint32_t)isect_ids[mid] < dbits) lo = mid + 1; else hi = mid;
    }
    while (lo < rg.end && flatten_ids[lo] != id) {
        if ((uint32_t)isect_ids[lo] != dbits) return;       // not in this list (clipped by an overflowed capacity)
        ++lo;
    }
    if (lo >= rg.end) return;
#pragma unroll
    for (int q = 0; q < UNITS_PER_TILE; ++q) {
        const int64_t row = (int64_t)q * det_cap + lo;
        if ((det_mask[row >> 5] >> (row & 31)) & 1u) {
            const float4* src = det_rows + row * 3;
            row_add(acc, src[0], src[1], src[2]);
        }
    }
}

// The (Gaussian, tile) pairs of a warp's 32 Gaussians are dealt to the lanes in flattened order, 32 per
// round, so the dependent loads of the 32 searches overlap and a Gaussian that covers hundreds of tiles
// costs more rounds, not one lane's serial walk (measured at 200 k Gaussians / 1.19 M intersections: the
// one-lane-per-Gaussian form of this kernel took 250 us -- its slowest lane walked 48 tiles at ~7 us each).
// A round's 32 rows go through shared memory; every owner then adds ITS rows of the round in pair order,
// and the rounds are sequential, so each Gaussian's sum runs over its tiles in row-major order: fixed.
__global__ void __launch_bounds__(256)
composite_bwd_gather_kernel(const float* __restrict__ rec, const int32_t* __restrict__ radii, int64_t CN, int N,
                            int tile_w, int tile_h, const int32_t* __restrict__ offsets, int n_lin,
                            int64_t n_isects, const int64_t* __restrict__ isect_ids,
                            const int32_t* __restrict__ flatten_ids, const float4* __restrict__ det_rows,
                            const unsigned* __restrict__ det_mask, int64_t det_cap, float4* __restrict__ v_rec) {
    __shared__ float4 s_rows[8][32][3];
    const unsigned full = 0xffffffffu;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = (i < CN) ? radii[i] : 0;
    int x0 = 0, x1 = 0, y0 = 0, y1 = 0, cnt = 0, cam = 0;
    uint32_t dbits = 0;
    if (r > 0) {
        const float4 r0 = reinterpret_cast<const float4*>(rec)[i * 3];
        tile_bbox(r0.x, r0.y, r, tile_w, tile_h, x0, x1, y0, y1);
        cnt = (x1 - x0) * (y1 - y0);
        cam = (int)(i / N);
        dbits = __float_as_uint(rec[i * GSB_REC + 9]);
    }
    const int n_tiles = tile_w * tile_h;
    const int bw = max(x1 - x0, 1);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(full, incl, o);
        if (lane >= o) incl += y;
    }
    const int excl = incl - cnt;
    const int total = __shfl_sync(full, incl, 31);
    Row12 acc;
    acc.a = acc.b = acc.c = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int base = 0; base < total; base += 32) {
        const int p = base + lane;
        // owner of pair p: the last lane whose first pair is <= p (zero-count lanes share their successor's start)
        int owner = 0;
#pragma unroll
        for (int st = 16; st > 0; st >>= 1) {
            const int cand = owner + st;
            const int e = __shfl_sync(full, excl, cand & 31);
            if (cand < 32 && e <= p) owner = cand;
        }
        const int t = p - __shfl_sync(full, excl, owner);
        const int o_x0 = __shfl_sync(full, x0, owner), o_y0 = __shfl_sync(full, y0, owner);
        const int o_bw = __shfl_sync(full, bw, owner), o_cam = __shfl_sync(full, cam, owner);
        const uint32_t o_d = __shfl_sync(full, dbits, owner);
        const int o_id = (int)__shfl_sync(full, (int)i, owner);
        Row12 part;
        part.a = part.b = part.c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p < total) {
            const int dy = t / o_bw;
            gather_tile(part, o_cam * n_tiles + (o_y0 + dy) * tile_w + o_x0 + (t - dy * o_bw), o_d, o_id, offsets,
                        n_lin, n_isects, isect_ids, flatten_ids, det_rows, det_mask, det_cap);
        }
        s_rows[warp][lane][0] = part.a; s_rows[warp][lane][1] = part.b; s_rows[warp][lane][2] = part.c;
        __syncwarp();
        const int qa = max(excl, base) - base, qb = min(excl + cnt, base + 32) - base;
        for (int q = qa; q < qb; ++q) row_add(acc, s_rows[warp][q][0], s_rows[warp][q][1], s_rows[warp][q][2]);
        __syncwarp();
    }
    if (i < CN) {
        float4* dst = v_rec + i * 3;
        dst[0] = acc.a; dst[1] = acc.b; dst[2] = acc.c;
    }
}

GSB_API size_t gsb_composite_bwd_det_bytes(int64_t capacity) {
    if (capacity <= 0) return 0;
    const size_t rows = (size_t)UNITS_PER_TILE * (size_t)capacity;
    return rows * 48 + ((rows + 31) / 32) * 4 + 16;
}

GSB_API int gsb_composite_bwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                              const float* backgrounds, int C, int N, int width, int height, int tile_w,
                              int tile_h, int64_t n_isects, const float* render_alphas,
                              const int32_t* last_ids, const float* v_render, const float* v_alphas,
                              float* v_rec, float* v_backgrounds, const int32_t* radii,
                              const int64_t* isect_ids, void* det_workspace, size_t det_workspace_bytes,
                              void* stream) {
    GSB_CHECK_ARG(C >= 0 && width > 0 && height > 0 && tile_w > 0 && tile_h > 0);
    const bool det = det_workspace != nullptr;
    if (C == 0) return 0;
    if (n_isects == 0) {
        if (det && N > 0) GSB_CUDA(cudaMemsetAsync(v_rec, 0, (size_t)C * N * GSB_REC * 4, gsb_stream(stream)));
        return 0;
    }
    // flatten_ids may be NULL only when it holds nothing (device-side count, capacity 0: every range is empty)
    GSB_CHECK_ARG(rec && (flatten_ids || n_isects == -1) && offsets && render_alphas && last_ids && v_render && v_rec);
    GSB_CHECK_ARG(n_isects < ((int64_t)1 << 31));
    GSB_CHECK_ARG(tile_w == (width + GSB_TILE - 1) / GSB_TILE && tile_h == (height + GSB_TILE - 1) / GSB_TILE);
    const int64_t n_lin64 = (int64_t)C * tile_w * tile_h;
    GSB_CHECK_ARG(n_lin64 * UNITS_PER_TILE < ((int64_t)1 << 31));
    const int n_lin = (int)n_lin64;
    const int grid = (UNITS_PER_TILE * n_lin + CMP_WARPS - 1) / CMP_WARPS;
    const float4* r4 = reinterpret_cast<const float4*>(rec);
    if (!det) {
        composite_bwd_kernel<false><<<grid, CMP_THREADS, 0, gsb_stream(stream)>>>(
            r4, flatten_ids, offsets, backgrounds, n_lin, width, height, tile_w, tile_h, n_isects,
            render_alphas, last_ids, v_render, v_alphas, v_rec, v_backgrounds, nullptr, nullptr, 0);
        GSB_LAUNCH_CHECK();
        return 0;
    }
    // deterministic accumulation: rows + mask in the caller's workspace, then the fixed-order gather
    const int64_t cap = n_isects > 0 ? n_isects : -n_isects - 1;
    GSB_CHECK_ARG(cap > 0 && radii && isect_ids && flatten_ids && v_backgrounds == nullptr);
    GSB_CHECK_ARG((reinterpret_cast<uintptr_t>(det_workspace) & 15) == 0);
    if (det_workspace_bytes < gsb_composite_bwd_det_bytes(cap)) {
        gsb_set_error("gsb_composite_bwd: deterministic workspace of %zu bytes, %zu needed", det_workspace_bytes,
                      gsb_composite_bwd_det_bytes(cap));
        return GSB_EWORKSPACE;
    }
    const size_t rows = (size_t)UNITS_PER_TILE * (size_t)cap;
    float4* det_rows = reinterpret_cast<float4*>(det_workspace);
    unsigned* det_mask = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(det_workspace) + rows * 48);
    GSB_CUDA(cudaMemsetAsync(det_mask, 0, ((rows + 31) / 32) * 4, gsb_stream(stream)));
    composite_bwd_kernel<true><<<grid, CMP_THREADS, 0, gsb_stream(stream)>>>(
        r4, flatten_ids, offsets, backgrounds, n_lin, width, height, tile_w, tile_h, n_isects, render_alphas,
        last_ids, v_render, v_alphas, v_rec, nullptr, det_rows, det_mask, cap);
    GSB_LAUNCH_CHECK();
    const int64_t CN = (int64_t)C * N;
    composite_bwd_gather_kernel<<<(unsigned)((CN + 255) / 256), 256, 0, gsb_stream(stream)>>>(
        rec, radii, CN, N, tile_w, tile_h, offsets, n_lin, n_isects, isect_ids, flatten_ids, det_rows, det_mask,
        cap, reinterpret_cast<float4*>(v_rec));
    GSB_LAUNCH_CHECK();
    return 0;
}
