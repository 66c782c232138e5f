// K7: compositing backward (gsplat rasterize_to_pixels bwd; autograd of the call at
// reference gslam/rasterization.py:325-339).
//
// Back-to-front over the tile's sorted list, like the forward pass reversed: the
// transmittance is rebuilt by dividing T_final by (1 - alpha), the colour behind
// each splat is a running suffix sum.
//
// Layout: one CTA of 2 warps per 16x16 tile, FOUR pixels per thread.  Warp w owns
// the 16x8 half-tile; pixel slot j of lane l is pixel (l & 7, l >> 3) of the j-th
// 8x4 block of that half.  Consequences:
//   * a thread sums the gradients of its 4 pixels in registers before any
//     cross-lane traffic, so the warp reduction runs once per 128 pixels;
//   * a slot is a compact 8x4 block, so small splats leave whole slots without a
//     hit and the warp skips their gradient math after one vote;
//   * the reduction is a transposing butterfly: 16 values over 32 lanes in
//     8+4+2+1+1 = 16 shuffles (a plain butterfly needs 11 x 5 = 55), after which
//     16 lanes each hold one finished sum and add it to shared memory in parallel;
//   * per (tile, Gaussian) the CTA issues three float4 atomics to v_rec.
#include "composite.cuh"

#define BWD_THREADS 64
#define BWD_BATCH 128
#define PPT 4

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

// 9 CTAs per SM (<= 113 registers): 148 x 9 = 1332 slots hold all 1200 tiles of a 640x480 view in
// one wave; at 8 per SM the last 16 tiles ran as a second wave on an otherwise idle GPU.
__global__ void __launch_bounds__(BWD_THREADS, 9)
composite_bwd_kernel(const float4* __restrict__ rec, const int32_t* __restrict__ flatten_ids,
                     const int32_t* __restrict__ offsets, const float* __restrict__ backgrounds, int C,
                     int width, int height, int tile_w, int tile_h, int64_t n_isects,
                     const float* __restrict__ render_alphas, const int32_t* __restrict__ last_ids,
                     const float* __restrict__ v_render, const float* __restrict__ v_alphas,
                     float* __restrict__ v_rec, float* __restrict__ v_backgrounds) {
    __shared__ float4 s0[BWD_BATCH];   // x, y, opacity, a'   (conic pre-scaled as in the forward pass)
    __shared__ float4 s1[BWD_BATCH];   // b', c', r, g
    __shared__ float4 s2[BWD_BATCH];   // b, depth, beta, -
    __shared__ int32_t s_id[BWD_BATCH];
    __shared__ __align__(16) float s_acc[BWD_BATCH][16];
    __shared__ int s_max[2];

    const int cam = blockIdx.z, tile = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // slot j -> pixel (bx + (j & 1) * 8, by + (j >> 1) * 4)
    const int bx = blockIdx.x * GSB_TILE + (lane & 7);
    const int by = blockIdx.y * GSB_TILE + warp * 8 + (lane >> 3);
    const float pxf = (float)bx + 0.5f, pyf = (float)by + 0.5f;

    TileRange rg = tile_range(offsets, cam, tile, tile_w * tile_h, C, n_isects);

    float T[PPT], tail[PPT], vr[PPT][GSB_CH], buf[PPT][GSB_CH];
    int bin_final[PPT];
    float bgv[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (backgrounds) {
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) bgv[k] = backgrounds[cam * GSB_CH + k];
    }
    float vbg[GSB_CH] = {0.f, 0.f, 0.f, 0.f, 0.f};
    int my_last = -1;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const int x = bx + (j & 1) * 8, y = by + (j >> 1) * 4;
        const bool inside = x < width && y < height;
        T[j] = 1.f; tail[j] = 0.f; bin_final[j] = -1;
#pragma unroll
        for (int k = 0; k < GSB_CH; ++k) { vr[j][k] = 0.f; buf[j][k] = 0.f; }
        if (inside) {
            const size_t pix = ((size_t)cam * height + y) * width + x;
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
            if (lane == 0 && s != 0.f) atomicAdd(v_backgrounds + cam * GSB_CH + k, s);
        }
    }
    const int warp_last = __reduce_max_sync(0xffffffffu, my_last);
    int slot_last[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) slot_last[j] = __reduce_max_sync(0xffffffffu, bin_final[j]);
    if (lane == 0) s_max[warp] = warp_last;
    __syncthreads();
    const int block_last = max(s_max[0], s_max[1]);
    if (block_last < rg.start || rg.end <= rg.start) return;

    const int range_end = min(rg.end, block_last + 1);
    const int n_batches = (range_end - rg.start + BWD_BATCH - 1) / BWD_BATCH;
    for (int b = n_batches - 1; b >= 0; --b) {
        const int b0 = rg.start + b * BWD_BATCH;
        const int bs = min(BWD_BATCH, range_end - b0);
        __syncthreads();
        for (int t = tid; t < bs; t += BWD_THREADS) {
            const int id = flatten_ids[b0 + t];
            stage_splat(rec + (size_t)id * 3, s0[t], s1[t], s2[t]);
            s_id[t] = id;
        }
        {
            float4* z = reinterpret_cast<float4*>(&s_acc[0][0]);
            for (int t = tid; t < bs * 4; t += BWD_THREADS) z[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
        for (int t = min(bs - 1, warp_last - b0); t >= 0; --t) {
            const int idx = b0 + t;
            const float4 g0 = s0[t];        // x, y, a', b'
            const float4 g1 = s1[t];        // c', thr, opacity, r
            // Phase 1, branch-free: sigma', exp and alpha of the four pixels (chains overlap)
            const float dx0 = g0.x - pxf, dx1 = dx0 - 8.0f;
            const float dy0 = g0.y - pyf, dy1 = dy0 - 4.0f;
            const float ax0 = g0.z * dx0 * dx0, ax1 = g0.z * dx1 * dx1;
            const float cy0 = g1.x * dy0 * dy0, cy1 = g1.x * dy1 * dy1;
            const float bx0 = g0.w * dx0, bx1 = g0.w * dx1;
            float sigma[PPT], vis[PPT], alpha[PPT];
            sigma[0] = fmaf(bx0, dy0, ax0 + cy0);
            sigma[1] = fmaf(bx1, dy0, ax1 + cy0);
            sigma[2] = fmaf(bx0, dy1, ax0 + cy1);
            sigma[3] = fmaf(bx1, dy1, ax1 + cy1);
            bool valid[PPT];
            unsigned votes[PPT];
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                vis[j] = fast_ex2(-sigma[j]);
                alpha[j] = fminf(ALPHA_MAX, g1.z * vis[j]);
                valid[j] = idx <= bin_final[j] && sigma[j] >= 0.f && alpha[j] >= ALPHA_MIN;
            }
#pragma unroll
            for (int j = 0; j < PPT; ++j) votes[j] = __ballot_sync(0xffffffffu, valid[j]);
            if ((votes[0] | votes[1] | votes[2] | votes[3]) == 0u) continue;
            const float4 g2 = s2[t];        // g, b, depth, beta
            const float col[GSB_CH] = {g1.w, g2.x, g2.y, g2.z, g2.w};
            float g[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) g[k] = 0.f;
            // Phase 2: per 8x4 block, a warp-uniform branch; inside, misses are predicated (alpha = 0)
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                if (votes[j] == 0u) continue;
                const float dx = (j & 1) ? dx1 : dx0, dy = (j >> 1) ? dy1 : dy0;
                const float al = valid[j] ? alpha[j] : 0.f;
                const float ra = fast_rcp(1.0f - al);
                const float Tn = T[j] * ra;
                const float fac = al * Tn;
                float v_alpha = tail[j] * ra;
#pragma unroll
                for (int k = 0; k < GSB_CH; ++k) {
                    g[6 + k] = fmaf(fac, vr[j][k], g[6 + k]);
                    v_alpha = fmaf(fmaf(col[k], Tn, -buf[j][k] * ra), vr[j][k], v_alpha);
                    buf[j][k] = fmaf(col[k], fac, buf[j][k]);
                }
                T[j] = Tn;
                const float ov = g1.z * vis[j];
                const float v_sigma = (valid[j] && ov <= ALPHA_MAX) ? -ov * v_alpha : 0.f;
                const float v_op = (valid[j] && ov <= ALPHA_MAX) ? vis[j] * v_alpha : 0.f;
                const float vsp = v_sigma * (1.0f / LOG2E);
                g[0] = fmaf(vsp, 2.f * g0.z * dx + g0.w * dy, g[0]);
                g[1] = fmaf(vsp, g0.w * dx + 2.f * g1.x * dy, g[1]);
                g[2] += v_op;
                g[3] = fmaf(0.5f * v_sigma * dx, dx, g[3]);
                g[4] = fmaf(v_sigma * dx, dy, g[4]);
                g[5] = fmaf(0.5f * v_sigma * dy, dy, g[5]);
            }
            int slot;
            const float total = transpose_reduce16(g, lane, slot);
            if (!(lane & 1) && slot < 11 && total != 0.f) atomicAdd(&s_acc[t][slot], total);
        }
        __syncthreads();
        for (int t = tid; t < bs; t += BWD_THREADS) {
            const float4* a = reinterpret_cast<const float4*>(&s_acc[t][0]);
            const float4 a0 = a[0], a1 = a[1], a2 = a[2];
            const bool nz = a0.x != 0.f || a0.y != 0.f || a0.z != 0.f || a0.w != 0.f || a1.x != 0.f ||
                            a1.y != 0.f || a1.z != 0.f || a1.w != 0.f || a2.x != 0.f || a2.y != 0.f ||
                            a2.z != 0.f;
            if (nz) {
                float4* dst = reinterpret_cast<float4*>(v_rec) + (size_t)s_id[t] * 3;
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
    composite_bwd_kernel<<<grid, BWD_THREADS, 0, gsb_stream(stream)>>>(
        reinterpret_cast<const float4*>(rec), flatten_ids, offsets, backgrounds, C, width, height, tile_w,
        tile_h, n_isects, render_alphas, last_ids, v_render, v_alphas, v_rec, v_backgrounds);
    GSB_LAUNCH_CHECK();
    return 0;
}
