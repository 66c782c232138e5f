// Definitions shared by the compositing forward and backward kernels.
#pragma once
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


// Shared-memory staging layout of one splat (three float4 planes):
//   s0 = {x, y, a', b'}   s1 = {c', thr, opacity, r}   s2 = {g, b, depth, beta}
// a', b', c' is the conic pre-scaled so that sigma' = a'dx^2 + b'dxdy + c'dy^2 = sigma * log2(e);
// thr = log2(255 * opacity) + margin: pairs with sigma' > thr cannot reach alpha >= 1/255, so the
// inner loops reject them on sigma' alone (no MUFU); pairs that pass still take the exact test.
#define THR_MARGIN 1e-3f
__device__ __forceinline__ void stage_splat(const float4* __restrict__ rp, float4& o0, float4& o1,
                                            float4& o2) {
    const float4 r0 = rp[0], r1 = rp[1], r2 = rp[2];
    o0 = make_float4(r0.x, r0.y, r0.w * (0.5f * LOG2E), r1.x * LOG2E);
    o1 = make_float4(r1.y * (0.5f * LOG2E), __log2f(255.0f * r0.z) + THR_MARGIN, r0.z, r1.z);
    o2 = make_float4(r1.w, r2.x, r2.y, r2.z);
}
