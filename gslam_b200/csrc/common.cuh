// Shared helpers for libgsb.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/gsb.h"

#define GSB_API extern "C" __attribute__((visibility("default")))

void gsb_set_error(const char* fmt, ...);

#define GSB_CHECK_ARG(cond)                                              \
    do {                                                                 \
        if (!(cond)) {                                                   \
            gsb_set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond); \
            return GSB_EINVAL;                                           \
        }                                                                \
    } while (0)

#define GSB_CUDA(expr)                                                   \
    do {                                                                 \
        cudaError_t _e = (expr);                                         \
        if (_e != cudaSuccess) {                                         \
            gsb_set_error("%s:%d: %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                              \
        }                                                                \
    } while (0)

#define GSB_LAUNCH_CHECK() GSB_CUDA(cudaGetLastError())

static inline cudaStream_t gsb_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 2^x on the SFU (MUFU.EX2), flush-to-zero.
__device__ __forceinline__ float fast_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// 1/x on the SFU (MUFU.RCP), about 1 ulp.
__device__ __forceinline__ float fast_rcp(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// Tile bounding box [x0,x1) x [y0,y1) of a splat, float32 arithmetic shared by
// the counting and emitting kernels (so they can never disagree).
__device__ __forceinline__ void tile_bbox(float mx, float my, int radius, int tile_w, int tile_h,
                                          int& x0, int& x1, int& y0, int& y1) {
    const float inv = 1.0f / GSB_TILE;
    float tr = (float)radius * inv, tx = mx * inv, ty = my * inv;
    x0 = min(max(0, (int)floorf(tx - tr)), tile_w);
    x1 = min(max(0, (int)ceilf(tx + tr)), tile_w);
    y0 = min(max(0, (int)floorf(ty - tr)), tile_h);
    y1 = min(max(0, (int)ceilf(ty + tr)), tile_h);
}

// Warp-balanced expansion of per-lane tile rectangles.  Every lane of a full warp
// brings a rectangle (x0, y0, width bw, cnt = bw*bh tiles, cnt may be 0); the
// warp's total work is dealt out one tile per lane per round, so one large splat
// cannot serialise its lane.  f(owner_lane, tx, ty, active) is called by ALL
// lanes each round (so it may shuffle from owner_lane); only `active` lanes own
// a tile.
template <typename F>
__device__ __forceinline__ void warp_expand_tiles(int x0, int y0, int bw, int cnt, F&& f) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(full, incl, o);
        if (lane >= o) incl += y;
    }
    const int total = __shfl_sync(full, incl, 31);
    const int excl = incl - cnt;
    for (int base = 0; base < total; base += 32) {
        const int k = base + lane;
        const bool active = k < total;
        // owner = last lane whose exclusive prefix is <= k (binary search over lanes)
        int lo = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            int probe = __shfl_sync(full, excl, lo + step);
            if (probe <= k) lo += step;
        }
        const int o_excl = __shfl_sync(full, excl, lo);
        const int o_x0 = __shfl_sync(full, x0, lo);
        const int o_y0 = __shfl_sync(full, y0, lo);
        const int o_bw = max(__shfl_sync(full, bw, lo), 1);
        const int local = active ? k - o_excl : 0;
        const int dy = local / o_bw;
        f(lo, o_x0 + (local - dy * o_bw), o_y0 + dy, active);
    }
}
