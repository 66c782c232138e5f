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

// dst[0] = sum of partials[0..n) in a fixed order (loss.cu); the finish of every deterministic-mode reduction
int gsb_det_sum_launch(const float* partials, int64_t n, float* dst, cudaStream_t st);

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

// Visit every tile of every lane's rectangle (x0, y0, width bw, cnt = bw*bh tiles, cnt may be 0).
// Lanes with up to TILES_DIRECT tiles (nearly all: the mean is about 8) walk their own rectangle
// with no cross-lane traffic -- a shuffle-based balanced expansion costs a five-deep dependent
// shuffle chain per 32 tiles, which made the binning kernels latency-bound.  A lane with a larger
// rectangle hands it to the whole warp afterwards, one tile per lane per step, so one huge splat
// cannot serialise its lane.  f(tx, ty, owner) is called for each tile by the lane that visits it;
// owner is the lane that owns the rectangle (== the calling lane in the direct part), so per-owner
// values can be fetched with __shfl_sync(..., owner) -- in the cooperative part ALL lanes call
// f_coop_begin(owner) first (warp-uniform) to let the caller broadcast what it needs.
#define TILES_DIRECT 48
template <typename F, typename B>
__device__ __forceinline__ void warp_for_tiles(int x0, int y0, int bw, int cnt, B&& begin_owner, F&& f) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool big = cnt > TILES_DIRECT;
    if (!big && cnt > 0) {
        begin_owner(lane, false);
        int tx = x0, ty = y0;
        const int xe = x0 + bw;
        for (int t = 0; t < cnt; ++t) {
            f(tx, ty);
            if (++tx == xe) { tx = x0; ++ty; }
        }
    }
    unsigned bigmask = __ballot_sync(full, big);
    while (bigmask) {
        const int src = __ffs(bigmask) - 1;
        bigmask &= bigmask - 1;
        const int o_x0 = __shfl_sync(full, x0, src), o_y0 = __shfl_sync(full, y0, src);
        const int o_bw = max(__shfl_sync(full, bw, src), 1), o_cnt = __shfl_sync(full, cnt, src);
        begin_owner(src, true);
        for (int t = lane; t < o_cnt; t += 32) {
            const int dy = t / o_bw;
            f(o_x0 + (t - dy * o_bw), o_y0 + dy);
        }
    }
}
