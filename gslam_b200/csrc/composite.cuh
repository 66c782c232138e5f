// Definitions shared by the compositing forward and backward kernels.
//
// Work decomposition (both directions): ONE WARP (= one CTA) owns one UNIT of a 16x16 tile and
// walks the tile's depth-sorted list on its own -- no CTA barrier anywhere.  The including .cu picks
// the unit with PPT (pixels per thread) before including this header:
//   PPT 2: 8x8 quarter-tile, slot j of lane l is pixel (l & 7, (l >> 3) + 4 j)   -- forward kernel
//   PPT 4: 16x8 half-tile, slots are the four 8x4 blocks                        -- backward kernel
// Small units keep many of them per SM (latency hiding; the hardware CTA scheduler hands out the
// tail of the grid dynamically) and make the culling below sharper; larger ones amortise the
// per-entry work over more pixels and need fewer slots.
//
// The list is consumed in rounds of 32 entries.  Splat records (48 B) are gathered
// from HBM/L2 into a three-stage shared-memory ring with cp.async (LDGSTS, 3 x 16 B per
// lane), two rounds ahead of use; the ids that address them are prefetched one round
// earlier still.  When a round lands, every lane takes one record, pre-scales the
// conic by log2(e), and runs an EXACT ellipse-vs-rectangle test: the minimum of the
// quadratic form over the unit's pixel-centre rectangle is compared with
// log2(255 * opacity), the largest exponent that can still reach alpha >= 1/255.
// Entries that cannot touch the unit (50-60 % of gsplat's 3-sigma bounding-
// square list) are dropped by a ballot/popc compaction, so the per-pixel loop only
// sees splats that matter.
#pragma once
#include "common.cuh"

#define ALPHA_MAX 0.999f
#define ALPHA_MIN (1.0f / 255.0f)
#define T_STOP 1e-4f
#define LOG2E 1.4426950408889634f

#ifndef CMP_WARPS
#define CMP_WARPS 1            // independent workers (warps) per CTA; set by the including .cu
#endif
#define CMP_THREADS (32 * CMP_WARPS)
#define RND 32                 // list entries per round (one per lane)
#define NSTAGE 3               // cp.async ring depth
#ifndef PPT
#define PPT 2                  // pixels per thread: 2 -> 8x8 quarter-tile units, 4 -> 16x8 half-tiles
#endif
#if PPT == 1
#define UNITS_PER_TILE 8
#define UNIT_W 8
#define UNIT_H 4
#define SLOT_DX(j) 0
#define SLOT_DY(j) 0
#elif PPT == 2
#define UNITS_PER_TILE 4
#define UNIT_W 8
#define UNIT_H 8
#define SLOT_DX(j) 0
#define SLOT_DY(j) ((j) * 4)
#else
#define UNITS_PER_TILE 2
#define UNIT_W 16
#define UNIT_H 8
#define SLOT_DX(j) (((j) & 1) * 8)
#define SLOT_DY(j) (((j) >> 1) * 4)
#endif

struct TileRange {
    int start, end;
};

__device__ __forceinline__ TileRange tile_range(const int32_t* offsets, int lin, int n_lin,
                                                int64_t n_isects) {
    TileRange r;
    r.start = offsets[lin];
    if (n_isects >= 0) {
        r.end = (lin == n_lin - 1) ? (int)n_isects : offsets[lin + 1];
    } else {
        // offsets has C*T+1 entries (device-side count, gsb_tile_scan); -n_isects - 1 is the
        // capacity of flatten_ids (0: unbounded).  A step whose count exceeds the capacity it was
        // captured with (the overflow flag of gsb_isect_scatter) must not read past the buffer.
        r.end = offsets[lin + 1];
        const int cap = (int)(-n_isects - 1);
        if (cap > 0) {
            r.start = min(r.start, cap);
            r.end = min(r.end, cap);
        }
    }
    return r;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// 16-byte shared-memory load from a 32-bit shared-window address (keeps the address arithmetic of the
// hot loops to one add: the generic-pointer path re-derives the window base every iteration).
__device__ __forceinline__ float4 lds128(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
#define K_PLANE_BYTES (RND * 16)   // the three planes of kept entries are contiguous
#define STAGE_BYTES (RND * 48)     // one ring slot

// Shared-window address of p, pinned in a register (the opaque move stops the compiler from
// re-deriving it from SR_CgaCtaId inside the hot loop).
__device__ __forceinline__ unsigned smem_addr(const void* p) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p), r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));
    return r;
}

// Per-warp staging area.  A ring slot first holds the gathered records, lane-major (48 B per lane);
// once every lane has its record in registers the SAME slot is overwritten with the compacted
// planes of the kept entries:
//   plane 0 [0, RND):       x, y, a', b'
//   plane 1 [RND, 2 RND):   c', list index (int bits), opacity, r
//   plane 2 [2 RND, 3 RND): g, b, depth, beta
// (in place: 4.6 KB per warp instead of 6.1 KB, which is what decides how many units fit an SM)
struct __align__(16) WarpStage {
    float4 raw[NSTAGE][RND * 3];
    int32_t gid[RND];              // flatten id (camera * N + gaussian) of the kept entries
};

// The unit's pixel-centre rectangle.
struct Rect {
    float x0, x1, y0, y1;
};

// One staged record -> compaction planes, plus the cull decision.
// a', b', c' is the conic pre-scaled so that sigma' = a'dx^2 + b'dxdy + c'dy^2 = sigma * log2(e).
__device__ __forceinline__ bool stage_and_cull(const float4 r0, const float4 r1, const float4 r2,
                                               const Rect rc, int list_index, float4& o0, float4& o1,
                                               float4& o2) {
    const float x = r0.x, y = r0.y, op = r0.z;
    const float A = r0.w * (0.5f * LOG2E), B = r1.x * LOG2E, Cc = r1.y * (0.5f * LOG2E);
    o0 = make_float4(x, y, A, B);
    o1 = make_float4(Cc, __int_as_float(list_index), op, r1.z);
    o2 = make_float4(r1.w, r2.x, r2.y, r2.z);
    // dx = x - px ranges over [dxl, dxh], dy likewise
    const float dxl = x - rc.x1, dxh = x - rc.x0, dyl = y - rc.y1, dyh = y - rc.y0;
    if (dxl <= 0.f && dxh >= 0.f && dyl <= 0.f && dyh >= 0.f) return true;   // centre inside
    // outside: the minimum of the (positive definite) form lies on the boundary; take each edge's
    // 1-D minimiser clamped to the edge
    const float hA = 0.5f * fast_rcp(A), hC = 0.5f * fast_rcp(Cc);
    float q = 3.0e38f;
    {
        const float dy0 = fminf(fmaxf(-B * dxl * hC, dyl), dyh);
        q = fminf(q, fmaf(dxl, fmaf(A, dxl, B * dy0), Cc * dy0 * dy0));
        const float dy1 = fminf(fmaxf(-B * dxh * hC, dyl), dyh);
        q = fminf(q, fmaf(dxh, fmaf(A, dxh, B * dy1), Cc * dy1 * dy1));
        const float dx0 = fminf(fmaxf(-B * dyl * hA, dxl), dxh);
        q = fminf(q, fmaf(dx0, fmaf(A, dx0, B * dyl), Cc * dyl * dyl));
        const float dx1 = fminf(fmaxf(-B * dyh * hA, dxl), dxh);
        q = fminf(q, fmaf(dx1, fmaf(A, dx1, B * dyh), Cc * dyh * dyh));
    }
    // alpha >= 1/255 needs sigma' <= log2(255 * opacity); the margin covers the rounding of the
    // per-pixel evaluation (cancellation in very anisotropic conics included)
    const float dxm = fmaxf(fabsf(dxl), fabsf(dxh)), dym = fmaxf(fabsf(dyl), fabsf(dyh));
    const float margin = 1e-3f + 4e-6f * (A * dxm * dxm + fabsf(B) * dxm * dym + Cc * dym * dym);
    return q <= __log2f(255.0f * op) + margin;
}

// Which pixels a warp owns.
struct Unit {
    int lin, cam;        // camera * T + tile, camera
    int bx, by;          // pixel of slot 0 of this lane
    Rect rc;
};

__device__ __forceinline__ Unit decode_unit(int unit, int lane, int tile_w, int n_tiles) {
    Unit u;
    u.lin = unit / UNITS_PER_TILE;
    const int q = unit - u.lin * UNITS_PER_TILE;
    u.cam = u.lin / n_tiles;
    const int tile = u.lin - u.cam * n_tiles;
    const int ty = tile / tile_w, tx = tile - ty * tile_w;
    const int ox = tx * GSB_TILE + (UNITS_PER_TILE >= 4 ? (q & 1) * 8 : 0);
    const int oy = ty * GSB_TILE + (UNITS_PER_TILE >= 4 ? (q >> 1) * UNIT_H : q * 8);
    u.bx = ox + (lane & 7);
    u.by = oy + (lane >> 3);
    u.rc.x0 = (float)ox + 0.5f; u.rc.x1 = u.rc.x0 + (float)(UNIT_W - 1);
    u.rc.y0 = (float)oy + 0.5f; u.rc.y1 = u.rc.y0 + (float)(UNIT_H - 1);
    return u;
}

// sigma' of this lane's PPT pixels for the staged splat (x, y, a', b') / c'.
__device__ __forceinline__ void eval_sigma(const float4 g0, float cc, float pxf, float pyf,
                                           float (&dx)[PPT], float (&dy)[PPT], float (&sigma)[PPT]) {
    const float dx0 = g0.x - pxf, dy0 = g0.y - pyf;
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        dx[j] = dx0 - (float)SLOT_DX(j);
        dy[j] = dy0 - (float)SLOT_DY(j);
        sigma[j] = fmaf(dx[j], fmaf(g0.w, dy[j], g0.z * dx[j]), cc * dy[j] * dy[j]);
    }
}

// Issue the gather of one round: lane copies the 48-byte record `id` into raw[stage][lane*3..].
__device__ __forceinline__ void issue_round(WarpStage& ws, int stage, int lane, const float4* rec, int id) {
    if (id >= 0) {
        const float4* src = rec + (size_t)id * 3;
        float4* dst = &ws.raw[stage][lane * 3];
        cp_async16(dst, src);
        cp_async16(dst + 1, src + 1);
        cp_async16(dst + 2, src + 2);
    }
    cp_async_commit();
}
