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

