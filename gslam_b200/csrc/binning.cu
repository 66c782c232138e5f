// K4/K5, B200-native form: counting sort by tile + per-tile depth sort in shared memory.
//
// gsplat sorts ALL (tile | depth) 64-bit keys with a global radix sort (7 passes
// over I keys) and then searches the sorted keys for the tile boundaries
// (reference call sites gslam/rasterization.py:261-274).  The keys are
// (camera, tile) major, so the same order is produced by
//   1. counting intersections per (camera, tile)          (atomics in project_fwd),
//   2. an exclusive scan over the C*T tile counters          -> isect_offsets, I,
//   3. scattering each (depth bits << 32 | flatten_id) into its tile's segment,
//   4. sorting every segment on that 64-bit word inside one CTA's shared memory.
// Ties in depth are broken by flatten_id, which is exactly the order a stable
// sort of the upstream emission order (c-major, n, tiles) gives, so the result
// is bit-identical to the global stable sort.  Nothing here needs the
// intersection count on the host: I lives in offsets[C*T].
#include "common.cuh"

// ---------------------------------------------------------------------------
// exclusive scan over the tile counters (one CTA; C*T is at most a few 10^4)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
tile_scan_kernel(const int32_t* __restrict__ counts, int n, int32_t* __restrict__ offsets,
                 int32_t* __restrict__ cursors, int64_t* __restrict__ n_isects_out) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        int i = base + tid;
        int v = (i < n) ? counts[(size_t)i * GSB_CSTRIDE] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;   // inclusive over warps
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + (warp ? warp_tot[warp - 1] : 0) + x - v;
        if (i < n) { offsets[i] = excl; cursors[(size_t)i * GSB_CSTRIDE] = excl; }
        __syncthreads();
        if (tid == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
    if (tid == 0) {
        offsets[n] = carry_s;
        if (n_isects_out) *n_isects_out = carry_s;
    }
}

GSB_API int gsb_tile_scan(const int32_t* tile_counts, int C, int tile_w, int tile_h, int32_t* offsets,
                          int32_t* cursors, int64_t* n_isects, void* stream) {
    int n = C * tile_w * tile_h;
    GSB_CHECK_ARG(n >= 0 && offsets && cursors);
    GSB_CHECK_ARG(n == 0 || tile_counts);
    tile_scan_kernel<<<1, 1024, 0, gsb_stream(stream)>>>(tile_counts, n, offsets, cursors, n_isects);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// scatter: one thread per (camera, Gaussian)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
isect_scatter_kernel(const float* __restrict__ means2d, const int32_t* __restrict__ radii,
                     const float* __restrict__ rec, int64_t CN, int N, int tile_w, int tile_h,
                     int32_t* __restrict__ cursors, int64_t capacity, uint64_t* __restrict__ dk,
                     int32_t* __restrict__ overflow) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // a warp never straddles two cameras when N % 32 == 0; otherwise each lane keeps its own camera
    int r = (i < CN) ? radii[i] : 0;
    int x0 = 0, x1 = 0, y0 = 0, y1 = 0, cnt = 0, cam = 0;
    uint32_t dbits = 0;
    if (r > 0) {
        float2 m = reinterpret_cast<const float2*>(means2d)[i];
        tile_bbox(m.x, m.y, r, tile_w, tile_h, x0, x1, y0, y1);
        cnt = (x1 - x0) * (y1 - y0);
        cam = (int)(i / N);
        dbits = __float_as_uint(rec[i * GSB_REC + 9]);
    }
    const int n_tiles = tile_w * tile_h;
    const uint32_t id = (uint32_t)i;
    uint64_t word = ((uint64_t)dbits << 32) | id;      // this lane's key, or the owner's in the cooperative part
    int32_t* cur = cursors + (size_t)cam * n_tiles * GSB_CSTRIDE;
    const uint64_t my_word = word;
    int32_t* const my_cur = cur;
    warp_for_tiles(x0, y0, x1 - x0, cnt,
        [&](int owner, bool coop) {
            if (coop) {
                word = __shfl_sync(0xffffffffu, my_word, owner);
                cur = reinterpret_cast<int32_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<uint64_t>(my_cur), owner));
            }
        },
        [&](int tx, int ty) {
            int pos = atomicAdd(cur + (size_t)(ty * tile_w + tx) * GSB_CSTRIDE, 1);
            if (pos < capacity) dk[pos] = word;
            else *overflow = 1;
        });
}

// Same scatter with the slot hand-out done per CTA: a CTA takes a contiguous chunk of one camera's
// Gaussians, histograms their tiles in shared memory, reserves each tile's share of the segment with
// ONE returning global atomic per (CTA, tile) -- all independent, issued side by side -- and then
// ranks its keys inside the reserved ranges with shared-memory atomics.  The per-key returning
// global atomic of the kernel above (whose round trip bounds it) is gone; the order inside a
// segment is arbitrary either way, the segment sort fixes it.
//
// The chunk a CTA takes is sized to the launch (gsb_isect_scatter): 2048 Gaussians at map sizes where
// that still gives a grid of 64+ CTAs, 256, 64 or 32 for the small maps of a starting sequence (a few
// thousand splats that each cover hundreds of tiles), where three 2048-chunks would leave the
// million shared-memory atomics of the expansion to three SMs.
template <int SCL_THREADS, int GPT>
__global__ void __launch_bounds__(SCL_THREADS)
isect_scatter_local_kernel(const float* __restrict__ means2d, const int32_t* __restrict__ radii,
                           const float* __restrict__ rec, int N, int tile_w, int tile_h,
                           int32_t* __restrict__ cursors, int64_t capacity, uint64_t* __restrict__ dk,
                           int32_t* __restrict__ overflow) {
    extern __shared__ int32_t s_tiles[];
    const int n_tiles = tile_w * tile_h;
    int32_t* const hist = s_tiles;              // count, later the running rank
    int32_t* const base = s_tiles + n_tiles;    // first reserved slot of this CTA in the tile's segment
    const int tid = threadIdx.x, cam = blockIdx.y;
    for (int t = tid; t < n_tiles; t += SCL_THREADS) hist[t] = 0;
    __syncthreads();
    int x0[GPT], y0[GPT], bw[GPT], cnt[GPT];
    uint64_t word[GPT];
#pragma unroll
    for (int g = 0; g < GPT; ++g) {
        const int n = (blockIdx.x * GPT + g) * SCL_THREADS + tid;
        const int64_t i = (int64_t)cam * N + n;
        const int r = (n < N) ? radii[i] : 0;
        x0[g] = y0[g] = bw[g] = cnt[g] = 0; word[g] = 0;
        if (r > 0) {
            const float2 m = reinterpret_cast<const float2*>(means2d)[i];
            int x1, y1;
            tile_bbox(m.x, m.y, r, tile_w, tile_h, x0[g], x1, y0[g], y1);
            bw[g] = x1 - x0[g];
            cnt[g] = bw[g] * (y1 - y0[g]);
            word[g] = ((uint64_t)__float_as_uint(rec[i * GSB_REC + 9]) << 32) | (uint32_t)i;
        }
    }
#pragma unroll
    for (int g = 0; g < GPT; ++g)
        warp_for_tiles(x0[g], y0[g], bw[g], cnt[g], [](int, bool) {},
                       [&](int tx, int ty) { atomicAdd(&hist[ty * tile_w + tx], 1); });
    __syncthreads();
    int32_t* const cur = cursors + (size_t)cam * n_tiles * GSB_CSTRIDE;
    for (int t = tid; t < n_tiles; t += SCL_THREADS) {
        const int c = hist[t];
        if (c) { base[t] = atomicAdd(cur + (size_t)t * GSB_CSTRIDE, c); hist[t] = 0; }
    }
    __syncthreads();
#pragma unroll
    for (int g = 0; g < GPT; ++g) {
        uint64_t w = word[g];
        warp_for_tiles(x0[g], y0[g], bw[g], cnt[g],
            [&](int owner, bool coop) { if (coop) w = __shfl_sync(0xffffffffu, word[g], owner); },
            [&](int tx, int ty) {
                const int t = ty * tile_w + tx;
                const int pos = base[t] + atomicAdd(&hist[t], 1);
                if (pos < capacity) dk[pos] = w;
                else *overflow = 1;
            });
    }
}

template <int THREADS, int GPT>
static int launch_scatter_local(const float* means2d, const int32_t* radii, const float* rec, int C, int N,
                                int tile_w, int tile_h, int32_t* cursors, int64_t capacity, uint64_t* dk,
                                int32_t* overflow, size_t smem, void* stream) {
    if (smem > 32 * 1024) {
        // the opt-in above 48 KB (static + dynamic) is a per-device, per-function attribute and setting it is idempotent and
        // cheap next to a launch, so it is simply set whenever it is needed: no process-wide flag
        GSB_CUDA(cudaFuncSetAttribute(isect_scatter_local_kernel<THREADS, GPT>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    }
    dim3 grid((N + THREADS * GPT - 1) / (THREADS * GPT), C);
    isect_scatter_local_kernel<THREADS, GPT><<<grid, THREADS, smem, gsb_stream(stream)>>>(
        means2d, radii, rec, N, tile_w, tile_h, cursors, capacity, dk, overflow);
    GSB_LAUNCH_CHECK();
    return 0;
}

GSB_API int gsb_isect_scatter(const float* means2d, const int32_t* radii, const float* rec, int C, int N,
                              int tile_w, int tile_h, int32_t* cursors, int64_t capacity,
                              int64_t* depth_id, int32_t* overflow, void* stream) {
    int64_t CN = (int64_t)C * N;
    if (CN == 0) return 0;
    GSB_CHECK_ARG(means2d && radii && rec && cursors && overflow);
    GSB_CHECK_ARG(capacity == 0 || depth_id);
    GSB_CHECK_ARG(CN < ((int64_t)1 << 31) && capacity < ((int64_t)1 << 31));
    const size_t smem = (size_t)tile_w * tile_h * 2 * sizeof(int32_t);
    if (smem <= 96 * 1024 && C <= 65535) {
        // chunk = 2048 Gaussians of one camera: ~10 keys per (CTA, tile) at SLAM densities
        // (measured at 200k: 1024 the same, 4096 slower -- too few CTAs); smaller chunks only when
        // 2048 would leave most SMs without a CTA
        uint64_t* dk = reinterpret_cast<uint64_t*>(depth_id);
        const int64_t want = 64;
        if ((int64_t)C * ((N + 2047) / 2048) >= want)
            return launch_scatter_local<1024, 2>(means2d, radii, rec, C, N, tile_w, tile_h, cursors, capacity, dk,
                                                 overflow, smem, stream);
        if ((int64_t)C * ((N + 255) / 256) >= want)
            return launch_scatter_local<256, 1>(means2d, radii, rec, C, N, tile_w, tile_h, cursors, capacity, dk,
                                                overflow, smem, stream);
        if ((int64_t)C * ((N + 63) / 64) >= 2 * want)
            return launch_scatter_local<64, 1>(means2d, radii, rec, C, N, tile_w, tile_h, cursors, capacity, dk,
                                               overflow, smem, stream);
        return launch_scatter_local<32, 1>(means2d, radii, rec, C, N, tile_w, tile_h, cursors, capacity, dk,
                                           overflow, smem, stream);
    }
    int blocks = (int)((CN + 255) / 256);
    isect_scatter_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(
        means2d, radii, rec, CN, N, tile_w, tile_h, cursors, capacity,
        reinterpret_cast<uint64_t*>(depth_id), overflow);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// per-tile segment sort.
//
// Keys are (depth bits << 32 | flatten_id).  Depth bits of positive floats are
// monotone in depth, so one CTA per tile does a bucket sort: min/max of the
// depth bits, a monotone linear map of every key to one of B ~ L buckets
// (shared-memory counters), an exclusive scan, a scatter, and a tiny insertion
// sort per bucket (about one key per bucket on average).  Buckets that stay
// large (equal or tightly clustered depths) are finished by a CTA-wide bitonic
// network.  Segments up to SEG_SMEM_ELEMS keys are staged in shared memory;
// longer ones run the same code through L2 with the counters still in shared
// memory.  O(L) work instead of the O(L log^2 L) of a pure bitonic sort.
// ---------------------------------------------------------------------------
#define SEG_THREADS 256
#define SEG_SMALL 24             // buckets up to this size are insertion-sorted by one thread
#define SEG_MAX_BIG 64

// all compare-exchanges ascending, so a non-power-of-two length needs no padding
// (out-of-range partners act as +inf)
__device__ __forceinline__ void bitonic_sort(uint64_t* a, int L) {
    int lg = 0;
    while ((1 << lg) < L) ++lg;
    const int half = (1 << lg) >> 1;
    for (int lk = 1; lk <= lg; ++lk) {
        const int hk = lk - 1;
        for (int t = threadIdx.x; t < half; t += SEG_THREADS) {
            int blk = t >> hk, off = t & ((1 << hk) - 1);
            int i = (blk << lk) + off, p = (blk << lk) + ((1 << lk) - 1 - off);
            if (p < L) {
                uint64_t x = a[i], y = a[p];
                if (x > y) { a[i] = y; a[p] = x; }
            }
        }
        __syncthreads();
        for (int lj = hk - 1; lj >= 0; --lj) {
            for (int t = threadIdx.x; t < half; t += SEG_THREADS) {
                int i = ((t >> lj) << (lj + 1)) + (t & ((1 << lj) - 1)), p = i + (1 << lj);
                if (p < L) {
                    uint64_t x = a[i], y = a[p];
                    if (x > y) { a[i] = y; a[p] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// Sorts in[0..L) into out[0..L).  cnt / fill: B ints each in shared memory.
__device__ __forceinline__ void bucket_sort(const uint64_t* in, uint64_t* out, int L, int* cnt,
                                            int* fill, int max_buckets) {
    __shared__ uint32_t red_mn[SEG_THREADS / 32], red_mx[SEG_THREADS / 32];
    __shared__ int warp_sum_s[SEG_THREADS / 32];
    __shared__ int big[SEG_MAX_BIG];
    __shared__ int n_big;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int lgB = 0;
    while ((1 << lgB) < L && (1 << lgB) < max_buckets) ++lgB;
    const int B = 1 << lgB;
    // 1. depth range
    uint32_t mn = 0xffffffffu, mx = 0u;
    for (int t = tid; t < L; t += SEG_THREADS) {
        uint32_t d = (uint32_t)(in[t] >> 32);
        mn = min(mn, d); mx = max(mx, d);
    }
    mn = __reduce_min_sync(0xffffffffu, mn);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0) { red_mn[warp] = mn; red_mx[warp] = mx; }
    for (int t = tid; t < B; t += SEG_THREADS) { cnt[t] = 0; fill[t] = 0; }
    if (tid == 0) n_big = 0;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < SEG_THREADS / 32; ++w) { mn = min(mn, red_mn[w]); mx = max(mx, red_mx[w]); }
    const float scale = (float)B / ((float)(mx - mn) + 1.0f);
    // 2. histogram
    for (int t = tid; t < L; t += SEG_THREADS) {
        uint32_t d = (uint32_t)(in[t] >> 32);
        int bkt = min(B - 1, (int)((float)(d - mn) * scale));
        atomicAdd(&cnt[bkt], 1);
    }
    __syncthreads();
    // 3. exclusive scan of cnt (B / 256 consecutive counters per thread), in place
    {
        const int per = max(B / SEG_THREADS, 1);
        const int b0 = tid * per;
        int local = 0;
        if (b0 < B)
            for (int k = 0; k < per; ++k) local += cnt[b0 + k];
        int x = local;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sum_s[warp] = x;
        __syncthreads();
        int base = 0;
        for (int w = 0; w < warp; ++w) base += warp_sum_s[w];
        int run = base + x - local;
        if (b0 < B)
            for (int k = 0; k < per; ++k) {
                int c = cnt[b0 + k];
                cnt[b0 + k] = run;
                run += c;
            }
    }
    __syncthreads();
    // 4. scatter into buckets
    for (int t = tid; t < L; t += SEG_THREADS) {
        uint64_t key = in[t];
        uint32_t d = (uint32_t)(key >> 32);
        int bkt = min(B - 1, (int)((float)(d - mn) * scale));
        int pos = cnt[bkt] + atomicAdd(&fill[bkt], 1);
        out[pos] = key;
    }
    __syncthreads();
    // 5. finish the buckets
    for (int bkt = tid; bkt < B; bkt += SEG_THREADS) {
        int s0 = cnt[bkt], m = fill[bkt];
        if (m < 2) continue;
        bool small = m <= SEG_SMALL;
        if (!small) {
            int slot = atomicAdd(&n_big, 1);
            if (slot < SEG_MAX_BIG) big[slot] = bkt;
            else small = true;          // list full: slow but correct
        }
        if (small) {
            for (int i = 1; i < m; ++i) {
                uint64_t key = out[s0 + i];
                int j = i - 1;
                while (j >= 0 && out[s0 + j] > key) { out[s0 + j + 1] = out[s0 + j]; --j; }
                out[s0 + j + 1] = key;
            }
        }
    }
    __syncthreads();
    const int nb = min(n_big, SEG_MAX_BIG);
    for (int k = 0; k < nb; ++k) bitonic_sort(out + cnt[big[k]], fill[big[k]]);
}

// Segments of up to KPT * 256 keys: the keys are read from global memory ONCE into registers
// (KPT per thread), bucketed through one shared-memory counter array (histogram -> exclusive
// scan -> the scanned counters double as scatter cursors, so afterwards cnt[b] is the END of
// bucket b) and finished in a shared-memory output array.  Shared memory is 12 bytes per key,
// so a 2048-key configuration keeps nine CTAs per SM resident.  Longer segments take the
// generic path through global scratch.
template <int KPT>
__global__ void __launch_bounds__(SEG_THREADS, KPT == 8 ? 6 : (KPT == 16 ? 4 : 2))   // what shared memory allows
segment_sort_kernel(uint64_t* __restrict__ dk, uint64_t* __restrict__ tmp,
                    const int32_t* __restrict__ offsets, int n_tiles, int tile_n_bits, int64_t capacity,
                    int64_t* __restrict__ isect_ids, int32_t* __restrict__ flatten_ids) {
    constexpr int ELEMS = KPT * SEG_THREADS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* out = reinterpret_cast<uint64_t*>(smem_raw);            // ELEMS keys
    int* cnt = reinterpret_cast<int*>(out + ELEMS);                   // ELEMS counters
    __shared__ uint32_t red_mn[SEG_THREADS / 32], red_mx[SEG_THREADS / 32];
    __shared__ int warp_sum_s[SEG_THREADS / 32];
    __shared__ int big[SEG_MAX_BIG];
    __shared__ int n_big;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lin = blockIdx.x;                 // camera * n_tiles + tile
    int64_t s = offsets[lin], e = offsets[lin + 1];
    if (e > capacity) e = capacity;
    if (s >= e) return;
    const int L = (int)(e - s);
    const int cam = lin / n_tiles, tile = lin % n_tiles;
    const int64_t hi = (((int64_t)cam << tile_n_bits) | tile) << 32;
    const uint64_t* sorted;
    if (L > ELEMS) {
        // generic path: global in / out, counters in shared memory (fill aliases the key array)
        bucket_sort(dk + s, tmp + s, L, cnt, reinterpret_cast<int*>(out), ELEMS);
        sorted = tmp + s;
    } else if (L == 1) {
        sorted = dk + s;
    } else {
        uint64_t key[KPT];
        uint32_t mn = 0xffffffffu, mx = 0u;
#pragma unroll
        for (int i = 0; i < KPT; ++i) {
            const int t = tid + i * SEG_THREADS;
            key[i] = t < L ? dk[s + t] : ~0ull;
            if (t < L) {
                const uint32_t d = (uint32_t)(key[i] >> 32);
                mn = min(mn, d); mx = max(mx, d);
            }
        }
        const int lgB = 32 - __clz(L - 1);      // L >= 2; L <= ELEMS, a power of two
        const int B = 1 << lgB;
        mn = __reduce_min_sync(0xffffffffu, mn);
        mx = __reduce_max_sync(0xffffffffu, mx);
        if (lane == 0) { red_mn[warp] = mn; red_mx[warp] = mx; }
        for (int t = tid; t < B; t += SEG_THREADS) cnt[t] = 0;
        if (tid == 0) n_big = 0;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < SEG_THREADS / 32; ++w) { mn = min(mn, red_mn[w]); mx = max(mx, red_mx[w]); }
        // monotone map of the depth bits to B buckets (about one key per bucket)
        const float scale = (float)B / ((float)(mx - mn) + 1.0f);
        auto bucket_of = [&](uint64_t k) -> int {
            return min(B - 1, (int)((float)((uint32_t)(k >> 32) - mn) * scale));
        };
#pragma unroll
        for (int i = 0; i < KPT; ++i)
            if (tid + i * SEG_THREADS < L) atomicAdd(&cnt[bucket_of(key[i])], 1);
        __syncthreads();
        {   // exclusive scan of cnt, in place (B / 256 consecutive counters per thread)
            const int per = max(B / SEG_THREADS, 1);
            const int b0 = tid * per;
            int local = 0;
            if (b0 < B)
                for (int k = 0; k < per; ++k) local += cnt[b0 + k];
            int x = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += y;
            }
            if (lane == 31) warp_sum_s[warp] = x;
            __syncthreads();
            // sum of the warps before this one (SEG_THREADS / 32 <= 32 of them)
            const int base = __reduce_add_sync(0xffffffffu, lane < warp ? warp_sum_s[lane] : 0);
            int run = base + x - local;
            if (b0 < B)
                for (int k = 0; k < per; ++k) {
                    int c = cnt[b0 + k];
                    cnt[b0 + k] = run;
                    run += c;
                }
        }
        __syncthreads();
        int slot[KPT];      // where the scatter put each key
#pragma unroll
        for (int i = 0; i < KPT; ++i) {
            slot[i] = -1;
            if (tid + i * SEG_THREADS < L) {
                slot[i] = atomicAdd(&cnt[bucket_of(key[i])], 1);
                out[slot[i]] = key[i];
            }
        }
        __syncthreads();
        // cnt[b] is now the end of bucket b.  Every key ranks itself inside its bucket (keys are
        // unique: the low word is the flatten id) -- m reads per key, all keys in parallel, instead of
        // one thread insertion-sorting a whole bucket while its warp waits.  Buckets above SEG_SMALL
        // keys (heavily clustered depths) are left as scattered and bitonic-sorted by the whole CTA.
#pragma unroll
        for (int i = 0; i < KPT; ++i) {
            if (slot[i] >= 0) {
                const int b = bucket_of(key[i]);
                const int s0 = b ? cnt[b - 1] : 0, m = cnt[b] - s0;
                if (m > SEG_SMALL) {
                    if (slot[i] == s0) {            // one key per bucket reports it
                        const int q = atomicAdd(&n_big, 1);
                        if (q < SEG_MAX_BIG) big[q] = b;
                    }
                } else if (m > 1) {
                    int r = 0;
                    for (int j = 0; j < m; ++j) r += out[s0 + j] < key[i];
                    slot[i] = s0 + r;       // from here on: the key's final position
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < KPT; ++i)
            if (slot[i] >= 0) out[slot[i]] = key[i];
        __syncthreads();
        const int nb = min(n_big, SEG_MAX_BIG);
        for (int k = 0; k < nb; ++k) {
            const int bkt = big[k], s0 = bkt ? cnt[bkt - 1] : 0;
            bitonic_sort(out + s0, cnt[bkt] - s0);
        }
        if (n_big > SEG_MAX_BIG) bitonic_sort(out, L);      // list full: slow but correct
        sorted = out;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < L; t += SEG_THREADS) {
        uint64_t w = sorted[t];
        isect_ids[s + t] = hi | (int64_t)(w >> 32);
        flatten_ids[s + t] = (int32_t)(uint32_t)w;
    }
}

template <int KPT>
static int launch_segment_sort(int n, uint64_t* dk, uint64_t* tmp, const int32_t* offsets, int n_tiles,
                               int tile_n_bits, int64_t capacity, int64_t* isect_ids, int32_t* flatten_ids,
                               cudaStream_t st) {
    const int smem = KPT * SEG_THREADS * 12;
    if (smem > 32 * 1024)    // per-device opt-in above 48 KB (static + dynamic); idempotent, so no process-wide flag is kept
        GSB_CUDA(cudaFuncSetAttribute(segment_sort_kernel<KPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    segment_sort_kernel<KPT><<<n, SEG_THREADS, smem, st>>>(dk, tmp, offsets, n_tiles, tile_n_bits, capacity,
                                                          isect_ids, flatten_ids);
    GSB_LAUNCH_CHECK();
    return 0;
}

GSB_API int gsb_segment_sort(int64_t* depth_id, int64_t* scratch, const int32_t* offsets, int C,
                             int tile_w, int tile_h, int tile_n_bits, int64_t capacity,
                             int64_t* isect_ids, int32_t* flatten_ids, void* stream) {
    int n = C * tile_w * tile_h;
    if (n == 0 || capacity == 0) return 0;
    GSB_CHECK_ARG(depth_id && scratch && offsets && isect_ids && flatten_ids);
    uint64_t* dk = reinterpret_cast<uint64_t*>(depth_id);
    uint64_t* tmp = reinterpret_cast<uint64_t*>(scratch);
    // the configuration follows the mean segment length (capacity bounds the intersection count):
    // 2048-key CTAs (24 KB, nine per SM) while the mean leaves 1.5x headroom, else 4096 / 8192
    const int64_t mean = capacity / n;
    if (mean * 3 <= 2048 * 2)
        return launch_segment_sort<8>(n, dk, tmp, offsets, tile_w * tile_h, tile_n_bits, capacity, isect_ids,
                                      flatten_ids, gsb_stream(stream));
    if (mean * 3 <= 4096 * 2)
        return launch_segment_sort<16>(n, dk, tmp, offsets, tile_w * tile_h, tile_n_bits, capacity, isect_ids,
                                       flatten_ids, gsb_stream(stream));
    return launch_segment_sort<32>(n, dk, tmp, offsets, tile_w * tile_h, tile_n_bits, capacity, isect_ids,
                                   flatten_ids, gsb_stream(stream));
}
