// The per-Gaussian tail of one mapping iteration as ONE kernel over a range of Gaussians:
//   union visibility (reduced over the window's views) -> isotropic-regulariser gradient
//   (reference gslam/backend.py:286-296) -> Adam on every parameter (:559-602, torch.optim.Adam
//   defaults) -> opacity decay of the Gaussians seen by more than one view (:356-359).
// The reference runs ~18 elementwise / reduction launches plus seven optimizer steps for this; round 1
// of this repo ran three kernels over all N on every rank.  Here the gradient arrives as row-major
// records [N, row_stride] (gsb_project_bwd's row layout: a Gaussian's whole gradient and its visibility
// count are one contiguous row), so a view-sharded step reduce-scatters the rows, each rank updates
// only its own Gaussian range [n0, n1) with this kernel, and the updated parameters are all-gathered:
// the O(N) work that does not scale with the number of GPUs shrinks by the number of ranks.
//
// Parameters and Adam moments stay in their own arrays ([N, width] each, the layout the rasterizer
// and the optimizer-state surgery use); only the gradient is row-major.
#include "common.cuh"

#define MU_MAX_TENSORS 8
#define MU_THREADS 256

struct MapUpdatePack {
    gsb_map_tensor t[MU_MAX_TENSORS];
    int n_tensors;
};

__device__ __forceinline__ float adam_elem(float p, float g, float* __restrict__ m_ptr, float* __restrict__ v_ptr,
                                           float beta1, float beta2, float eps, float step_size, float rsqrt_bc2) {
    const float m = beta1 * *m_ptr + (1.f - beta1) * g;
    const float v = beta2 * *v_ptr + (1.f - beta2) * g * g;
    *m_ptr = m;
    *v_ptr = v;
    return p - step_size * m / (sqrtf(v) * rsqrt_bc2 + eps);
}

// ---- NVLink multicast (NVLS) accessors: the address is a multicast mapping of a symmetric allocation ----
// A load-reduce returns the SUM over every GPU's copy (the addition happens in the NVSwitch), a store
// lands in every GPU's copy.  SASS: LDGMC.E.ADD.F32x4 / STG on the multicast address.
__device__ __forceinline__ float4 mc_ld_reduce4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ float mc_ld_reduce1(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st1(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void mc_st4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w) : "memory");
}

struct McPack {                 // multicast side of a fused reduce-scatter + update + all-gather launch
    float* param_mc[MU_MAX_TENSORS];   // multicast address of every parameter array
    float* vis_mc;              // multicast address of the int32 visibility counts (bit pattern through f32 stores)
    const float* stats_in;      // [2] local: this rank's (total, photometric) loss share
    float* stats_mc;            // multicast address of stats[world][4]
    float* scratch;             // local: per-block isotropic partials [gridDim.x], then an int ticket (zero between launches)
    int rank;
};

// MC = false: rows / parameters are this GPU's own arrays (one GPU, or between NCCL collectives).
// MC = true: ONE kernel is reduce-scatter + update + all-gather -- `rows` is the multicast address of the
// symmetric gradient rows (every load is summed over the GPUs in the switch), updated parameters, visibility
// counts and this rank's loss scalars are multicast-stored into every GPU's copy.  The caller brackets the
// launch with two cross-GPU barriers (rows complete before, stores visible after).
template <bool MC>
__global__ void __launch_bounds__(MU_THREADS)
map_update_kernel(const MapUpdatePack pk, const McPack mc, const float* __restrict__ rows, int64_t row_stride,
                  int vis_col, int64_t n0, int64_t n1, int scales_tensor, float iso_weight, int decay_tensor,
                  float decay, float beta1, float beta2, float eps, float bc1, float rsqrt_bc2, int32_t* step_dev,
                  int32_t* __restrict__ vis_count_out, float* __restrict__ iso_loss_sum,
                  float* __restrict__ det_partials) {
    __shared__ float s_bc[2];
    __shared__ float red[MU_THREADS / 32];
    if (step_dev) {     // device-side step count (CUDA-graph replays): read by every block, bumped by the last
        if (threadIdx.x == 0) {
            const int t = *reinterpret_cast<volatile int32_t*>(step_dev) + 1;
            s_bc[0] = (float)(1.0 - pow((double)beta1, (double)t));
            s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)t)));
        }
        __syncthreads();
        bc1 = s_bc[0];
        rsqrt_bc2 = s_bc[1];
    }
    const int lane = threadIdx.x & 31;
    const int64_t n = n0 + (int64_t)blockIdx.x * MU_THREADS + threadIdx.x;
    const bool live = n < n1;
    const float* row = rows + (live ? n : n0) * row_stride;
    // the narrow columns of a row (quats | means opac | scales unc | first colours .. ) in registers: with
    // multicast each 16-byte load is one in-switch reduction over the GPUs
    float head[16];
    float vis = 0.f;
    if (MC) {
        if (live) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                if (v * 4 < row_stride) {
                    const float4 q = mc_ld_reduce4(row + v * 4);
                    head[v * 4] = q.x; head[v * 4 + 1] = q.y; head[v * 4 + 2] = q.z; head[v * 4 + 3] = q.w;
                }
            }
            vis = vis_col < 16 ? head[vis_col] : mc_ld_reduce1(row + vis_col);
        }
        if (live && mc.vis_mc) mc_st1(mc.vis_mc + n, __int_as_float((int32_t)vis));
    } else {
        vis = live ? row[vis_col] : 0.f;
        if (live && vis_count_out) vis_count_out[n] = (int32_t)vis;
    }
    float iso_part = 0.f;
#pragma unroll 1
    for (int ti = 0; ti < pk.n_tensors; ++ti) {
        const gsb_map_tensor T = pk.t[ti];
        const float step_size = T.lr / bc1;
        if (T.width <= 4) {
            if (!live) continue;
            float p[4], g[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (k < T.width) { p[k] = T.param[n * T.width + k]; g[k] = MC ? head[(T.col + k) & 15] : row[T.col + k]; }
            if (ti == scales_tensor && T.width == 3 && vis > 0.f) {
                // d/ds of  w * sum_axis | exp(s) - exp(mean_axis s).detach() |  on the union visibility
                const float m = expf((p[0] + p[1] + p[2]) * (1.0f / 3.0f));
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float e = expf(p[k]), d = e - m;
                    iso_part += fabsf(d);
                    g[k] += iso_weight * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) * e;
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (k < T.width) {
                    const int64_t i = n * T.width + k;
                    float q = adam_elem(p[k], g[k], T.exp_avg + i, T.exp_avg_sq + i, beta1, beta2, eps, step_size,
                                        rsqrt_bc2);
                    if (ti == decay_tensor && vis > 1.f) q *= decay;
                    if (MC) mc_st1(mc.param_mc[ti] + i, q);
                    else T.param[i] = q;
                }
            }
        } else {
            // wide rows (SH coefficients): the warp walks its 32 Gaussians' elements together, so the
            // parameter and moment accesses are contiguous
            const int64_t nb = n - lane;                                   // first Gaussian of this warp
            const int cnt = (int)max((int64_t)0, min((int64_t)32, n1 - nb));
            const int total = cnt * T.width;
            for (int e = lane; e < total; e += 32) {
                const int gi = e / T.width, col = e - gi * T.width;
                const int64_t i = (nb + gi) * T.width + col;
                const float* gp = rows + (nb + gi) * row_stride + T.col + col;
                const float g = MC ? mc_ld_reduce1(gp) : *gp;
                float q = adam_elem(T.param[i], g, T.exp_avg + i, T.exp_avg_sq + i, beta1, beta2, eps, step_size,
                                    rsqrt_bc2);
                if (MC) mc_st1(mc.param_mc[ti] + i, q);
                else T.param[i] = q;
            }
        }
    }
    if (MC || iso_loss_sum) {
        iso_part = warp_sum(iso_part);
        if (lane == 0) red[threadIdx.x >> 5] = iso_part;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < MU_THREADS / 32; ++w) s += red[w];
            if (MC) {
                // the block that finishes last adds the partials in block order (fixed order: reproducible) and
                // multicasts this rank's (total share, photometric share, isotropic sum, 0) to every GPU
                mc.scratch[blockIdx.x] = s;
                __threadfence();
                int* ticket = reinterpret_cast<int*>(mc.scratch + gridDim.x);
                if (atomicAdd(ticket, 1) == (int)gridDim.x - 1) {
                    __threadfence();
                    float tot = 0.f;
                    for (unsigned b = 0; b < gridDim.x; ++b) tot += reinterpret_cast<volatile float*>(mc.scratch)[b];
                    *ticket = 0;
                    if (mc.stats_mc)
                        mc_st4(mc.stats_mc + mc.rank * 4, make_float4(mc.stats_in ? mc.stats_in[0] : 0.f,
                                                                      mc.stats_in ? mc.stats_in[1] : 0.f, tot, 0.f));
                }
            } else if (det_partials) det_partials[blockIdx.x] = s;
            else if (s != 0.f) atomicAdd(iso_loss_sum, s);
        }
    }
    if (step_dev) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(step_dev + 1, 1) == (int)gridDim.x - 1) {
                step_dev[0] += 1;
                step_dev[1] = 0;
            }
        }
    }
}

GSB_API int gsb_map_update(const gsb_map_tensor* tensors_host, int n_tensors, const float* grad_rows,
                           int64_t row_stride, int vis_col, int64_t n0, int64_t n1, int scales_tensor,
                           float iso_weight, int decay_tensor, float decay, float beta1, float beta2, float eps,
                           int step, int32_t* step_dev, int32_t* vis_count_out, float* iso_loss_sum,
                           float* det_partials, void* stream) {
    GSB_CHECK_ARG(n_tensors >= 0 && n_tensors <= MU_MAX_TENSORS && (step >= 1 || step_dev));
    GSB_CHECK_ARG(n0 >= 0 && n1 >= n0 && row_stride > 0 && vis_col >= 0 && vis_col < row_stride);
    if (n1 == n0 && !step_dev) return 0;       // (with a device-side step count an empty range still counts a step)
    GSB_CHECK_ARG(tensors_host && (grad_rows || n1 == n0));
    GSB_CHECK_ARG(scales_tensor < n_tensors && decay_tensor < n_tensors);
    MapUpdatePack pk;
    pk.n_tensors = n_tensors;
    for (int k = 0; k < n_tensors; ++k) {
        const gsb_map_tensor& t = tensors_host[k];
        GSB_CHECK_ARG(t.param && t.exp_avg && t.exp_avg_sq && t.width >= 1 && t.col >= 0 && t.col + t.width <= row_stride);
        pk.t[k] = t;
    }
    GSB_CHECK_ARG(scales_tensor < 0 || pk.t[scales_tensor].width == 3);
    GSB_CHECK_ARG(decay_tensor < 0 || pk.t[decay_tensor].width <= 4);
    if (step < 1) step = 1;     // unused when step_dev is given
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    const int64_t blocks = n1 > n0 ? (n1 - n0 + MU_THREADS - 1) / MU_THREADS : 1;
    GSB_CHECK_ARG(blocks < ((int64_t)1 << 31));
    McPack mc = {};
    map_update_kernel<false><<<(unsigned)blocks, MU_THREADS, 0, gsb_stream(stream)>>>(
        pk, mc, grad_rows, row_stride, vis_col, n0, n1, scales_tensor, iso_weight, decay_tensor, decay, beta1, beta2,
        eps, (float)bc1, (float)(1.0 / sqrt(bc2)), step_dev, vis_count_out, iso_loss_sum,
        iso_loss_sum ? det_partials : nullptr);
    GSB_LAUNCH_CHECK();
    if (iso_loss_sum && det_partials) return gsb_det_sum_launch(det_partials, blocks, iso_loss_sum, gsb_stream(stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// Fused reduce-scatter + update + all-gather over NVLink multicast, the throughput form.  The generic
// kernel above (MC = true) is correct but issues one 4-byte multicast store per parameter element and four
// strided 16-byte load-reduces per Gaussian -- NVLink moves those as tiny packets (measured: 333 us for
// 250 k Gaussians on two GPUs, slower than NCCL).  Here a CTA of 256 Gaussians
//   1. load-reduces the first 16 columns of its 256 rows with fully coalesced 16-byte multimem.ld_reduce
//      (a warp instruction = 512 contiguous bytes) into shared memory,
//   2. runs the update one thread per Gaussian out of shared memory (rows at a 20-word stride: the four
//      LDS.128 of a row are conflict-free), staging the new parameters per tensor in shared memory,
//   3. multicast-stores every tensor's 256-Gaussian chunk -- contiguous in the parameter arrays -- with
//      coalesced 16-byte multimem.st.
// Wide tensors (SH coefficients) are walked by the warp over contiguous elements as before.
// ---------------------------------------------------------------------------------------------------
#define MC_ROW_S 20             // shared-memory row stride in words (16 columns + pad, 16-byte aligned)
#define MC_OUT_MAX 16           // staged output floats per Gaussian: narrow tensors (<= 15) + visibility

__global__ void __launch_bounds__(MU_THREADS)
map_update_mc_kernel(const MapUpdatePack pk, const McPack mc, const float* __restrict__ rows, int64_t row_stride,
                     int vis_col, int64_t n0, int64_t n1, int scales_tensor, float iso_weight, int decay_tensor,
                     float decay, float beta1, float beta2, float eps, float bc1, float rsqrt_bc2,
                     int32_t* step_dev) {
    __shared__ __align__(16) float s_rows[MU_THREADS * MC_ROW_S];
    __shared__ __align__(16) float s_out[MU_THREADS * MC_OUT_MAX];
    __shared__ float s_bc[2];
    __shared__ float red[MU_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31;
    if (step_dev) {
        if (tid == 0) {
            const int t = *reinterpret_cast<volatile int32_t*>(step_dev) + 1;
            s_bc[0] = (float)(1.0 - pow((double)beta1, (double)t));
            s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)t)));
        }
    }
    const int64_t nb0 = n0 + (int64_t)blockIdx.x * MU_THREADS;       // first Gaussian of this CTA
    const int cnt = (int)max((int64_t)0, min((int64_t)MU_THREADS, n1 - nb0));
    // 1. coalesced in-switch reduction of columns 0..15 of the CTA's rows
    const int quads = row_stride == 16 ? cnt * 4 : 0;
    if (row_stride == 16) {
        const float* src = rows + nb0 * 16;
        for (int q = tid; q < quads; q += MU_THREADS) {
            const float4 v = mc_ld_reduce4(src + q * 4);
            *reinterpret_cast<float4*>(&s_rows[(q >> 2) * MC_ROW_S + (q & 3) * 4]) = v;
        }
    } else {
        for (int q = tid; q < cnt * 4; q += MU_THREADS) {
            const float4 v = mc_ld_reduce4(rows + (nb0 + (q >> 2)) * row_stride + (q & 3) * 4);
            *reinterpret_cast<float4*>(&s_rows[(q >> 2) * MC_ROW_S + (q & 3) * 4]) = v;
        }
    }
    __syncthreads();
    if (step_dev) { bc1 = s_bc[0]; rsqrt_bc2 = s_bc[1]; }
    const int64_t n = nb0 + tid;
    const bool live = tid < cnt;
    float head[16];
#pragma unroll
    for (int v = 0; v < 4; ++v) {
        const float4 q = *reinterpret_cast<const float4*>(&s_rows[tid * MC_ROW_S + v * 4]);
        head[v * 4] = q.x; head[v * 4 + 1] = q.y; head[v * 4 + 2] = q.z; head[v * 4 + 3] = q.w;
    }
    const float vis = !live ? 0.f : (vis_col < 16 ? head[vis_col] : mc_ld_reduce1(rows + n * row_stride + vis_col));
    float iso_part = 0.f;
    int out_off = 0;            // running offset (floats per Gaussian) of the narrow tensors in s_out
#pragma unroll 1
    for (int ti = 0; ti < pk.n_tensors; ++ti) {
        const gsb_map_tensor T = pk.t[ti];
        const float step_size = T.lr / bc1;
        if (T.width <= 4) {
            if (live) {
                float p[4], g[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (k < T.width) { p[k] = T.param[n * T.width + k]; g[k] = head[(T.col + k) & 15]; }
                if (ti == scales_tensor && T.width == 3 && vis > 0.f) {
                    const float m = expf((p[0] + p[1] + p[2]) * (1.0f / 3.0f));
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const float e = expf(p[k]), d = e - m;
                        iso_part += fabsf(d);
                        g[k] += iso_weight * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) * e;
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (k < T.width) {
                        const int64_t i = n * T.width + k;
                        float q = adam_elem(p[k], g[k], T.exp_avg + i, T.exp_avg_sq + i, beta1, beta2, eps, step_size,
                                            rsqrt_bc2);
                        if (ti == decay_tensor && vis > 1.f) q *= decay;
                        s_out[out_off * MU_THREADS + tid * T.width + k] = q;
                    }
                }
            }
            out_off += T.width;
        } else {
            const int64_t nb = n - lane;
            const int wcnt = (int)max((int64_t)0, min((int64_t)32, n1 - nb));
            const int total = wcnt * T.width;
            for (int e = lane; e < total; e += 32) {
                const int gi = e / T.width, col = e - gi * T.width;
                const int64_t i = (nb + gi) * T.width + col;
                const float g = mc_ld_reduce1(rows + (nb + gi) * row_stride + T.col + col);
                const float q = adam_elem(T.param[i], g, T.exp_avg + i, T.exp_avg_sq + i, beta1, beta2, eps, step_size,
                                          rsqrt_bc2);
                mc_st1(mc.param_mc[ti] + i, q);
            }
        }
    }
    if (live) s_out[out_off * MU_THREADS + tid] = __int_as_float((int32_t)vis);
    __syncthreads();
    // 3. coalesced multicast stores of every narrow tensor's chunk (+ the visibility counts)
    out_off = 0;
#pragma unroll 1
    for (int ti = 0; ti <= pk.n_tensors; ++ti) {
        const bool is_vis = ti == pk.n_tensors;
        const int w = is_vis ? 1 : pk.t[ti].width;
        if (!is_vis && w > 4) continue;
        float* dst = is_vis ? mc.vis_mc : mc.param_mc[ti];
        if (dst) {
            dst += nb0 * w;
            const float* src = s_out + out_off * MU_THREADS;
            const int valid = cnt * w;
            for (int q = tid; q * 4 < valid; q += MU_THREADS) {
                if (q * 4 + 3 < valid) mc_st4(dst + q * 4, *reinterpret_cast<const float4*>(src + q * 4));
                else
                    for (int k = q * 4; k < valid; ++k) mc_st1(dst + k, src[k]);
            }
        }
        out_off += w;
    }
    // this rank's loss scalars: block partials, the last block adds them in block order and multicasts
    iso_part = warp_sum(iso_part);
    if (lane == 0) red[tid >> 5] = iso_part;
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < MU_THREADS / 32; ++w) s += red[w];
        mc.scratch[blockIdx.x] = s;
        __threadfence();
        int* ticket = reinterpret_cast<int*>(mc.scratch + gridDim.x);
        if (atomicAdd(ticket, 1) == (int)gridDim.x - 1) {
            __threadfence();
            float tot = 0.f;
            for (unsigned b = 0; b < gridDim.x; ++b) tot += reinterpret_cast<volatile float*>(mc.scratch)[b];
            *ticket = 0;
            if (mc.stats_mc)
                mc_st4(mc.stats_mc + mc.rank * 4, make_float4(mc.stats_in ? mc.stats_in[0] : 0.f,
                                                              mc.stats_in ? mc.stats_in[1] : 0.f, tot, 0.f));
            if (step_dev) step_dev[0] += 1;
        }
    }
}

// Fused reduce-scatter + update + all-gather over NVLink multicast (see map_update_kernel<true>).
GSB_API size_t gsb_map_update_mc_scratch_floats(int64_t n0, int64_t n1) {
    const int64_t blocks = n1 > n0 ? (n1 - n0 + MU_THREADS - 1) / MU_THREADS : 1;
    return (size_t)blocks + 4;
}

GSB_API int gsb_map_update_mc(const gsb_map_tensor* tensors_host, const uint64_t* param_mc_host, int n_tensors,
                              const float* grad_rows_mc, int64_t row_stride, int vis_col, int64_t n0, int64_t n1,
                              int scales_tensor, float iso_weight, int decay_tensor, float decay, float beta1,
                              float beta2, float eps, int step, int32_t* step_dev, int32_t* vis_count_mc,
                              const float* stats_in, float* stats_mc, int rank, float* scratch, void* stream) {
    GSB_CHECK_ARG(n_tensors >= 1 && n_tensors <= MU_MAX_TENSORS && (step >= 1 || step_dev));
    GSB_CHECK_ARG(n0 >= 0 && n1 >= n0 && row_stride > 0 && row_stride % 4 == 0 && vis_col >= 0 && vis_col < row_stride);
    GSB_CHECK_ARG(tensors_host && param_mc_host && grad_rows_mc && scratch && rank >= 0);
    GSB_CHECK_ARG((reinterpret_cast<uintptr_t>(grad_rows_mc) & 15) == 0 && (reinterpret_cast<uintptr_t>(stats_mc) & 15) == 0);
    GSB_CHECK_ARG(scales_tensor < n_tensors && decay_tensor < n_tensors);
    MapUpdatePack pk;
    McPack mc = {};
    pk.n_tensors = n_tensors;
    for (int k = 0; k < n_tensors; ++k) {
        const gsb_map_tensor& t = tensors_host[k];
        GSB_CHECK_ARG(t.param && t.exp_avg && t.exp_avg_sq && t.width >= 1 && t.col >= 0 && t.col + t.width <= row_stride);
        GSB_CHECK_ARG(t.width > 4 || t.col + t.width <= 16);      // narrow tensors live in the first 16 columns
        GSB_CHECK_ARG(param_mc_host[k] != 0);
        pk.t[k] = t;
        mc.param_mc[k] = reinterpret_cast<float*>(param_mc_host[k]);
    }
    GSB_CHECK_ARG(scales_tensor < 0 || pk.t[scales_tensor].width == 3);
    GSB_CHECK_ARG(decay_tensor < 0 || pk.t[decay_tensor].width <= 4);
    mc.vis_mc = reinterpret_cast<float*>(vis_count_mc);
    mc.stats_in = stats_in;
    mc.stats_mc = stats_mc;
    mc.scratch = scratch;
    mc.rank = rank;
    if (step < 1) step = 1;
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    const int64_t blocks = n1 > n0 ? (n1 - n0 + MU_THREADS - 1) / MU_THREADS : 1;
    GSB_CHECK_ARG(blocks < ((int64_t)1 << 31));
    // 16-byte multicast stores of every tensor's chunk: the range must start on a multiple of 4 Gaussians and the
    // parameter arrays on 16 bytes (FlatSplats(align = 4 * ranks) gives both)
    int narrow = 1;
    bool aligned = (n0 % 4) == 0 && (reinterpret_cast<uintptr_t>(vis_count_mc) & 15) == 0;
    for (int k = 0; k < n_tensors; ++k) {
        if (pk.t[k].width <= 4) narrow += pk.t[k].width;
        aligned = aligned && (param_mc_host[k] & 15) == 0;
    }
    if (aligned && narrow <= MC_OUT_MAX) {
        map_update_mc_kernel<<<(unsigned)blocks, MU_THREADS, 0, gsb_stream(stream)>>>(
            pk, mc, grad_rows_mc, row_stride, vis_col, n0, n1, scales_tensor, iso_weight, decay_tensor, decay, beta1,
            beta2, eps, (float)bc1, (float)(1.0 / sqrt(bc2)), step_dev);
        GSB_LAUNCH_CHECK();
        return 0;
    }
    map_update_kernel<true><<<(unsigned)blocks, MU_THREADS, 0, gsb_stream(stream)>>>(
        pk, mc, grad_rows_mc, row_stride, vis_col, n0, n1, scales_tensor, iso_weight, decay_tensor, decay, beta1,
        beta2, eps, (float)bc1, (float)(1.0 / sqrt(bc2)), step_dev, nullptr, nullptr, nullptr);
    GSB_LAUNCH_CHECK();
    return 0;
}
