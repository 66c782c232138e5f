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

__global__ void __launch_bounds__(MU_THREADS)
map_update_kernel(const MapUpdatePack pk, const float* __restrict__ rows, int64_t row_stride, int vis_col,
                  int64_t n0, int64_t n1, int scales_tensor, float iso_weight, int decay_tensor, float decay,
                  float beta1, float beta2, float eps, float bc1, float rsqrt_bc2, int32_t* step_dev,
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
    const float vis = live ? row[vis_col] : 0.f;
    if (live && vis_count_out) vis_count_out[n] = (int32_t)vis;
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
                if (k < T.width) { p[k] = T.param[n * T.width + k]; g[k] = row[T.col + k]; }
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
                    T.param[i] = q;
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
                const float g = rows[(nb + gi) * row_stride + T.col + col];
                float q = adam_elem(T.param[i], g, T.exp_avg + i, T.exp_avg_sq + i, beta1, beta2, eps, step_size,
                                    rsqrt_bc2);
                if (ti == decay_tensor && rows[(nb + gi) * row_stride + vis_col] > 1.f) q *= decay;
                T.param[i] = q;
            }
        }
    }
    if (iso_loss_sum) {
        iso_part = warp_sum(iso_part);
        if (lane == 0) red[threadIdx.x >> 5] = iso_part;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < MU_THREADS / 32; ++w) s += red[w];
            if (det_partials) det_partials[blockIdx.x] = s;
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
    if (step < 1) step = 1;     // unused when step_dev is given
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    const int64_t blocks = n1 > n0 ? (n1 - n0 + MU_THREADS - 1) / MU_THREADS : 1;
    GSB_CHECK_ARG(blocks < ((int64_t)1 << 31));
    map_update_kernel<<<(unsigned)blocks, MU_THREADS, 0, gsb_stream(stream)>>>(
        pk, grad_rows, row_stride, vis_col, n0, n1, scales_tensor, iso_weight, decay_tensor, decay, beta1, beta2,
        eps, (float)bc1, (float)(1.0 / sqrt(bc2)), step_dev, vis_count_out, iso_loss_sum,
        iso_loss_sum ? det_partials : nullptr);
    GSB_LAUNCH_CHECK();
    if (iso_loss_sum && det_partials) return gsb_det_sum_launch(det_partials, blocks, iso_loss_sum, gsb_stream(stream));
    return 0;
}
