// Small per-Gaussian pieces of the mapping step that sit between the big kernels
// (reference gslam/backend.py:287-296, :356-359).  As torch expressions they are ~18 elementwise
// and reduction launches over [N,3] per step (about 110 us at N = 500k, a sixth of a one-view
// step); here they are two launches.
#include "common.cuh"

// vis_count[n] = number of views in which Gaussian n is visible: (radii > 0).sum(dim=0)
// (gslam/backend.py:287 uses radii.sum(0) > 0 as the mask, :357 counts the views).
__global__ void __launch_bounds__(256)
vis_count_kernel(const int32_t* __restrict__ radii, int C, int64_t N, int32_t* __restrict__ out) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int c = 0;
    for (int k = 0; k < C; ++k) c += radii[(int64_t)k * N + n] > 0;
    out[n] = c;
}

// Isotropic regulariser on the visible Gaussians (gslam/backend.py:288-296):
//   L = weight * sum_{visible} sum_axis | exp(s) - exp(mean_axis s).detach() |
// adds dL/ds = weight * sign(e - m) * e to v_log_scales and the UNWEIGHTED sum of |e - m| to *loss_sum.
__global__ void __launch_bounds__(256)
isotropic_reg_kernel(const float* __restrict__ log_scales, const int32_t* __restrict__ vis_count,
                     int64_t N, float weight, float* __restrict__ v_log_scales, float* __restrict__ loss_sum,
                     float* __restrict__ det_partials) {
    __shared__ float red[8];
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    float part = 0.f;
    if (n < N && vis_count[n] > 0) {
        const float s0 = log_scales[n * 3], s1 = log_scales[n * 3 + 1], s2 = log_scales[n * 3 + 2];
        const float m = expf((s0 + s1 + s2) * (1.0f / 3.0f));
        const float e[3] = {expf(s0), expf(s1), expf(s2)};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float d = e[k] - m;
            part += fabsf(d);
            const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
            v_log_scales[n * 3 + k] += weight * sg * e[k];
        }
    }
    part = warp_sum(part);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0 && loss_sum) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[w];
        if (det_partials) det_partials[blockIdx.x] = s;     // deterministic mode: fixed-order sum afterwards
        else if (s != 0.f) atomicAdd(loss_sum, s);
    }
}

GSB_API int gsb_vis_count(const int32_t* radii, int C, int64_t N, int32_t* vis_count, void* stream) {
    GSB_CHECK_ARG(C >= 0 && N >= 0);
    if (N == 0) return 0;
    GSB_CHECK_ARG(vis_count && (radii || C == 0));
    vis_count_kernel<<<(unsigned)((N + 255) / 256), 256, 0, gsb_stream(stream)>>>(radii, C, N, vis_count);
    GSB_LAUNCH_CHECK();
    return 0;
}

GSB_API int gsb_isotropic_reg(const float* log_scales, const int32_t* vis_count, int64_t N, float weight,
                              float* v_log_scales, float* loss_sum, float* det_partials, void* stream) {
    GSB_CHECK_ARG(N >= 0);
    if (N == 0) return 0;
    GSB_CHECK_ARG(log_scales && vis_count && v_log_scales);
    const unsigned blocks = (unsigned)((N + 255) / 256);
    isotropic_reg_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(
        log_scales, vis_count, N, weight, v_log_scales, loss_sum, loss_sum ? det_partials : nullptr);
    GSB_LAUNCH_CHECK();
    if (loss_sum && det_partials) return gsb_det_sum_launch(det_partials, blocks, loss_sum, gsb_stream(stream));
    return 0;
}
