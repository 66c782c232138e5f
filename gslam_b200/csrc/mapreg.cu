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

// ---------------------------------------------------------------------------
// Allocation-free variant of gsb_isotropic_reg's deterministic mode (no counterpart in the reference;
// added for callers that cannot allocate device memory: code
// under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The per-block
// partial sums live in a file-scope device array instead of caller scratch: 64 Ki floats = 256 KiB,
// held by every GPU that loads the library until the process exits -> N <= 256 * 65536 = 16 777 216.
// In host code the array's NAME is a host-side stand-in, not a device address; the address is taken
// once per device with cudaGetSymbolAddress (with that device current) and cached.
// Every call on a device shares the array and finds the previous call's partials in it: the kernel
// overwrites partials[0 .. ceil(N/256)) (every block stores its sum, zero included) and the fixed-order
// sum reads exactly that range, so nothing has to be cleared -- but two calls that OVERLAP on one GPU
// (different streams, no ordering between them) race on it.  Calls on one stream are safe.
// ---------------------------------------------------------------------------
#define GSB_ISO_STATIC_MAX_BLOCKS 65536
static __device__ __align__(256) float g_iso_static_scratch[GSB_ISO_STATIC_MAX_BLOCKS];
#define GSB_MAX_DEVICES 64
static float* g_iso_static_ptr[GSB_MAX_DEVICES];        // device address of the array, per device (0: not taken yet)

GSB_API int64_t gsb_isotropic_reg_static_max_n(void) { return (int64_t)GSB_ISO_STATIC_MAX_BLOCKS * 256; }

GSB_API int gsb_isotropic_reg_static(const float* log_scales, const int32_t* vis_count, int64_t N, float weight,
                                     float* v_log_scales, float* loss_sum, void* stream) {
    GSB_CHECK_ARG(N >= 0);
    if (N > gsb_isotropic_reg_static_max_n()) {
        gsb_set_error("gsb_isotropic_reg_static: N = %lld exceeds the static scratch (at most %lld Gaussians); "
                      "use gsb_isotropic_reg with caller scratch", (long long)N,
                      (long long)gsb_isotropic_reg_static_max_n());
        return GSB_EINVAL;
    }
    GSB_CHECK_ARG(loss_sum);
    if (N == 0) {
        GSB_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), gsb_stream(stream)));
        return 0;
    }
    GSB_CHECK_ARG(log_scales && vis_count && v_log_scales);
    int dev = -1;
    GSB_CUDA(cudaGetDevice(&dev));
    GSB_CHECK_ARG(dev >= 0 && dev < GSB_MAX_DEVICES);
    float* scratch = __atomic_load_n(&g_iso_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (!scratch) {
        // set-up, once per device, with that device current (racing threads store the same value)
        GSB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_iso_static_scratch));
        GSB_CHECK_ARG(scratch != nullptr);
        __atomic_store_n(&g_iso_static_ptr[dev], scratch, __ATOMIC_RELEASE);
    }
    const unsigned blocks = (unsigned)((N + 255) / 256);
    isotropic_reg_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(log_scales, vis_count, N, weight, v_log_scales,
                                                                loss_sum, scratch);
    GSB_LAUNCH_CHECK();
    return gsb_det_sum_launch(scratch, blocks, loss_sum, gsb_stream(stream));
}
