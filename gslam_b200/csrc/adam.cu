// K12/K13: single-launch multi-tensor Adam with the opacity-decay epilogue.
// Replaces the seven torch.optim.Adam(fused=True) steps of reference
// gslam/backend.py:559-602 (one optimizer per parameter tensor) and the masked
// `opacities *= 0.995` at gslam/backend.py:356-359.  Semantics are those of
// torch.optim.Adam defaults (betas (0.9, 0.999), eps 1e-8, no weight decay, no
// amsgrad): m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
// p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
#include "common.cuh"

#define ADAM_MAX_TENSORS 8
#define ADAM_PER_BLOCK 1024

struct AdamPack {
    gsb_adam_tensor t[ADAM_MAX_TENSORS];
    int first_block[ADAM_MAX_TENSORS + 1];
    int n_tensors;
};

__global__ void __launch_bounds__(256)
adam_multi_kernel(const AdamPack pk, float beta1, float beta2, float eps, float bc1, float rsqrt_bc2,
                  int decay_tensor, const int32_t* __restrict__ vis_count, int vis_div, float decay,
                  int32_t* step_dev) {
    // Device-side step count (CUDA-graph replays cannot take it as a launch argument): every block
    // reads step_dev[0] before doing anything; the block that finishes LAST bumps it.
    __shared__ float s_bc[2];
    if (step_dev) {
        if (threadIdx.x == 0) {
            const int t = *reinterpret_cast<volatile int32_t*>(step_dev) + 1;
            s_bc[0] = (float)(1.0 - pow((double)beta1, (double)t));
            s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)t)));
        }
        __syncthreads();
        bc1 = s_bc[0];
        rsqrt_bc2 = s_bc[1];
    }
    int ti = 0;
#pragma unroll
    for (int k = 1; k < ADAM_MAX_TENSORS; ++k)
        if (k < pk.n_tensors && (int)blockIdx.x >= pk.first_block[k]) ti = k;
    const gsb_adam_tensor T = pk.t[ti];
    const int64_t base = (int64_t)(blockIdx.x - pk.first_block[ti]) * ADAM_PER_BLOCK;
    const float step_size = T.lr / bc1;
#pragma unroll
    for (int u = 0; u < ADAM_PER_BLOCK / 256; ++u) {
        const int64_t i = base + u * 256 + threadIdx.x;
        if (i >= T.n) break;
        const float g = T.grad[i];
        const float m = beta1 * T.exp_avg[i] + (1.f - beta1) * g;
        const float v = beta2 * T.exp_avg_sq[i] + (1.f - beta2) * g * g;
        T.exp_avg[i] = m;
        T.exp_avg_sq[i] = v;
        float p = T.param[i] - step_size * m / (sqrtf(v) * rsqrt_bc2 + eps);
        if (ti == decay_tensor && vis_count[i / vis_div] > 1) p *= decay;
        T.param[i] = p;
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

GSB_API int gsb_adam_multi(const gsb_adam_tensor* tensors_host, int n_tensors, float beta1, float beta2,
                           float eps, int step, int decay_tensor, const int32_t* vis_count,
                           int vis_div, float decay, int32_t* step_dev, void* stream) {
    GSB_CHECK_ARG(n_tensors >= 0 && n_tensors <= ADAM_MAX_TENSORS && (step >= 1 || step_dev));
    if (n_tensors == 0) return 0;
    GSB_CHECK_ARG(tensors_host);
    GSB_CHECK_ARG(decay_tensor < n_tensors && (decay_tensor < 0 || (vis_count && vis_div >= 1)));
    AdamPack pk;
    pk.n_tensors = n_tensors;
    int blocks = 0;
    for (int k = 0; k < n_tensors; ++k) {
        const gsb_adam_tensor& t = tensors_host[k];
        GSB_CHECK_ARG(t.n >= 0 && (t.n == 0 || (t.param && t.grad && t.exp_avg && t.exp_avg_sq)));
        pk.t[k] = t;
        pk.first_block[k] = blocks;
        blocks += (int)((t.n + ADAM_PER_BLOCK - 1) / ADAM_PER_BLOCK);
    }
    for (int k = n_tensors; k <= ADAM_MAX_TENSORS; ++k) pk.first_block[k] = blocks;
    if (blocks == 0) return 0;
    if (step < 1) step = 1;     // unused when step_dev is given
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    adam_multi_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(pk, beta1, beta2, eps, (float)bc1,
                                                             (float)(1.0 / sqrt(bc2)), decay_tensor,
                                                             vis_count, vis_div, decay, step_dev);
    GSB_LAUNCH_CHECK();
    return 0;
}
