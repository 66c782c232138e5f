// K11: camera pose chain, forward and backward in one launch each.
// Replaces PoseZhou.forward + rotation_6d_to_matrix (reference gslam/primitives.py:15-36, 82-92):
//   viewmat = Rt . [[GS(dR + (1,0,0,0,1,0)), dt], [0 0 0 1]]
// where GS is the Gram-Schmidt map of Zhou et al.'s 6-D rotation representation (rows b1, b2, b3).
// The reference runs ~25 autograd nodes per view per iteration for this; here one thread per view
// does the whole 16 -> 9 chain, so the gradient sink of v_viewmats[C,4,4] costs one tiny launch.
#include "common.cuh"

struct Gs {
    float a1[3], a2[3], b1[3], u[3], b2[3], b3[3];
    float n1, nu, d12;      // |a1|, |u|, b1 . a2
};

__device__ __forceinline__ float dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ void cross3(const float* a, const float* b, float* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

__device__ __forceinline__ void gram_schmidt(const float* dR, Gs& g) {
    g.a1[0] = dR[0] + 1.f; g.a1[1] = dR[1]; g.a1[2] = dR[2];
    g.a2[0] = dR[3]; g.a2[1] = dR[4] + 1.f; g.a2[2] = dR[5];
    g.n1 = fmaxf(sqrtf(dot3(g.a1, g.a1)), 1e-12f);      // torch.nn.functional.normalize: eps = 1e-12
#pragma unroll
    for (int k = 0; k < 3; ++k) g.b1[k] = g.a1[k] / g.n1;
    g.d12 = dot3(g.b1, g.a2);
#pragma unroll
    for (int k = 0; k < 3; ++k) g.u[k] = g.a2[k] - g.d12 * g.b1[k];
    g.nu = fmaxf(sqrtf(dot3(g.u, g.u)), 1e-12f);
#pragma unroll
    for (int k = 0; k < 3; ++k) g.b2[k] = g.u[k] / g.nu;
    cross3(g.b1, g.b2, g.b3);
}

__global__ void pose_fwd_kernel(const float* __restrict__ Rt, const float* __restrict__ dR,
                                const float* __restrict__ dt, int C, float* __restrict__ viewmats) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    Gs g;
    gram_schmidt(dR + c * 6, g);
    // D = [[b1; b2; b3], dt], rows are the basis vectors
    float D[4][4] = {{g.b1[0], g.b1[1], g.b1[2], dt[c * 3]},
                     {g.b2[0], g.b2[1], g.b2[2], dt[c * 3 + 1]},
                     {g.b3[0], g.b3[1], g.b3[2], dt[c * 3 + 2]},
                     {0.f, 0.f, 0.f, 1.f}};
    const float* R = Rt + c * 16;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < 4; ++k) s = fmaf(R[i * 4 + k], D[k][j], s);
            viewmats[c * 16 + i * 4 + j] = s;
        }
}

__global__ void pose_bwd_kernel(const float* __restrict__ Rt, const float* __restrict__ dR,
                                const float* __restrict__ v_viewmats, int C, float* __restrict__ v_dR,
                                float* __restrict__ v_dt) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    Gs g;
    gram_schmidt(dR + c * 6, g);
    // dL/dD = Rt^T . dL/dV  (only its first three rows matter)
    const float* R = Rt + c * 16;
    const float* G = v_viewmats + c * 16;
    float gD[3][4];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) s = fmaf(R[i * 4 + k], G[i * 4 + j], s);
            gD[k][j] = s;
        }
    if (v_dt) {
#pragma unroll
        for (int k = 0; k < 3; ++k) v_dt[c * 3 + k] = gD[k][3];
    }
    if (!v_dR) return;
    float g1[3] = {gD[0][0], gD[0][1], gD[0][2]};
    float g2[3] = {gD[1][0], gD[1][1], gD[1][2]};
    const float g3[3] = {gD[2][0], gD[2][1], gD[2][2]};
    // b3 = b1 x b2
    float t[3];
    cross3(g.b2, g3, t);
#pragma unroll
    for (int k = 0; k < 3; ++k) g1[k] += t[k];
    cross3(g3, g.b1, t);
#pragma unroll
    for (int k = 0; k < 3; ++k) g2[k] += t[k];
    // b2 = u / |u|
    float gu[3];
    const float p2 = dot3(g2, g.b2);
#pragma unroll
    for (int k = 0; k < 3; ++k) gu[k] = (g2[k] - p2 * g.b2[k]) / g.nu;
    // u = a2 - (b1 . a2) b1
    const float pu = dot3(gu, g.b1);
    float ga2[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        ga2[k] = gu[k] - pu * g.b1[k];
        g1[k] -= g.d12 * gu[k] + pu * g.a2[k];
    }
    // b1 = a1 / |a1|
    const float p1 = dot3(g1, g.b1);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        v_dR[c * 6 + k] = (g1[k] - p1 * g.b1[k]) / g.n1;
        v_dR[c * 6 + 3 + k] = ga2[k];
    }
}

GSB_API int gsb_pose_fwd(const float* Rt, const float* dR, const float* dt, int C, float* viewmats,
                         void* stream) {
    GSB_CHECK_ARG(C >= 0);
    if (C == 0) return 0;
    GSB_CHECK_ARG(Rt && dR && dt && viewmats);
    pose_fwd_kernel<<<(C + 63) / 64, 64, 0, gsb_stream(stream)>>>(Rt, dR, dt, C, viewmats);
    GSB_LAUNCH_CHECK();
    return 0;
}

GSB_API int gsb_pose_bwd(const float* Rt, const float* dR, const float* v_viewmats, int C, float* v_dR,
                         float* v_dt, void* stream) {
    GSB_CHECK_ARG(C >= 0);
    if (C == 0) return 0;
    GSB_CHECK_ARG(Rt && dR && v_viewmats && (v_dR || v_dt));
    pose_bwd_kernel<<<(C + 63) / 64, 64, 0, gsb_stream(stream)>>>(Rt, dR, v_viewmats, C, v_dR, v_dt);
    GSB_LAUNCH_CHECK();
    return 0;
}
