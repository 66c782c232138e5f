// Backward of the spherical-harmonics colour extension (sh.cuh; forward is fused into
// gsb_project_fwd).  One thread owns one Gaussian and loops over the cameras, so the coefficient
// gradients accumulate in registers and the contribution to v_means is a plain read-modify-write.
#include "sh.cuh"

template <int DEG>
__global__ void __launch_bounds__(128)
sh_bwd_kernel(const float* __restrict__ means, const float* __restrict__ coeffs, int sh_stride,
              const float* __restrict__ viewmats, const int32_t* __restrict__ radii,
              const float* __restrict__ v_rec, int C, int N, float* __restrict__ v_sh,
              float* __restrict__ v_means, int64_t row_stride) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float mu[3] = {means[n * 3], means[n * 3 + 1], means[n * 3 + 2]};
    const float* cf = coeffs + (size_t)n * sh_stride * 3;
    float c[K][3], vc[K][3];
    // 16-byte accesses when the row allows it (K = 4, 16 with a matching stride): see sh_colour
    // (row_stride != 0: v_sh / v_means are columns of a row-major [N, row_stride] gradient record, a multiple
    // of 4 floats wide -- the mapping step's layout, see gsb_project_bwd)
    const bool vec = (K * 3) % 4 == 0 && (sh_stride * 3) % 4 == 0 &&
                     ((reinterpret_cast<uintptr_t>(coeffs) | reinterpret_cast<uintptr_t>(v_sh)) & 15) == 0;
    if (vec) {
        const float4* c4 = reinterpret_cast<const float4*>(cf);
#pragma unroll
        for (int v = 0; v < K * 3 / 4; ++v) {
            const float4 q = c4[v];
            const float e[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) c[(v * 4 + j) / 3][(v * 4 + j) % 3] = e[j];
        }
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) c[k][ch] = cf[k * 3 + ch];
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) vc[k][ch] = 0.f;
    float vmu[3] = {0.f, 0.f, 0.f};
    for (int cam = 0; cam < C; ++cam) {
        const size_t idx = (size_t)cam * N + n;
        if (radii[idx] <= 0) continue;
        const float* V = viewmats + (size_t)cam * 16;
        const float R[9] = {V[0], V[1], V[2], V[4], V[5], V[6], V[8], V[9], V[10]};
        const float t[3] = {V[3], V[7], V[11]};
        float centre[3];
        cam_centre(R, t, centre);
        const float dx = mu[0] - centre[0], dy = mu[1] - centre[1], dz = mu[2] - centre[2];
        const float inv = rsqrtf(fmaxf(dx * dx + dy * dy + dz * dz, 1e-24f));
        const float x = dx * inv, y = dy * inv, z = dz * inv;
        float Y[SH_MAX_K];
        sh_basis(DEG, x, y, z, Y);
        float pre[3] = {0.5f, 0.5f, 0.5f};
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) pre[ch] = fmaf(Y[k], c[k][ch], pre[ch]);
        // colour gradient of this (camera, Gaussian): rec floats 6, 7, 8; zero where the clamp was active
        const float* g = v_rec + idx * GSB_REC;
        float vr[3] = {g[6], g[7], g[8]};
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) vr[ch] = pre[ch] > 0.f ? vr[ch] : 0.f;
        float w[SH_MAX_K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            w[k] = c[k][0] * vr[0] + c[k][1] * vr[1] + c[k][2] * vr[2];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) vc[k][ch] = fmaf(Y[k], vr[ch], vc[k][ch]);
        }
        float gx, gy, gz;
        sh_basis_bwd(DEG, x, y, z, w, gx, gy, gz);
        // through the normalisation d = v / |v|
        const float dotp = gx * x + gy * y + gz * z;
        vmu[0] += (gx - dotp * x) * inv;
        vmu[1] += (gy - dotp * y) * inv;
        vmu[2] += (gz - dotp * z) * inv;
    }
    float* out = v_sh + (size_t)n * (row_stride ? (size_t)row_stride : (size_t)sh_stride * 3);
    if (vec) {
        float4* o4 = reinterpret_cast<float4*>(out);
#pragma unroll
        for (int v = 0; v < K * 3 / 4; ++v)
            o4[v] = make_float4(vc[(v * 4) / 3][(v * 4) % 3], vc[(v * 4 + 1) / 3][(v * 4 + 1) % 3],
                                vc[(v * 4 + 2) / 3][(v * 4 + 2) % 3], vc[(v * 4 + 3) / 3][(v * 4 + 3) % 3]);
    } else {
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) out[k * 3 + ch] = vc[k][ch];
    }
    for (int k = K; k < sh_stride; ++k)      // coefficients above the active degree get no gradient
        for (int ch = 0; ch < 3; ++ch) out[k * 3 + ch] = 0.f;
    if (v_means) {
#pragma unroll
        for (int j = 0; j < 3; ++j) v_means[n * (row_stride ? row_stride : 3) + j] += vmu[j];
    }
}

GSB_API int gsb_sh_bwd(const float* means, const float* sh_coeffs, int sh_degree, int sh_stride,
                       const float* viewmats, const int32_t* radii, const float* v_rec, int C, int N,
                       float* v_sh, float* v_means, int64_t row_stride, void* stream) {
    GSB_CHECK_ARG(C >= 0 && N >= 0 && row_stride >= 0 && row_stride % 4 == 0);
    if (N == 0) return 0;
    GSB_CHECK_ARG(means && sh_coeffs && viewmats && radii && v_rec && v_sh);
    GSB_CHECK_ARG(sh_degree >= 0 && sh_degree <= 3 && sh_stride >= (sh_degree + 1) * (sh_degree + 1));
    const int blocks = (N + 127) / 128;
#define GSB_SH(D)                                                                                   \
    sh_bwd_kernel<D><<<blocks, 128, 0, gsb_stream(stream)>>>(means, sh_coeffs, sh_stride, viewmats, \
                                                             radii, v_rec, C, N, v_sh, v_means, row_stride)
    switch (sh_degree) {
        case 0: GSB_SH(0); break;
        case 1: GSB_SH(1); break;
        case 2: GSB_SH(2); break;
        default: GSB_SH(3); break;
    }
#undef GSB_SH
    GSB_LAUNCH_CHECK();
    return 0;
}
