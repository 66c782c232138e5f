// Real spherical harmonics up to degree 3 (16 coefficients per channel), the view-dependent
// colour extension of SURVEY 8(a) "SH note" (NOT a reference feature: gslam's colours are
// sigmoid(logit_colors), gslam/rasterization.py:146).  Semantics follow upstream gsplat's
// rendering path: d = normalize(mu - camera centre), rgb = max(sum_k Y_k(d) c_k + 0.5, 0).
#pragma once
#include "common.cuh"

#define SH_MAX_K 16

__device__ __forceinline__ int sh_num_coeffs(int degree) { return (degree + 1) * (degree + 1); }

// Basis values Y[0..K) at the unit vector (x, y, z); 3DGS sign convention.
__device__ __forceinline__ void sh_basis(int degree, float x, float y, float z, float* Y) {
    Y[0] = 0.28209479177387814f;
    if (degree < 1) return;
    const float C1 = 0.4886025119029199f;
    Y[1] = -C1 * y; Y[2] = C1 * z; Y[3] = -C1 * x;
    if (degree < 2) return;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    Y[4] = 1.0925484305920792f * xy;
    Y[5] = -1.0925484305920792f * yz;
    Y[6] = 0.31539156525252005f * (2.f * zz - xx - yy);
    Y[7] = -1.0925484305920792f * xz;
    Y[8] = 0.5462742152960396f * (xx - yy);
    if (degree < 3) return;
    Y[9] = -0.5900435899266435f * y * (3.f * xx - yy);
    Y[10] = 2.890611442640554f * xy * z;
    Y[11] = -0.4570457994644658f * y * (4.f * zz - xx - yy);
    Y[12] = 0.3731763325901154f * z * (2.f * zz - 3.f * xx - 3.f * yy);
    Y[13] = -0.4570457994644658f * x * (4.f * zz - xx - yy);
    Y[14] = 1.445305721320277f * z * (xx - yy);
    Y[15] = -0.5900435899266435f * x * (xx - 3.f * yy);
}

// Given w[k] = dL/dY_k, accumulate dL/d(x, y, z) (as if x, y, z were free variables).
__device__ __forceinline__ void sh_basis_bwd(int degree, float x, float y, float z, const float* w,
                                             float& gx, float& gy, float& gz) {
    gx = gy = gz = 0.f;
    if (degree < 1) return;
    const float C1 = 0.4886025119029199f;
    gy += -C1 * w[1]; gz += C1 * w[2]; gx += -C1 * w[3];
    if (degree < 2) return;
    const float a = 1.0925484305920792f, b = 0.31539156525252005f, c = 0.5462742152960396f;
    gx += a * y * w[4];            gy += a * x * w[4];
    gy += -a * z * w[5];           gz += -a * y * w[5];
    gx += -2.f * b * x * w[6];     gy += -2.f * b * y * w[6];   gz += 4.f * b * z * w[6];
    gx += -a * z * w[7];           gz += -a * x * w[7];
    gx += 2.f * c * x * w[8];      gy += -2.f * c * y * w[8];
    if (degree < 3) return;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    const float d = -0.5900435899266435f, e = 2.890611442640554f, f = -0.4570457994644658f;
    const float g = 0.3731763325901154f, h = 1.445305721320277f;
    gx += d * 6.f * xy * w[9];                 gy += d * (3.f * xx - 3.f * yy) * w[9];
    gx += e * yz * w[10];                      gy += e * xz * w[10];                 gz += e * xy * w[10];
    gx += f * -2.f * xy * w[11];               gy += f * (4.f * zz - xx - 3.f * yy) * w[11];  gz += f * 8.f * yz * w[11];
    gx += g * -6.f * xz * w[12];               gy += g * -6.f * yz * w[12];          gz += g * (6.f * zz - 3.f * xx - 3.f * yy) * w[12];
    gx += f * (4.f * zz - 3.f * xx - yy) * w[13];  gy += f * -2.f * xy * w[13];      gz += f * 8.f * xz * w[13];
    gx += h * 2.f * xz * w[14];                gy += h * -2.f * yz * w[14];          gz += h * (xx - yy) * w[14];
    gx += d * (3.f * xx - 3.f * yy) * w[15];   gy += d * -6.f * xy * w[15];
}

// Camera centre of a world-to-camera transform (R, t): -R^T t.
__device__ __forceinline__ void cam_centre(const float* R, const float* t, float* o) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o[j] = -(R[j] * t[0] + R[3 + j] * t[1] + R[6 + j] * t[2]);
}

// Colour of Gaussian n seen from `centre`: coeffs is [N, K, 3].
__device__ __forceinline__ void sh_colour(int degree, const float* __restrict__ coeffs, int K, const float* mu,
                                          const float* centre, float* rgb) {
    float dx = mu[0] - centre[0], dy = mu[1] - centre[1], dz = mu[2] - centre[2];
    const float inv = rsqrtf(fmaxf(dx * dx + dy * dy + dz * dz, 1e-24f));
    float Y[SH_MAX_K];
    sh_basis(degree, dx * inv, dy * inv, dz * inv, Y);
    float acc[3] = {0.5f, 0.5f, 0.5f};
    if ((reinterpret_cast<uintptr_t>(coeffs) & 15) == 0) {
        // the 3K floats of one Gaussian as 16-byte loads: a quarter of the memory instructions of the
        // scalar walk (each of which touches 32 different lines across the warp)
        const float4* c4 = reinterpret_cast<const float4*>(coeffs);
#pragma unroll
        for (int v = 0; v < SH_MAX_K * 3 / 4; ++v) {
            if (v * 4 < K * 3) {
                const float4 q = c4[v];
                const float e[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int f = v * 4 + j;            // compile-time: coefficient f / 3, channel f % 3
                    if (f < K * 3) acc[f % 3] = fmaf(Y[f / 3], e[j], acc[f % 3]);
                }
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < SH_MAX_K; ++k) {
            if (k < K) {
                acc[0] = fmaf(Y[k], coeffs[k * 3], acc[0]);
                acc[1] = fmaf(Y[k], coeffs[k * 3 + 1], acc[1]);
                acc[2] = fmaf(Y[k], coeffs[k * 3 + 2], acc[2]);
            }
        }
    }
    const float r = acc[0], g = acc[1], b = acc[2];
    rgb[0] = fmaxf(r, 0.f); rgb[1] = fmaxf(g, 0.f); rgb[2] = fmaxf(b, 0.f);
}
