// K1/K2/K3/K8: fused activations + EWA projection (+ tile counting), forward and backward.
// Replaces the chain at reference gslam/rasterization.py:145-170, :187, :225-256.
//
// One thread owns one Gaussian and loops over the cameras: the 60 B of
// parameters and the 3-D covariance are read / built once, per-Gaussian
// gradients accumulate in registers over cameras (no atomics), and only the
// per-camera pose gradient needs a block reduction.
#include "common.cuh"
#include "sh.cuh"

#define MAX_CAMS_SMEM 64

struct Cam {
    float R[9];
    float t[3];
    float fx, fy, cx, cy;
};

__device__ __forceinline__ void load_cams(Cam* cams, const float* viewmats, const float* Ks, int c0,
                                          int nc) {
    for (int i = threadIdx.x; i < nc; i += blockDim.x) {
        const float* V = viewmats + (size_t)(c0 + i) * 16;
        const float* K = Ks + (size_t)(c0 + i) * 9;
        Cam cm;
        cm.R[0] = V[0]; cm.R[1] = V[1]; cm.R[2] = V[2];  cm.t[0] = V[3];
        cm.R[3] = V[4]; cm.R[4] = V[5]; cm.R[5] = V[6];  cm.t[1] = V[7];
        cm.R[6] = V[8]; cm.R[7] = V[9]; cm.R[8] = V[10]; cm.t[2] = V[11];
        cm.fx = K[0]; cm.cx = K[2]; cm.fy = K[4]; cm.cy = K[5];
        cams[i] = cm;
    }
}

// Row n of a [N,4] tensor.  One 16-byte access when the base allows it: as views of a flat [15 N]
// buffer (mapping.FlatSplats, the flat gradient of project_bwd) the quaternion rows start 12 N bytes
// in, which is 16-byte aligned only when N is a multiple of 4.
__device__ __forceinline__ float4 load_row4(const float* base, int n) {
    if ((reinterpret_cast<uintptr_t>(base) & 15) == 0) return reinterpret_cast<const float4*>(base)[n];
    return make_float4(base[n * 4], base[n * 4 + 1], base[n * 4 + 2], base[n * 4 + 3]);
}
__device__ __forceinline__ void store_row4(float* base, int n, float4 v) {
    if ((reinterpret_cast<uintptr_t>(base) & 15) == 0) { reinterpret_cast<float4*>(base)[n] = v; return; }
    base[n * 4] = v.x; base[n * 4 + 1] = v.y; base[n * 4 + 2] = v.z; base[n * 4 + 3] = v.w;
}

// Rotation of the normalised quaternion (wxyz), row-major.
__device__ __forceinline__ void quat_to_R(float w, float x, float y, float z, float* R) {
    R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
    R[3] = 2.f * (x * y + w * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - w * x);
    R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
}

// S = A * B for row-major 3x3.
__device__ __forceinline__ void mat3_mul(const float* A, const float* B, float* S) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            S[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}

// S = A * B^T.
__device__ __forceinline__ void mat3_mul_bt(const float* A, const float* B, float* S) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            S[i * 3 + j] = A[i * 3] * B[j * 3] + A[i * 3 + 1] * B[j * 3 + 1] + A[i * 3 + 2] * B[j * 3 + 2];
}

// S = A^T * B.
__device__ __forceinline__ void mat3_mul_at(const float* A, const float* B, float* S) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            S[i * 3 + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}

struct Proj {
    float x, y, z;          // camera-space mean
    float rz, tx, ty;       // 1/z and the clamped x,y used in the Jacobian
    bool clamp_x, clamp_y;
    float Sc[9];            // camera-space covariance
    float c00, c01, c11;    // blurred 2-D covariance
    float det;
    float mx, my;
    float radius;
    bool ok;
};

__device__ __forceinline__ void project_one(const Cam& cm, const float* mu, const float* Sig, int width,
                                            int height, float eps2d, float near_plane, float far_plane,
                                            float radius_clip, Proj& p) {
    p.ok = false;
    p.x = cm.R[0] * mu[0] + cm.R[1] * mu[1] + cm.R[2] * mu[2] + cm.t[0];
    p.y = cm.R[3] * mu[0] + cm.R[4] * mu[1] + cm.R[5] * mu[2] + cm.t[1];
    p.z = cm.R[6] * mu[0] + cm.R[7] * mu[1] + cm.R[8] * mu[2] + cm.t[2];
    if (!(p.z >= near_plane && p.z <= far_plane)) return;
    float tmp[9];
    mat3_mul(cm.R, Sig, tmp);
    mat3_mul_bt(tmp, cm.R, p.Sc);
    float tanx = 0.5f * (float)width / cm.fx, tany = 0.5f * (float)height / cm.fy;
    float lim_xp = ((float)width - cm.cx) / cm.fx + 0.3f * tanx;
    float lim_xn = cm.cx / cm.fx + 0.3f * tanx;
    float lim_yp = ((float)height - cm.cy) / cm.fy + 0.3f * tany;
    float lim_yn = cm.cy / cm.fy + 0.3f * tany;
    p.rz = 1.0f / p.z;
    float xz = p.x * p.rz, yz = p.y * p.rz;
    p.clamp_x = !(xz <= lim_xp && xz >= -lim_xn);
    p.clamp_y = !(yz <= lim_yp && yz >= -lim_yn);
    p.tx = p.z * fminf(lim_xp, fmaxf(-lim_xn, xz));
    p.ty = p.z * fminf(lim_yp, fmaxf(-lim_yn, yz));
    float rz2 = p.rz * p.rz;
    float a = cm.fx * p.rz, b = -cm.fx * p.tx * rz2, d = cm.fy * p.rz, e = -cm.fy * p.ty * rz2;
    const float* S = p.Sc;
    p.c00 = a * a * S[0] + 2.f * a * b * S[2] + b * b * S[8] + eps2d;
    p.c01 = a * d * S[1] + a * e * S[2] + b * d * S[5] + b * e * S[8];
    p.c11 = d * d * S[4] + 2.f * d * e * S[5] + e * e * S[8] + eps2d;
    p.det = p.c00 * p.c11 - p.c01 * p.c01;
    if (!(p.det > 0.f)) return;
    p.mx = cm.fx * p.x * p.rz + cm.cx;
    p.my = cm.fy * p.y * p.rz + cm.cy;
    float bb = 0.5f * (p.c00 + p.c11);
    float v1 = bb + sqrtf(fmaxf(0.01f, bb * bb - p.det));
    p.radius = ceilf(3.f * sqrtf(v1));
    if (p.radius <= radius_clip) return;
    if (p.mx + p.radius <= 0.f || p.mx - p.radius >= (float)width || p.my + p.radius <= 0.f ||
        p.my - p.radius >= (float)height)
        return;
    p.ok = true;
}

__device__ __forceinline__ void build_sigma(const float* q, const float* ls, float* Rq, float* s,
                                            float* Sig, float& qinv) {
    qinv = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    quat_to_R(q[0] * qinv, q[1] * qinv, q[2] * qinv, q[3] * qinv, Rq);
    s[0] = expf(ls[0]); s[1] = expf(ls[1]); s[2] = expf(ls[2]);
    float M[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) M[i * 3 + j] = Rq[i * 3 + j] * s[j];
    mat3_mul_bt(M, M, Sig);
}

#ifndef PF_THREADS
#define PF_THREADS 64       // measured 23.1 us (64) / 24.8 (128) / 24.6 (256) at 200 k Gaussians: small CTAs shorten the tail
#endif
#ifdef PF_MINB
#define PF_BOUNDS __launch_bounds__(PF_THREADS, PF_MINB)
#else
#define PF_BOUNDS __launch_bounds__(PF_THREADS)     // (a second argument, even 1, makes ptxas spend 12 more registers)
#endif
__global__ void PF_BOUNDS
project_fwd_kernel(const float* __restrict__ means, const float* __restrict__ quats,
                   const float* __restrict__ log_scales, const float* __restrict__ logit_opac,
                   const float* __restrict__ logit_colors, const float* __restrict__ log_unc,
                   const float* __restrict__ viewmats, const float* __restrict__ Ks, int C, int N,
                   int width, int height, float eps2d, float near_plane, float far_plane,
                   float radius_clip, int tile_w, int tile_h, int32_t* __restrict__ radii,
                   float* __restrict__ means2d, float* __restrict__ rec,
                   int32_t* __restrict__ tiles_per_gauss, int32_t* __restrict__ tile_counts,
                   const float* __restrict__ sh_coeffs, int sh_degree, int sh_stride) {
    __shared__ Cam cams[MAX_CAMS_SMEM];
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    float mu[3], q[4], ls[3], Rq[9], s[3], Sig[9], qinv;
    float opac = 0.f, col[3] = {0.f, 0.f, 0.f}, beta = 0.f;
    if (n < N) {
        mu[0] = means[n * 3]; mu[1] = means[n * 3 + 1]; mu[2] = means[n * 3 + 2];
        float4 qq = load_row4(quats, n);
        q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
        ls[0] = log_scales[n * 3]; ls[1] = log_scales[n * 3 + 1]; ls[2] = log_scales[n * 3 + 2];
        build_sigma(q, ls, Rq, s, Sig, qinv);
        opac = sigmoidf_(logit_opac[n]);
        if (!sh_coeffs) {
            col[0] = sigmoidf_(logit_colors[n * 3]);
            col[1] = sigmoidf_(logit_colors[n * 3 + 1]);
            col[2] = sigmoidf_(logit_colors[n * 3 + 2]);
        }
        if (log_unc) beta = fmaxf(expf(log_unc[n]), 0.01f);
    }
    for (int c0 = 0; c0 < C; c0 += MAX_CAMS_SMEM) {
        int nc = min(MAX_CAMS_SMEM, C - c0);
        __syncthreads();
        load_cams(cams, viewmats, Ks, c0, nc);
        __syncthreads();
        for (int ci = 0; ci < nc; ++ci) {
            size_t idx = (size_t)(c0 + ci) * N + n;
            Proj p;
            p.ok = false;
            if (n < N)
                project_one(cams[ci], mu, Sig, width, height, eps2d, near_plane, far_plane, radius_clip, p);
            float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
            int rad = 0, cnt = 0, x0 = 0, x1 = 0, y0 = 0, y1 = 0;
            float2 m2 = make_float2(0.f, 0.f);
            if (p.ok) {
                if (sh_coeffs) {    // view-dependent colour (extension; the reference path is the sigmoid above)
                    float centre[3];
                    cam_centre(cams[ci].R, cams[ci].t, centre);
                    sh_colour(sh_degree, sh_coeffs + (size_t)n * sh_stride * 3, sh_num_coeffs(sh_degree), mu, centre, col);
                }
                float idet = 1.0f / p.det;
                rad = (int)p.radius;
                m2 = make_float2(p.mx, p.my);
                r0 = make_float4(p.mx, p.my, opac, p.c11 * idet);
                r1 = make_float4(-p.c01 * idet, p.c00 * idet, col[0], col[1]);
                r2 = make_float4(col[2], p.z, beta, 0.f);
                tile_bbox(p.mx, p.my, rad, tile_w, tile_h, x0, x1, y0, y1);
                cnt = (x1 - x0) * (y1 - y0);
            }
            if (n < N) {
                radii[idx] = rad;
                tiles_per_gauss[idx] = cnt;
                reinterpret_cast<float2*>(means2d)[idx] = m2;
                float4* rp = reinterpret_cast<float4*>(rec) + idx * 3;
                rp[0] = r0; rp[1] = r1; rp[2] = r2;
            }
            if (tile_counts) {
                int32_t* tc = tile_counts + (size_t)(c0 + ci) * tile_w * tile_h * GSB_CSTRIDE;
                warp_for_tiles(x0, y0, x1 - x0, cnt, [](int, bool) {}, [&](int tx, int ty) {
                    atomicAdd(tc + (size_t)(ty * tile_w + tx) * GSB_CSTRIDE, 1);
                });
            }
        }
    }
}

GSB_API int gsb_project_fwd(const float* means, const float* quats, const float* log_scales,
                            const float* logit_opacities, const float* logit_colors,
                            const float* log_unc, const float* viewmats, const float* Ks, int C,
                            int N, int width, int height, float eps2d, float near_plane,
                            float far_plane, float radius_clip, int tile_w, int tile_h,
                            int32_t* radii, float* means2d, float* rec, int32_t* tiles_per_gauss,
                            int32_t* tile_counts, const float* sh_coeffs, int sh_degree, int sh_stride,
                            void* stream) {
    GSB_CHECK_ARG(C >= 0 && N >= 0 && width > 0 && height > 0);
    if (C == 0 || N == 0) return 0;
    GSB_CHECK_ARG(means && quats && log_scales && logit_opacities && viewmats && Ks);
    GSB_CHECK_ARG(logit_colors || sh_coeffs);
    GSB_CHECK_ARG(!sh_coeffs || (sh_degree >= 0 && sh_degree <= 3 && sh_stride >= (sh_degree + 1) * (sh_degree + 1)));
    GSB_CHECK_ARG(radii && means2d && rec && tiles_per_gauss);
    int blocks = (N + PF_THREADS - 1) / PF_THREADS;
    project_fwd_kernel<<<blocks, PF_THREADS, 0, gsb_stream(stream)>>>(
        means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks, C, N, width,
        height, eps2d, near_plane, far_plane, radius_clip, tile_w, tile_h, radii, means2d, rec,
        tiles_per_gauss, tile_counts, sh_coeffs, sh_degree, sh_stride);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// Backward.  v_rec slots: [2] opacity [3..5] conic [6..8] rgb [9] depth [10] beta;
// the x,y gradient comes from v_means2d (reference: out.means2d.retain_grad(),
// gslam/backend.py:326, makes means2d a separate autograd edge).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
project_bwd_kernel(const float* __restrict__ means, const float* __restrict__ quats,
                   const float* __restrict__ log_scales, const float* __restrict__ logit_opac,
                   const float* __restrict__ logit_colors, const float* __restrict__ log_unc,
                   const float* __restrict__ viewmats, const float* __restrict__ Ks, int C, int N,
                   int width, int height, float eps2d, const int32_t* __restrict__ radii,
                   const float* __restrict__ v_means2d, int64_t v_m2_stride,
                   const float* __restrict__ v_rec, int need_gauss, float* __restrict__ v_means,
                   float* __restrict__ v_quats, float* __restrict__ v_log_scales,
                   float* __restrict__ v_logit_opac, float* __restrict__ v_logit_colors,
                   float* __restrict__ v_log_unc, float* __restrict__ v_viewmats,
                   float* __restrict__ det_partials, int64_t row_stride, float* __restrict__ v_vis) {
    __shared__ Cam cams[MAX_CAMS_SMEM];
    __shared__ float red[8][12];
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float mu[3] = {0.f, 0.f, 0.f}, q[4], ls[3], Rq[9], s[3], Sig[9], qinv = 1.f;
    if (n < N) {
        mu[0] = means[n * 3]; mu[1] = means[n * 3 + 1]; mu[2] = means[n * 3 + 2];
        float4 qq = load_row4(quats, n);
        q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
        ls[0] = log_scales[n * 3]; ls[1] = log_scales[n * 3 + 1]; ls[2] = log_scales[n * 3 + 2];
        build_sigma(q, ls, Rq, s, Sig, qinv);
    }
    float v_mu[3] = {0.f, 0.f, 0.f}, v_Sig[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float v_opac = 0.f, v_col[3] = {0.f, 0.f, 0.f}, v_beta = 0.f;
    int n_vis = 0;          // views in which this Gaussian is visible (radii > 0)

    for (int c0 = 0; c0 < C; c0 += MAX_CAMS_SMEM) {
        int nc = min(MAX_CAMS_SMEM, C - c0);
        __syncthreads();
        load_cams(cams, viewmats, Ks, c0, nc);
        __syncthreads();
        for (int ci = 0; ci < nc; ++ci) {
            const Cam& cm = cams[ci];
            size_t idx = (size_t)(c0 + ci) * N + n;
            float vR[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, vt[3] = {0.f, 0.f, 0.f};
            bool active = (n < N) && radii[idx] > 0;
            n_vis += active ? 1 : 0;
            if (active) {
                Proj p;
                project_one(cm, mu, Sig, width, height, eps2d, -1e30f, 1e30f, -1.f, p);
                // culls were decided in the forward pass (radii > 0); only det can still fail here
                active = p.det > 0.f;
                if (active) {
                    const float4* vr = reinterpret_cast<const float4*>(v_rec) + idx * 3;
                    float4 g0 = vr[0], g1 = vr[1], g2 = vr[2];
                    float vmx = v_means2d[idx * v_m2_stride], vmy = v_means2d[idx * v_m2_stride + 1];
                    float vA = g0.w, vB = g1.x, vC = g1.y, v_depth = g2.y;
                    v_opac += g0.z;
                    v_col[0] += g1.z; v_col[1] += g1.w; v_col[2] += g2.x;
                    v_beta += g2.z;
                    float idet = 1.0f / p.det;
                    float A = p.c11 * idet, B = -p.c01 * idet, Cc = p.c00 * idet;
                    float h = 0.5f * vB;
                    float T00 = vA * A + h * B, T01 = vA * B + h * Cc;
                    float T10 = h * A + vC * B, T11 = h * B + vC * Cc;
                    float G00 = -(A * T00 + B * T10), G01 = -(A * T01 + B * T11);
                    float G11 = -(B * T01 + Cc * T11);
                    float rz = p.rz, rz2 = rz * rz, rz3 = rz2 * rz;
                    float a = cm.fx * rz, b = -cm.fx * p.tx * rz2, d = cm.fy * rz, e = -cm.fy * p.ty * rz2;
                    // v_Sc = J^T G J
                    float JtG[6] = {a * G00, a * G01, d * G01, d * G11, b * G00 + e * G01, b * G01 + e * G11};
                    float J0[3] = {a, 0.f, b}, J1[3] = {0.f, d, e};
                    float vSc[9];
#pragma unroll
                    for (int i = 0; i < 3; ++i)
#pragma unroll
                        for (int j = 0; j < 3; ++j) vSc[i * 3 + j] = JtG[i * 2] * J0[j] + JtG[i * 2 + 1] * J1[j];
                    // v_J = 2 G J Sc
                    const float* S = p.Sc;
                    float W0[3], W1[3];
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        W0[j] = a * S[j] + b * S[6 + j];
                        W1[j] = d * S[3 + j] + e * S[6 + j];
                    }
                    float v_a = 2.f * (G00 * W0[0] + G01 * W1[0]);
                    float v_b = 2.f * (G00 * W0[2] + G01 * W1[2]);
                    float v_d = 2.f * (G01 * W0[1] + G11 * W1[1]);
                    float v_e = 2.f * (G01 * W0[2] + G11 * W1[2]);
                    float vp[3];
                    vp[0] = cm.fx * rz * vmx;
                    vp[1] = cm.fy * rz * vmy;
                    vp[2] = -(cm.fx * p.x * vmx + cm.fy * p.y * vmy) * rz2 + v_depth;
                    if (!p.clamp_x) vp[0] += -cm.fx * rz2 * v_b; else vp[2] += -cm.fx * rz3 * v_b * p.tx;
                    if (!p.clamp_y) vp[1] += -cm.fy * rz2 * v_e; else vp[2] += -cm.fy * rz3 * v_e * p.ty;
                    vp[2] += -cm.fx * rz2 * v_a - cm.fy * rz2 * v_d + 2.f * cm.fx * p.tx * rz3 * v_b +
                             2.f * cm.fy * p.ty * rz3 * v_e;
                    // to world space
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        v_mu[j] += cm.R[j] * vp[0] + cm.R[3 + j] * vp[1] + cm.R[6 + j] * vp[2];
                    float RS[9], t2[9];
                    mat3_mul(cm.R, Sig, RS);        // Rc * Sigma
                    mat3_mul(vSc, RS, t2);          // v_Sc * Rc * Sigma
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
#pragma unroll
                        for (int j = 0; j < 3; ++j) vR[i * 3 + j] = vp[i] * mu[j] + 2.f * t2[i * 3 + j];
                        vt[i] = vp[i];
                    }
                    if (need_gauss) {
                        float t3[9], t4[9];
                        mat3_mul_at(cm.R, vSc, t3);     // Rc^T v_Sc
                        mat3_mul(t3, cm.R, t4);         // Rc^T v_Sc Rc
#pragma unroll
                        for (int k = 0; k < 9; ++k) v_Sig[k] += t4[k];
                    }
                }
            }
            if (v_viewmats) {
                // block reduction of the 12 pose-gradient entries of this camera
                bool any = __any_sync(0xffffffffu, active);
                if (any) {
#pragma unroll
                    for (int k = 0; k < 9; ++k) vR[k] = warp_sum(vR[k]);
#pragma unroll
                    for (int k = 0; k < 3; ++k) vt[k] = warp_sum(vt[k]);
                }
                if (lane == 0) {
#pragma unroll
                    for (int k = 0; k < 9; ++k) red[warp][k] = vR[k];
#pragma unroll
                    for (int k = 0; k < 3; ++k) red[warp][9 + k] = vt[k];
                }
                __syncthreads();
                if (threadIdx.x < 12) {
                    float acc = 0.f;
#pragma unroll
                    for (int w = 0; w < 8; ++w) acc += red[w][threadIdx.x];
                    int k = threadIdx.x;
                    int slot = (k < 9) ? (k / 3) * 4 + (k % 3) : (k - 9) * 4 + 3;
                    if (det_partials)       // deterministic mode: per-block partials, summed in a fixed order below
                        det_partials[((size_t)(c0 + ci) * 12 + k) * gridDim.x + blockIdx.x] = acc;
                    else if (acc != 0.f) atomicAdd(v_viewmats + (size_t)(c0 + ci) * 16 + slot, acc);
                }
                __syncthreads();
            }
        }
    }
    if (!need_gauss || n >= N) return;
    // Sigma = M M^T, M = Rq diag(s)
    float M[9], vM[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) M[i * 3 + j] = Rq[i * 3 + j] * s[j];
    float vS2[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) vS2[i * 3 + j] = v_Sig[i * 3 + j] + v_Sig[j * 3 + i];
    mat3_mul(vS2, M, vM);
    float vRq[9], v_ls[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        float vs = Rq[j] * vM[j] + Rq[3 + j] * vM[3 + j] + Rq[6 + j] * vM[6 + j];
        v_ls[j] = vs * s[j];
#pragma unroll
        for (int i = 0; i < 3; ++i) vRq[i * 3 + j] = vM[i * 3 + j] * s[j];
    }
    float w = q[0] * qinv, x = q[1] * qinv, y = q[2] * qinv, z = q[3] * qinv;
    float vqn[4];
    vqn[0] = 2.f * (-z * vRq[1] + y * vRq[2] + z * vRq[3] - x * vRq[5] - y * vRq[6] + x * vRq[7]);
    vqn[1] = 2.f * (y * vRq[1] + z * vRq[2] + y * vRq[3] - 2.f * x * vRq[4] - w * vRq[5] + z * vRq[6] +
                    w * vRq[7] - 2.f * x * vRq[8]);
    vqn[2] = 2.f * (-2.f * y * vRq[0] + x * vRq[1] + w * vRq[2] + x * vRq[3] + z * vRq[5] - w * vRq[6] +
                    z * vRq[7] - 2.f * y * vRq[8]);
    vqn[3] = 2.f * (-2.f * z * vRq[0] - w * vRq[1] + x * vRq[2] + w * vRq[3] - 2.f * z * vRq[4] +
                    y * vRq[5] + x * vRq[6] + y * vRq[7]);
    float dotp = vqn[0] * w + vqn[1] * x + vqn[2] * y + vqn[3] * z;
    float4 vq = make_float4((vqn[0] - dotp * w) * qinv, (vqn[1] - dotp * x) * qinv,
                            (vqn[2] - dotp * y) * qinv, (vqn[3] - dotp * z) * qinv);
    // Output addressing: natural ([N,3], [N,4], ... one array per parameter) when row_stride == 0, else
    // every pointer is the first column of its parameter inside ONE row-major [N, row_stride] gradient
    // record (the mapping step's layout: a Gaussian's whole gradient is one contiguous row, so the
    // rows of a Gaussian range are one contiguous block for a reduce-scatter).
    const int64_t sm = row_stride ? row_stride : 3, sq = row_stride ? row_stride : 4, so = row_stride ? row_stride : 1;
    v_means[n * sm] = v_mu[0]; v_means[n * sm + 1] = v_mu[1]; v_means[n * sm + 2] = v_mu[2];
    if (row_stride) *reinterpret_cast<float4*>(v_quats + n * sq) = vq;     // 16-byte aligned by contract
    else store_row4(v_quats, n, vq);
    v_log_scales[n * sm] = v_ls[0]; v_log_scales[n * sm + 1] = v_ls[1]; v_log_scales[n * sm + 2] = v_ls[2];
    float o = sigmoidf_(logit_opac[n]);
    v_logit_opac[n * so] = v_opac * o * (1.f - o);
    if (logit_colors) {     // NULL with SH colours: gsb_sh_bwd owns that gradient
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float cc = sigmoidf_(logit_colors[n * 3 + k]);
            v_logit_colors[n * sm + k] = v_col[k] * cc * (1.f - cc);
        }
    }
    if (v_log_unc) {
        float gu = 0.f;
        if (log_unc) {
            float bexp = expf(log_unc[n]);
            gu = (bexp >= 0.01f) ? v_beta * bexp : 0.f;
        }
        if (log_unc || row_stride) v_log_unc[n * so] = gu;
    }
    if (v_vis) v_vis[n * so] = (float)n_vis;
}

// Fixed-order sum of the per-block pose-gradient partials [C*12][n_blocks]: one warp per entry, lanes
// stride over the blocks, a fixed butterfly joins them.  grid C, block 12 * 32.
__global__ void __launch_bounds__(384)
viewmat_partials_sum_kernel(const float* __restrict__ partials, int n_blocks, float* __restrict__ v_viewmats) {
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31, cam = blockIdx.x;
    const float* p = partials + ((size_t)cam * 12 + k) * n_blocks;
    float s = 0.f;
    for (int b = lane; b < n_blocks; b += 32) s += p[b];
    s = warp_sum(s);
    const int slot = (k < 9) ? (k / 3) * 4 + (k % 3) : (k - 9) * 4 + 3;
    if (lane == 0) v_viewmats[(size_t)cam * 16 + slot] = s;
}

GSB_API size_t gsb_project_bwd_det_floats(int C, int N) {
    return (size_t)C * 12 * (size_t)((N + 255) / 256);
}

GSB_API int gsb_project_bwd(const float* means, const float* quats, const float* log_scales,
                            const float* logit_opacities, const float* logit_colors,
                            const float* log_unc, const float* viewmats, const float* Ks, int C,
                            int N, int width, int height, float eps2d, const int32_t* radii,
                            const float* rec, const float* v_means2d, int64_t v_means2d_stride,
                            const float* v_rec, int need_gauss_grads, float* v_means, float* v_quats,
                            float* v_log_scales, float* v_logit_opacities, float* v_logit_colors,
                            float* v_log_unc, float* v_viewmats, float* det_partials,
                            int64_t row_stride, float* v_vis, void* stream) {
    (void)rec;
    GSB_CHECK_ARG(row_stride == 0 || (row_stride % 4 == 0 && need_gauss_grads &&
                                      (reinterpret_cast<uintptr_t>(v_quats) & 15) == 0));
    GSB_CHECK_ARG(!v_vis || need_gauss_grads);
    GSB_CHECK_ARG(C >= 0 && N >= 0);
    if (C == 0 || N == 0) return 0;
    GSB_CHECK_ARG(means && quats && log_scales && logit_opacities && viewmats && Ks);
    GSB_CHECK_ARG(radii && v_means2d && v_rec && v_means2d_stride >= 2);
    if (need_gauss_grads)   // logit_colors == NULL: SH colours, their gradient comes from gsb_sh_bwd
        GSB_CHECK_ARG(v_means && v_quats && v_log_scales && v_logit_opacities && (v_logit_colors || !logit_colors));
    int blocks = (N + 255) / 256;
    project_bwd_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(
        means, quats, log_scales, logit_opacities, logit_colors, log_unc, viewmats, Ks, C, N, width,
        height, eps2d, radii, v_means2d, v_means2d_stride, v_rec, need_gauss_grads, v_means, v_quats,
        v_log_scales, v_logit_opacities, v_logit_colors, v_log_unc, v_viewmats,
        v_viewmats ? det_partials : nullptr, row_stride, v_vis);
    GSB_LAUNCH_CHECK();
    if (v_viewmats && det_partials) {
        viewmat_partials_sum_kernel<<<C, 384, 0, gsb_stream(stream)>>>(det_partials, blocks, v_viewmats);
        GSB_LAUNCH_CHECK();
    }
    return 0;
}

// ---------------------------------------------------------------------------
// K3: compaction of the visible rows for the packed projection
// (reference get_new_splat_depth, gslam/rasterization.py:390-419).
// ---------------------------------------------------------------------------
__global__ void pack_visible_kernel(const int32_t* __restrict__ radii, const float* __restrict__ means2d,
                                    const float* __restrict__ rec, const int64_t* __restrict__ cum,
                                    int64_t CN, int N, int64_t* __restrict__ camera_ids,
                                    int64_t* __restrict__ gaussian_ids, int32_t* __restrict__ radii_out,
                                    float* __restrict__ means2d_out, float* __restrict__ depths_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    if (r <= 0) return;
    int64_t o = cum[i] - 1;
    camera_ids[o] = i / N;
    gaussian_ids[o] = i % N;
    radii_out[o] = r;
    means2d_out[o * 2] = means2d[i * 2];
    means2d_out[o * 2 + 1] = means2d[i * 2 + 1];
    depths_out[o] = rec[i * GSB_REC + 9];
}

GSB_API int gsb_pack_visible(const int32_t* radii, const float* means2d, const float* rec,
                             const int64_t* cum, int C, int N, int64_t* camera_ids,
                             int64_t* gaussian_ids, int32_t* radii_out, float* means2d_out,
                             float* depths_out, void* stream) {
    int64_t CN = (int64_t)C * N;
    if (CN == 0) return 0;
    GSB_CHECK_ARG(radii && means2d && rec && cum);
    int blocks = (int)((CN + 255) / 256);
    pack_visible_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(radii, means2d, rec, cum, CN, N,
                                                                camera_ids, gaussian_ids, radii_out,
                                                                means2d_out, depths_out);
    GSB_LAUNCH_CHECK();
    return 0;
}
