// K9/K10: fused SSIM and photometric loss.
//
// K9 replaces fused_ssim (rahul-goel/fused-ssim@30fb258) as called at reference
// gslam/backend.py:303-307: 11-tap sigma=1.5 separable Gaussian window with zero
// padding, C1 = 0.01^2, C2 = 0.03^2, per channel; padding='valid' averages the
// map without its 5-pixel border.  Images are addressed through element strides
// so the reference's `rgbs.permute(0, 3, 1, 2)` view needs no copy.
//
// K10 replaces the exposure-affine + active-NeRF terms (gslam/backend.py:273-283,
// gslam/frontend.py:121-129): because every term is a mean over pixels the
// gradient is local, so ONE pass produces the loss sums and the gradients.
#include "common.cuh"

__constant__ float G11[11] = {1.0283800845e-03f, 7.5987581352e-03f, 3.6000772128e-02f, 1.0936068951e-01f,
                              2.1300553771e-01f, 2.6601172486e-01f, 2.1300553771e-01f, 1.0936068951e-01f,
                              3.6000772128e-02f, 7.5987581352e-03f, 1.0283800845e-03f};
#define SSIM_C1 (0.01f * 0.01f)
#define SSIM_C2 (0.03f * 0.03f)
// Tiling of the two SSIM kernels: a 256-thread CTA produces a 32 x 32 tile of one (image, channel)
// plane.  The 11-tap window is separable; both passes are register-blocked so that a thread loads
// 14 neighbours once for FOUR outputs (horizontal: 4 consecutive columns of one row; vertical: 4
// consecutive rows of one column) -- 3.5 shared-memory loads per tap-sum instead of 11.  Row
// strides are = 1 (mod 4) words, which makes the 4-row x 8-group access pattern of a warp hit 32
// different banks.
#define SB 32                   // tile edge
#define SH (SB + 10)            // tile + halo
#define SXS 45                  // row stride of the input tiles  (>= SH, = 1 mod 4)
#define HCS 33                  // row stride of the row-filtered tiles (>= SB, = 1 mod 4)

struct Strides4 {
    int64_t n, c, h, w;
};

__device__ __forceinline__ float block_sum_256(float v, float* red) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x < 8) s = red[threadIdx.x];
    if (threadIdx.x < 32) {
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
    }
    return s;   // valid in thread 0
}

// grid (ceil(W/32), ceil(H/32), N*C), block 256
__global__ void __launch_bounds__(256)
ssim_fwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, Strides4 s1, Strides4 s2,
                int C, int H, int W, int valid, float* __restrict__ ssim_sum,
                float* __restrict__ dm_dmu1, float* __restrict__ dm_ds1, float* __restrict__ dm_ds12,
                float* __restrict__ det_partials) {
    __shared__ float sx[SH][SXS], sy[SH][SXS];
    __shared__ float hc[5][SH][HCS];
    __shared__ float red[8];
    const int nc = blockIdx.z, n = nc / C, c = nc % C;
    const int tid = threadIdx.x;
    const int h0 = blockIdx.y * SB, w0 = blockIdx.x * SB;
    const float* p1 = img1 + n * s1.n + c * s1.c;
    const float* p2 = img2 + n * s2.n + c * s2.c;
    for (int i = tid; i < SH * SH; i += 256) {
        const int r = i / SH, q = i - r * SH;
        const int gh = h0 + r - 5, gw = w0 + q - 5;
        const bool in = gh >= 0 && gh < H && gw >= 0 && gw < W;
        sx[r][q] = in ? p1[gh * s1.h + gw * s1.w] : 0.f;
        sy[r][q] = in ? p2[gh * s2.h + gw * s2.w] : 0.f;
    }
    __syncthreads();
    // horizontal pass: item = (row r, group of 4 columns)
    for (int i = tid; i < SH * (SB / 4); i += 256) {
        const int r = i / (SB / 4), q0 = (i - r * (SB / 4)) * 4;
        float m1[4] = {0.f, 0.f, 0.f, 0.f}, m2[4] = {0.f, 0.f, 0.f, 0.f}, xx[4] = {0.f, 0.f, 0.f, 0.f};
        float yy[4] = {0.f, 0.f, 0.f, 0.f}, xy[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float x = sx[r][q0 + k], y = sy[r][q0 + k];
            const float x2 = x * x, y2 = y * y, xyv = x * y;
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int tap = k - o;          // compile-time after unrolling
                if (tap >= 0 && tap < 11) {
                    const float w = G11[tap];
                    m1[o] = fmaf(w, x, m1[o]); m2[o] = fmaf(w, y, m2[o]);
                    xx[o] = fmaf(w, x2, xx[o]); yy[o] = fmaf(w, y2, yy[o]); xy[o] = fmaf(w, xyv, xy[o]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            hc[0][r][q0 + o] = m1[o]; hc[1][r][q0 + o] = m2[o]; hc[2][r][q0 + o] = xx[o];
            hc[3][r][q0 + o] = yy[o]; hc[4][r][q0 + o] = xy[o];
        }
    }
    __syncthreads();
    // vertical pass: thread = (column tx, group of 4 rows)
    const int tx = tid & 31, ty0 = (tid >> 5) * 4;
    float acc[5][4];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[m][o] = 0.f;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float v = hc[m][ty0 + k][tx];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int tap = k - o;
                if (tap >= 0 && tap < 11) acc[m][o] = fmaf(G11[tap], v, acc[m][o]);
            }
        }
    }
    float part = 0.f;
    const int w = w0 + tx;
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int h = h0 + ty0 + o;
        const float mu1 = acc[0][o], mu2 = acc[1][o], e11 = acc[2][o], e22 = acc[3][o], e12 = acc[4][o];
        const bool inside = h < H && w < W;
        const bool counted = inside && (!valid || (h >= 5 && h < H - 5 && w >= 5 && w < W - 5));
        const float sg1 = e11 - mu1 * mu1, sg2 = e22 - mu2 * mu2, sg12 = e12 - mu1 * mu2;
        const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * sg12 + SSIM_C2;
        const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = sg1 + sg2 + SSIM_C2;
        const float inv = 1.0f / (B1 * B2);
        const float m = A1 * A2 * inv;
        if (dm_dmu1 && inside) {
            // m as a function of (mu1, E[x^2], E[xy]) with img2 fixed
            float d_mu1 = 0.f, d_s1 = 0.f, d_s12 = 0.f;
            if (counted) {
                d_mu1 = (2.f * mu2 * (A2 - A1) * inv) - m * 2.f * mu1 * (B2 - B1) * inv;
                d_s1 = -m / B2;
                d_s12 = 2.f * A1 * inv;
            }
            const size_t o_ = ((size_t)nc * H + h) * W + w;
            dm_dmu1[o_] = d_mu1; dm_ds1[o_] = d_s1; dm_ds12[o_] = d_s12;
        }
        part += counted ? m : 0.f;
    }
    const float tot = block_sum_256(part, red);
    if (tid == 0) {
        if (det_partials)   // deterministic mode: summed in a fixed order by det_sum_kernel
            det_partials[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = tot;
        else if (tot != 0.f) atomicAdd(ssim_sum, tot);
    }
}

// dst[0] = sum of p[0..n) in a fixed order (one CTA of 256: strided per-thread sums, then the block tree)
__global__ void __launch_bounds__(256) det_sum_kernel(const float* __restrict__ p, int64_t n, float* __restrict__ dst) {
    __shared__ float red[8];
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < n; i += 256) s += p[i];
    s = block_sum_256(s, red);
    if (threadIdx.x == 0) dst[0] = s;
}

int gsb_det_sum_launch(const float* partials, int64_t n, float* dst, cudaStream_t st) {
    det_sum_kernel<<<1, 256, 0, st>>>(partials, n, dst);
    GSB_LAUNCH_CHECK();
    return 0;
}

// v_img1[p] (+)= scale * sum_q w(q-p) [ dmu1(q) + 2 x(p) ds1(q) + y(p) ds12(q) ],  scale = *grad * coef
__global__ void __launch_bounds__(256)
ssim_bwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, Strides4 s1, Strides4 s2,
                int C, int H, int W, const float* __restrict__ dm_dmu1, const float* __restrict__ dm_ds1,
                const float* __restrict__ dm_ds12, const float* __restrict__ grad, float coef,
                float* __restrict__ v_img1, Strides4 so, int accumulate) {
    __shared__ float sa[3][SH][SXS];
    __shared__ float hc[3][SH][HCS];
    const int nc = blockIdx.z, n = nc / C, c = nc % C;
    const int tid = threadIdx.x;
    const int h0 = blockIdx.y * SB, w0 = blockIdx.x * SB;
    const size_t plane = (size_t)nc * H * W;
    for (int i = tid; i < SH * SH; i += 256) {
        const int r = i / SH, q = i - r * SH;
        const int gh = h0 + r - 5, gw = w0 + q - 5;
        const bool in = gh >= 0 && gh < H && gw >= 0 && gw < W;
        const size_t o = plane + (size_t)gh * W + gw;
        sa[0][r][q] = in ? dm_dmu1[o] : 0.f;
        sa[1][r][q] = in ? dm_ds1[o] : 0.f;
        sa[2][r][q] = in ? dm_ds12[o] : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < SH * (SB / 4); i += 256) {
        const int r = i / (SB / 4), q0 = (i - r * (SB / 4)) * 4;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 14; ++k) {
                const float v = sa[m][r][q0 + k];
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    const int tap = k - o;
                    if (tap >= 0 && tap < 11) a[o] = fmaf(G11[tap], v, a[o]);
                }
            }
#pragma unroll
            for (int o = 0; o < 4; ++o) hc[m][r][q0 + o] = a[o];
        }
    }
    __syncthreads();
    const int tx = tid & 31, ty0 = (tid >> 5) * 4;
    float acc[3][4];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[m][o] = 0.f;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float v = hc[m][ty0 + k][tx];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int tap = k - o;
                if (tap >= 0 && tap < 11) acc[m][o] = fmaf(G11[tap], v, acc[m][o]);
            }
        }
    }
    const int w = w0 + tx;
    if (w >= W) return;
    const float scale = (grad ? *grad : 1.0f) * coef;
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int h = h0 + ty0 + o;
        if (h >= H) continue;
        const float x = img1[n * s1.n + c * s1.c + h * s1.h + w * s1.w];
        const float y = img2[n * s2.n + c * s2.c + h * s2.h + w * s2.w];
        const float g = scale * (acc[0][o] + 2.f * x * acc[1][o] + y * acc[2][o]);
        float* dst = v_img1 + n * so.n + c * so.c + h * so.h + w * so.w;
        *dst = accumulate ? *dst + g : g;
    }
}

static Strides4 mk(const int64_t* s) { return Strides4{s[0], s[1], s[2], s[3]}; }

GSB_API int gsb_ssim_fwd(const float* img1, const float* img2, const int64_t* strides1_host,
                         const int64_t* strides2_host, int N, int C, int H, int W, int valid,
                         float* ssim_sum, float* dm_dmu1, float* dm_ds1, float* dm_ds12, float* det_partials,
                         void* stream) {
    GSB_CHECK_ARG(N >= 0 && C >= 0 && H > 0 && W > 0);
    if (N * C == 0) return 0;
    GSB_CHECK_ARG(img1 && img2 && strides1_host && strides2_host && ssim_sum && N * C <= 65535);
    GSB_CHECK_ARG((dm_dmu1 == nullptr) == (dm_ds1 == nullptr) && (dm_ds1 == nullptr) == (dm_ds12 == nullptr));
    dim3 grid((W + SB - 1) / SB, (H + SB - 1) / SB, N * C);
    ssim_fwd_kernel<<<grid, 256, 0, gsb_stream(stream)>>>(img1, img2, mk(strides1_host), mk(strides2_host), C, H, W,
                                                          valid, ssim_sum, dm_dmu1, dm_ds1, dm_ds12, det_partials);
    GSB_LAUNCH_CHECK();
    if (det_partials) return gsb_det_sum_launch(det_partials, (int64_t)grid.x * grid.y * grid.z, ssim_sum, gsb_stream(stream));
    return 0;
}

GSB_API size_t gsb_ssim_fwd_det_floats(int N, int C, int H, int W) {
    return (size_t)((W + SB - 1) / SB) * ((H + SB - 1) / SB) * (size_t)N * C;
}

GSB_API int gsb_ssim_bwd(const float* img1, const float* img2, const int64_t* strides1_host,
                         const int64_t* strides2_host, int N, int C, int H, int W, const float* dm_dmu1,
                         const float* dm_ds1, const float* dm_ds12, const float* grad, float coef,
                         float* v_img1, const int64_t* strides_out_host, int accumulate, void* stream) {
    GSB_CHECK_ARG(N >= 0 && C >= 0 && H > 0 && W > 0);
    if (N * C == 0) return 0;
    GSB_CHECK_ARG(img1 && img2 && strides1_host && strides2_host && dm_dmu1 && dm_ds1 && dm_ds12 && v_img1 &&
                  strides_out_host && N * C <= 65535);
    dim3 grid((W + SB - 1) / SB, (H + SB - 1) / SB, N * C);
    ssim_bwd_kernel<<<grid, 256, 0, gsb_stream(stream)>>>(img1, img2, mk(strides1_host), mk(strides2_host), C, H, W,
                                                          dm_dmu1, dm_ds1, dm_ds12, grad, coef, v_img1,
                                                          mk(strides_out_host), accumulate);
    GSB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// K10: photometric loss, value and gradient in one pass.
//   mode 0  mapping active-NeRF NLL   mean(sum_ch e^2 / (2 beta^2)) + mean(0.5 ln^2 beta)
//   mode 1  tracking 'active-nerf'    mean(sum_ch e^2 beta^-2)
//   mode 2  'l1'                      mean over pixels and channels of |e|
//   mode 3  'mse'                     mean over pixels and channels of e^2
// e = rgb * exp(a_c) + b_c - gt (exposure[C,2] = (a, b); NULL = identity).
// grid (min(ceil(H*W/256), ~4 CTAs per SM / C), C), chunks of 256 pixels strided over the CTAs
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
photo_loss_kernel(const float* __restrict__ render, const float* __restrict__ gt,
                  const float* __restrict__ exposure, int mode, int HW, float inv_count, float coef,
                  float* __restrict__ loss_sum, float* __restrict__ loss_out, float* __restrict__ v_render,
                  float* __restrict__ v_exposure, float* __restrict__ det_partials) {
    __shared__ float red[8];
    // the CTA's 256 pixels pass through shared memory so that global traffic is 16-byte and
    // coalesced; a thread then reads its own pixel at stride 5 / 3 floats (conflict-free)
    __shared__ __align__(16) float s_r[256 * GSB_CH];
    __shared__ __align__(16) float s_g[256 * 3];
    const int cam = blockIdx.y;
    float ea = 1.f, eb = 0.f;
    if (exposure) { ea = expf(exposure[cam * 2]); eb = exposure[cam * 2 + 1]; }
    float t1 = 0.f, t2 = 0.f, ga = 0.f, gb = 0.f;
    // a CTA walks several 256-pixel chunks and keeps its sums in registers: the loss sums are ONE
    // address for the whole grid, and a few hundred CTAs queueing on it cost more than the image traffic
    for (int p0 = blockIdx.x * 256; p0 < HW; p0 += gridDim.x * 256) {
    const int p = p0 + threadIdx.x;
    const int n_px = min(256, HW - p0);
    const size_t pix0 = (size_t)cam * HW + p0;
    const bool vec = (HW & 3) == 0 && n_px == 256 &&
                     ((reinterpret_cast<uintptr_t>(render) | reinterpret_cast<uintptr_t>(gt) |
                       reinterpret_cast<uintptr_t>(v_render)) & 15) == 0;
    {
        // pix0 * 20 B and pix0 * 12 B are multiples of 16 when p0 and HW are multiples of 4
        const float* rsrc = render + pix0 * GSB_CH;
        const float* gsrc = gt + pix0 * 3;
        if (vec) {
            for (int i = threadIdx.x; i < 256 * GSB_CH / 4; i += 256)
                reinterpret_cast<float4*>(s_r)[i] = reinterpret_cast<const float4*>(rsrc)[i];
            if (threadIdx.x < 256 * 3 / 4)
                reinterpret_cast<float4*>(s_g)[threadIdx.x] = reinterpret_cast<const float4*>(gsrc)[threadIdx.x];
        } else {
            for (int i = threadIdx.x; i < n_px * GSB_CH; i += 256) s_r[i] = rsrc[i];
            for (int i = threadIdx.x; i < n_px * 3; i += 256) s_g[i] = gsrc[i];
        }
    }
    __syncthreads();
    if (p < HW) {
        float* r = s_r + threadIdx.x * GSB_CH;
        const float* g = s_g + threadIdx.x * 3;
        const float beta = r[4];
        float e[3], vr[3], vbeta = 0.f, se = 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) { e[k] = r[k] * ea + eb - g[k]; se += e[k] * e[k]; }
        if (mode == 0) {
            const float ib2 = 1.0f / (beta * beta), lb = logf(beta);
            t1 += 0.5f * se * ib2; t2 += 0.5f * lb * lb;
#pragma unroll
            for (int k = 0; k < 3; ++k) vr[k] = e[k] * ib2;
            vbeta = (-se * ib2 + lb) / beta;
        } else if (mode == 1) {
            const float ib2 = 1.0f / (beta * beta);
            t1 += se * ib2;
#pragma unroll
            for (int k = 0; k < 3; ++k) vr[k] = 2.f * e[k] * ib2;
            vbeta = -2.f * se * ib2 / beta;
        } else if (mode == 2) {
            t1 += (fabsf(e[0]) + fabsf(e[1]) + fabsf(e[2])) * (1.f / 3.f);
#pragma unroll
            for (int k = 0; k < 3; ++k) vr[k] = (e[k] > 0.f ? 1.f : (e[k] < 0.f ? -1.f : 0.f)) * (1.f / 3.f);
        } else {
            t1 += se * (1.f / 3.f);
#pragma unroll
            for (int k = 0; k < 3; ++k) vr[k] = 2.f * e[k] * (1.f / 3.f);
        }
        const float sc = inv_count * coef;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float rk = r[k];
            r[k] = sc * vr[k] * ea;             // the gradient replaces the pixel in shared memory
            ga += sc * vr[k] * rk * ea;
            gb += sc * vr[k];
        }
        r[3] = 0.f;
        r[4] = sc * vbeta;
    }
    __syncthreads();
    {
        float* dst = v_render + pix0 * GSB_CH;
        if (vec) {
            for (int i = threadIdx.x; i < 256 * GSB_CH / 4; i += 256)
                reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(s_r)[i];
        } else {
            for (int i = threadIdx.x; i < n_px * GSB_CH; i += 256) dst[i] = s_r[i];
        }
    }
    __syncthreads();        // the next chunk overwrites the staging buffers
    }
    float s;
    if (det_partials) {
        // deterministic mode: the four block sums go to partials[k][block]; photo_finish_kernel adds them up
        const size_t nb = (size_t)gridDim.x * gridDim.y, b = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
        s = block_sum_256(t1, red); if (threadIdx.x == 0) det_partials[b] = s;
        __syncthreads();
        s = block_sum_256(t2, red); if (threadIdx.x == 0) det_partials[nb + b] = s;
        __syncthreads();
        s = block_sum_256(ga, red); if (threadIdx.x == 0) det_partials[2 * nb + b] = s;
        __syncthreads();
        s = block_sum_256(gb, red); if (threadIdx.x == 0) det_partials[3 * nb + b] = s;
        return;
    }
    s = block_sum_256(t1, red); if (threadIdx.x == 0 && s != 0.f) atomicAdd(loss_sum, s);
    __syncthreads();
    s = block_sum_256(t2, red); if (threadIdx.x == 0 && s != 0.f) atomicAdd(loss_sum + 1, s);
    if (v_exposure) {
        __syncthreads();
        s = block_sum_256(ga, red); if (threadIdx.x == 0 && s != 0.f) atomicAdd(v_exposure + cam * 2, s);
        __syncthreads();
        s = block_sum_256(gb, red); if (threadIdx.x == 0 && s != 0.f) atomicAdd(v_exposure + cam * 2 + 1, s);
    }
    // the block that finishes last turns the two sums into the loss values, so the caller needs no
    // scalar arithmetic kernels: loss_out = (coef * photometric, photometric)
    if (loss_out && threadIdx.x == 0) {
        __threadfence();
        const unsigned total = gridDim.x * gridDim.y;
        if (atomicAdd(reinterpret_cast<unsigned*>(loss_sum + 2), 1u) == total - 1) {
            __threadfence();
            const float photo = (atomicAdd(loss_sum, 0.f) + atomicAdd(loss_sum + 1, 0.f)) * inv_count;
            loss_out[0] = coef * photo;
            loss_out[1] = photo;
        }
    }
}

// Deterministic finish: block 0 sums the two loss partials over all CTAs, block 1 + c the exposure
// gradient of camera c over that camera's CTAs; fixed order (strided per-thread sums + the block tree).
__global__ void __launch_bounds__(256)
photo_finish_kernel(const float* __restrict__ partials, int gx, int C, float inv_count, float coef,
                    float* __restrict__ loss_sum, float* __restrict__ loss_out, float* __restrict__ v_exposure) {
    __shared__ float red[8];
    const size_t nb = (size_t)gx * C;
    auto fixed_sum = [&](const float* p, size_t n) -> float {
        float s = 0.f;
        for (size_t i = threadIdx.x; i < n; i += 256) s += p[i];
        __syncthreads();
        return block_sum_256(s, red);
    };
    if (blockIdx.x == 0) {
        const float a = fixed_sum(partials, nb);
        const float b = fixed_sum(partials + nb, nb);
        if (threadIdx.x == 0) {
            loss_sum[0] = a; loss_sum[1] = b;
            if (loss_out) { const float photo = (a + b) * inv_count; loss_out[0] = coef * photo; loss_out[1] = photo; }
        }
    } else if (v_exposure) {
        const int cam = blockIdx.x - 1;
        const float a = fixed_sum(partials + 2 * nb + (size_t)cam * gx, gx);
        const float b = fixed_sum(partials + 3 * nb + (size_t)cam * gx, gx);
        if (threadIdx.x == 0) { v_exposure[cam * 2] = a; v_exposure[cam * 2 + 1] = b; }
    }
}

static dim3 photo_grid(int C, int H, int W) {
    const int HW = H * W;
    const int chunks = (HW + 255) / 256, per_cam = (148 * 4 + C - 1) / C;
    return dim3(chunks < per_cam ? chunks : per_cam, C);
}

GSB_API size_t gsb_photo_loss_det_floats(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return 0;
    const dim3 g = photo_grid(C, H, W);
    return (size_t)4 * g.x * g.y;
}

GSB_API int gsb_photo_loss(const float* render, const float* gt, const float* exposure, int mode, int C,
                           int H, int W, float coef, float* loss_sum, float* loss_out, float* v_render,
                           float* v_exposure, float* det_partials, void* stream) {
    GSB_CHECK_ARG(C >= 0 && H > 0 && W > 0 && mode >= 0 && mode <= 3);
    if (C == 0) return 0;
    GSB_CHECK_ARG(render && gt && loss_sum && v_render && C <= 65535);
    const int HW = H * W;
    const dim3 grid = photo_grid(C, H, W);
    const float inv_count = 1.0f / ((float)C * (float)HW);
    photo_loss_kernel<<<grid, 256, 0, gsb_stream(stream)>>>(render, gt, exposure, mode, HW, inv_count, coef, loss_sum,
                                                            loss_out, v_render, v_exposure, det_partials);
    GSB_LAUNCH_CHECK();
    if (det_partials) {
        photo_finish_kernel<<<1 + (v_exposure ? C : 0), 256, 0, gsb_stream(stream)>>>(
            det_partials, (int)grid.x, C, inv_count, coef, loss_sum, loss_out, v_exposure);
        GSB_LAUNCH_CHECK();
    }
    return 0;
}

// v_render[C,H,W,5] *= *grad  (device scalar), used when the fused loss is not the graph root
__global__ void scale_kernel(float* __restrict__ x, int64_t n, const float* __restrict__ grad) {
    const float g = *grad;
    if (g == 1.0f) return;      // loss.backward() on the fused loss itself: nothing to do, no traffic
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] *= g;
}

GSB_API int gsb_scale_by_scalar(float* x, int64_t n, const float* grad, void* stream) {
    if (n == 0) return 0;
    GSB_CHECK_ARG(x && grad);
    // a few CTAs per SM, grid-stride: the common case (grad == 1) then costs one small launch
    const int64_t want = (n + 255) / 256;
    const unsigned blocks = (unsigned)(want < 148 * 8 ? want : 148 * 8);
    scale_kernel<<<blocks, 256, 0, gsb_stream(stream)>>>(x, n, grad);
    GSB_LAUNCH_CHECK();
    return 0;
}
