// Bodies of the fused L1 + SSIM loss (loss.cu; include/msgs_loss.h; utils/loss_utils.py:17-64), separable version:
// a CTA owns a 32 x 16 tile of one channel of one image; it stages the tile plus a 5-pixel halo (zero padding outside
// the image, as conv2d(padding=5)) in shared memory, runs the 11-tap window along x into a second shared buffer and then
// along y — 11 + 11 taps per pixel and quantity instead of 121.  gaussian(11, 1.5)'s 2-D window is the outer product of
// the 1-D taps (create_window), so the two passes give the same sums up to fp32 summation order.
// The phases are __host__ __device__ functions of (thread id, thread count): tests/loss_host_harness.cu runs them on the
// host, phase by phase over every thread id, against golden vectors from the reference's own loss_utils + autograd.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace msgs {

constexpr int LT_W = 32, LT_H = 16;                 // output tile
constexpr int LHALO = 5;                            // window_size // 2
constexpr int LIN_W = LT_W + 2 * LHALO, LIN_H = LT_H + 2 * LHALO;       // 42 x 26 staged inputs
constexpr int LIN_S = LIN_W + 1, LH_S = LT_W + 1;   // row strides (odd: conflict-free column walks)

struct LossFwdSmem {
    float a[LIN_H * LIN_S], b[LIN_H * LIN_S];       // img, gt (zero outside the image)
    float h[5][LIN_H * LH_S];                       // row sums of a, b, a a, b b, a b
};
struct LossBwdSmem {
    float m[3][LIN_H * LIN_S];                      // dmu, de11, de12 (zero outside the image)
    float h[3][LIN_H * LH_S];
};

// stage `n` maps (channel plane pointers src[k], H x W) around tile origin (x0, y0) into dst[k]
template <int N>
__host__ __device__ __forceinline__ void loss_stage(int tid, int nthr, const float* const (&src)[N], float* const (&dst)[N],
                                                    int H, int W, int x0, int y0) {
    for (int i = tid; i < LIN_H * LIN_W; i += nthr) {
        const int r = i / LIN_W, c = i - r * LIN_W;
        const int yy = y0 - LHALO + r, xx = x0 - LHALO + c;
        const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
#pragma unroll
        for (int k = 0; k < N; ++k) dst[k][r * LIN_S + c] = in ? src[k][(long long)yy * W + xx] : 0.f;
    }
}

// forward, pass along x: h[q][r][x] = sum_k w[k] * f_q(a, b)[r][x + k]
__host__ __device__ __forceinline__ void loss_fwd_rows(int tid, int nthr, const float* w, LossFwdSmem& sm) {
    for (int i = tid; i < LIN_H * LT_W; i += nthr) {
        const int r = i / LT_W, x = i - r * LT_W;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float p = sm.a[r * LIN_S + x + k], q = sm.b[r * LIN_S + x + k], wk = w[k];
            s0 += wk * p; s1 += wk * q; s2 += wk * p * p; s3 += wk * q * q; s4 += wk * p * q;
        }
        sm.h[0][r * LH_S + x] = s0; sm.h[1][r * LH_S + x] = s1; sm.h[2][r * LH_S + x] = s2;
        sm.h[3][r * LH_S + x] = s3; sm.h[4][r * LH_S + x] = s4;
    }
}

// forward, pass along y + the SSIM map and its derivative maps; returns this thread's sum of
// (1 - lam) |a - b| + lam (1 - ssim) over its pixels inside the image
__host__ __device__ __forceinline__ float loss_fwd_cols(int tid, int nthr, const float* w, const LossFwdSmem& sm, float lam,
                                                        int H, int W, int x0, int y0, float* __restrict__ loss_px,
                                                        float* __restrict__ dmu, float* __restrict__ de11,
                                                        float* __restrict__ de12) {
    float total = 0.f;
    for (int i = tid; i < LT_H * LT_W; i += nthr) {
        const int y = i / LT_W, x = i - y * LT_W;
        const int yy = y0 + y, xx = x0 + x;
        if (yy >= H || xx >= W) continue;
        float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float wk = w[k];
            const int o = (y + k) * LH_S + x;
            mu1 += wk * sm.h[0][o]; mu2 += wk * sm.h[1][o]; e11 += wk * sm.h[2][o]; e22 += wk * sm.h[3][o]; e12 += wk * sm.h[4][o];
        }
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;
        const float A = 2.f * mu1 * mu2 + C1, B = 2.f * s12 + C2, Cc = mu1 * mu1 + mu2 * mu2 + C1, D = s1 + s2 + C2;
        const float S = (A * B) / (Cc * D);
        const float dS_ds1 = -(A * B) / (Cc * D * D);            // = dS/dE[img^2]
        const float dS_ds12 = 2.f * A / (Cc * D);                // = dS/dE[img gt]
        // total derivative w.r.t. mu1 (s1 = e11 - mu1^2 and s12 = e12 - mu1 mu2 depend on it)
        const float dS_dmu1 = 2.f * mu2 * B / (Cc * D) - 2.f * mu1 * A * B / (Cc * Cc * D) - 2.f * mu1 * dS_ds1 - mu2 * dS_ds12;
        const float p = sm.a[(y + LHALO) * LIN_S + x + LHALO], q = sm.b[(y + LHALO) * LIN_S + x + LHALO];
        const float l = (1.f - lam) * fabsf(p - q) + lam * (1.f - S);
        const long long o = (long long)yy * W + xx;
        if (loss_px) loss_px[o] = l;
        dmu[o] = dS_dmu1; de11[o] = dS_ds1; de12[o] = dS_ds12;
        total += l;
    }
    return total;
}

// backward, pass along x over the three derivative maps
__host__ __device__ __forceinline__ void loss_bwd_rows(int tid, int nthr, const float* w, LossBwdSmem& sm) {
    for (int i = tid; i < LIN_H * LT_W; i += nthr) {
        const int r = i / LT_W, x = i - r * LT_W;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float wk = w[k];
            s0 += wk * sm.m[0][r * LIN_S + x + k]; s1 += wk * sm.m[1][r * LIN_S + x + k]; s2 += wk * sm.m[2][r * LIN_S + x + k];
        }
        sm.h[0][r * LH_S + x] = s0; sm.h[1][r * LH_S + x] = s1; sm.h[2][r * LH_S + x] = s2;
    }
}

// backward, pass along y + the gradient:  dimg = scale * ((1 - lam) sign(p - q) - lam (W*dmu + 2 p W*de11 + q W*de12)),
// scale = weight * upstream / (C H W)   (the window is symmetric, so the adjoint of the convolution is the convolution)
__host__ __device__ __forceinline__ void loss_bwd_cols(int tid, int nthr, const float* w, const LossBwdSmem& sm, float lam,
                                                       float scale, int H, int W, int x0, int y0, const float* __restrict__ img,
                                                       const float* __restrict__ gt, float* __restrict__ dimg) {
    for (int i = tid; i < LT_H * LT_W; i += nthr) {
        const int y = i / LT_W, x = i - y * LT_W;
        const int yy = y0 + y, xx = x0 + x;
        if (yy >= H || xx >= W) continue;
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) {
            const float wk = w[k];
            const int o = (y + k) * LH_S + x;
            c0 += wk * sm.h[0][o]; c1 += wk * sm.h[1][o]; c2 += wk * sm.h[2][o];
        }
        const long long o = (long long)yy * W + xx;
        const float p = img[o], q = gt[o];
        const float d = p - q, sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        dimg[o] = scale * ((1.f - lam) * sgn - lam * (c0 + 2.f * p * c1 + q * c2));
    }
}

}  // namespace msgs
