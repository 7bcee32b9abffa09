// Per-thread bodies of the fused L1 + SSIM loss (loss.cu; include/msgs_loss.h; utils/loss_utils.py:17-64).
// __host__ __device__: the same code runs on the host in tests/test_zz_loss.py against golden vectors from the reference.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace msgs {

// one thread per (c, y, x): window sums over the zero-padded 11x11 neighbourhood -> ssim value and its derivatives
__host__ __device__ __forceinline__ void
l1_ssim_forward_thread(long long i, int C, int H, int W, float lam, const float* __restrict__ img,
                       const float* __restrict__ gt, const float* __restrict__ w1d, float* __restrict__ loss_px,
                       float* __restrict__ dmu, float* __restrict__ de11, float* __restrict__ de12) {
    const long long HW = (long long)H * W;
    if (i >= C * HW) return;
    const int c = (int)(i / HW), y = (int)((i - c * HW) / W), x = (int)(i - c * HW - (long long)y * W);
    const float* a = img + c * HW;
    const float* b = gt + c * HW;
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
    for (int dy = 0; dy < 11; ++dy) {
        const int yy = y + dy - 5;
        if (yy < 0 || yy >= H) continue;
        const float wy = w1d[dy];
        for (int dx = 0; dx < 11; ++dx) {
            const int xx = x + dx - 5;
            if (xx < 0 || xx >= W) continue;
            const float w = wy * w1d[dx], p = a[(long long)yy * W + xx], q = b[(long long)yy * W + xx];
            mu1 += w * p; mu2 += w * q; e11 += w * p * p; e22 += w * q * q; e12 += w * p * q;
        }
    }
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;
    const float A = 2.f * mu1 * mu2 + C1, B = 2.f * s12 + C2, Cc = mu1 * mu1 + mu2 * mu2 + C1, D = s1 + s2 + C2;
    const float S = (A * B) / (Cc * D);
    const float dS_ds1 = -(A * B) / (Cc * D * D);            // = dS/dE[img^2]
    const float dS_ds12 = 2.f * A / (Cc * D);                // = dS/dE[img gt]
    // total derivative w.r.t. mu1 (s1 = e11 - mu1^2 and s12 = e12 - mu1 mu2 depend on it)
    const float dS_dmu1 = 2.f * mu2 * B / (Cc * D) - 2.f * mu1 * A * B / (Cc * Cc * D) - 2.f * mu1 * dS_ds1 - mu2 * dS_ds12;
    loss_px[i] = (1.f - lam) * fabsf(a[(long long)y * W + x] - b[(long long)y * W + x]) + lam * (1.f - S);
    dmu[i] = dS_dmu1; de11[i] = dS_ds1; de12[i] = dS_ds12;
}

// one thread per (c, y, x): gathers the derivative maps of the 11x11 pixels whose windows contain this one
__host__ __device__ __forceinline__ void
l1_ssim_backward_thread(long long i, int C, int H, int W, float lam, const float* __restrict__ img,
                        const float* __restrict__ gt, const float* __restrict__ w1d, const float* __restrict__ dmu,
                        const float* __restrict__ de11, const float* __restrict__ de12, const float* __restrict__ up,
                        float* __restrict__ dimg) {
    const long long HW = (long long)H * W;
    if (i >= C * HW) return;
    const int c = (int)(i / HW), y = (int)((i - c * HW) / W), x = (int)(i - c * HW - (long long)y * W);
    const long long o = c * HW;
    const float p = img[i], q = gt[i];
    float acc = 0.f;
    for (int dy = 0; dy < 11; ++dy) {
        const int yy = y + dy - 5;
        if (yy < 0 || yy >= H) continue;
        const float wy = w1d[dy];
        for (int dx = 0; dx < 11; ++dx) {
            const int xx = x + dx - 5;
            if (xx < 0 || xx >= W) continue;
            const long long j = o + (long long)yy * W + xx;
            acc += wy * w1d[dx] * (dmu[j] + 2.f * p * de11[j] + q * de12[j]);
        }
    }
    const float d = p - q, sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    const float g = (up ? *up : 1.f) / (float)(C * HW);
    dimg[i] = g * ((1.f - lam) * sgn - lam * acc);
}

}  // namespace msgs
