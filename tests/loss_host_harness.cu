// Test-only: the CTA phases of ms-gs_b200/csrc/loss_body.cuh run on the HOST — for every tile of every channel, each phase
// over every thread id of a 256-thread CTA, in the order the kernels synchronise them.  Built by tests/test_zz_loss.py.
#include "../ms-gs_b200/csrc/loss_body.cuh"

using namespace msgs;
static const int NT = 256;

extern "C" double host_l1_ssim_forward(int C, int H, int W, float lam, const float* img, const float* gt, const float* w1d,
                                       float* loss_px, float* dmu, float* de11, float* de12) {
    static LossFwdSmem sm;
    double total = 0.0;
    const long long HW = (long long)H * W;
    for (int c = 0; c < C; ++c)
        for (int y0 = 0; y0 < H; y0 += LT_H)
            for (int x0 = 0; x0 < W; x0 += LT_W) {
                const float* const src[2] = {img + c * HW, gt + c * HW};
                float* const dst[2] = {sm.a, sm.b};
                for (int t = 0; t < NT; ++t) loss_stage<2>(t, NT, src, dst, H, W, x0, y0);
                for (int t = 0; t < NT; ++t) loss_fwd_rows(t, NT, w1d, sm);
                for (int t = 0; t < NT; ++t)
                    total += loss_fwd_cols(t, NT, w1d, sm, lam, H, W, x0, y0, loss_px ? loss_px + c * HW : nullptr,
                                           dmu + c * HW, de11 + c * HW, de12 + c * HW);
            }
    return total / (double)(C * HW);
}

extern "C" void host_l1_ssim_backward(int C, int H, int W, float lam, const float* img, const float* gt, const float* w1d,
                                      const float* dmu, const float* de11, const float* de12, float up, float weight, float* dimg) {
    static LossBwdSmem sm;
    const long long HW = (long long)H * W;
    const float scale = up * weight / (float)(C * HW);
    for (int c = 0; c < C; ++c)
        for (int y0 = 0; y0 < H; y0 += LT_H)
            for (int x0 = 0; x0 < W; x0 += LT_W) {
                const float* const src[3] = {dmu + c * HW, de11 + c * HW, de12 + c * HW};
                float* const dst[3] = {sm.m[0], sm.m[1], sm.m[2]};
                for (int t = 0; t < NT; ++t) loss_stage<3>(t, NT, src, dst, H, W, x0, y0);
                for (int t = 0; t < NT; ++t) loss_bwd_rows(t, NT, w1d, sm);
                for (int t = 0; t < NT; ++t)
                    loss_bwd_cols(t, NT, w1d, sm, lam, scale, H, W, x0, y0, img + c * HW, gt + c * HW, dimg + c * HW);
            }
}
