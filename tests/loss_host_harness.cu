// Test-only: the per-thread bodies of ms-gs_b200/csrc/loss_body.cuh run on the HOST for every thread index of the launch
// grid (256-thread blocks, overshoot included).  Built by tests/test_zz_loss.py.
#include "../ms-gs_b200/csrc/loss_body.cuh"

extern "C" void host_l1_ssim_forward(int C, int H, int W, float lam, const float* img, const float* gt, const float* w1d,
                                     float* loss_px, float* dmu, float* de11, float* de12) {
    const long long threads = ((long long)C * H * W + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i) msgs::l1_ssim_forward_thread(i, C, H, W, lam, img, gt, w1d, loss_px, dmu, de11, de12);
}

extern "C" void host_l1_ssim_backward(int C, int H, int W, float lam, const float* img, const float* gt, const float* w1d,
                                      const float* dmu, const float* de11, const float* de12, const float* up, float* dimg) {
    const long long threads = ((long long)C * H * W + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i) msgs::l1_ssim_backward_thread(i, C, H, W, lam, img, gt, w1d, dmu, de11, de12, up, dimg);
}
