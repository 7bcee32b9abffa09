/* msgs_loss.h — C ABI of the fused training loss ("loss (L1 + SSIM 11x11 conv) fused").
 *
 * Replaces train.py:210-211  `loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))`
 * (utils/loss_utils.py:17-18 l1_loss, :24-64 gaussian / create_window / ssim / _ssim: five grouped 11x11 conv2d with zero
 * padding 5 plus ~15 element-wise launches forward, and their autograd backward) by ONE forward and ONE backward launch.
 * Images are (C,H,W) fp32 contiguous device memory; w1d = the 11 fp32 taps of gaussian(11, 1.5) (device memory; the 2-D
 * window is their outer product, as create_window builds it).  Exported by libmsgs_raster.so; MSGS_OK or MSGS_ERR_*.
 * First step: correctness-first gather kernels (each thread reads its 11x11 window through L1/L2); the separable
 * shared-memory version is the follow-up.
 */
#ifndef MSGS_LOSS_H
#define MSGS_LOSS_H
#ifdef __cplusplus
extern "C" {
#endif

/* Forward: loss_px[c,y,x] = (1 - lambda) |img - gt| + lambda (1 - ssim_map[c,y,x])   (loss = mean of loss_px), and the
 * three per-pixel derivative maps of ssim_map w.r.t. the window sums of img (mu1, E[img^2], E[img gt]) for the backward. */
int msgs_l1_ssim_forward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt, const float* w1d,
                         float* loss_px, float* dmu, float* de11, float* de12, void* stream);

/* Backward: dL_dimg[c,y,x] for loss = mean(loss_px), times the upstream scalar *dL_dloss (device pointer, may be NULL = 1). */
int msgs_l1_ssim_backward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt, const float* w1d,
                          const float* dmu, const float* de11, const float* de12, const float* dL_dloss, float* dL_dimg,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif
