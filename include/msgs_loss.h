/* msgs_loss.h — C ABI of the fused training loss ("loss (L1 + SSIM 11x11 conv) fused for the 7 small
 * images").
 *
 * Replaces train.py:210-214  `loss = ((1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))) * multiplier`
 * (utils/loss_utils.py:17-18 l1_loss, :24-64 gaussian / create_window / ssim / _ssim: five grouped 11x11 conv2d with zero
 * padding 5 plus ~15 element-wise launches forward, and their autograd backward; multiplier = 1 at the finest scale, 0.1 at
 * the coarse ones, train.py:212-214) by ONE forward and ONE backward launch for ALL images of a step (the scales of a view).
 * Separable: a CTA stages a 32 x 16 tile + halo of one channel in shared memory and applies the 11-tap window along x, then
 * along y (11 + 11 taps per pixel and quantity instead of 121); the per-image loss is reduced inside the kernel.
 * Images are (C,H,W) fp32 contiguous device memory; w1d = the 11 fp32 taps of gaussian(11, 1.5) (device memory; the 2-D
 * window is their outer product, as create_window builds it).  Exported by libmsgs_raster.so; MSGS_OK or MSGS_ERR_*.
 */
#ifndef MSGS_LOSS_H
#define MSGS_LOSS_H
#ifdef __cplusplus
extern "C" {
#endif

/* One image of a batch.  Forward reads img, gt and writes the three per-pixel derivative maps of the SSIM map w.r.t. the
 * window sums of img (dmu: mu1, total; de11: E[img^2]; de12: E[img gt]) for the backward, and loss_px (optional, may be
 * NULL): (1 - lambda) |img - gt| + lambda (1 - ssim_map).  Backward reads img, gt and the maps and writes dL_dimg.
 * weight: the image's loss multiplier (train.py:212-214). */
typedef struct {
    int C, H, W;
    float weight;
    const float* img;
    const float* gt;
    float* dmu;
    float* de11;
    float* de12;
    float* loss_px;
    float* dL_dimg;
} msgs_loss_image;

#define MSGS_LOSS_MAX_IMAGES 8

/* Forward for n_images <= 8 images in one launch: loss_out[i] = weight_i * mean over image i of loss_px (loss_out: n_images
 * floats of device memory, overwritten).  The step's loss is their sum. */
int msgs_l1_ssim_forward_batch(int n_images, const msgs_loss_image* images, float lambda_dssim, const float* w1d,
                               float* loss_out, void* stream);

/* Backward for the same images in one launch: dL_dimg_i = (*dL_dloss) * weight_i * d(mean loss_px_i)/d img_i
 * (dL_dloss: device scalar, may be NULL = 1). */
int msgs_l1_ssim_backward_batch(int n_images, const msgs_loss_image* images, float lambda_dssim, const float* w1d,
                                const float* dL_dloss, void* stream);

/* Single-image forms (same kernels, n_images = 1, weight 1).  Forward also writes the loss_px map (loss = its mean). */
int msgs_l1_ssim_forward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt, const float* w1d,
                         float* loss_px, float* dmu, float* de11, float* de12, void* stream);
int msgs_l1_ssim_backward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt, const float* w1d,
                          const float* dmu, const float* de11, const float* de12, const float* dL_dloss, float* dL_dimg,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif
