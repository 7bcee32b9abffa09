// Fused L1 + SSIM training loss (include/msgs_loss.h): train.py:210-211 with utils/loss_utils.py:17-64
// as one forward and one backward launch.  First step (gather kernels, one thread per pixel-channel, the 121-tap window
// read through L1/L2): L1/LSU-bound; algorithmic HBM bytes 8 read + 16 written per pixel-channel forward, 20 read + 4
// written backward.  The separable shared-memory version is the follow-up.
#include "common.cuh"
#include "../../include/msgs_raster.h"
#include "../../include/msgs_loss.h"
#include "loss_body.cuh"

namespace msgs {
__global__ void __launch_bounds__(256)
l1_ssim_forward_kernel(int C, int H, int W, float lam, const float* __restrict__ img, const float* __restrict__ gt,
                       const float* __restrict__ w1d, float* __restrict__ loss_px, float* __restrict__ dmu,
                       float* __restrict__ de11, float* __restrict__ de12) {
    l1_ssim_forward_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, C, H, W, lam, img, gt, w1d, loss_px, dmu, de11, de12);
}
__global__ void __launch_bounds__(256)
l1_ssim_backward_kernel(int C, int H, int W, float lam, const float* __restrict__ img, const float* __restrict__ gt,
                        const float* __restrict__ w1d, const float* __restrict__ dmu, const float* __restrict__ de11,
                        const float* __restrict__ de12, const float* __restrict__ up, float* __restrict__ dimg) {
    l1_ssim_backward_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, C, H, W, lam, img, gt, w1d, dmu, de11, de12, up, dimg);
}
}  // namespace msgs

extern "C" int msgs_l1_ssim_forward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt,
                                    const float* w1d, float* loss_px, float* dmu, float* de11, float* de12, void* stream) {
    using namespace msgs;
    if (C < 0 || H < 0 || W < 0) return MSGS_ERR_ARG;
    const long long n = (long long)C * H * W;
    if (n == 0) return MSGS_OK;
    if (!img || !gt || !w1d || !loss_px || !dmu || !de11 || !de12) return MSGS_ERR_ARG;
    l1_ssim_forward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(C, H, W, lambda_dssim, img, gt, w1d,
                                                                                          loss_px, dmu, de11, de12);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_l1_ssim_backward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt,
                                     const float* w1d, const float* dmu, const float* de11, const float* de12,
                                     const float* dL_dloss, float* dL_dimg, void* stream) {
    using namespace msgs;
    if (C < 0 || H < 0 || W < 0) return MSGS_ERR_ARG;
    const long long n = (long long)C * H * W;
    if (n == 0) return MSGS_OK;
    if (!img || !gt || !w1d || !dmu || !de11 || !de12 || !dL_dimg) return MSGS_ERR_ARG;
    l1_ssim_backward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(C, H, W, lambda_dssim, img, gt, w1d,
                                                                                           dmu, de11, de12, dL_dloss, dL_dimg);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
