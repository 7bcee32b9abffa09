// Fused L1 + SSIM training loss (include/msgs_loss.h): train.py:210-214 with utils/loss_utils.py:17-64 as
// ONE forward and ONE backward launch for all images of a step.  Separable shared-memory kernels (loss_body.cuh): per
// pixel-channel 8 B read + 12 B written forward, 20 B read + 4 B written backward; 11 + 11 taps per quantity from shared
// memory.  HBM/L2-bound on large images, launch-bound on the coarse scales — hence one launch for all of them.
#include "common.cuh"
#include "../../include/msgs_raster.h"
#include "../../include/msgs_loss.h"
#include "loss_body.cuh"

namespace msgs {

constexpr int LOSS_THREADS = 256;

struct LossBatch {
    int n;
    int cta_begin[MSGS_LOSS_MAX_IMAGES + 1];        // first CTA of image i (CTA = (channel, tile row, tile column))
    int tiles_x[MSGS_LOSS_MAX_IMAGES], tiles_y[MSGS_LOSS_MAX_IMAGES];
    msgs_loss_image im[MSGS_LOSS_MAX_IMAGES];
};

__device__ __forceinline__ bool loss_locate(const LossBatch& B, int cta, int& img, int& c, int& x0, int& y0) {
    img = 0;
    while (img + 1 < B.n && cta >= B.cta_begin[img + 1]) ++img;
    const int local = cta - B.cta_begin[img];
    const int per = B.tiles_x[img] * B.tiles_y[img];
    c = local / per;
    const int t = local - c * per;
    const int ty = t / B.tiles_x[img], tx = t - ty * B.tiles_x[img];
    x0 = tx * LT_W; y0 = ty * LT_H;
    return c < B.im[img].C;
}

__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_forward_kernel(const __grid_constant__ LossBatch B, float lam, const float* __restrict__ w1d, float* __restrict__ loss_out) {
    __shared__ LossFwdSmem sm;
    __shared__ float s_w[11];
    __shared__ float s_red[LOSS_THREADS / 32];
    const int tid = threadIdx.x;
    int img, c, x0, y0;
    if (!loss_locate(B, blockIdx.x, img, c, x0, y0)) return;
    const msgs_loss_image& I = B.im[img];
    const long long HW = (long long)I.H * I.W;
    if (tid < 11) s_w[tid] = w1d[tid];
    const float* const src[2] = {I.img + c * HW, I.gt + c * HW};
    float* const dst[2] = {sm.a, sm.b};
    loss_stage<2>(tid, LOSS_THREADS, src, dst, I.H, I.W, x0, y0);
    __syncthreads();
    loss_fwd_rows(tid, LOSS_THREADS, s_w, sm);
    __syncthreads();
    float part = loss_fwd_cols(tid, LOSS_THREADS, s_w, sm, lam, I.H, I.W, x0, y0, I.loss_px ? I.loss_px + c * HW : nullptr,
                               I.dmu + c * HW, I.de11 + c * HW, I.de12 + c * HW);
    // per-image mean, reduced here: warp shuffle -> one atomic per CTA
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
    if ((tid & 31) == 0) s_red[tid >> 5] = part;
    __syncthreads();
    if (tid == 0 && loss_out) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < LOSS_THREADS / 32; ++k) t += s_red[k];
        atomicAdd(loss_out + img, t * (I.weight / (float)(I.C * HW)));
    }
}

__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_backward_kernel(const __grid_constant__ LossBatch B, float lam, const float* __restrict__ w1d, const float* __restrict__ up) {
    __shared__ LossBwdSmem sm;
    __shared__ float s_w[11];
    const int tid = threadIdx.x;
    int img, c, x0, y0;
    if (!loss_locate(B, blockIdx.x, img, c, x0, y0)) return;
    const msgs_loss_image& I = B.im[img];
    const long long HW = (long long)I.H * I.W;
    if (tid < 11) s_w[tid] = w1d[tid];
    const float* const src[3] = {I.dmu + c * HW, I.de11 + c * HW, I.de12 + c * HW};
    float* const dst[3] = {sm.m[0], sm.m[1], sm.m[2]};
    loss_stage<3>(tid, LOSS_THREADS, src, dst, I.H, I.W, x0, y0);
    __syncthreads();
    loss_bwd_rows(tid, LOSS_THREADS, s_w, sm);
    __syncthreads();
    const float scale = (up ? *up : 1.f) * I.weight / (float)(I.C * HW);
    loss_bwd_cols(tid, LOSS_THREADS, s_w, sm, lam, scale, I.H, I.W, x0, y0, I.img + c * HW, I.gt + c * HW, I.dL_dimg + c * HW);
}

static int build_batch(int n_images, const msgs_loss_image* images, bool backward, LossBatch& B) {
    if (n_images < 0 || n_images > MSGS_LOSS_MAX_IMAGES || (n_images > 0 && !images)) return MSGS_ERR_ARG;
    B.n = 0;
    int cta = 0;
    for (int i = 0; i < n_images; ++i) {
        const msgs_loss_image& I = images[i];
        if (I.C < 0 || I.H < 0 || I.W < 0) return MSGS_ERR_ARG;
        B.cta_begin[i] = cta;
        B.im[i] = I;
        B.tiles_x[i] = (I.W + LT_W - 1) / LT_W; B.tiles_y[i] = (I.H + LT_H - 1) / LT_H;
        const long long n = (long long)I.C * B.tiles_x[i] * B.tiles_y[i];
        if (n > 0) {
            if (!I.img || !I.gt || !I.dmu || !I.de11 || !I.de12 || (backward && !I.dL_dimg)) return MSGS_ERR_ARG;
            if (cta + n > (1 << 30)) return MSGS_ERR_ARG;
        }
        if (B.tiles_x[i] == 0 || B.tiles_y[i] == 0) { B.tiles_x[i] = B.tiles_y[i] = 1; B.im[i].C = 0; }   // empty image: no CTA
        cta += (int)((long long)B.im[i].C * B.tiles_x[i] * B.tiles_y[i]);
    }
    B.n = n_images;
    B.cta_begin[n_images] = cta;
    return cta;
}

}  // namespace msgs

using namespace msgs;

extern "C" int msgs_l1_ssim_forward_batch(int n_images, const msgs_loss_image* images, float lambda_dssim, const float* w1d,
                                          float* loss_out, void* stream) {
    LossBatch B;
    const int ctas = build_batch(n_images, images, false, B);
    if (ctas < 0) return ctas;
    if (n_images == 0) return MSGS_OK;
    if (!loss_out || !w1d) return MSGS_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(loss_out, 0, (size_t)n_images * sizeof(float), st) != cudaSuccess) return MSGS_ERR_CUDA;
    if (ctas == 0) return MSGS_OK;
    note_launches(1);
    l1_ssim_forward_kernel<<<(unsigned)ctas, LOSS_THREADS, 0, st>>>(B, lambda_dssim, w1d, loss_out);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_l1_ssim_backward_batch(int n_images, const msgs_loss_image* images, float lambda_dssim, const float* w1d,
                                           const float* dL_dloss, void* stream) {
    LossBatch B;
    const int ctas = build_batch(n_images, images, true, B);
    if (ctas < 0) return ctas;
    if (ctas == 0) return MSGS_OK;
    if (!w1d) return MSGS_ERR_ARG;
    note_launches(1);
    l1_ssim_backward_kernel<<<(unsigned)ctas, LOSS_THREADS, 0, (cudaStream_t)stream>>>(B, lambda_dssim, w1d, dL_dloss);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

// single-image form: the contract of this entry point is the loss_px map (loss = its mean); the in-kernel reduction is off
extern "C" int msgs_l1_ssim_forward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt,
                                    const float* w1d, float* loss_px, float* dmu, float* de11, float* de12, void* stream) {
    if (C < 0 || H < 0 || W < 0) return MSGS_ERR_ARG;
    if ((long long)C * H * W == 0) return MSGS_OK;
    if (!loss_px || !w1d) return MSGS_ERR_ARG;
    msgs_loss_image I = {C, H, W, 1.0f, img, gt, dmu, de11, de12, loss_px, nullptr};
    LossBatch B;
    const int ctas = build_batch(1, &I, false, B);
    if (ctas < 0) return ctas;
    note_launches(1);
    l1_ssim_forward_kernel<<<(unsigned)ctas, LOSS_THREADS, 0, (cudaStream_t)stream>>>(B, lambda_dssim, w1d, nullptr);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_l1_ssim_backward(int C, int H, int W, float lambda_dssim, const float* img, const float* gt,
                                     const float* w1d, const float* dmu, const float* de11, const float* de12,
                                     const float* dL_dloss, float* dL_dimg, void* stream) {
    if (C < 0 || H < 0 || W < 0) return MSGS_ERR_ARG;
    if ((long long)C * H * W == 0) return MSGS_OK;
    msgs_loss_image I = {C, H, W, 1.0f, img, gt, const_cast<float*>(dmu), const_cast<float*>(de11), const_cast<float*>(de12),
                         nullptr, dL_dimg};
    return msgs_l1_ssim_backward_batch(1, &I, lambda_dssim, w1d, dL_dloss, stream);
}
