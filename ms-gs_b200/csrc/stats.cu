// Fused per-step statistics update (include/msgs_stats.h): update_pixel_sizes, max_radii2D,
// add_densification_stats and update_base_gaussian_mask of one rendered unit as ONE launch instead of ~20 masked-index
// PyTorch launches with host synchronisations.  HBM-bound, one thread per Gaussian, coalesced scalar accesses; algorithmic
// bytes per Gaussian: 4 (radii) always; visible ones add <= 4 + 8 + 16 (pixel sizes) + 12 + 8 + 16 (densify) + 1.
#include "common.cuh"
#include "../../include/msgs_raster.h"
#include "../../include/msgs_stats.h"
#include "stats_body.cuh"

namespace msgs {
__global__ void __launch_bounds__(256)
update_stats_kernel(int P, int L, int lvl, int what, const int32_t* __restrict__ radii, const float* __restrict__ px,
                    const int64_t* __restrict__ target, const float* __restrict__ g2d, float* __restrict__ max_px,
                    float* __restrict__ min_px, float* __restrict__ max_r, float* __restrict__ accum,
                    float* __restrict__ denom, uint8_t* __restrict__ base) {
    update_stats_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, L, lvl, what, radii, px, target, g2d, max_px,
                        min_px, max_r, accum, denom, base);
}
}  // namespace msgs

extern "C" int msgs_update_stats(int P, int L, int reso_lvl, int what, const int32_t* radii, const float* pixel_sizes,
                                 const int64_t* target_reso_lvl, const float* dL_dmeans2D, float* max_pixel_sizes,
                                 float* min_pixel_sizes, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                                 uint8_t* base_mask, void* stream) {
    using namespace msgs;
    if (P < 0 || L < 1 || reso_lvl < 0 || reso_lvl >= L || (what & ~7)) return MSGS_ERR_ARG;
    if (P == 0 || what == 0) return MSGS_OK;
    if (!radii) return MSGS_ERR_ARG;
    if ((what & MSGS_STATS_PIXEL_SIZES) && (!pixel_sizes || !target_reso_lvl || !max_pixel_sizes || !min_pixel_sizes)) return MSGS_ERR_ARG;
    if ((what & MSGS_STATS_DENSIFY) && (!dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom)) return MSGS_ERR_ARG;
    if ((what & MSGS_STATS_BASE_MASK) && !base_mask) return MSGS_ERR_ARG;
    update_stats_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, L, reso_lvl, what, radii, pixel_sizes, target_reso_lvl, dL_dmeans2D, max_pixel_sizes, min_pixel_sizes, max_radii2D,
        xyz_gradient_accum, denom, base_mask);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
