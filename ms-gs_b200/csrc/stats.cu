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

__global__ void __launch_bounds__(256)
update_stats_units_kernel(int P, int L, const __grid_constant__ StatsUnits U, const int64_t* __restrict__ target,
                          float* __restrict__ max_px, float* __restrict__ min_px, float* __restrict__ max_r,
                          float* __restrict__ accum, float* __restrict__ denom, uint8_t* __restrict__ base) {
    update_stats_units_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, L, U, target, max_px, min_px, max_r, accum,
                              denom, base);
}

// sum lanes: dacc | ddenom; max lanes: radius (padded to 4 words) | obs_hi | base_seen; min lanes: obs_lo.  The lanes the
// NEXT render depends on (pixel-size observations, base mask) are the contiguous tail [fast_off, end): a sharded step may
// reduce only those every step and keep the densification sums rank-local until they are consumed (msgs_stats.h).
struct DeltaLanes { float *dacc, *ddenom, *radius, *obs_hi, *base_seen, *obs_lo; long long n_sum, n_max, n_min, fast_off, fast_max; };
static inline long long pad4(long long n) { return (n + 3) / 4 * 4; }
static DeltaLanes delta_lanes(int P, int L, float* delta) {
    DeltaLanes d;
    d.n_sum = pad4(2LL * P * L); d.fast_max = pad4(2LL * P); d.n_max = pad4(P) + d.fast_max; d.n_min = pad4(P);
    d.fast_off = d.n_sum + pad4(P);
    d.dacc = delta; d.ddenom = delta + (long long)P * L;
    d.radius = delta + d.n_sum; d.obs_hi = delta + d.fast_off; d.base_seen = d.obs_hi + P;
    d.obs_lo = delta + d.n_sum + d.n_max;
    return d;
}

__global__ void __launch_bounds__(256)
stats_delta_reset_kernel(float4* __restrict__ delta, long long zero4, long long total4) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
        const float v = i < zero4 ? 0.f : INFINITY;
        delta[i] = make_float4(v, v, v, v);
    }
}

__global__ void __launch_bounds__(256)
stats_delta_units_kernel(int P, int L, const __grid_constant__ StatsUnits U, const int64_t* __restrict__ target,
                         float* __restrict__ dacc, float* __restrict__ ddenom, float* __restrict__ radius,
                         float* __restrict__ obs_hi, float* __restrict__ obs_lo, float* __restrict__ base_seen) {
    stats_delta_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, L, U, target, dacc, ddenom, radius, obs_hi, obs_lo,
                       base_seen);
}

__global__ void __launch_bounds__(256)
stats_delta_apply_kernel(int P, int L, const int64_t* __restrict__ target, const float* __restrict__ dacc,
                         const float* __restrict__ ddenom, const float* __restrict__ radius, const float* __restrict__ obs_hi,
                         const float* __restrict__ obs_lo, const float* __restrict__ base_seen, float* __restrict__ max_px,
                         float* __restrict__ min_px, float* __restrict__ max_r, float* __restrict__ accum,
                         float* __restrict__ denom, uint8_t* __restrict__ base) {
    stats_apply_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, L, target, dacc, ddenom, radius, obs_hi, obs_lo,
                       base_seen, max_px, min_px, max_r, accum, denom, base);
}

// validates the units and packs them for the kernels; returns the OR of the units' `what`, or an error
static int pack_units(int P, int L, int n_units, const msgs_stats_unit* units, StatsUnits& U) {
    if (P < 0 || L < 1 || n_units < 0 || n_units > MSGS_STATS_MAX_UNITS || (n_units > 0 && !units)) return MSGS_ERR_ARG;
    int any = 0;
    U.n = n_units;
    for (int u = 0; u < n_units; ++u) {
        const msgs_stats_unit& s = units[u];
        if (s.reso_lvl < 0 || s.reso_lvl >= L || (s.what & ~7)) return MSGS_ERR_ARG;
        if (P > 0 && s.what) {
            if (!s.radii) return MSGS_ERR_ARG;
            if ((s.what & MSGS_STATS_PIXEL_SIZES) && !s.pixel_sizes) return MSGS_ERR_ARG;
            if ((s.what & MSGS_STATS_DENSIFY) && !s.dL_dmeans2D) return MSGS_ERR_ARG;
        }
        U.lvl[u] = s.reso_lvl; U.what[u] = s.what; U.radii[u] = s.radii; U.px[u] = s.pixel_sizes; U.g2d[u] = s.dL_dmeans2D;
        any |= s.what;
    }
    return any;
}
}  // namespace msgs

extern "C" int msgs_update_stats_units(int P, int L, int n_units, const msgs_stats_unit* units, const int64_t* target_reso_lvl,
                                       float* max_pixel_sizes, float* min_pixel_sizes, float* max_radii2D,
                                       float* xyz_gradient_accum, float* denom, uint8_t* base_mask, void* stream) {
    using namespace msgs;
    StatsUnits U;
    const int any = pack_units(P, L, n_units, units, U);
    if (any < 0) return any;
    if (P == 0 || any == 0) return MSGS_OK;
    if ((any & MSGS_STATS_PIXEL_SIZES) && (!target_reso_lvl || !max_pixel_sizes || !min_pixel_sizes)) return MSGS_ERR_ARG;
    if ((any & MSGS_STATS_DENSIFY) && (!max_radii2D || !xyz_gradient_accum || !denom)) return MSGS_ERR_ARG;
    if ((any & MSGS_STATS_BASE_MASK) && !base_mask) return MSGS_ERR_ARG;
    note_launches(1);
    update_stats_units_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, L, U, target_reso_lvl, max_pixel_sizes, min_pixel_sizes, max_radii2D, xyz_gradient_accum, denom, base_mask);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" void msgs_stats_delta_words(int P, int L, long long* n_sum, long long* n_max, long long* n_min) {
    const msgs::DeltaLanes d = msgs::delta_lanes(P < 0 ? 0 : P, L < 1 ? 1 : L, nullptr);
    if (n_sum) *n_sum = d.n_sum;
    if (n_max) *n_max = d.n_max;
    if (n_min) *n_min = d.n_min;
}

extern "C" void msgs_stats_delta_fast_words(int P, int L, long long* offset, long long* n_max, long long* n_min) {
    const msgs::DeltaLanes d = msgs::delta_lanes(P < 0 ? 0 : P, L < 1 ? 1 : L, nullptr);
    if (offset) *offset = d.fast_off;
    if (n_max) *n_max = d.fast_max;
    if (n_min) *n_min = d.n_min;
}

extern "C" int msgs_stats_delta_reset(int P, int L, float* delta, void* stream) {
    using namespace msgs;
    if (P < 0 || L < 1) return MSGS_ERR_ARG;
    if (P == 0) return MSGS_OK;
    if (!delta || ((uintptr_t)delta & 15)) return MSGS_ERR_ARG;
    const DeltaLanes d = delta_lanes(P, L, delta);
    const long long zero4 = (d.n_sum + d.n_max) / 4, total4 = zero4 + d.n_min / 4;
    const long long want = (total4 + 255) / 256;
    note_launches(1);
    stats_delta_reset_kernel<<<(unsigned)(want < 148 * 8 ? want : 148 * 8), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(delta), zero4, total4);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_stats_delta_units(int P, int L, int n_units, const msgs_stats_unit* units, const int64_t* target_reso_lvl,
                                      float* delta, void* stream) {
    using namespace msgs;
    StatsUnits U;
    const int any = pack_units(P, L, n_units, units, U);
    if (any < 0) return any;
    if (P == 0 || any == 0) return MSGS_OK;
    if (!delta || ((any & MSGS_STATS_PIXEL_SIZES) && !target_reso_lvl)) return MSGS_ERR_ARG;
    const DeltaLanes d = delta_lanes(P, L, delta);
    note_launches(1);
    stats_delta_units_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, L, U, target_reso_lvl, d.dacc, d.ddenom, d.radius, d.obs_hi, d.obs_lo, d.base_seen);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_stats_delta_apply(int P, int L, const int64_t* target_reso_lvl, const float* delta, float* max_pixel_sizes,
                                      float* min_pixel_sizes, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                                      uint8_t* base_mask, void* stream) {
    using namespace msgs;
    if (P < 0 || L < 1) return MSGS_ERR_ARG;
    if (P == 0) return MSGS_OK;
    if (!delta || !target_reso_lvl || !max_pixel_sizes || !min_pixel_sizes || !max_radii2D || !xyz_gradient_accum || !denom)
        return MSGS_ERR_ARG;
    const DeltaLanes d = delta_lanes(P, L, const_cast<float*>(delta));
    note_launches(1);
    stats_delta_apply_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, L, target_reso_lvl, d.dacc, d.ddenom, d.radius, d.obs_hi, d.obs_lo, d.base_seen, max_pixel_sizes, min_pixel_sizes,
        max_radii2D, xyz_gradient_accum, denom, base_mask);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

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
    note_launches(1);
    update_stats_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        P, L, reso_lvl, what, radii, pixel_sizes, target_reso_lvl, dL_dmeans2D, max_pixel_sizes, min_pixel_sizes, max_radii2D,
        xyz_gradient_accum, denom, base_mask);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
