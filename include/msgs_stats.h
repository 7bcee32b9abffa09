/* msgs_stats.h — C ABI of the fused per-step statistics update ("per-level stat updates").
 *
 * After every render of a (view, scale) unit train.py updates per-Gaussian bookkeeping with ~20 small PyTorch launches,
 * most of them boolean-mask indexing (each `x[mask]` is a nonzero + gather and a host synchronisation):
 *   gaussians.update_pixel_sizes(visibility_filter, pixel_sizes, reso_idx, iteration)   train.py:245; scene/gaussian_model.py:662-686
 *   gaussians.max_radii2D[vis] = max(max_radii2D[vis], radii[vis])                      train.py:249
 *   gaussians.add_densification_stats(viewspace_point_tensor, vis, reso_lvl)            train.py:250; gaussian_model.py:698-701
 *   gaussians.update_base_gaussian_mask(vis)                                            train.py:241; gaussian_model.py:703-705
 * with vis = radii > 0.  ONE launch, no host synchronisation, every array read and written once, in place.
 * Device pointers, `stream` is a cudaStream_t.  Exported by libmsgs_raster.so.  Returns MSGS_OK or MSGS_ERR_* (msgs_raster.h).
 */
#ifndef MSGS_STATS_H
#define MSGS_STATS_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSGS_STATS_PIXEL_SIZES 1 /* update_pixel_sizes: needs pixel_sizes, target_reso_lvl, max/min_pixel_sizes        */
#define MSGS_STATS_DENSIFY     2 /* max_radii2D + add_densification_stats: needs dL_dmeans2D, max_radii2D, accum, denom */
#define MSGS_STATS_BASE_MASK   4 /* update_base_gaussian_mask: needs base_mask                                         */

/* P Gaussians, L resolution levels (reso_lvls), reso_lvl in [0, L).  radii (P) int32; pixel_sizes (P) f32;
 * target_reso_lvl (P) int64; dL_dmeans2D (P,3) f32 (= viewspace_point_tensor.grad).  In place: max_pixel_sizes (P),
 * min_pixel_sizes (P), max_radii2D (P) f32; xyz_gradient_accum (P,L,1), denom (P,L,1) f32; base_mask (P) bool (1 byte). */
int msgs_update_stats(int P, int L, int reso_lvl, int what, const int32_t* radii, const float* pixel_sizes,
                      const int64_t* target_reso_lvl, const float* dL_dmeans2D, float* max_pixel_sizes,
                      float* min_pixel_sizes, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                      uint8_t* base_mask, void* stream);

/* ---- every unit of a step in ONE launch ----
 * A step renders up to 8 units (the scales of a view); unit u = (reso_lvl, what, radii, pixel_sizes, dL_dmeans2D).
 * msgs_update_stats_units applies them in order, exactly as n_units consecutive msgs_update_stats calls would. */
#define MSGS_STATS_MAX_UNITS 8
typedef struct {
    int reso_lvl, what;
    const int32_t* radii;
    const float* pixel_sizes;
    const float* dL_dmeans2D;
} msgs_stats_unit;

int msgs_update_stats_units(int P, int L, int n_units, const msgs_stats_unit* units, const int64_t* target_reso_lvl,
                            float* max_pixel_sizes, float* min_pixel_sizes, float* max_radii2D, float* xyz_gradient_accum,
                            float* denom, uint8_t* base_mask, void* stream);

/* ---- the sharded step (one process per GPU, each rank renders its own views) ----
 * The state must end the step identical on every rank, so a rank does not update it directly: its units write their
 * observations into a DELTA buffer, the ranks' deltas are reduced with ONE typed-lane collective
 * (msgs_stats_allreduce_multimem, msgs_optim.h) and every rank applies the reduced delta to its replica of the state.
 * Delta layout in 4-byte words (msgs_stats_delta_words fills n_sum / n_max / n_min, each a multiple of 4; the buffer is
 * their sum):
 *   sum lanes  [0, n_sum):          dacc (P,L) f32 | ddenom (P,L) f32
 *   max lanes  [n_sum, +n_max):     radius (P, padded to 4) | obs_hi (P) | base_seen (P)   non-negative f32, reduced as u32 max
 *   min lanes  [.., +n_min):        obs_lo (P)                                   positive f32 or +inf, reduced as u32 min
 * msgs_stats_delta_reset: sum / max lanes 0, min lanes +inf.  msgs_stats_delta_units: accumulate up to 8 units' observations.
 * msgs_stats_delta_apply: state update from the (reduced) delta; with ONE rank and one view per step the result is
 * bit-identical to msgs_update_stats_units; several views' observations of one Gaussian in one step act as one observation
 * with the largest / smallest size (the 0.95 / 1.05 ageing of gaussian_model.py:669-672 is applied once per step, not once
 * per view) and their gradient norms are summed before they are added (fp32 rounding only). */
void msgs_stats_delta_words(int P, int L, long long* n_sum, long long* n_max, long long* n_min);
/* The lanes the NEXT render depends on — obs_hi, base_seen (max) and obs_lo (min): they feed max / min_pixel_sizes and the base
 * mask, which the MS-GS filters of the rasterizer read — are the contiguous tail of the buffer: words [offset, offset + n_max +
 * n_min).  A sharded step may reduce only that tail every step (msgs_stats_allreduce_multimem(delta_mc + offset, 0, n_max,
 * n_min, ...): 3 words per Gaussian instead of 2 L + 4) and apply the rest — gradient-norm sums, visibility counts, radii: pure
 * sums / maxima that only densification reads (train.py:253-262, every densification_interval iterations) — to RANK-LOCAL
 * accumulators that are summed over the ranks when they are consumed. */
void msgs_stats_delta_fast_words(int P, int L, long long* offset, long long* n_max, long long* n_min);
int msgs_stats_delta_reset(int P, int L, float* delta, void* stream);
int msgs_stats_delta_units(int P, int L, int n_units, const msgs_stats_unit* units, const int64_t* target_reso_lvl,
                           float* delta, void* stream);
int msgs_stats_delta_apply(int P, int L, const int64_t* target_reso_lvl, const float* delta, float* max_pixel_sizes,
                           float* min_pixel_sizes, float* max_radii2D, float* xyz_gradient_accum, float* denom,
                           uint8_t* base_mask, void* stream);

#ifdef __cplusplus
}
#endif
#endif
