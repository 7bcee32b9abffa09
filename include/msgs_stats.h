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

#ifdef __cplusplus
}
#endif
#endif
