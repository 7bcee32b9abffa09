/* msgs_activations.h — C ABI of the fused parameter-activation kernels.
 *
 * Replaces, on the caller side of the rasterizer, what MS-GS's `gaussian_renderer.render()` runs as separate PyTorch
 * kernels over all P Gaussians on every call (gaussian_renderer/__init__.py:57-89):
 *   pc.get_scaling   = exp(_scaling)                               scene/gaussian_model.py:127-129, 37
 *   pc.get_rotation  = torch.nn.functional.normalize(_rotation)    scene/gaussian_model.py:131-133, 47   (x / max(|x|, 1e-12))
 *   pc.get_opacity   = sigmoid(_opacity)                           scene/gaussian_model.py:153-155, 44
 *   pc.get_features  = cat((_features_dc, _features_rest), dim=1)  scene/gaussian_model.py:143-150
 * and their backward (exp / sigmoid / normalize backward, the split of dL_dshs into the two feature tensors).
 * ONE launch forward and ONE backward; plain pointers, fp32, row-major contiguous tensors, device memory, `stream` is a
 * cudaStream_t.  Exported by libmsgs_raster.so.  Return MSGS_OK (0) or a negative MSGS_ERR_* of msgs_raster.h.
 * No CPU fallback.
 */
#ifndef MSGS_ACTIVATIONS_H
#define MSGS_ACTIVATIONS_H

#ifdef __cplusplus
extern "C" {
#endif

/* Forward.  D = number of "rest" SH coefficients per Gaussian (15 for degree 3); shs is (P, 1 + D, 3).
 * features_dc (P,1,3), features_rest (P,D,3), scaling_raw (P,3), rotation_raw (P,4), opacity_raw (P,1)
 * -> shs (P,1+D,3), scales (P,3), rotations (P,4), opacities (P,1).  Any D >= 0 (SH degree 0-3: D = 0, 3, 8, 15). */
int msgs_activate(int P, int D, const float* features_dc, const float* features_rest, const float* scaling_raw,
                  const float* rotation_raw, const float* opacity_raw, float* shs, float* scales, float* rotations,
                  float* opacities, void* stream);

/* Backward.  Takes the forward OUTPUTS scales / opacities (exp' = scales, sigmoid' = o (1 - o)), the raw rotation, and
 * the gradients of the four outputs (any of them may be NULL = zero gradient); writes the gradients of the five inputs
 * (every element written once; an output pointer may be NULL to skip it). */
int msgs_activate_backward(int P, int D, const float* scales, const float* rotation_raw, const float* opacities,
                           const float* dL_dshs, const float* dL_dscales, const float* dL_drotations,
                           const float* dL_dopacities, float* dL_dfeatures_dc, float* dL_dfeatures_rest,
                           float* dL_dscaling_raw, float* dL_drotation_raw, float* dL_dopacity_raw, void* stream);

#ifdef __cplusplus
}
#endif
#endif
