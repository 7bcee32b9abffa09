/*
 * msgs_raster.h — C ABI of the B200-native MS-GS differentiable Gaussian rasterizer (libmsgs_raster.so).
 *
 * Drop-in boundary for the reference's C++ core API
 *   CudaRasterizer::Rasterizer::{forward, backward, markVisible}
 *   (/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:24-104)
 * which the reference's torch binding (rasterize_points.cu:35-262, ext.cpp:15-19) calls.  Differences from
 * the C++ signature, all forced by the C ABI or by stream-awareness:
 *   - std::function<char*(size_t)> allocators  ->  (msgs_alloc_fn, void* ctx) pairs;
 *   - `bool` arguments -> int (0/1); `const bool* base_mask` -> const uint8_t* (1 byte per Gaussian, as torch.bool);
 *   - an explicit cudaStream_t (passed as void*) on every call: all work is enqueued on that stream;
 *   - errors are returned (negative) instead of thrown; msgs_last_error() returns the message.
 * Optional inputs are signalled by NULL exactly as the reference signals them by empty tensors
 * (forward.cu:221,375; rasterizer_impl.cu:345,430,462; backward.cu:455,467).
 *
 * All pointers except the allocator contexts are DEVICE pointers (fp32 unless stated).  No torch types.
 * The library has no CPU fallback: every entry point fails with MSGS_ERR_CUDA when no device is usable.
 */
#ifndef MSGS_RASTER_H
#define MSGS_RASTER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSGS_ABI_VERSION 1

#define MSGS_OK 0
#define MSGS_ERR_ARG  (-1) /* bad argument (e.g. NUM_CHANNELS!=3 analogue, NULL required pointer)       */
#define MSGS_ERR_CUDA (-2) /* a CUDA call or launch failed; see msgs_last_error()                        */
#define MSGS_ERR_ALLOC (-3)/* an allocator callback returned NULL                                         */

/* Replaces std::function<char*(size_t)> (rasterizer.h:33-35).  Must return device memory of at least
 * `bytes` bytes, 128-byte aligned, valid until the matching backward has run.  Called once per forward for
 * the geometry and the image buffer, then once or twice for the binning buffer (rasterizer_impl.cu:240, 253,
 * 308): the library may first ask for a speculative size (so that it can enqueue placement and blend before it
 * reads num_rendered back) and a second time, with the exact size, if the guess was too small; the buffer of the
 * LAST call is the one the backward must be given. */
typedef void* (*msgs_alloc_fn)(void* ctx, size_t bytes);

/* A ready-made msgs_alloc_fn for bindings that know the size in advance (msgs_geom_bytes, msgs_image_bytes,
 * msgs_binning_hint_bytes): pass msgs_fixed_alloc with a msgs_fixed_buffer* as ctx.  A request that fits `bytes` is served
 * from `ptr` (served_fixed = 1), any other goes to `fallback` (served_fixed = 0); `requested` = the size asked for. */
typedef struct msgs_fixed_buffer {
    void* ptr; size_t bytes;
    msgs_alloc_fn fallback; void* fallback_ctx;
    size_t requested; int served_fixed;
} msgs_fixed_buffer;
void* msgs_fixed_alloc(void* ctx, size_t bytes);
/* bytes of the speculative binning buffer the next forward of this (P, width, height) will request on this host thread
 * (0 = no history yet: the forward will ask for the exact size after its read-back) */
size_t msgs_binning_hint_bytes(int P, int width, int height);

int msgs_abi_version(void);
const char* msgs_last_error(void);

/* Replaces Rasterizer::markVisible (rasterizer.h:26-31; kernel checkFrustum rasterizer_impl.cu:54-66):
 * present[i] = (view-space z of means3D[i] > 0.2). */
int msgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream);

/* Replaces Rasterizer::forward (rasterizer.h:33-68; rasterizer_impl.cu:200-367).
 * Returns num_rendered (>= 0: number of (Gaussian, tile) instances) or a negative MSGS_ERR_*.
 * Outputs: out_color (3,H,W), out_acc_pixel_size (H,W), out_depth (H,W), radii int32 (P), pixel_sizes (P);
 * every element of every output is written (no pre-zeroing needed).  occ_multiplier / dc_delta / fade_size
 * are accepted and ignored exactly as the reference ignores them (compile-time dc_occ_mult=false,
 * forward.cu:291,469).  `debug` != 0 synchronises and checks after every stage (auxiliary.h:171-178). */
int msgs_forward(msgs_alloc_fn geom_alloc, void* geom_ctx,
                 msgs_alloc_fn binning_alloc, void* binning_ctx,
                 msgs_alloc_fn img_alloc, void* img_ctx,
                 int P, int D, int M,
                 const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* max_pixel_sizes, const float* min_pixel_sizes,
                 const float* opacities, const float* occ_multiplier, const float* dc_delta,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const uint8_t* base_mask,
                 const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                 float tan_fovx, float tan_fovy,
                 int prefiltered, int filter_small, int filter_large, float fade_size,
                 float* out_color, float* out_acc_pixel_size, float* out_depth,
                 int* radii, float* pixel_sizes, int debug, void* stream);

/* Extension (several independent (view, scale) units of a step in flight): msgs_forward in two halves.
 * msgs_forward_begin takes the arguments of msgs_forward, enqueues every stage on `stream` — placement + blend into a
 * speculatively sized binning buffer included — and returns a handle (>= 0, at most 64 per host thread) WITHOUT waiting
 * for the device; msgs_forward_finish(handle), called from the same host thread, waits for that call's 4-byte
 * read-back only, launches placement + blend again with an exactly sized buffer if the speculation missed (or there
 * was no history), releases the handle and returns num_rendered as msgs_forward does.  msgs_forward is begin + finish.
 * The reference's single blocking cudaMemcpy (rasterizer_impl.cu:305) thereby stops serialising independent calls:
 * a caller begins all units of a step on their own streams and then finishes them. */
int msgs_forward_begin(msgs_alloc_fn geom_alloc, void* geom_ctx,
                       msgs_alloc_fn binning_alloc, void* binning_ctx,
                       msgs_alloc_fn img_alloc, void* img_ctx,
                       int P, int D, int M,
                       const float* background, int width, int height,
                       const float* means3D, const float* shs, const float* colors_precomp,
                       const float* max_pixel_sizes, const float* min_pixel_sizes,
                       const float* opacities, const float* occ_multiplier, const float* dc_delta,
                       const float* scales, float scale_modifier, const float* rotations,
                       const float* cov3D_precomp, const uint8_t* base_mask,
                       const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                       float tan_fovx, float tan_fovy,
                       int prefiltered, int filter_small, int filter_large, float fade_size,
                       float* out_color, float* out_acc_pixel_size, float* out_depth,
                       int* radii, float* pixel_sizes, int debug, void* stream);
int msgs_forward_finish(int handle);

/* Extension (SURVEY.md §8e "the scale-independent per-view preprocess can be shared across that view's scales"):
 * the forward of a VIEW GROUP — n_units (<= 8) units that share the camera (viewmatrix, projmatrix, cam_pos, tan_fov),
 * the Gaussians and every flag, and differ only in the image size (the scales of one view of train.py --ms_train,
 * utils/camera_utils.py:38-39).  The per-Gaussian loads, frustum test, projection, cov3D, SH colour and depth key are
 * computed ONCE by one preprocess launch that writes all units' records, and the Gaussians are depth-sorted ONCE (the
 * depth order of a view does not depend on the resolution); counting, placement and blend then run per unit on that
 * unit's stream.  Every unit's outputs and state buffers are bit-identical in meaning to a separate msgs_forward_begin
 * of the unit (the state buffers feed msgs_backward* unchanged), except that unit 0's geometry buffer also holds the
 * group's depth order, which the other units' placement reads: keep it alive until every unit is finished.
 * units[i].stream: preprocess + sort run on units[0].stream, the other streams wait for them by an event.
 * On success handles[0..n_units) receive one handle per unit, each to be passed to msgs_forward_finish. */
typedef struct msgs_group_unit {
    int width, height;
    const float* background;
    float* out_color; float* out_acc_pixel_size; float* out_depth;
    int* radii; float* pixel_sizes;
    msgs_alloc_fn geom_alloc; void* geom_ctx;
    msgs_alloc_fn binning_alloc; void* binning_ctx;
    msgs_alloc_fn img_alloc; void* img_ctx;
    void* stream;
} msgs_group_unit;
int msgs_forward_group_begin(int n_units, const msgs_group_unit* units, int* handles,
                             int P, int D, int M,
                             const float* means3D, const float* shs, const float* colors_precomp,
                             const float* max_pixel_sizes, const float* min_pixel_sizes,
                             const float* opacities, const float* occ_multiplier, const float* dc_delta,
                             const float* scales, float scale_modifier, const float* rotations,
                             const float* cov3D_precomp, const uint8_t* base_mask,
                             const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                             float tan_fovx, float tan_fovy,
                             int prefiltered, int filter_small, int filter_large, float fade_size, int debug);

/* Extension: the preprocess backward of a view group in one pass (see msgs_forward_group_begin).  For every unit the
 * caller first runs the blend backward alone — msgs_backward_accumulate_part(..., part = 1, 0, -1) on the unit's stream —
 * and makes `stream` wait for them; this call then adds the group's training gradients (dL_dmean3D, dL_dsh, dL_dopacity,
 * dL_dscale, dL_drot; caller-zeroed accumulators as for msgs_backward_accumulate) for the Gaussians any unit rendered and
 * writes every unit's dL_dmean2D (P,3; may be NULL).  The resolution-independent linear maps (projection, SH and cov3D
 * backward) run once on the per-unit sums: one SH row read and one gradient row reduction per Gaussian and view.  Equals
 * the sum of the units' msgs_backward_accumulate calls up to fp32 summation order.  SH colours with M == 16 only. */
typedef struct msgs_group_bwd_unit {
    int width, height;
    const int* radii; const float* pixel_sizes;   /* the unit's forward outputs */
    void* geom_buffer;                            /* the unit's geometry state buffer */
    float* dL_dmean2D;                            /* (P,3) or NULL */
} msgs_group_bwd_unit;
int msgs_backward_group_accumulate(int n_units, const msgs_group_bwd_unit* units, int P, int D, int M,
                                   const float* means3D, const float* shs,
                                   const float* max_pixel_sizes, const float* min_pixel_sizes,
                                   const float* scales, float scale_modifier, const float* rotations,
                                   const float* cov3D_precomp, const uint8_t* base_mask,
                                   const float* viewmatrix, const float* projmatrix, const float* campos,
                                   float tan_fovx, float tan_fovy, int filter_small, int filter_large,
                                   float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                   int debug, void* stream);

/* Extension (the gradient all-reduce of the sharded training step): two-shot all-reduce (fp32 sum, in place)
 * of a symmetric buffer over the NVSwitch multicast mapping, hand-written with multimem.ld_reduce / multimem.st
 * (exchange.cu).  multicast_ptr: the buffer's address inside the multicast mapping (16-byte aligned); n_floats: a
 * multiple of 4; rank r reduces and broadcasts the r-th of world_size slices.  ctas <= 0: one CTA per SM.  The caller
 * brackets the call with stream-ordered barriers across the ranks (msgs_sharding.SymmetricGradientExchange.allreduce). */
int msgs_allreduce_multimem(void* multicast_ptr, size_t n_floats, int rank, int world_size, int ctas, void* stream);

/* Replaces Rasterizer::backward (rasterizer.h:70-104; rasterizer_impl.cu:371-494).
 * geom/binning/image buffers are the ones the allocators returned in the matching forward (same P, R, W, H).
 * Gradient outputs: dL_dmean2D (P,3; z==0), dL_dconic (P,4; [2] never accumulated, reference Q10),
 * dL_dopacity (P), dL_docc_multiplier (P,4) == 0, dL_ddc_delta (P,12) == 0, dL_dcolor (P,3), dL_dmean3D (P,3),
 * dL_dcov3D (P,6), dL_dsh (P,M,3), dL_dscale (P,3), dL_drot (P,4).  Every element of every non-NULL output
 * is written by the call (zeros where the reference leaves its torch::zeros untouched,
 * rasterize_points.cu:184-194), so callers may pass uninitialised memory.  dL_docc_multiplier,
 * dL_ddc_delta, dL_dconic, dL_dmean2D, dL_dcolor and dL_dcov3D may be NULL (then skipped: a binding that knows which
 * inputs need a gradient — autograd's needs_input_grad — saves their 12-64 B per Gaussian of writes). */
int msgs_backward(int P, int D, int M, int R,
                  const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const uint8_t* base_mask,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tan_fovx, float tan_fovy, const int* radii,
                  void* geom_buffer, void* binning_buffer, void* image_buffer,
                  int filter_small, int filter_large, float fade_size,
                  const float* dL_dpix,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int debug, void* stream);

/* Extension for steps made of several (view, scale) units (no reference counterpart: the reference renders one unit
 * per optimiser step and lets autograd sum): same arguments as msgs_backward, but the five gradients that training
 * sums over the units of a step — dL_dmean3D, dL_dsh, dL_dopacity, dL_dscale, dL_drot — are ADDED to the contents of
 * the output arrays, and only for the Gaussians this unit rendered (radii > 0); the caller zeroes them once per step.
 * dL_dmean2D is written for every Gaussian as in msgs_backward (per-view densification statistic) unless NULL;
 * dL_dcolor, dL_dcov3D, dL_dconic, dL_docc_multiplier, dL_ddc_delta may be NULL (skipped) and are written, not added. */
int msgs_backward_accumulate(int P, int D, int M, int R,
                  const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const uint8_t* base_mask,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tan_fovx, float tan_fovy, const int* radii,
                  void* geom_buffer, void* binning_buffer, void* image_buffer,
                  int filter_small, int filter_large, float fade_size,
                  const float* dL_dpix,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int debug, void* stream);

/* msgs_backward_accumulate in pieces, so that a multi-GPU step can overlap its gradient exchange with the backward
 * (no reference counterpart).  part bit 0: run the blend backward of this unit (fills the per-Gaussian accumulators in
 * the geometry buffer); part bit 1: run the per-Gaussian preprocess backward over the Gaussians [first, first + count)
 * only (first a multiple of 128; count < 0 = up to P) and add their training gradients as msgs_backward_accumulate does.
 * The view-group path calls part 1 for every unit of a view and then ONE msgs_backward_group_accumulate for the group;
 * part 3 over [0, P) is msgs_backward_accumulate. */
int msgs_backward_accumulate_part(int P, int D, int M, int R,
                  const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const uint8_t* base_mask,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tan_fovx, float tan_fovy, const int* radii,
                  void* geom_buffer, void* binning_buffer, void* image_buffer,
                  int filter_small, int filter_large, float fade_size,
                  const float* dL_dpix,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int debug, void* stream, int part, int first, int count);

/* Raw-parameter entry points (gaussian_renderer.render()'s activations fused into the op).  Same
 * arguments as their namesakes, but `scales`, `rotations` and `opacities` are GaussianModel's RAW parameters (_scaling,
 * _rotation, _opacity; scene/gaussian_model.py:127-155) and `shs` is the (P,16,3) concatenation of _features_dc and
 * _features_rest: exp / F.normalize / sigmoid (gaussian_renderer/__init__.py:57-89) run in registers inside the forward
 * preprocess, and the backward chains their derivatives, so the gradients written / accumulated are those of the raw
 * parameters — no activated copies, no separate activation passes (2 x 448 B per Gaussian of traffic and 2 launches less
 * than msgs_activate + msgs_forward; results bit-identical to that composition).  The backward needs the raw opacities
 * (`opacity_raw`, P floats) for sigmoid'.  Scale/rotation inputs only (cov3D_precomp must be NULL). */
int msgs_forward_begin_raw(msgs_alloc_fn geom_alloc, void* geom_ctx,
                 msgs_alloc_fn binning_alloc, void* binning_ctx,
                 msgs_alloc_fn img_alloc, void* img_ctx,
                 int P, int D, int M,
                 const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* max_pixel_sizes, const float* min_pixel_sizes,
                 const float* opacities, const float* occ_multiplier, const float* dc_delta,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const uint8_t* base_mask,
                 const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                 float tan_fovx, float tan_fovy,
                 int prefiltered, int filter_small, int filter_large, float fade_size,
                 float* out_color, float* out_acc_pixel_size, float* out_depth,
                 int* radii, float* pixel_sizes, int debug, void* stream);
int msgs_forward_group_begin_raw(int n_units, const msgs_group_unit* units, int* handles, int P, int D, int M,
                                 const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* max_pixel_sizes, const float* min_pixel_sizes,
                                 const float* opacities, const float* occ_multiplier, const float* dc_delta,
                                 const float* scales, float scale_modifier, const float* rotations,
                                 const float* cov3D_precomp, const uint8_t* base_mask,
                                 const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                                 float tan_fovx, float tan_fovy,
                                 int prefiltered, int filter_small, int filter_large, float fade_size, int debug);
/* accumulate: 0 = every gradient element written once (msgs_backward), 1 = training gradients added (msgs_backward_accumulate);
 * part: bit 0 = blend backward, bit 1 = preprocess backward (3 = both). */
int msgs_backward_raw(const float* opacity_raw, int accumulate, int part, int P, int D, int M, int R,
                  const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const uint8_t* base_mask,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tan_fovx, float tan_fovy, const int* radii,
                  void* geom_buffer, void* binning_buffer, void* image_buffer,
                  int filter_small, int filter_large, float fade_size,
                  const float* dL_dpix,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int debug, void* stream);
int msgs_backward_group_accumulate_raw(const float* opacity_raw, int n_units, const msgs_group_bwd_unit* units, int P, int D, int M,
                                   const float* means3D, const float* shs,
                                   const float* max_pixel_sizes, const float* min_pixel_sizes,
                                   const float* scales, float scale_modifier, const float* rotations,
                                   const float* cov3D_precomp, const uint8_t* base_mask,
                                   const float* viewmatrix, const float* projmatrix, const float* campos,
                                   float tan_fovx, float tan_fovy, int filter_small, int filter_large,
                                   float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                   int debug, void* stream);

/* Extension: the WHOLE backward of a view group in one call.  blends[i] belongs to units[i]: the unit's blend backward
 * (msgs_backward_raw(..., part = 1)) is launched on blends[i].stream behind what `stream` holds at the call, coarsest unit
 * first, and `stream` waits for all of them before the group's preprocess backward (msgs_backward_group_accumulate[_raw])
 * runs on it.  opacity_raw NULL: activated-parameter gradients; else raw-parameter gradients.  One host call per view
 * instead of one per unit plus one. */
typedef struct msgs_group_bwd_blend {
    const float* background;      /* (3) */
    void* binning_buffer;         /* the unit's forward state buffers */
    void* image_buffer;
    int R;                        /* the unit's num_rendered */
    const float* dL_dpix;         /* (3, height, width) */
    void* stream;                 /* cudaStream_t of the unit */
} msgs_group_bwd_blend;
int msgs_backward_group_full(const float* opacity_raw, const msgs_group_bwd_blend* blends,
                             int n_units, const msgs_group_bwd_unit* units, int P, int D, int M,
                             const float* means3D, const float* shs,
                             const float* max_pixel_sizes, const float* min_pixel_sizes,
                             const float* scales, float scale_modifier, const float* rotations,
                             const float* cov3D_precomp, const uint8_t* base_mask,
                             const float* viewmatrix, const float* projmatrix, const float* campos,
                             float tan_fovx, float tan_fovy, int filter_small, int filter_large,
                             float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                             int debug, void* stream);

/* The two halves of msgs_backward_group_full on their own (sharded training step: the gradient exchange of a slice of the
 * Gaussians runs while the next slice's preprocess backward is being computed).  msgs_backward_group_blend: the blend half —
 * `stream` ends up waiting for every unit's blend backward.  msgs_backward_group_accumulate_part: Gaussians
 * [first, first + count) of the group's preprocess backward (first a multiple of 128; count < 0: to the end); opacity_raw NULL:
 * activated-parameter gradients. */
int msgs_backward_group_blend(const float* opacity_raw, const msgs_group_bwd_blend* blends,
                              int n_units, const msgs_group_bwd_unit* units, int P, int D, int M,
                              const float* means3D, const float* shs,
                              const float* max_pixel_sizes, const float* min_pixel_sizes,
                              const float* scales, float scale_modifier, const float* rotations,
                              const float* cov3D_precomp, const uint8_t* base_mask,
                              const float* viewmatrix, const float* projmatrix, const float* campos,
                              float tan_fovx, float tan_fovy, int filter_small, int filter_large,
                              float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                              int debug, void* stream);
int msgs_backward_group_accumulate_part(const float* opacity_raw, int first, int count,
                              int n_units, const msgs_group_bwd_unit* units, int P, int D, int M,
                              const float* means3D, const float* shs,
                              const float* max_pixel_sizes, const float* min_pixel_sizes,
                              const float* scales, float scale_modifier, const float* rotations,
                              const float* cov3D_precomp, const uint8_t* base_mask,
                              const float* viewmatrix, const float* projmatrix, const float* campos,
                              float tan_fovx, float tan_fovy, int filter_small, int filter_large,
                              float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                              int debug, void* stream);

/* Speculation statistics of this host thread (msgs_forward_begin sizes the binning buffer from recent num_rendered values
 * of the same (P, W, H) before the read-back): out[0] = forwards finished, out[1] = those whose speculative buffer was
 * large enough (placement + blend ran once), out[2] = those that had to run them again.  reset != 0 clears the counters. */
void msgs_speculation_stats(long long out[3], int reset);

/* Buffer sizes the allocators will be asked for (analogue of required<T>(), rasterizer_impl.h:68-74). */
size_t msgs_geom_bytes(int P);
size_t msgs_image_bytes(int width, int height, int P);
size_t msgs_binning_bytes(int R);

/* Introspection of the private state layout, for parity tests only (the reference's equivalent is
 * GeometryState/ImageState/BinningState::fromChunk, rasterizer_impl.cu:155-196).
 * geom   out[0..7]: splat records (P x 4 float4: {x,y,depth,blend_px},{conic.x,conic.y,conic.z,opacity},
 *                   {r,g,b,bwd_skip},{cull hx,hy,thr,-}), cov3D (P x 6), clamped (P bytes, bits 0..2), tiles_touched u32,
 *                   tile rectangles (P x 4 u16: xmin,ymin,xmax,ymax), grad accumulators (P x 12 f32),
 *                   depth order (P u32 Gaussian indices), meta (16 u32, [0] = num_rendered)
 * image  out[0..3]: final_T f32 (W*H), n_contrib u32 (W*H), ranges uint2 (tiles), tile_max_contrib u32 (tiles)
 * binning out[0]:   point_list (R u32, tile-major, depth order inside a tile); out[1..3] = 0.  The library never
 *                   materialises the reference's 64-bit keys: the sorted key of list position i is
 *                   (tile of i from `ranges`) << 32 | float bits of the depth of point_list[i]. */
void msgs_geom_layout(void* chunk, int P, uint64_t out[8]);
void msgs_image_layout(void* chunk, int width, int height, int P, uint64_t out[4]);
void msgs_binning_layout(void* chunk, int R, uint64_t out[4]);

/* Stand-alone entry to the hand-written onesweep radix sort (replaces cub::DeviceRadixSort::SortPairs at
 * rasterizer_impl.cu:327-332): stable ascending sort of n (u64 key, u32 value) pairs on key bits
 * [0, end_bit).  temp must hold msgs_sort_temp_bytes(n) bytes.  The forward uses the same kernels (u32 keys) to
 * put the Gaussians in depth order; this entry exposes them for callers and tests. */
size_t msgs_sort_temp_bytes(int n);
int msgs_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                    uint64_t* keys_alt, uint32_t* vals_alt,
                    int n, int end_bit, void* temp, void* stream);

/* msgs_sort_pairs without any allocation, by the library or by the caller: the onesweep scratch (digit histograms, tile
 * tickets, look-back words) lives in a static __device__ array inside the library (540 KiB, held by every GPU that
 * loads the library until the process exits), so code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc can sort.  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, result and errors as msgs_sort_pairs minus `temp`; the caller still passes the input, output and
 * ping-pong (`_alt`) buffers.  Limit: n <= MSGS_SORT_STATIC_MAX_N pairs, beyond it MSGS_ERR_ARG with a message in
 * msgs_last_error() (use msgs_sort_pairs with your own temp).  Calls no cudaMalloc / cudaFree / allocating routine:
 * two async D2D copies, one memset of the part of the scratch this call reads (the array keeps the last call's
 * contents), then the histogram, scan and pass kernels on `stream`.  Every call on a device shares the one array:
 * calls that overlap on the GPU (different streams, or several host threads without ordering) RACE — order them on
 * one stream or with events.  The scratch's device address is looked up once per device (cudaGetSymbolAddress, with
 * that device current) on the first call there. */
#define MSGS_SORT_STATIC_MAX_N (1 << 21)
int msgs_sort_pairs_static(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                           uint64_t* keys_alt, uint32_t* vals_alt, int n, int end_bit, void* stream);

/* Measurement hooks (bench.py): number of kernels launched by this library so far, and optional per-stage
 * CUDA-event timing on the caller's stream.  With profiling enabled every stage is bracketed by two events;
 * msgs_profile_collect() waits for them and ADDS elapsed ms / launch counts per stage into the arrays. */
long long msgs_launch_count(void);
int msgs_profile_num_stages(void);
const char* msgs_profile_stage_name(int i);
void msgs_profile_enable(int on);
int msgs_profile_collect(double* ms, long long* counts, int n);

#ifdef __cplusplus
}
#endif
#endif /* MSGS_RASTER_H */
