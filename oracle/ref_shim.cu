// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI shim around the *unmodified* reference CUDA rasterizer
// (/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu), compiled in place
// for sm_100a by oracle/Makefile into oracle/_ref/libref_rasterizer.so.  It lets tests/ and the
// bench's reference-CUDA leg drive CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (rasterizer.h:24-104) through ctypes without the 9-minute torch binding build, and exposes the
// private buffer carving (rasterizer_impl.h:29-66, rasterizer_impl.cu:155-196) so intermediate
// state (sorted keys, ranges, n_contrib, ...) can be compared bit-for-bit.
//
// Nothing here is copied from the reference: this file only *calls* it.
#include <cstdint>
#include <cstddef>
#include <functional>
#include <stdexcept>
#include <cstdio>
#include <cuda_runtime.h>
#include "rasterizer_impl.h"   // found via -I$(REF)/cuda_rasterizer

using namespace CudaRasterizer;

extern "C" {

typedef char* (*ref_alloc_fn)(void* ctx, size_t bytes);

int ref_forward(
    ref_alloc_fn geom_alloc, void* geom_ctx,
    ref_alloc_fn binning_alloc, void* binning_ctx,
    ref_alloc_fn img_alloc, void* img_ctx,
    int P, int D, int M,
    const float* background, int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* max_pixel_sizes, const float* min_pixel_sizes,
    const float* opacities, const float* occ_multiplier, const float* dc_delta,
    const float* scales, float scale_modifier, const float* rotations,
    const float* cov3D_precomp, const bool* base_mask,
    const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy,
    int prefiltered, int filter_small, int filter_large, float fade_size,
    float* out_color, float* out_acc_pixel_size, float* out_depth,
    int* radii, float* pixel_sizes, int debug)
{
    std::function<char*(size_t)> g = [=](size_t n) { return geom_alloc(geom_ctx, n); };
    std::function<char*(size_t)> b = [=](size_t n) { return binning_alloc(binning_ctx, n); };
    std::function<char*(size_t)> i = [=](size_t n) { return img_alloc(img_ctx, n); };
    try {
        return Rasterizer::forward(g, b, i, P, D, M, background, width, height, means3D, shs,
            colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities, occ_multiplier, dc_delta,
            scales, scale_modifier, rotations, cov3D_precomp, base_mask, viewmatrix, projmatrix,
            cam_pos, tan_fovx, tan_fovy, prefiltered != 0, filter_small != 0, filter_large != 0,
            fade_size, out_color, out_acc_pixel_size, out_depth, radii, pixel_sizes, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_forward] %s\n", e.what());
        return -1;
    }
}

int ref_backward(
    int P, int D, int M, int R,
    const float* background, int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
    const float* scales, float scale_modifier, const float* rotations,
    const float* cov3D_precomp, const bool* base_mask,
    const float* viewmatrix, const float* projmatrix, const float* campos,
    float tan_fovx, float tan_fovy, const int* radii,
    char* geom_buffer, char* binning_buffer, char* img_buffer,
    int filter_small, int filter_large, float fade_size,
    const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
    float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor, float* dL_dmean3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug)
{
    try {
        Rasterizer::backward(P, D, M, R, background, width, height, means3D, shs, colors_precomp,
            pixel_sizes, max_pixel_sizes, min_pixel_sizes, scales, scale_modifier, rotations,
            cov3D_precomp, base_mask, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
            geom_buffer, binning_buffer, img_buffer, filter_small != 0, filter_large != 0, fade_size,
            dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_docc_multiplier, dL_ddc_delta, dL_dcolor,
            dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_backward] %s\n", e.what());
        return -1;
    }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present)
{
    Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

// Addresses of the reference's private sub-arrays inside its three byte buffers.
// geom: depths, clamped, internal_radii, means2D, cov3D, conic_opacity, occ_mult_interp,
//       dc_delta_interp, rgb, tiles_touched, scanning_space, point_offsets
void ref_geom_layout(char* chunk, size_t P, uint64_t out[12])
{
    GeometryState g = GeometryState::fromChunk(chunk, P);
    out[0] = (uint64_t)g.depths;        out[1] = (uint64_t)g.clamped;
    out[2] = (uint64_t)g.internal_radii; out[3] = (uint64_t)g.means2D;
    out[4] = (uint64_t)g.cov3D;         out[5] = (uint64_t)g.conic_opacity;
    out[6] = (uint64_t)g.occ_mult_interp; out[7] = (uint64_t)g.dc_delta_interp;
    out[8] = (uint64_t)g.rgb;           out[9] = (uint64_t)g.tiles_touched;
    out[10] = (uint64_t)g.scanning_space; out[11] = (uint64_t)g.point_offsets;
}

// img: accum_alpha (final_T), n_contrib, ranges
void ref_image_layout(char* chunk, size_t N, uint64_t out[3])
{
    ImageState s = ImageState::fromChunk(chunk, N);
    out[0] = (uint64_t)s.accum_alpha; out[1] = (uint64_t)s.n_contrib; out[2] = (uint64_t)s.ranges;
}

// binning: point_list (sorted values), point_list_unsorted, point_list_keys (sorted), point_list_keys_unsorted
void ref_binning_layout(char* chunk, size_t R, uint64_t out[4])
{
    BinningState s = BinningState::fromChunk(chunk, R);
    out[0] = (uint64_t)s.point_list; out[1] = (uint64_t)s.point_list_unsorted;
    out[2] = (uint64_t)s.point_list_keys; out[3] = (uint64_t)s.point_list_keys_unsorted;
}

} // extern "C"
