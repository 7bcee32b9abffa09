// Shared declarations of libmsgs_raster.so (B200 / sm_100a).  Private to ms-gs_b200/csrc.
//
// State layout (opaque to callers; replaces GeometryState / ImageState / BinningState of the reference,
// submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.h:29-66 — "private layout, free
// to redesign").  Redesign for HBM3e/L2:
//   * per-Gaussian blend inputs live in ONE 48-byte "splat record" (3 x float4) so that the per-tile
//     gathers of the blend kernels touch 2 sectors instead of 5 separate arrays;
//   * the 9 atomically accumulated gradients of the blend backward live in one 48-byte record
//     (one L2 line region per Gaussian) inside the geometry buffer;
//   * sort keys are compacted: (tile << depth_bits) | (depth_bits_as_uint - min_depth_bits), an
//     order-preserving remap of the reference key (tile << 32 | depth_bits, rasterizer_impl.cu:98-108)
//     that needs fewer radix passes.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

namespace msgs {

constexpr int TILE_X = 16;              // config.h:16-17 — baked into the key/range contract
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;

// ---- splat record -------------------------------------------------------------------------------
// rec[3*i+0] = {mean2D.x, mean2D.y, depth (view z), blend_pixel_size (Q6, forward.cu:499-506)}
// rec[3*i+1] = {conic.x, conic.y, conic.z, opacity}                       (forward.cu:388)
// rec[3*i+2] = {r, g, b, bwd_skip (1.0f if the backward's per-pair filter re-check skips it, Q2)}

struct GeomState {
    float4*   rec;            // P x 3
    float*    cov3D;          // P x 6   (computed from scale/rot when not precomputed)
    uint8_t*  clamped;        // P       bits 0..2 = channel clamped at 0 (forward.cu:71-73)
    uint32_t* tiles_touched;  // P
    uint32_t* point_offsets;  // P       inclusive scan of tiles_touched
    float*    gacc;           // P x 12  blend-backward accumulators: color[3], mean2D[2], conic[3], opacity, pad[3]
    uint32_t* scan_state;     // decoupled look-back tile descriptors (u64 each) + counters
    uint32_t* meta;           // 16 words: [0]=num_rendered [1]=min depth bits [2]=max depth bits [3..] = scratch
};

struct ImageState {
    float*    final_T;        // W*H
    uint32_t* n_contrib;      // W*H
    uint2*    ranges;         // tiles
    uint32_t* tile_max_contrib; // tiles: max n_contrib over the tile's pixels (backward start point)
};

struct BinningState {
    uint32_t* vals[2];        // [0] = point_list (sorted result), [1] = alternate
    uint64_t* keys[2];        // [0] = sorted compacted keys, [1] = alternate
    void*     sort_temp;
};

constexpr size_t SCAN_TILE = 2048;      // Gaussians per scan tile (256 threads x 8)

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
    p = (char*)align_up((size_t)p, 128);
    out = (T*)p;
    p += count * sizeof(T);
}

size_t sort_temp_bytes(size_t n);

inline GeomState geom_from_chunk(char* chunk, size_t P) {
    GeomState g;
    char* p = chunk;
    carve(p, g.rec, 3 * P);
    carve(p, g.cov3D, 6 * P);
    carve(p, g.clamped, P);
    carve(p, g.tiles_touched, P);
    carve(p, g.point_offsets, P);
    carve(p, g.gacc, 12 * P);
    carve(p, g.scan_state, 2 * ((P + SCAN_TILE - 1) / SCAN_TILE + 1) + 32);
    carve(p, g.meta, 16);
    return g;
}
inline size_t geom_bytes(size_t P) {
    GeomState g = geom_from_chunk((char*)0, P);
    return (size_t)((char*)(g.meta + 16) - (char*)0) + 128;
}

inline ImageState image_from_chunk(char* chunk, int W, int H) {
    ImageState s;
    char* p = chunk;
    const size_t N = (size_t)W * H;
    const size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    carve(p, s.final_T, N);
    carve(p, s.n_contrib, N);
    carve(p, s.ranges, T);
    carve(p, s.tile_max_contrib, T);
    return s;
}
inline size_t image_bytes(int W, int H) {
    ImageState s = image_from_chunk((char*)0, W, H);
    const size_t T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    return (size_t)((char*)(s.tile_max_contrib + T) - (char*)0) + 128;
}

inline BinningState binning_from_chunk(char* chunk, size_t R) {
    BinningState b;
    char* p = chunk;
    carve(p, b.vals[0], R);
    carve(p, b.vals[1], R);
    carve(p, b.keys[0], R);
    carve(p, b.keys[1], R);
    char* t = nullptr;
    carve(p, t, sort_temp_bytes(R));
    b.sort_temp = t;
    return b;
}
inline size_t binning_bytes(size_t R) {
    BinningState b = binning_from_chunk((char*)0, R);
    return (size_t)((char*)b.sort_temp - (char*)0) + sort_temp_bytes(R) + 128;
}

// ---- per-call constants ---------------------------------------------------------------------------
struct FwdParams {
    int P, D, M, W, H, gx, gy;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int prefiltered, filter_small, filter_large;
    const float *view, *proj, *campos;   // device pointers (4x4, 4x4, 3)
};

// ---- stage launchers (each enqueues on `st`; definitions in the matching .cu) -----------------------
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);

void launch_preprocess_fwd(const FwdParams& prm, const float* means3D, const float* scales, const float* rotations,
                           const float* opacities, const float* shs, const float* cov3D_precomp,
                           const float* colors_precomp, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, int* radii, float* pixel_sizes, GeomState g, cudaStream_t st);

void launch_init_state(GeomState g, int P, cudaStream_t st);
void launch_scan(GeomState g, int P, cudaStream_t st);

void launch_duplicate(GeomState g, int P, int gx, int gy, const int* radii, uint32_t min_depth_bits, int depth_bits,
                      uint64_t* keys, uint32_t* vals, cudaStream_t st);

// stable LSD onesweep sort of n pairs on key bits [0,end_bit); result ends in (keys_out, vals_out).
// The input must already be in the buffer sort_input_index(end_bit) says (0 = *_out, 1 = *_alt).
int  sort_num_passes(int end_bit);
int  sort_input_index(int end_bit);
void launch_sort_histogram(const uint64_t* keys_in, size_t n, int end_bit, void* temp, cudaStream_t st);
void launch_sort_pass(uint64_t* keys_out, uint32_t* vals_out, uint64_t* keys_alt, uint32_t* vals_alt,
                      size_t n, int end_bit, int p, void* temp, cudaStream_t st);
void launch_sort(uint64_t* keys_out, uint32_t* vals_out, uint64_t* keys_alt, uint32_t* vals_alt,
                 size_t n, int end_bit, void* temp, cudaStream_t st);

void launch_tile_ranges(const uint64_t* keys_sorted, size_t R, int depth_bits, uint2* ranges, int tiles, cudaStream_t st);

void launch_blend_fwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, float* final_T, uint32_t* n_contrib,
                      uint32_t* tile_max_contrib, float* out_color, float* out_acc_px, float* out_depth,
                      cudaStream_t st);

void launch_blend_bwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, const float* final_T, const uint32_t* n_contrib,
                      const uint32_t* tile_max_contrib, const float* dL_dpix, float* gacc, cudaStream_t st);

struct BwdParams {
    int P, D, M, W, H;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int filter_small, filter_large;
    const float *view, *proj, *campos;   // device pointers
};
void launch_preprocess_bwd(const BwdParams& prm, const float* means3D, const float* shs, const float* scales,
                           const float* rotations, const float* cov3D_precomp, const int* radii,
                           const float* pixel_sizes, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, GeomState g,
                           float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_docc, float* dL_ddc,
                           float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                           float* dL_dscale, float* dL_drot, cudaStream_t st);

}  // namespace msgs
