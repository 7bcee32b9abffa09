// Shared declarations of libmsgs_raster.so (B200 / sm_100a).  Private to ms-gs_b200/csrc.
//
// State layout (opaque to callers; replaces GeometryState / ImageState / BinningState of the reference,
// submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.h:29-66 — "private layout, free
// to redesign").  Redesign for HBM3e/L2:
//   * per-Gaussian blend inputs live in ONE 48-byte "splat record" (3 x float4) so that the per-tile
//     gathers of the blend kernels touch 2 sectors instead of 5 separate arrays;
//   * the 9 atomically accumulated gradients of the blend backward live in one 48-byte record
//     (one L2 line region per Gaussian) inside the geometry buffer;
//   * binning does NOT sort R (Gaussian, tile) instances on a 64-bit (tile | depth) key as the reference
//     does (rasterizer_impl.cu:98-108, 327-332).  The same order — stable by (tile, depth bits, Gaussian
//     index) — is produced by (1) a radix sort of the P Gaussians by their 32 depth bits, (2) emission of
//     the instances in that depth order, (3) a stable radix partition of the instances by tile id only
//     (1-2 digits instead of 5-6), with the per-tile ranges and the digit histograms taken from per-tile
//     counts that come from a 4-corner scatter + 2-D prefix sum instead of from the instances.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

namespace msgs {

constexpr int TILE_X = 16;              // config.h:16-17 — baked into the key/range contract
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int SORT_MAX_PASSES = 7;

struct SortPlan { int passes; int shift[SORT_MAX_PASSES]; uint32_t mask[SORT_MAX_PASSES]; };
SortPlan make_sort_plan(int begin_bit, int end_bit);
size_t sort_temp_bytes(size_t n);
uint32_t* sort_temp_hist(void* temp);
void sort_temp_clear(void* temp, size_t n, cudaStream_t st);

// ---- splat record -------------------------------------------------------------------------------
// rec[3*i+0] = {mean2D.x, mean2D.y, depth (view z), blend_pixel_size (Q6, forward.cu:499-506)}
// rec[3*i+1] = {conic.x, conic.y, conic.z, opacity}                       (forward.cu:388)
// rec[3*i+2] = {r, g, b, bwd_skip (1.0f if the backward's per-pair filter re-check skips it, Q2)}

struct GeomState {
    float4*   rec;            // P x 3
    float*    cov3D;          // P x 6   (computed from scale/rot when not precomputed)
    uint8_t*  clamped;        // P       bits 0..2 = channel clamped at 0 (forward.cu:71-73)
    uint32_t* tiles_touched;  // P
    ushort4*  rect;           // P       tile rectangle {xmin, ymin, xmax, ymax} of rendered Gaussians
    uint32_t* dkeys[2];       // P each  depth sort keys: float bits of view z, 0xFFFFFFFF if not rendered
    uint32_t* dvals[2];       // P each  depth sort values: Gaussian index; dvals[0] ends up depth-ordered
    uint32_t* point_offsets;  // P       inclusive scan of tiles_touched IN DEPTH ORDER
    float*    gacc;           // P x 12  blend-backward accumulators: color[3], mean2D[2], conic[3], opacity, pad[3]
    uint32_t* scan_state;     // decoupled look-back tile descriptors (u64 each) + ticket
    uint32_t* meta;           // 16 words: [0] = num_rendered
    void*     dsort_temp;     // onesweep temp for the depth sort
};

struct ImageState {
    float*    final_T;        // W*H
    uint32_t* n_contrib;      // W*H
    uint2*    ranges;         // tiles
    uint32_t* tile_max_contrib; // tiles: max n_contrib over the tile's pixels (backward start point)
    int*      corners;        // (gx+1) x (gy+1): 4-corner scatter of the tile rectangles
    uint32_t* part_hist;      // SORT_MAX_PASSES x 256: exclusive digit offsets of the tile partition passes
};

struct BinningState {
    uint32_t* vals[2];        // [0] = point_list (sorted result), [1] = alternate
    uint32_t* tkeys[2];       // tile ids travelling with the values through the partition passes
    void*     sort_temp;      // tickets + look-back words of the partition passes
};

constexpr size_t SCAN_TILE = 2048;      // Gaussians per scan tile (256 threads x 8)

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
    p = (char*)align_up((size_t)p, 128);
    out = (T*)p;
    p += count * sizeof(T);
}

inline size_t scan_state_words(size_t P) { return 2 * ((P + SCAN_TILE - 1) / SCAN_TILE + 1) + 32; }

inline GeomState geom_from_chunk(char* chunk, size_t P, char** end = nullptr) {
    GeomState g;
    char* p = chunk;
    carve(p, g.rec, 3 * P);
    carve(p, g.cov3D, 6 * P);
    carve(p, g.clamped, P);
    carve(p, g.tiles_touched, P);
    carve(p, g.rect, P);
    carve(p, g.dkeys[0], P);
    carve(p, g.dkeys[1], P);
    carve(p, g.dvals[0], P);
    carve(p, g.dvals[1], P);
    carve(p, g.point_offsets, P);
    carve(p, g.gacc, 12 * P);
    carve(p, g.scan_state, scan_state_words(P));
    carve(p, g.meta, 16);
    char* t = nullptr;
    carve(p, t, sort_temp_bytes(P));
    g.dsort_temp = t;
    if (end) *end = p;
    return g;
}
inline size_t geom_bytes(size_t P) {
    char* end = nullptr;
    geom_from_chunk((char*)0, P, &end);
    return (size_t)(end - (char*)0) + 128;
}

inline ImageState image_from_chunk(char* chunk, int W, int H, char** end = nullptr) {
    ImageState s;
    char* p = chunk;
    const size_t N = (size_t)W * H;
    const size_t gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    carve(p, s.final_T, N);
    carve(p, s.n_contrib, N);
    carve(p, s.ranges, gx * gy);
    carve(p, s.tile_max_contrib, gx * gy);
    carve(p, s.corners, (gx + 1) * (gy + 1));
    carve(p, s.part_hist, (size_t)SORT_MAX_PASSES * 256);
    if (end) *end = p;
    return s;
}
inline size_t image_bytes(int W, int H) {
    char* end = nullptr;
    image_from_chunk((char*)0, W, H, &end);
    return (size_t)(end - (char*)0) + 128;
}

inline BinningState binning_from_chunk(char* chunk, size_t R, char** end = nullptr) {
    BinningState b;
    char* p = chunk;
    carve(p, b.vals[0], R);
    carve(p, b.vals[1], R);
    carve(p, b.tkeys[0], R);
    carve(p, b.tkeys[1], R);
    char* t = nullptr;
    carve(p, t, sort_temp_bytes(R));
    b.sort_temp = t;
    if (end) *end = p;
    return b;
}
inline size_t binning_bytes(size_t R) {
    char* end = nullptr;
    binning_from_chunk((char*)0, R, &end);
    return (size_t)(end - (char*)0) + 128;
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

void launch_init_state(GeomState g, ImageState img, int P, int gx, int gy, cudaStream_t st);

void launch_preprocess_fwd(const FwdParams& prm, const float* means3D, const float* scales, const float* rotations,
                           const float* opacities, const float* shs, const float* cov3D_precomp,
                           const float* colors_precomp, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, int* radii, float* pixel_sizes, GeomState g, int* corners,
                           cudaStream_t st);

// inclusive scan of tiles_touched[order[k]] over k -> point_offsets[k]; total -> meta[0]
void launch_scan_depth_order(GeomState g, int P, cudaStream_t st);

// per-tile counts from the corner grid -> ranges (empty tiles {0,0}), exclusive digit offsets of every partition pass
void launch_tile_scan(ImageState img, int gx, int gy, SortPlan plan, cudaStream_t st);

// instances in depth order: tile id + Gaussian index
void launch_emit(GeomState g, int P, int gx, uint32_t* tkeys, uint32_t* vals, cudaStream_t st);

template <typename KeyT>
void launch_sort_histogram(const KeyT* keys_in, size_t n, const SortPlan& pl, void* temp, cudaStream_t st);
// one onesweep pass; keys_out may be null (keys not needed after the last pass); hist = exclusive digit offsets
template <typename KeyT>
void launch_sort_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out, size_t n,
                      const SortPlan& pl, int p, const uint32_t* hist, void* temp, cudaStream_t st);

void launch_blend_fwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, float* final_T, uint32_t* n_contrib,
                      uint32_t* tile_max_contrib, float* out_color, float* out_acc_px, float* out_depth,
                      cudaStream_t st);

void launch_bwd_prepare(int P, GeomState g, const float* pixel_sizes, const float* max_px, const float* min_px,
                        const uint8_t* base_mask, int filter_small, int filter_large, cudaStream_t st);

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
