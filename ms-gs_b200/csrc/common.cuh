// Shared declarations of libmsgs_raster.so (B200 / sm_100a).  Private to ms-gs_b200/csrc.
//
// State layout (opaque to callers; replaces GeometryState / ImageState / BinningState of the reference,
// submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.h:29-66 — "private layout, free
// to redesign").  Redesign for HBM3e/L2:
//   * per-Gaussian blend inputs live in ONE 64-byte "splat record" (4 x float4, two whole sectors) so that the
//     per-tile gathers of the blend kernels touch 2 sectors instead of 5 separate arrays;
//   * the 9 atomically accumulated gradients of the blend backward live in one 48-byte record
//     (one L2 line region per Gaussian) inside the geometry buffer;
//   * binning does NOT sort R (Gaussian, tile) instances on a 64-bit (tile | depth) key as the reference
//     does (rasterizer_impl.cu:98-108, 327-332), and it never materialises keys at all.  The same order —
//     stable by (tile, depth bits, Gaussian index) — is produced by (1) a radix sort of the P Gaussians by
//     their 32 depth bits, (2) COUNTING: the depth-ordered Gaussians are cut into chunks; per chunk the number
//     of its Gaussians covering every tile comes from a 4-corner scatter of the tile rectangles + a 2-D
//     prefix sum in shared memory; an exclusive scan of that (chunk x tile) matrix down the chunks plus a
//     scan of the tile totals gives every chunk its first slot in every tile's list (and the ranges, and
//     num_rendered); (3) PLACEMENT: each chunk walks its Gaussians in depth order and appends the index to
//     the lists of the tiles it covers.  HBM traffic per instance: one 4-byte store.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstdlib>
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
size_t sort_temp_clear_bytes(size_t n);
void sort_temp_clear(void* temp, size_t n, cudaStream_t st);

// ---- splat record -------------------------------------------------------------------------------
// rec[4*i+0] = {mean2D.x, mean2D.y, depth (view z), blend_pixel_size (Q6, forward.cu:499-506)}
// rec[4*i+1] = {conic.x, conic.y, conic.z, opacity}                       (forward.cu:388)
// rec[4*i+2] = {r, g, b, filter bits as a float: 1 = fails the backward's filter_small test, 2 = fails its filter_large
//              test (backward.cu:649-660 re-checks them INDEPENDENTLY, quirk Q2; evaluated once per Gaussian here)}
// rec[4*i+3] = {hx, hy, thr, -}: footprint of alpha >= 1/255 for the blend kernels' per-warp cull (cull.cuh):
//              -2 * power <= thr inside it, |dx| <= hx, |dy| <= hy
constexpr int REC_F4 = 4;               // float4 per splat record

struct GeomState {
    float4*   rec;            // P x REC_F4
    float*    cov3D;          // P x 6   (computed from scale/rot when not precomputed)
    uint8_t*  clamped;        // P       bits 0..2 = channel clamped at 0 (forward.cu:71-73)
    uint32_t* tiles_touched;  // P       (kept for parity introspection; binning uses the rectangles)
    ushort4*  rect;           // P       tile rectangle {xmin, ymin, xmax, ymax}; zero area if not rendered
    ushort4*  rect_sorted;    // P       the rectangles gathered into depth order (binning.cu)
    uint32_t* dkeys[2];       // P each  depth sort keys: float bits of view z, 0xFFFFFFFF if not rendered
    uint32_t* dvals[2];       // P each  depth sort values: Gaussian index; dvals[0] ends up depth-ordered
    float*    gacc;           // P x 12  blend-backward accumulators: color[3], mean2D[2], conic[3], opacity, pad[3];
                              //         zero for every rendered Gaussian outside a backward call (the forward zeroes them,
                              //         the backward's preprocess re-zeroes what it has consumed)
    uint32_t* meta;           // 32 words (directly in front of dsort_temp): [0] = num_rendered, [1] = rendered Gaussians
                              //          (view group: by any unit of the group)
    void*     dsort_temp;     // onesweep temp for the depth sort
};

// Binning plan (binning.cu).  The V rendered Gaussians (they come first in depth order; V is only known on the
// device, meta[1]) are cut into NC chunk groups.  Counting pass: tile rows are cut into count-bands so that one band's
// corner grid fits the 48 KB of the counting kernel; with few (group, band) pairs a group's Gaussians are split over
// several CTAs.
//
// Two-level placement (r02).  Level 1: the tile grid is cut into super-blocks of sb x sb tiles (8 x 8; 16 x 16 on very
// large grids); the depth-ordered Gaussians are cut into LIST CHUNKS of sub = 256 * items Gaussians, and every
// (super-block, list chunk) pair gets a contiguous, depth-ordered run of 8-byte entries {column mask | row mask << 16,
// Gaussian index} — one entry per (Gaussian, super-block it meets), found by counting (l1_count), an exclusive scan over
// the (super-block, list chunk) matrix in that order (l1_scan) and an ordered emission (l1_emit).  A chunk GROUP is K
// consecutive list chunks, so for a fixed super-block the entries of a group are contiguous.  Level 2 (place2): one CTA
// per (group, super-block) walks only ITS OWN entries, 32 at a time (lane = entry); per tile a ballot of the covering
// entries gives every one of them its slot, and the warp writes them as ONE run of consecutive 4-byte words — nobody
// re-walks a chunk per super-block and no lane-per-tile scattered partial-sector store is left.
struct BinPlan {
    int NC, band_rows, nbands, split;   // split: CTAs per (group, band) of the counting pass
    int sb, sb_shift, sbx, sby, S;      // super-block edge in tiles (8 / 16), its log2, the super-block grid
    int items, sub;                     // Gaussians per thread / per list chunk of the level-1 kernels (sub = 256 * items)
    int nl1;                            // list chunks that P Gaussians can span = ceil(P / sub): row stride of l1off
};
// Gaussians per chunk group, from the device-side V: K list chunks, K = ceil(ceil(V / sub) / NC) (at least one)
__host__ __device__ inline int bin_group_chunks(uint32_t V, const BinPlan& bp) {
    const int n = (int)((V + (uint32_t)bp.sub - 1u) / (uint32_t)bp.sub);
    const int k = (n + bp.NC - 1) / bp.NC;
    return k < 1 ? 1 : k;
}
__host__ __device__ inline int bin_chunk_size(uint32_t V, const BinPlan& bp) {
    return bin_group_chunks(V, bp) * bp.sub;
}
constexpr int BIN_SMEM_WORDS = 12288;   // 48 KB
constexpr int L1_SEG = 4096;            // elements per CTA of the list-offset scan
inline int bin_nc_cap() {       // MSGS_NC: experiment knob for the chunk count
    static const int cap = [] { const char* e = getenv("MSGS_NC"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= 256 ? v : 0; }();
    return cap;
}
inline BinPlan make_bin_plan(size_t P, int gx, int gy) {
    BinPlan b;
    const size_t tiles = (size_t)gx * gy;
    // super-blocks: 8 x 8 tiles unless the level-1 ballot tables (one word per super-block column and row per 32
    // Gaussians) would outgrow shared memory
    b.sb = 8; b.sb_shift = 3;
    if ((gx + 7) / 8 + (gy + 7) / 8 > 64) { b.sb = 16; b.sb_shift = 4; }
    b.sbx = (gx + b.sb - 1) / b.sb; b.sby = (gy + b.sb - 1) / b.sb; b.S = b.sbx * b.sby;
    b.items = P > (size_t)2000000 ? 16 : 4;
    b.sub = 256 * b.items;
    b.nl1 = (int)((P + b.sub - 1) / b.sub);
    if (b.nl1 < 1) b.nl1 = 1;
    {
        // enough (group, super-block) CTAs to fill the GPU a few times over, no more groups than list chunks
        int nc = (8 * 148 + b.S - 1) / b.S;
        if (nc > 64) nc = 64;
        if (bin_nc_cap()) nc = bin_nc_cap();
        if (nc > b.nl1) nc = b.nl1;
        if (nc < 1) nc = 1;
        b.NC = nc;
    }
    int rows = BIN_SMEM_WORDS / (gx + 2) - 1;          // (rows + 1) x (gx + 1 [+1 pad]) corner cells
    if (rows < 1) rows = 1;
    if (rows > gy) rows = gy;
    b.band_rows = rows;
    b.nbands = (gy + rows - 1) / rows;
    // counting pass: with few (group, band) pairs one CTA per pair leaves most SMs idle and every CTA walks a long
    // chunk; split the group's Gaussians over several CTAs that add into the (zeroed) counts
    int split = 1;
    const int split_cap = 32;
    if (b.NC * b.nbands < 148)
        while (b.NC * b.nbands * split < 2 * 148 && split < split_cap) ++split;
    b.split = split;
    return b;
}
// the two-level placement's limits: column + row ballot words per 32 Gaussians
inline bool bin_plan_supported(const BinPlan& b) { return b.sbx + b.sby <= 64; }

struct ImageState {
    float*    final_T;        // W*H
    uint32_t* n_contrib;      // W*H
    uint2*    ranges;         // tiles
    uint32_t* tile_max_contrib; // tiles: max n_contrib over the tile's pixels (backward start point)
    uint32_t* tile_total;     // tiles: instances per tile                       (forward only)
    uint32_t* tile_start;     // tiles: first slot of the tile's list             (forward only)
    uint32_t* counts;         // NC x tiles: per chunk, Gaussians covering the tile -> exclusive prefix down the chunks
    uint32_t* l1off;          // S x nl1 (+1): entries per (super-block, list chunk) -> exclusive scan in that order; the
                              //               extra word holds the total                                (forward only)
    uint32_t* l1seg;          // per L1_SEG elements of l1off: their sum (scan helper)
};

struct BinningState {
    uint32_t* point_list;     // R: Gaussian indices, tile-major, depth order inside a tile
    uint2*    l1;             // <= R level-1 entries {column mask | row mask << 16, Gaussian index}: every (Gaussian,
                              //      super-block) pair holds at least one instance, so R entries always suffice
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
    p = (char*)align_up((size_t)p, 128);
    out = (T*)p;
    p += count * sizeof(T);
}

inline GeomState geom_from_chunk(char* chunk, size_t P, char** end = nullptr) {
    GeomState g;
    char* p = chunk;
    carve(p, g.rec, REC_F4 * P);
    carve(p, g.cov3D, 6 * P);
    carve(p, g.clamped, P);
    carve(p, g.tiles_touched, P);
    carve(p, g.rect, P);
    carve(p, g.rect_sorted, P);
    carve(p, g.dkeys[0], P);
    carve(p, g.dkeys[1], P);
    carve(p, g.dvals[0], P);
    carve(p, g.dvals[1], P);
    carve(p, g.gacc, 12 * P);
    carve(p, g.meta, 32);
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

inline ImageState image_from_chunk(char* chunk, int W, int H, size_t P, char** end = nullptr) {
    ImageState s;
    char* p = chunk;
    const size_t N = (size_t)W * H;
    const size_t gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const BinPlan bp = make_bin_plan(P, (int)gx, (int)gy);
    carve(p, s.final_T, N);
    carve(p, s.n_contrib, N);
    carve(p, s.ranges, gx * gy);
    carve(p, s.tile_max_contrib, gx * gy);
    carve(p, s.tile_total, gx * gy);
    carve(p, s.tile_start, gx * gy);
    carve(p, s.counts, (size_t)bp.NC * gx * gy);
    const size_t nl1off = (size_t)bp.S * bp.nl1 + 1;
    carve(p, s.l1off, nl1off + 3);
    carve(p, s.l1seg, (nl1off + L1_SEG - 1) / L1_SEG + 1);
    if (end) *end = p;
    return s;
}
inline size_t image_bytes(int W, int H, size_t P) {
    char* end = nullptr;
    image_from_chunk((char*)0, W, H, P, &end);
    return (size_t)(end - (char*)0) + 128;
}

inline BinningState binning_from_chunk(char* chunk, size_t R, char** end = nullptr) {
    BinningState b;
    char* p = chunk;
    carve(p, b.point_list, R);
    carve(p, b.l1, R);
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
    int raw;                             // scales / rotations / opacities are GaussianModel's RAW parameters: exp / normalize /
                                         // sigmoid run in registers here (act.cuh; msgs_*_raw entry points)
    const float *view, *proj, *campos;   // device pointers (4x4, 4x4, 3)
};

// ---- stage launchers (each enqueues on `st`; definitions in the matching .cu) -----------------------
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);

void launch_preprocess_fwd(const FwdParams& prm, const float* means3D, const float* scales, const float* rotations,
                           const float* opacities, const float* shs, const float* cov3D_precomp,
                           const float* colors_precomp, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, int* radii, float* pixel_sizes, GeomState g, cudaStream_t st);

// View group (msgs_forward_group_begin): per-unit outputs of the shared preprocess kernel (preprocess_fwd.cu)
constexpr int MAX_GROUP = 8;
struct GroupScale {
    int W, H, gx, gy;
    float focal_x, focal_y;
    int* radii; float* pixel_sizes;                          // the unit's API outputs
    float4* rec; float* cov3D; uint8_t* clamped;             // the unit's GeomState arrays
    uint32_t* tiles_touched; ushort4* rect; float* gacc; uint32_t* meta;
};
struct GroupParams { int S; GroupScale sc[MAX_GROUP]; };
void launch_preprocess_fwd_group(const FwdParams& prm, const GroupParams& gp, const float* means3D, const float* scales,
                                 const float* rotations, const float* opacities, const float* shs,
                                 const float* cov3D_precomp, const float* colors_precomp, const float* max_px,
                                 const float* min_px, const uint8_t* base_mask, uint32_t* dkeys, uint32_t* dvals,
                                 cudaStream_t st);

// binning (binning.cu): per-chunk tile counts -> scan down the chunks -> tile starts / ranges / num_rendered -> placement
void launch_chunk_count(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp, cudaStream_t st);
void launch_chunk_scan(const ImageState& img, int tiles, const BinPlan& bp, cudaStream_t st);
// host_out: mapped pinned host word that also receives num_rendered (may be null)
void launch_tile_scan(const ImageState& img, int tiles, uint32_t* meta, uint32_t* host_out, cudaStream_t st);
// `capacity`: instances the binning buffer can hold; if num_rendered (meta[0]) exceeds it the kernels return without writing
// two-level placement: list counting + scan (no dependence on the binning buffer), then list emission + placement
void launch_l1_count_scan(const GeomState& g, const ImageState& img, int P, const BinPlan& bp, cudaStream_t st);
void launch_place2(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp,
                   const BinningState& bin, uint32_t capacity, cudaStream_t st);

template <typename KeyT>
void launch_sort_histogram(const KeyT* keys_in, size_t n, const SortPlan& pl, void* temp, cudaStream_t st);
// one onesweep pass; keys_out may be null (keys not needed after the last pass); hist = exclusive digit offsets
template <typename KeyT>
void launch_sort_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out, size_t n,
                      const SortPlan& pl, int p, const uint32_t* hist, void* temp, cudaStream_t st);

void launch_blend_fwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, float* final_T, uint32_t* n_contrib,
                      uint32_t* tile_max_contrib, float* out_color, float* out_acc_px, float* out_depth,
                      const uint32_t* meta, uint32_t capacity, cudaStream_t st);

void launch_blend_bwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, const float* final_T, const uint32_t* n_contrib,
                      const uint32_t* tile_max_contrib, const float* dL_dpix, float* gacc, int filter_small,
                      int filter_large, cudaStream_t st);

struct BwdParams {
    int P, D, M, W, H;    // P: one past the last Gaussian this launch handles
    int first;            // first Gaussian this launch handles (a multiple of 128; msgs_backward_accumulate_part)
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int filter_small, filter_large;
    int raw;                             // as FwdParams.raw: the gradients written are those of the RAW parameters
    const float* opacity_raw;            // raw mode: the raw opacities (sigmoid' needs them; the reference's backward gets none)
    const float *view, *proj, *campos;   // device pointers
};
void launch_preprocess_bwd(const BwdParams& prm, const float* means3D, const float* shs, const float* scales,
                           const float* rotations, const float* cov3D_precomp, const int* radii,
                           const float* pixel_sizes, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, GeomState g,
                           float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_docc, float* dL_ddc,
                           float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                           float* dL_dscale, float* dL_drot, int accumulate, cudaStream_t st);   // 0 write, 1 add

// View group backward (msgs_backward_group_accumulate, preprocess_bwd.cu)
struct GroupBwdScale {
    float focal_x, focal_y;
    const int* radii; const float* pixel_sizes;              // the unit's forward outputs
    const float* cov3D; const uint8_t* clamped; float* gacc; // the unit's GeomState arrays
    float* dL_dmean2D;                                       // (P,3) per unit, may be null
};
struct GroupBwdParams { int S; GroupBwdScale sc[MAX_GROUP]; };
void launch_preprocess_bwd_group(const BwdParams& prm, const GroupBwdParams& gp, const float* means3D, const float* shs,
                                 const float* scales, const float* rotations, const float* cov3D_precomp,
                                 const float* max_px, const float* min_px, const uint8_t* base_mask, float* dL_dopacity,
                                 float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot, cudaStream_t st);

// launch counter shared by every translation unit of the library (msgs_launch_count, api.cu)
void note_launches(int kernels);

}  // namespace msgs
