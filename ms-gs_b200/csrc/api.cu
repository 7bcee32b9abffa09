// extern "C" surface of libmsgs_raster.so (include/msgs_raster.h): host orchestration of the stages.
// Mirrors CudaRasterizer::Rasterizer::{forward,backward,markVisible} (rasterizer_impl.cu:141-153, 200-367,
// 371-494).  All work is enqueued on the caller's stream; the single host synchronisation per forward is
// the 4-byte read-back of num_rendered that sizes the binning buffer and is the call's return value (the reference
// blocks at the same place: cudaMemcpy at rasterizer_impl.cu:305) — taken with placement + blend already enqueued.
#include "common.cuh"
#include "../../include/msgs_raster.h"
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>


using namespace msgs;

static thread_local std::string g_err;

// Forward calls in flight between msgs_forward_begin and msgs_forward_finish (per host thread): what the second half
// needs to read num_rendered back and, if the speculative binning buffer was too small, to launch placement + blend again.
constexpr int MAX_PENDING = 64;
struct Pending {
    bool used = false, immediate = false;
    GeomState g; ImageState img; BinPlan bp;
    int P = 0, W = 0, H = 0, gx = 0, gy = 0, debug = 0;
    const float* background = nullptr;
    float *out_color = nullptr, *out_acc = nullptr, *out_depth = nullptr;
    msgs_alloc_fn binning_alloc = nullptr; void* binning_ctx = nullptr;
    cudaStream_t st = nullptr;
    uint32_t guess = 0;
    cudaEvent_t copied = nullptr;
    int copied_dev = -1;         // device the event was created on (a host thread may drive several GPUs one after the other)
};
static thread_local Pending g_pending[MAX_PENDING];
static thread_local uint32_t* g_pinned = nullptr;     // MAX_PENDING words of mapped pinned host memory, one per slot
static thread_local uint32_t* g_pinned_dev = nullptr; // the device's address of the same words

// Recent num_rendered per (P, W, H): the speculative size of the next binning buffer (see msgs_forward).
struct CapEntry { int P, W, H; uint32_t r[8]; int n; };       // the last 8 forwards: a training loop rotates over views
static thread_local CapEntry g_caps[16];
static thread_local int g_ncaps = 0, g_cap_rr = 0;
static thread_local long long g_spec[3] = {0, 0, 0};      // forwards finished / speculation hits / misses (this host thread)
static uint32_t capacity_hint(int P, int W, int H) {
    for (int i = 0; i < g_ncaps; ++i) {
        const CapEntry& e = g_caps[i];
        if (e.P == P && e.W == W && e.H == H) {
            uint32_t m = 0;
            for (int k = 0; k < 8; ++k) m = e.r[k] > m ? e.r[k] : m;
            const uint64_t c = (uint64_t)m + m / 8 + 65536;       // 12.5 % + 64 Ki slack over the recent maximum
            return c >= (1ull << 31) ? 0u : (uint32_t)c;
        }
    }
    return 0u;
}
static void capacity_note(int P, int W, int H, uint32_t R) {
    for (int i = 0; i < g_ncaps; ++i) {
        CapEntry& e = g_caps[i];
        if (e.P == P && e.W == W && e.H == H) { e.r[e.n++ & 7] = R; return; }
    }
    CapEntry& e = g_caps[g_ncaps < 16 ? g_ncaps++ : (g_cap_rr++ & 15)];
    e.P = P; e.W = W; e.H = H; e.n = 1; e.r[0] = R; for (int k = 1; k < 8; ++k) e.r[k] = 0;
}

static int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    g_err = what;
    if (e != cudaSuccess) { g_err += ": "; g_err += cudaGetErrorString(e); }
    return code;
}
#define CUDA_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) return fail(MSGS_ERR_CUDA, #expr, _e); } while (0)
#define STAGE_CHECK(name) do { \
    cudaError_t _e = cudaPeekAtLastError(); \
    if (_e == cudaSuccess && debug) _e = cudaStreamSynchronize(st); \
    if (_e != cudaSuccess) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, name, _e); } } while (0)

// ---- optional per-stage CUDA-event timing (bench.py's live roofline measurement) and launch counting ----
enum Stage { ST_PREPROCESS_FWD = 0, ST_DEPTH_SORT, ST_CHUNK_COUNT, ST_CHUNK_SCAN, ST_TILE_SCAN, ST_PLACE,
             ST_BLEND_FWD, ST_BLEND_BWD, ST_PREPROCESS_BWD, ST_LIST_COUNT, ST_COUNT };
static const char* kStageNames[ST_COUNT] = {"preprocess_fwd", "depth_sort", "chunk_count", "chunk_scan", "tile_scan",
                                            "place", "blend_fwd", "blend_bwd", "preprocess_bwd", "list_count"};
static std::atomic<long long> g_launches{0};
namespace msgs { void note_launches(int kernels) { g_launches += kernels; } }
static bool g_prof_on = false;
struct ProfRec { int stage; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;
static cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
struct StageScope {
    int stage; cudaStream_t st; cudaEvent_t a = nullptr;
    StageScope(int stage_, cudaStream_t st_, int kernels) : stage(stage_), st(st_) {
        g_launches += kernels;
        if (g_prof_on) { a = prof_event(); cudaEventRecord(a, st); }
    }
    ~StageScope() {
        if (a) { cudaEvent_t b = prof_event(); cudaEventRecord(b, st); g_prof_recs.push_back({stage, a, b}); }
    }
};

// placement + forward blend of one pending forward into the binning buffer (capacity entries; device-side guard)
static void launch_tail(const Pending& pd, const BinningState& bin, uint32_t capacity) {
    cudaStream_t st = pd.st;
    uint32_t* point_list = bin.point_list;
    { StageScope sc(ST_PLACE, st, 2);      // list emission + placement
      launch_place2(pd.g, pd.img, pd.P, pd.gx, pd.gy, pd.bp, bin, capacity, st); }
    { StageScope sc(ST_BLEND_FWD, st, 1);
      launch_blend_fwd(pd.W, pd.H, pd.gx, pd.gy, pd.img.ranges, point_list, pd.g.rec, pd.background, pd.img.final_T,
                       pd.img.n_contrib, pd.img.tile_max_contrib, pd.out_color, pd.out_acc, pd.out_depth, pd.g.meta,
                       capacity, st); }
}

// Second half of a forward's enqueue (after the depth sort): counting, scans, the num_rendered read-back and the
// speculative placement + blend.  pd is filled by the caller; marks the slot used on success.
static int begin_tail(Pending& pd, int slot) {
    cudaStream_t st = pd.st;
    const int debug = pd.debug;
    const int tiles = pd.gx * pd.gy;
    // (2) counting: per-chunk tile counts -> scan down the chunks -> tile starts, ranges, num_rendered
    { StageScope sc(ST_CHUNK_COUNT, st, 1); launch_chunk_count(pd.g, pd.img, pd.P, pd.gx, pd.gy, pd.bp, st); }
    STAGE_CHECK("chunk count");
    { StageScope sc(ST_CHUNK_SCAN, st, 1); launch_chunk_scan(pd.img, tiles, pd.bp, st); }
    STAGE_CHECK("chunk scan");
    // num_rendered goes to the device copy (meta[0]) and, by a direct store of the kernel into mapped pinned memory, to
    // the host: no cudaMemcpy, no copy engine
    if (!g_pinned) {
        CUDA_TRY(cudaHostAlloc((void**)&g_pinned, MAX_PENDING * sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable));
        CUDA_TRY(cudaHostGetDevicePointer((void**)&g_pinned_dev, g_pinned, 0));
    }
    { StageScope sc(ST_TILE_SCAN, st, 1); launch_tile_scan(pd.img, tiles, pd.g.meta, g_pinned_dev + slot, st); }
    STAGE_CHECK("tile scan");
    // two-level placement: the (super-block, list chunk) entry counts and their scan need no binning buffer
    { StageScope sc(ST_LIST_COUNT, st, 2);
      launch_l1_count_scan(pd.g, pd.img, pd.P, pd.bp, st); }
    STAGE_CHECK("list count / scan");

    // num_rendered sizes the binning buffer and is part of the return value, so the host must read it (the reference
    // blocks at the same place, rasterizer_impl.cu:305).  To keep the GPU busy across that read the placement and the
    // blend are enqueued FIRST into a speculatively sized buffer (the largest recent R of this (P, W, H) plus slack)
    // with a device-side guard (R > capacity -> both kernels return at once); the host then waits only for the
    // 4-byte copy that was enqueued in front of them.  If the guess was too small (or there is no history) the buffer
    // is allocated with the exact size and the two kernels are launched (again).
    {   // events belong to the device that was current when they were created
        int cur_dev = 0;
        CUDA_TRY(cudaGetDevice(&cur_dev));
        if (pd.copied && pd.copied_dev != cur_dev) { cudaEventDestroy(pd.copied); pd.copied = nullptr; }
        if (!pd.copied) { CUDA_TRY(cudaEventCreateWithFlags(&pd.copied, cudaEventDisableTiming)); pd.copied_dev = cur_dev; }
    }
    CUDA_TRY(cudaEventRecord(pd.copied, st));        // behind the tile scan that stored num_rendered into g_pinned[slot]
    static const bool speculate = !(getenv("MSGS_SPECULATE") && getenv("MSGS_SPECULATE")[0] == '0');
    pd.guess = (debug || !speculate) ? 0u : capacity_hint(pd.P, pd.W, pd.H);
    if (pd.guess) {
        char* bin_chunk = (char*)pd.binning_alloc(pd.binning_ctx, binning_bytes((size_t)pd.guess));
        if (!bin_chunk) return fail(MSGS_ERR_ALLOC, "binning buffer allocation failed");
        launch_tail(pd, binning_from_chunk(bin_chunk, (size_t)pd.guess), pd.guess);
    }
    STAGE_CHECK("place / blend forward (speculative)");
    pd.used = true;
    return 0;
}

extern "C" {

long long msgs_launch_count(void) { return g_launches.load(); }
int msgs_profile_num_stages(void) { return ST_COUNT; }
const char* msgs_profile_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }
void msgs_profile_enable(int on) { g_prof_on = on != 0; }
// Adds the elapsed ms / launch counts of every stage recorded since the last call to ms[] / counts[].
int msgs_profile_collect(double* ms, long long* counts, int n) {
    for (auto& r : g_prof_recs) {
        cudaEventSynchronize(r.b);
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess && r.stage < n) { ms[r.stage] += t; counts[r.stage] += 1; }
        g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b);
    }
    g_prof_recs.clear();
    return ST_COUNT;
}

void msgs_speculation_stats(long long out[3], int reset) {
    if (out) { out[0] = g_spec[0]; out[1] = g_spec[1]; out[2] = g_spec[2]; }
    if (reset) g_spec[0] = g_spec[1] = g_spec[2] = 0;
}

int msgs_abi_version(void) { return MSGS_ABI_VERSION; }
const char* msgs_last_error(void) { return g_err.c_str(); }

// Fixed-buffer allocator: an msgs_alloc_fn a binding can pass instead of its own callback when it already knows the
// size (msgs_geom_bytes / msgs_image_bytes / msgs_binning_hint_bytes): serves the request from the caller's buffer when it
// fits, otherwise forwards it to the binding's real callback.  Saves one host-language call-back per buffer and forward.
void* msgs_fixed_alloc(void* ctx, size_t bytes) {
    msgs_fixed_buffer* b = (msgs_fixed_buffer*)ctx;
    if (!b) return nullptr;
    b->requested = bytes;
    if (b->ptr && bytes <= b->bytes) { b->served_fixed = 1; return b->ptr; }
    b->served_fixed = 0;
    return b->fallback ? b->fallback(b->fallback_ctx, bytes) : nullptr;
}
// bytes of the speculative binning buffer the next forward of this shape will ask for on this host thread (0: no history)
size_t msgs_binning_hint_bytes(int P, int width, int height) {
    const uint32_t c = capacity_hint(P, width, height);
    return c ? binning_bytes((size_t)c) : 0;
}
size_t msgs_geom_bytes(int P) { return geom_bytes((size_t)(P > 0 ? P : 0)); }
size_t msgs_image_bytes(int width, int height, int P) { return image_bytes(width, height, (size_t)(P > 0 ? P : 0)); }
size_t msgs_binning_bytes(int R) { return binning_bytes((size_t)(R > 0 ? R : 0)); }
size_t msgs_sort_temp_bytes(int n) { return sort_temp_bytes((size_t)(n > 0 ? n : 0)); }

void msgs_geom_layout(void* chunk, int P, uint64_t out[8]) {
    GeomState g = geom_from_chunk((char*)chunk, (size_t)P);
    out[0] = (uint64_t)g.rec; out[1] = (uint64_t)g.cov3D; out[2] = (uint64_t)g.clamped;
    out[3] = (uint64_t)g.tiles_touched; out[4] = (uint64_t)g.rect; out[5] = (uint64_t)g.gacc;
    out[6] = (uint64_t)g.dvals[0]; out[7] = (uint64_t)g.meta;
}
void msgs_image_layout(void* chunk, int width, int height, int P, uint64_t out[4]) {
    ImageState s = image_from_chunk((char*)chunk, width, height, (size_t)P);
    out[0] = (uint64_t)s.final_T; out[1] = (uint64_t)s.n_contrib; out[2] = (uint64_t)s.ranges;
    out[3] = (uint64_t)s.tile_max_contrib;
}
void msgs_binning_layout(void* chunk, int R, uint64_t out[4]) {
    BinningState b = binning_from_chunk((char*)chunk, (size_t)R);
    out[0] = (uint64_t)b.point_list; out[1] = 0; out[2] = 0; out[3] = 0;
}

int msgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream) {
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return fail(MSGS_ERR_ARG, "msgs_mark_visible: null argument");
    launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, "mark_visible launch", e); }
    return MSGS_OK;
}

int msgs_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                    uint64_t* keys_alt, uint32_t* vals_alt, int n, int end_bit, void* temp, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || end_bit < 0 || end_bit > 56 || n >= (1 << 28)) return fail(MSGS_ERR_ARG, "msgs_sort_pairs: need 0<=end_bit<=56, n<2^28");
    if (n == 0) return MSGS_OK;
    const SortPlan pl = make_sort_plan(0, end_bit);
    // ping-pong between (out, alt) so that the last pass writes into `out`
    uint64_t* kb[2] = {keys_out, keys_alt};
    uint32_t* vb[2] = {vals_out, vals_alt};
    int cur = (pl.passes & 1) ? 1 : 0;
    CUDA_TRY(cudaMemcpyAsync(kb[cur], keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(vb[cur], vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (pl.passes == 0) return MSGS_OK;
    sort_temp_clear(temp, (size_t)n, st);
    launch_sort_histogram<uint64_t>(kb[cur], (size_t)n, pl, temp, st);
    for (int p = 0; p < pl.passes; ++p) {
        launch_sort_pass<uint64_t>(kb[cur], kb[cur ^ 1], vb[cur], vb[cur ^ 1], (size_t)n, pl, p, sort_temp_hist(temp), temp, st);
        cur ^= 1;
    }
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, "sort launch", e); }
    return MSGS_OK;
}

}  // extern "C"

// ---- msgs_sort_pairs_static: the same sort with its scratch in a static device array (no allocation anywhere) -----------
// The onesweep temp (digit histograms, tile tickets, look-back words) of up to MSGS_SORT_STATIC_MAX_N pairs:
// 7 x 256 x 4 + 32 + (512 + 1) x 256 x 4 + 256 = 532 768 bytes; every GPU that loads the library holds the array until exit.
// Host code may NOT pass this name to a kernel (on the host it is a stand-in, not a device address): the address comes
// from cudaGetSymbolAddress, once per device with that device current (sort_static_temp below).
constexpr size_t SORT_STATIC_BYTES = 540 * 1024;
static __device__ __align__(256) unsigned char g_sort_static_scratch[SORT_STATIC_BYTES];

constexpr int SORT_STATIC_MAX_DEVICES = 64;
static std::mutex g_sort_static_mu;
static void* g_sort_static_ptr[SORT_STATIC_MAX_DEVICES];

// device address of the scratch array on the CURRENT device (looked up on the first call per device); 0 + error on failure
static int sort_static_temp(void** out) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= SORT_STATIC_MAX_DEVICES) return fail(MSGS_ERR_ARG, "msgs_sort_pairs_static: device ordinal >= 64");
    std::lock_guard<std::mutex> lock(g_sort_static_mu);
    if (!g_sort_static_ptr[dev]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_sort_static_scratch);
        if (e != cudaSuccess || !ptr) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, "msgs_sort_pairs_static: cudaGetSymbolAddress(scratch)", e); }
        if (sort_temp_bytes((size_t)MSGS_SORT_STATIC_MAX_N) > SORT_STATIC_BYTES)
            return fail(MSGS_ERR_ARG, "msgs_sort_pairs_static: static scratch smaller than the sort's temp layout (rebuild)");
        g_sort_static_ptr[dev] = ptr;
    }
    *out = g_sort_static_ptr[dev];
    return MSGS_OK;
}

extern "C" {

int msgs_sort_pairs_static(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                           uint64_t* keys_alt, uint32_t* vals_alt, int n, int end_bit, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || end_bit < 0 || end_bit > 56) return fail(MSGS_ERR_ARG, "msgs_sort_pairs_static: need n>=0, 0<=end_bit<=56");
    if (n > MSGS_SORT_STATIC_MAX_N)
        return fail(MSGS_ERR_ARG, "msgs_sort_pairs_static: n exceeds MSGS_SORT_STATIC_MAX_N (2097152 pairs): the static scratch "
                                  "cannot hold it; use msgs_sort_pairs with a caller-provided temp buffer");
    if (n == 0) return MSGS_OK;
    if (!keys_in || !keys_out || !vals_in || !vals_out || !keys_alt || !vals_alt) return fail(MSGS_ERR_ARG, "msgs_sort_pairs_static: null buffer");
    void* temp = nullptr;
    if (int rc = sort_static_temp(&temp)) return rc;
    const SortPlan pl = make_sort_plan(0, end_bit);
    uint64_t* kb[2] = {keys_out, keys_alt};
    uint32_t* vb[2] = {vals_out, vals_alt};
    int cur = (pl.passes & 1) ? 1 : 0;
    CUDA_TRY(cudaMemcpyAsync(kb[cur], keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(vb[cur], vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (pl.passes == 0) return MSGS_OK;
    // the array keeps the previous call's histograms, tickets and look-back words: everything this call reads is cleared first
    sort_temp_clear(temp, (size_t)n, st);
    launch_sort_histogram<uint64_t>(kb[cur], (size_t)n, pl, temp, st);
    for (int p = 0; p < pl.passes; ++p) {
        launch_sort_pass<uint64_t>(kb[cur], kb[cur ^ 1], vb[cur], vb[cur ^ 1], (size_t)n, pl, p, sort_temp_hist(temp), temp, st);
        cur ^= 1;
    }
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, "msgs_sort_pairs_static: sort launch", e); }
    return MSGS_OK;
}

#define MSGS_FWD_PARAMS msgs_alloc_fn geom_alloc, void* geom_ctx, msgs_alloc_fn binning_alloc, void* binning_ctx, \
                 msgs_alloc_fn img_alloc, void* img_ctx, int P, int D, int M, \
                 const float* background, int width, int height, \
                 const float* means3D, const float* shs, const float* colors_precomp, \
                 const float* max_pixel_sizes, const float* min_pixel_sizes, \
                 const float* opacities, const float* occ_multiplier, const float* dc_delta, \
                 const float* scales, float scale_modifier, const float* rotations, \
                 const float* cov3D_precomp, const uint8_t* base_mask, \
                 const float* viewmatrix, const float* projmatrix, const float* cam_pos, \
                 float tan_fovx, float tan_fovy, int prefiltered, int filter_small, int filter_large, float fade_size, \
                 float* out_color, float* out_acc_pixel_size, float* out_depth, \
                 int* radii, float* pixel_sizes, int debug, void* stream
#define MSGS_FWD_ARGS geom_alloc, geom_ctx, binning_alloc, binning_ctx, img_alloc, img_ctx, P, D, M, background, width, height, \
                 means3D, shs, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities, occ_multiplier, dc_delta, scales, \
                 scale_modifier, rotations, cov3D_precomp, base_mask, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, \
                 prefiltered, filter_small, filter_large, fade_size, out_color, out_acc_pixel_size, out_depth, radii, \
                 pixel_sizes, debug, stream

}  // extern "C"

static int forward_begin_impl(int raw, MSGS_FWD_PARAMS) {
    (void)occ_multiplier; (void)dc_delta; (void)fade_size;   // dead inputs in the reference too (quirk Q9)
    cudaStream_t st = (cudaStream_t)stream;
    if (P < 0 || width <= 0 || height <= 0) return fail(MSGS_ERR_ARG, "msgs_forward: bad P / image size");
    int slot = 0;
    while (slot < MAX_PENDING && g_pending[slot].used) ++slot;
    if (slot == MAX_PENDING) return fail(MSGS_ERR_ARG, "msgs_forward_begin: too many forward calls in flight (finish some first)");
    if (!geom_alloc || !binning_alloc || !img_alloc) return fail(MSGS_ERR_ARG, "msgs_forward: null allocator");
    if (!background || !out_color || !out_acc_pixel_size || !out_depth) return fail(MSGS_ERR_ARG, "msgs_forward: null output/background");
    const int gx = (width + TILE_X - 1) / TILE_X, gy = (height + TILE_Y - 1) / TILE_Y;

    if (P == 0) {   // rasterize_points.cu:93 — nothing to render: outputs stay at their fill value (0)
        CUDA_TRY(cudaMemsetAsync(out_color, 0, (size_t)3 * width * height * 4, st));
        CUDA_TRY(cudaMemsetAsync(out_acc_pixel_size, 0, (size_t)width * height * 4, st));
        CUDA_TRY(cudaMemsetAsync(out_depth, 0, (size_t)width * height * 4, st));
        g_pending[slot].used = true; g_pending[slot].immediate = true;
        return slot;
    }
    if (!means3D || !opacities || !max_pixel_sizes || !min_pixel_sizes || !base_mask || !viewmatrix || !projmatrix ||
        !cam_pos || !radii || !pixel_sizes)
        return fail(MSGS_ERR_ARG, "msgs_forward: null required input");
    if (!colors_precomp && !shs) return fail(MSGS_ERR_ARG, "msgs_forward: need shs or colors_precomp");
    if (!cov3D_precomp && (!scales || !rotations)) return fail(MSGS_ERR_ARG, "msgs_forward: need scales+rotations or cov3D_precomp");
    if (!colors_precomp && (M <= 0 || M < (D + 1) * (D + 1) || D < 0 || D > 3))
        return fail(MSGS_ERR_ARG, "msgs_forward: SH degree / coefficient count mismatch");

    // rasterizer_impl.cu:236-254
    FwdParams prm;
    prm.P = P; prm.D = D; prm.M = M; prm.W = width; prm.H = height; prm.gx = gx; prm.gy = gy;
    prm.scale_modifier = scale_modifier; prm.tan_fovx = tan_fovx; prm.tan_fovy = tan_fovy;
    prm.focal_y = height / (2.0f * tan_fovy);
    prm.focal_x = width / (2.0f * tan_fovx);
    prm.prefiltered = prefiltered; prm.filter_small = filter_small; prm.filter_large = filter_large;
    prm.raw = raw;
    prm.view = viewmatrix; prm.proj = projmatrix; prm.campos = cam_pos;

    char* geom_chunk = (char*)geom_alloc(geom_ctx, geom_bytes((size_t)P));
    if (!geom_chunk) return fail(MSGS_ERR_ALLOC, "geometry buffer allocation failed");
    GeomState g = geom_from_chunk(geom_chunk, (size_t)P);
    char* img_chunk = (char*)img_alloc(img_ctx, image_bytes(width, height, (size_t)P));
    if (!img_chunk) return fail(MSGS_ERR_ALLOC, "image buffer allocation failed");
    ImageState img = image_from_chunk(img_chunk, width, height, (size_t)P);
    const BinPlan bp = make_bin_plan((size_t)P, gx, gy);
    if (!bin_plan_supported(bp)) return fail(MSGS_ERR_ARG, "msgs_forward: image too large for the placement tables (width + height above 16384 tiles edge sum)");

    // one memset: counters (meta) + the depth sort's histogram / tickets / look-back words right behind them
    CUDA_TRY(cudaMemsetAsync(g.meta, 0, (size_t)((char*)g.dsort_temp - (char*)g.meta) + sort_temp_clear_bytes((size_t)P), st));
    { StageScope sc(ST_PREPROCESS_FWD, st, 1);
    launch_preprocess_fwd(prm, means3D, scales, rotations, opacities, shs, cov3D_precomp, colors_precomp,
                          max_pixel_sizes, min_pixel_sizes, base_mask, radii, pixel_sizes, g, st); }
    STAGE_CHECK("preprocess");
    // (1) Gaussians in depth order: onesweep sort of P (u32 depth bits, index) pairs, 4 passes -> dvals[0]
    {
        const SortPlan dpl = make_sort_plan(0, 32);
        StageScope sc(ST_DEPTH_SORT, st, 3 + dpl.passes);
        launch_sort_histogram<uint32_t>(g.dkeys[0], (size_t)P, dpl, g.dsort_temp, st);
        int cur = 0;
        for (int p = 0; p < dpl.passes; ++p) {
            // values of the first pass = the Gaussian's index (preprocess writes no value array); nobody reads the sorted
            // keys, so the last pass writes none
            launch_sort_pass<uint32_t>(g.dkeys[cur], p + 1 < dpl.passes ? g.dkeys[cur ^ 1] : nullptr, p == 0 ? nullptr : g.dvals[cur],
                                       g.dvals[cur ^ 1], (size_t)P, dpl, p, sort_temp_hist(g.dsort_temp), g.dsort_temp, st);
            cur ^= 1;
        }
    }
    STAGE_CHECK("depth sort");
    Pending& pd = g_pending[slot];
    pd.immediate = false;
    pd.g = g; pd.img = img; pd.bp = bp;
    pd.P = P; pd.W = width; pd.H = height; pd.gx = gx; pd.gy = gy; pd.debug = debug;
    pd.background = background; pd.out_color = out_color; pd.out_acc = out_acc_pixel_size; pd.out_depth = out_depth;
    pd.binning_alloc = binning_alloc; pd.binning_ctx = binning_ctx; pd.st = st;
    const int rc = begin_tail(pd, slot);
    return rc < 0 ? rc : slot;
}

extern "C" {

int msgs_forward_begin(MSGS_FWD_PARAMS) { return forward_begin_impl(0, MSGS_FWD_ARGS); }
/* raw-parameter forward: scales / rotations / opacities are GaussianModel's _scaling / _rotation / _opacity */
int msgs_forward_begin_raw(MSGS_FWD_PARAMS) { return forward_begin_impl(1, MSGS_FWD_ARGS); }

int msgs_forward_finish(int handle) {
    if (handle < 0 || handle >= MAX_PENDING || !g_pending[handle].used) return fail(MSGS_ERR_ARG, "msgs_forward_finish: bad handle");
    Pending& pd = g_pending[handle];
    pd.used = false;
    if (pd.immediate) return 0;
    const int debug = pd.debug;
    cudaStream_t st = pd.st;
    CUDA_TRY(cudaEventSynchronize(pd.copied));
    const uint32_t R = *reinterpret_cast<volatile uint32_t*>(g_pinned + handle);
    if (R >= (1u << 31)) return fail(MSGS_ERR_ARG, "msgs_forward: more than 2^31 (Gaussian, tile) instances");
    capacity_note(pd.P, pd.W, pd.H, R);
    ++g_spec[0];
    if (pd.guess && R <= pd.guess) ++g_spec[1]; else ++g_spec[2];
    if (!pd.guess || R > pd.guess) {
        char* bin_chunk = (char*)pd.binning_alloc(pd.binning_ctx, binning_bytes((size_t)R));
        if (!bin_chunk) return fail(MSGS_ERR_ALLOC, "binning buffer allocation failed");
        launch_tail(pd, binning_from_chunk(bin_chunk, (size_t)R), R);
    }
    STAGE_CHECK("place / blend forward");
    return (int)R;
}

#define MSGS_GRP_PARAMS int n_units, const msgs_group_unit* units, int* handles, int P, int D, int M, \
                             const float* means3D, const float* shs, const float* colors_precomp, \
                             const float* max_pixel_sizes, const float* min_pixel_sizes, \
                             const float* opacities, const float* occ_multiplier, const float* dc_delta, \
                             const float* scales, float scale_modifier, const float* rotations, \
                             const float* cov3D_precomp, const uint8_t* base_mask, \
                             const float* viewmatrix, const float* projmatrix, const float* cam_pos, \
                             float tan_fovx, float tan_fovy, \
                             int prefiltered, int filter_small, int filter_large, float fade_size, int debug
#define MSGS_GRP_ARGS n_units, units, handles, P, D, M, means3D, shs, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities, \
                             occ_multiplier, dc_delta, scales, scale_modifier, rotations, cov3D_precomp, base_mask, viewmatrix, \
                             projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, filter_small, filter_large, fade_size, debug

}  // extern "C"

static int forward_group_begin_impl(int raw, MSGS_GRP_PARAMS) {
    (void)occ_multiplier; (void)dc_delta; (void)fade_size;   // dead inputs in the reference too (quirk Q9)
    if (n_units < 1 || n_units > MAX_GROUP || !units || !handles) return fail(MSGS_ERR_ARG, "msgs_forward_group_begin: need 1..8 units");
    if (P <= 0) return fail(MSGS_ERR_ARG, "msgs_forward_group_begin: P must be positive (use msgs_forward for P == 0)");
    if (!means3D || !opacities || !max_pixel_sizes || !min_pixel_sizes || !base_mask || !viewmatrix || !projmatrix || !cam_pos)
        return fail(MSGS_ERR_ARG, "msgs_forward: null required input");
    if (!colors_precomp && !shs) return fail(MSGS_ERR_ARG, "msgs_forward: need shs or colors_precomp");
    if (!cov3D_precomp && (!scales || !rotations)) return fail(MSGS_ERR_ARG, "msgs_forward: need scales+rotations or cov3D_precomp");
    if (!colors_precomp && (M <= 0 || M < (D + 1) * (D + 1) || D < 0 || D > 3))
        return fail(MSGS_ERR_ARG, "msgs_forward: SH degree / coefficient count mismatch");
    int slots[MAX_GROUP], found = 0;
    for (int s = 0; s < MAX_PENDING && found < n_units; ++s) if (!g_pending[s].used) slots[found++] = s;
    if (found < n_units) return fail(MSGS_ERR_ARG, "msgs_forward_begin: too many forward calls in flight (finish some first)");
    for (int i = 0; i < n_units; ++i) {
        const msgs_group_unit& u = units[i];
        if (u.width <= 0 || u.height <= 0) return fail(MSGS_ERR_ARG, "msgs_forward: bad P / image size");
        if (!u.geom_alloc || !u.binning_alloc || !u.img_alloc) return fail(MSGS_ERR_ARG, "msgs_forward: null allocator");
        if (!u.background || !u.out_color || !u.out_acc_pixel_size || !u.out_depth || !u.radii || !u.pixel_sizes)
            return fail(MSGS_ERR_ARG, "msgs_forward: null output/background");
    }
    cudaStream_t st = (cudaStream_t)units[0].stream;       // the group's stream: shared preprocess + depth sort

    FwdParams prm;
    prm.P = P; prm.D = D; prm.M = M; prm.W = units[0].width; prm.H = units[0].height;
    prm.gx = (prm.W + TILE_X - 1) / TILE_X; prm.gy = (prm.H + TILE_Y - 1) / TILE_Y;
    prm.scale_modifier = scale_modifier; prm.tan_fovx = tan_fovx; prm.tan_fovy = tan_fovy;
    prm.focal_y = prm.H / (2.0f * tan_fovy);
    prm.focal_x = prm.W / (2.0f * tan_fovx);
    prm.prefiltered = prefiltered; prm.filter_small = filter_small; prm.filter_large = filter_large;
    prm.raw = raw;
    prm.view = viewmatrix; prm.proj = projmatrix; prm.campos = cam_pos;

    GroupParams gp;
    gp.S = n_units;
    GeomState gs[MAX_GROUP]; ImageState imgs[MAX_GROUP]; BinPlan bps[MAX_GROUP];
    for (int i = 0; i < n_units; ++i) {
        const msgs_group_unit& u = units[i];
        const int gx = (u.width + TILE_X - 1) / TILE_X, gy = (u.height + TILE_Y - 1) / TILE_Y;
        char* geom_chunk = (char*)u.geom_alloc(u.geom_ctx, geom_bytes((size_t)P));
        if (!geom_chunk) return fail(MSGS_ERR_ALLOC, "geometry buffer allocation failed");
        gs[i] = geom_from_chunk(geom_chunk, (size_t)P);
        char* img_chunk = (char*)u.img_alloc(u.img_ctx, image_bytes(u.width, u.height, (size_t)P));
        if (!img_chunk) return fail(MSGS_ERR_ALLOC, "image buffer allocation failed");
        imgs[i] = image_from_chunk(img_chunk, u.width, u.height, (size_t)P);
        bps[i] = make_bin_plan((size_t)P, gx, gy);
        if (!bin_plan_supported(bps[i])) return fail(MSGS_ERR_ARG, "msgs_forward: image too large for the placement tables");
        GroupScale& o = gp.sc[i];
        o.W = u.width; o.H = u.height; o.gx = gx; o.gy = gy;
        o.focal_y = u.height / (2.0f * tan_fovy);           // rasterizer_impl.cu:236-237, per unit
        o.focal_x = u.width / (2.0f * tan_fovx);
        o.radii = u.radii; o.pixel_sizes = u.pixel_sizes;
        o.rec = gs[i].rec; o.cov3D = gs[i].cov3D; o.clamped = gs[i].clamped; o.tiles_touched = gs[i].tiles_touched;
        o.rect = gs[i].rect; o.gacc = gs[i].gacc; o.meta = gs[i].meta;
        // counters: unit 0 also clears the depth sort's histogram / tickets / look-back words behind its meta
        const size_t clr = (i == 0) ? (size_t)((char*)gs[0].dsort_temp - (char*)gs[0].meta) + sort_temp_clear_bytes((size_t)P)
                                    : (size_t)32 * sizeof(uint32_t);
        CUDA_TRY(cudaMemsetAsync(gs[i].meta, 0, clr, st));
    }
    { StageScope sc(ST_PREPROCESS_FWD, st, 1);
      launch_preprocess_fwd_group(prm, gp, means3D, scales, rotations, opacities, shs, cov3D_precomp, colors_precomp,
                                  max_pixel_sizes, min_pixel_sizes, base_mask, gs[0].dkeys[0], gs[0].dvals[0], st); }
    STAGE_CHECK("preprocess (group)");
    {
        const SortPlan dpl = make_sort_plan(0, 32);
        StageScope sc(ST_DEPTH_SORT, st, 3 + dpl.passes);
        launch_sort_histogram<uint32_t>(gs[0].dkeys[0], (size_t)P, dpl, gs[0].dsort_temp, st);
        int cur = 0;
        for (int p = 0; p < dpl.passes; ++p) {
            launch_sort_pass<uint32_t>(gs[0].dkeys[cur], p + 1 < dpl.passes ? gs[0].dkeys[cur ^ 1] : nullptr,
                                       p == 0 ? nullptr : gs[0].dvals[cur], gs[0].dvals[cur ^ 1], (size_t)P,
                                       dpl, p, sort_temp_hist(gs[0].dsort_temp), gs[0].dsort_temp, st);
            cur ^= 1;
        }
    }
    STAGE_CHECK("depth sort (group)");
    static thread_local cudaEvent_t sorted_ev[64] = {};          // one per device this host thread has used
    int cur_dev = 0;
    CUDA_TRY(cudaGetDevice(&cur_dev));
    cudaEvent_t& sorted = sorted_ev[cur_dev & 63];
    if (!sorted) CUDA_TRY(cudaEventCreateWithFlags(&sorted, cudaEventDisableTiming));
    CUDA_TRY(cudaEventRecord(sorted, st));
    // unit tails coarsest first: a coarse unit is a few CTAs with very long lists (its blend runs as long as the
    // full-resolution one), so it must be in flight before the full-resolution kernels fill the GPU
    int order[MAX_GROUP];
    for (int i = 0; i < n_units; ++i) order[i] = i;
    for (int a = 1; a < n_units; ++a)
        for (int b = a; b > 0 && (long long)units[order[b]].width * units[order[b]].height <
                                 (long long)units[order[b - 1]].width * units[order[b - 1]].height; --b) {
            const int t = order[b]; order[b] = order[b - 1]; order[b - 1] = t;
        }
    for (int oi = 0; oi < n_units; ++oi) {
        const int i = order[oi];
        const msgs_group_unit& u = units[i];
        cudaStream_t ust = (cudaStream_t)u.stream;
        if (ust != st) CUDA_TRY(cudaStreamWaitEvent(ust, sorted, 0));
        Pending& pd = g_pending[slots[i]];
        pd.immediate = false;
        pd.g = gs[i];
        pd.g.dvals[0] = gs[0].dvals[0];                     // the group's depth order (lives in unit 0's buffer)
        pd.img = imgs[i]; pd.bp = bps[i];
        pd.P = P; pd.W = u.width; pd.H = u.height; pd.gx = gp.sc[i].gx; pd.gy = gp.sc[i].gy; pd.debug = debug;
        pd.background = u.background; pd.out_color = u.out_color; pd.out_acc = u.out_acc_pixel_size; pd.out_depth = u.out_depth;
        pd.binning_alloc = u.binning_alloc; pd.binning_ctx = u.binning_ctx; pd.st = ust;
        const int rc = begin_tail(pd, slots[i]);
        if (rc < 0) {
            for (int k = 0; k < oi; ++k) g_pending[slots[order[k]]].used = false;
            return rc;
        }
        handles[i] = slots[i];
    }
    return MSGS_OK;
}

extern "C" {

int msgs_forward_group_begin(MSGS_GRP_PARAMS) { return forward_group_begin_impl(0, MSGS_GRP_ARGS); }
int msgs_forward_group_begin_raw(MSGS_GRP_PARAMS) { return forward_group_begin_impl(1, MSGS_GRP_ARGS); }

int msgs_forward(msgs_alloc_fn geom_alloc, void* geom_ctx, msgs_alloc_fn binning_alloc, void* binning_ctx,
                 msgs_alloc_fn img_alloc, void* img_ctx, int P, int D, int M,
                 const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* max_pixel_sizes, const float* min_pixel_sizes,
                 const float* opacities, const float* occ_multiplier, const float* dc_delta,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const uint8_t* base_mask,
                 const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                 float tan_fovx, float tan_fovy, int prefiltered, int filter_small, int filter_large, float fade_size,
                 float* out_color, float* out_acc_pixel_size, float* out_depth,
                 int* radii, float* pixel_sizes, int debug, void* stream) {
    const int h = msgs_forward_begin(geom_alloc, geom_ctx, binning_alloc, binning_ctx, img_alloc, img_ctx, P, D, M, background,
                                     width, height, means3D, shs, colors_precomp, max_pixel_sizes, min_pixel_sizes, opacities,
                                     occ_multiplier, dc_delta, scales, scale_modifier, rotations, cov3D_precomp, base_mask,
                                     viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, filter_small,
                                     filter_large, fade_size, out_color, out_acc_pixel_size, out_depth, radii, pixel_sizes,
                                     debug, stream);
    return h < 0 ? h : msgs_forward_finish(h);
}

}  // extern "C"

static int backward_impl(int accumulate, int part, int first, int count, const float* opacity_raw, int P, int D, int M, int R, const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* pixel_sizes, const float* max_pixel_sizes, const float* min_pixel_sizes,
                  const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const uint8_t* base_mask,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tan_fovx, float tan_fovy, const int* radii,
                  void* geom_buffer, void* binning_buffer, void* image_buffer,
                  int filter_small, int filter_large, float fade_size, const float* dL_dpix,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                  float* dL_docc_multiplier, float* dL_ddc_delta, float* dL_dcolor,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int debug, void* stream) {
    (void)fade_size; (void)colors_precomp;
    cudaStream_t st = (cudaStream_t)stream;
    if (P < 0 || R < 0 || width <= 0 || height <= 0) return fail(MSGS_ERR_ARG, "msgs_backward: bad sizes");
    if (P == 0) return MSGS_OK;   // rasterize_points.cu:196
    if (!geom_buffer || !image_buffer || (R > 0 && !binning_buffer)) return fail(MSGS_ERR_ARG, "msgs_backward: null state buffer");
    if (!means3D || !radii || !pixel_sizes || !max_pixel_sizes || !min_pixel_sizes || !base_mask || !viewmatrix ||
        !projmatrix || !campos || !dL_dpix || !background)
        return fail(MSGS_ERR_ARG, "msgs_backward: null required input");
    if (!dL_dopacity || !dL_dmean3D || !dL_dscale || !dL_drot || (M > 0 && !dL_dsh))
        return fail(MSGS_ERR_ARG, "msgs_backward: null gradient output");
    // dL_dmean2D, dL_dconic, dL_dcolor, dL_dcov3D, dL_docc_multiplier, dL_ddc_delta may be NULL (not wanted: skipped)
    const int gx = (width + TILE_X - 1) / TILE_X, gy = (height + TILE_Y - 1) / TILE_Y;

    GeomState g = geom_from_chunk((char*)geom_buffer, (size_t)P);
    ImageState img = image_from_chunk((char*)image_buffer, width, height, (size_t)P);

    // part: bit 0 = blend backward, bit 1 = preprocess backward over the Gaussians [first, first + count)
    if (first < 0 || (first % 128) != 0 || first > P) return fail(MSGS_ERR_ARG, "msgs_backward: first must be a multiple of 128 in [0, P]");
    const int end = (count < 0 || count > P - first) ? P : first + count;
    if ((part & 1) && R > 0) {
        BinningState bin = binning_from_chunk((char*)binning_buffer, (size_t)R);
        { StageScope sc(ST_BLEND_BWD, st, 1);
          launch_blend_bwd(width, height, gx, gy, img.ranges, bin.point_list, g.rec, background, img.final_T, img.n_contrib,
                           img.tile_max_contrib, dL_dpix, g.gacc, filter_small, filter_large, st); }
        STAGE_CHECK("blend backward");
    }
    if (!(part & 2)) return MSGS_OK;
    BwdParams prm;
    prm.P = end; prm.first = first; prm.D = D; prm.M = M; prm.W = width; prm.H = height;
    prm.scale_modifier = scale_modifier; prm.tan_fovx = tan_fovx; prm.tan_fovy = tan_fovy;
    prm.focal_y = height / (2.0f * tan_fovy);
    prm.focal_x = width / (2.0f * tan_fovx);
    prm.filter_small = filter_small; prm.filter_large = filter_large;
    prm.raw = opacity_raw != nullptr; prm.opacity_raw = opacity_raw;
    prm.view = viewmatrix; prm.proj = projmatrix; prm.campos = campos;
    { StageScope sc(ST_PREPROCESS_BWD, st, 1);
    launch_preprocess_bwd(prm, means3D, shs, scales, rotations, cov3D_precomp, radii, pixel_sizes, max_pixel_sizes,
                          min_pixel_sizes, base_mask, g, dL_dmean2D, dL_dconic, dL_dopacity, dL_docc_multiplier,
                          dL_ddc_delta, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, accumulate, st); }
    STAGE_CHECK("preprocess backward");
    return MSGS_OK;
}

#define MSGS_BWD_ARGS P, D, M, R, background, width, height, means3D, shs, colors_precomp, pixel_sizes, max_pixel_sizes, \
    min_pixel_sizes, scales, scale_modifier, rotations, cov3D_precomp, base_mask, viewmatrix, projmatrix, campos, \
    tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, image_buffer, filter_small, filter_large, fade_size, dL_dpix, \
    dL_dmean2D, dL_dconic, dL_dopacity, dL_docc_multiplier, dL_ddc_delta, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, \
    dL_dscale, dL_drot, debug, stream
#define MSGS_BWD_PARAMS int P, int D, int M, int R, const float* background, int width, int height, \
    const float* means3D, const float* shs, const float* colors_precomp, const float* pixel_sizes, \
    const float* max_pixel_sizes, const float* min_pixel_sizes, const float* scales, float scale_modifier, \
    const float* rotations, const float* cov3D_precomp, const uint8_t* base_mask, const float* viewmatrix, \
    const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii, void* geom_buffer, \
    void* binning_buffer, void* image_buffer, int filter_small, int filter_large, float fade_size, const float* dL_dpix, \
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_docc_multiplier, float* dL_ddc_delta, \
    float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug, \
    void* stream

extern "C" {
#define MSGS_GBWD_PARAMS int n_units, const msgs_group_bwd_unit* units, int P, int D, int M, \
                                   const float* means3D, const float* shs, \
                                   const float* max_pixel_sizes, const float* min_pixel_sizes, \
                                   const float* scales, float scale_modifier, const float* rotations, \
                                   const float* cov3D_precomp, const uint8_t* base_mask, \
                                   const float* viewmatrix, const float* projmatrix, const float* campos, \
                                   float tan_fovx, float tan_fovy, int filter_small, int filter_large, \
                                   float* dL_dopacity, float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot, \
                                   int debug, void* stream
#define MSGS_GBWD_ARGS n_units, units, P, D, M, means3D, shs, max_pixel_sizes, min_pixel_sizes, scales, scale_modifier, rotations, \
                                   cov3D_precomp, base_mask, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, filter_small, \
                                   filter_large, dL_dopacity, dL_dmean3D, dL_dsh, dL_dscale, dL_drot, debug, stream
}  // extern "C"

static int backward_group_impl(const float* opacity_raw, MSGS_GBWD_PARAMS, int first = 0, int count = -1) {
    cudaStream_t st = (cudaStream_t)stream;
    if (first < 0 || (first % 128) != 0 || first > (P > 0 ? P : 0))
        return fail(MSGS_ERR_ARG, "msgs_backward_group_accumulate: first must be a multiple of 128 in [0, P]");
    if (n_units < 1 || n_units > MAX_GROUP || !units) return fail(MSGS_ERR_ARG, "msgs_backward_group_accumulate: need 1..8 units");
    if (P < 0) return fail(MSGS_ERR_ARG, "msgs_backward: bad sizes");
    if (P == 0) return MSGS_OK;
    if (!shs || M != 16 || D < 0 || D > 3) return fail(MSGS_ERR_ARG, "msgs_backward_group_accumulate: SH colours with M == 16 only");
    if (!cov3D_precomp && (!scales || !rotations)) return fail(MSGS_ERR_ARG, "msgs_backward: need scales+rotations or cov3D_precomp");
    if (!means3D || !max_pixel_sizes || !min_pixel_sizes || !base_mask || !viewmatrix || !projmatrix || !campos)
        return fail(MSGS_ERR_ARG, "msgs_backward: null required input");
    if (!dL_dopacity || !dL_dmean3D || !dL_dsh || !dL_dscale || !dL_drot) return fail(MSGS_ERR_ARG, "msgs_backward: null gradient output");
    GroupBwdParams gp;
    gp.S = n_units;
    for (int i = 0; i < n_units; ++i) {
        const msgs_group_bwd_unit& u = units[i];
        if (u.width <= 0 || u.height <= 0 || !u.radii || !u.pixel_sizes || !u.geom_buffer)
            return fail(MSGS_ERR_ARG, "msgs_backward_group_accumulate: bad unit");
        const GeomState g = geom_from_chunk((char*)u.geom_buffer, (size_t)P);
        GroupBwdScale& o = gp.sc[i];
        o.focal_y = u.height / (2.0f * tan_fovy);
        o.focal_x = u.width / (2.0f * tan_fovx);
        o.radii = u.radii; o.pixel_sizes = u.pixel_sizes;
        o.cov3D = g.cov3D; o.clamped = g.clamped; o.gacc = g.gacc; o.dL_dmean2D = u.dL_dmean2D;
    }
    BwdParams prm;
    prm.P = (count < 0 || count > P - first) ? P : first + count; prm.first = first;
    prm.D = D; prm.M = M; prm.W = units[0].width; prm.H = units[0].height;
    prm.scale_modifier = scale_modifier; prm.tan_fovx = tan_fovx; prm.tan_fovy = tan_fovy;
    prm.focal_y = gp.sc[0].focal_y; prm.focal_x = gp.sc[0].focal_x;
    prm.filter_small = filter_small; prm.filter_large = filter_large;
    prm.raw = opacity_raw != nullptr; prm.opacity_raw = opacity_raw;
    prm.view = viewmatrix; prm.proj = projmatrix; prm.campos = campos;
    { StageScope sc(ST_PREPROCESS_BWD, st, 1);
      launch_preprocess_bwd_group(prm, gp, means3D, shs, scales, rotations, cov3D_precomp, max_pixel_sizes, min_pixel_sizes,
                                  base_mask, dL_dopacity, dL_dmean3D, dL_dsh, dL_dscale, dL_drot, st); }
    STAGE_CHECK("preprocess backward (group)");
    return MSGS_OK;
}
// The whole backward of a view group in ONE call: every unit's blend backward on the unit's stream (coarsest first, see
// forward_group_begin_impl), then the group's preprocess backward on `stream` behind all of them.
static int backward_group_full_impl(int part, const float* opacity_raw, const msgs_group_bwd_blend* blends, MSGS_GBWD_PARAMS) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_units < 1 || n_units > MAX_GROUP || !units || !blends) return fail(MSGS_ERR_ARG, "msgs_backward_group_full: need 1..8 units");
    if (P < 0) return fail(MSGS_ERR_ARG, "msgs_backward: bad sizes");
    if (P == 0) return MSGS_OK;
    static thread_local cudaEvent_t evs[64][MAX_GROUP + 1] = {};      // per device this host thread has used
    int cur_dev = 0;
    CUDA_TRY(cudaGetDevice(&cur_dev));
    cudaEvent_t* ev = evs[cur_dev & 63];
    for (int i = 0; i <= MAX_GROUP; ++i)
        if (!ev[i]) CUDA_TRY(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    CUDA_TRY(cudaEventRecord(ev[MAX_GROUP], st));                     // what the caller enqueued on `stream` (dL_dpix, zeroed sums)
    int order[MAX_GROUP];
    for (int i = 0; i < n_units; ++i) order[i] = i;
    for (int a = 1; a < n_units; ++a)
        for (int b = a; b > 0 && (long long)units[order[b]].width * units[order[b]].height <
                                 (long long)units[order[b - 1]].width * units[order[b - 1]].height; --b) {
            const int t = order[b]; order[b] = order[b - 1]; order[b - 1] = t;
        }
    for (int oi = 0; oi < n_units; ++oi) {
        const int i = order[oi];
        const msgs_group_bwd_unit& u = units[i];
        const msgs_group_bwd_blend& bl = blends[i];
        if (u.width <= 0 || u.height <= 0 || !u.geom_buffer || !bl.image_buffer || !bl.dL_dpix || !bl.background || bl.R < 0 ||
            (bl.R > 0 && !bl.binning_buffer))
            return fail(MSGS_ERR_ARG, "msgs_backward_group_full: bad unit");
        cudaStream_t ust = (cudaStream_t)bl.stream;
        if (ust != st) CUDA_TRY(cudaStreamWaitEvent(ust, ev[MAX_GROUP], 0));
        if (bl.R > 0) {
            const int gx = (u.width + TILE_X - 1) / TILE_X, gy = (u.height + TILE_Y - 1) / TILE_Y;
            const GeomState g = geom_from_chunk((char*)u.geom_buffer, (size_t)P);
            const ImageState img = image_from_chunk((char*)bl.image_buffer, u.width, u.height, (size_t)P);
            const BinningState bin = binning_from_chunk((char*)bl.binning_buffer, (size_t)bl.R);
            StageScope sc(ST_BLEND_BWD, ust, 1);
            launch_blend_bwd(u.width, u.height, gx, gy, img.ranges, bin.point_list, g.rec, bl.background, img.final_T, img.n_contrib,
                             img.tile_max_contrib, bl.dL_dpix, g.gacc, filter_small, filter_large, ust);
        }
        {   // STAGE_CHECK on the unit's stream
            cudaError_t e_ = cudaPeekAtLastError();
            if (e_ == cudaSuccess && debug) e_ = cudaStreamSynchronize(ust);
            if (e_ != cudaSuccess) { cudaGetLastError(); return fail(MSGS_ERR_CUDA, "blend backward (group)", e_); }
        }
        if (ust != st) CUDA_TRY(cudaEventRecord(ev[i], ust));
    }
    // only now: the group's stream carries one of the blend kernels itself (the full-resolution unit's, launched last)
    for (int i = 0; i < n_units; ++i)
        if ((cudaStream_t)blends[i].stream != st) CUDA_TRY(cudaStreamWaitEvent(st, ev[i], 0));
    if (!(part & 2)) return MSGS_OK;
    return backward_group_impl(opacity_raw, MSGS_GBWD_ARGS);
}
extern "C" {
int msgs_backward_group_full(const float* opacity_raw, const msgs_group_bwd_blend* blends, MSGS_GBWD_PARAMS) {
    return backward_group_full_impl(3, opacity_raw, blends, MSGS_GBWD_ARGS);
}
/* the blend half only: `stream` ends up waiting for every unit's blend backward; the preprocess half follows as one
 * msgs_backward_group_accumulate[_raw] call or as msgs_backward_group_accumulate_part slices */
int msgs_backward_group_blend(const float* opacity_raw, const msgs_group_bwd_blend* blends, MSGS_GBWD_PARAMS) {
    return backward_group_full_impl(1, opacity_raw, blends, MSGS_GBWD_ARGS);
}
/* Gaussians [first, first + count) of the group's preprocess backward (first a multiple of 128; count < 0: to the end): the
 * slices of one pass may be exchanged / optimised while the next slice is being computed (msgs_train, sharded step) */
int msgs_backward_group_accumulate_part(const float* opacity_raw, int first, int count, MSGS_GBWD_PARAMS) {
    return backward_group_impl(opacity_raw, MSGS_GBWD_ARGS, first, count);
}
int msgs_backward_group_accumulate(MSGS_GBWD_PARAMS) { return backward_group_impl(nullptr, MSGS_GBWD_ARGS); }
/* raw-parameter variants: scales / rotations are the raw parameters, opacity_raw (P) the raw opacities; the
 * gradients written / accumulated are those of the raw parameters (exp / normalize / sigmoid backward chained in registers) */
int msgs_backward_group_accumulate_raw(const float* opacity_raw, MSGS_GBWD_PARAMS) {
    if (!opacity_raw) return fail(MSGS_ERR_ARG, "msgs_backward_group_accumulate_raw: null opacity_raw");
    return backward_group_impl(opacity_raw, MSGS_GBWD_ARGS);
}
int msgs_backward(MSGS_BWD_PARAMS) { return backward_impl(0, 3, 0, -1, nullptr, MSGS_BWD_ARGS); }
int msgs_backward_accumulate(MSGS_BWD_PARAMS) { return backward_impl(1, 3, 0, -1, nullptr, MSGS_BWD_ARGS); }
int msgs_backward_accumulate_part(MSGS_BWD_PARAMS, int part, int first, int count) { return backward_impl(1, part, first, count, nullptr, MSGS_BWD_ARGS); }
/* accumulate: 0 = write every gradient element once (msgs_backward), 1 = add the training gradients (msgs_backward_accumulate);
 * part: bit 0 = blend backward, bit 1 = preprocess backward (3 = both) */
int msgs_backward_raw(const float* opacity_raw, int accumulate, int part, MSGS_BWD_PARAMS) {
    if (!opacity_raw) return fail(MSGS_ERR_ARG, "msgs_backward_raw: null opacity_raw");
    if (part < 1 || part > 3) return fail(MSGS_ERR_ARG, "msgs_backward_raw: part must be 1, 2 or 3");
    return backward_impl(accumulate ? 1 : 0, part, 0, -1, opacity_raw, MSGS_BWD_ARGS);
}
}  // extern "C"
