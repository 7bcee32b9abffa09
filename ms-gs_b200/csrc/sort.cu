// Hand-written onesweep LSD radix sort of (key, u32 value) pairs — replaces cub::DeviceRadixSort::SortPairs
// (rasterizer_impl.cu:190-193, 327-332).  Stable, ascending.  Keys are u32 (the depth sort of the forward) or u64 (public
// msgs_sort_pairs).  Traffic: one up-front histogram read of the keys plus, per pass, one read and one write of every
// pair; each pass is ONE kernel whose tiles chain their per-digit prefixes with a decoupled look-back (no separate
// scan / scatter launches, no per-tile histogram round trip through HBM).  At the sizes of the rendering path (1-6 M
// keys, resident in L2) the sort is a latency chain, and the structure below is what the per-phase measurements of
// tools/sort_microbench.cu led to (profiles/r02y_sort_microbench.md).
//
// Pass kernel, per 512-thread CTA and tile of SORT_TILE = 8192 pairs (one CTA per SM):
//   0. take a ticket and clear the counters, THEN wait for the previous kernel (programmatic dependent launch);
//   1. the ticket is the tile (predecessors are then resident or finished); load keys warp-striped (coalesced);
//   2. rank them inside each warp — lanes with the same digit found with 8 ballots (stable: item order, then lane
//      order), per-warp digit counters in shared memory;
//   3. scan the counters across warps and digits (thread d < 256 owns digit d), publish the tile's digit counts,
//      scatter keys and values to their tile-local sorted slot in shared memory, and only then look back: the counts
//      of the earlier tiles of the group of 16, plus a chain of group words (two-level look-back);
//   4. stream the sorted tile out so that each digit's run is written with coalesced stores.  A pass without an output
//      key array writes none (the forward's last pass); a pass without input values takes the pair's position.
#include "common.cuh"

namespace msgs {

constexpr int RADIX_BITS = 8;
constexpr int BINS = 1 << RADIX_BITS;
// Tunables.  The defaults are the configuration measured fastest on the depth sort (1.2 M u32 keys, B200) with
// tools/sort_microbench.cu, which builds this file with other values (-DMSGS_SORT_...): profiles/r02y_sort_microbench.md.
#ifndef MSGS_SORT_THREADS
#define MSGS_SORT_THREADS 512        // threads per pass CTA (>= BINS: thread d < BINS owns digit d in the scan / look-back steps)
#endif
#ifndef MSGS_SORT_ITEMS
#define MSGS_SORT_ITEMS 16           // pairs per thread: tiles of 8192 pairs (1.2 M keys = 147 tiles = one CTA per SM)
#endif
#ifndef MSGS_SORT_WINDOW
#define MSGS_SORT_WINDOW 4           // look-back: group words read per round trip (independent loads in flight)
#endif
#ifndef MSGS_SORT_HIST_BATCH
#define MSGS_SORT_HIST_BATCH 8       // histogram: keys loaded per thread before any of them is counted
#endif
#ifndef MSGS_SORT_HIST_THREADS
#define MSGS_SORT_HIST_THREADS 1024  // histogram: one fat CTA per SM (fewer same-address flushes into the global bins)
#endif
#ifndef MSGS_SORT_HIST_CPS
#define MSGS_SORT_HIST_CPS 1         // histogram CTAs per SM
#endif
#ifndef MSGS_SORT_FUSE_SCAN
#define MSGS_SORT_FUSE_SCAN 1        // every pass CTA scans the pass's 256 digit totals itself (no scan launch)
#endif
#ifndef MSGS_SORT_BALLOT
#define MSGS_SORT_BALLOT 1           // in-warp matching with 8 ballots per item instead of match.any
#endif
#ifndef MSGS_SORT_LATE_LOOKBACK
#define MSGS_SORT_LATE_LOOKBACK 1    // scatter into shared memory BEFORE the look-back (predecessors publish meanwhile)
#endif
#ifndef MSGS_SORT_TWO_LEVEL
#define MSGS_SORT_TWO_LEVEL 1        // look-back over groups of 16 tiles (in-group aggregates + a group-level chain)
#endif
#ifndef MSGS_SORT_PDL
#define MSGS_SORT_PDL 1              // pass kernels launched with programmatic stream serialization: the prologue (ticket,
#endif                               // shared-memory clear) overlaps the previous kernel's tail; griddepcontrol.wait before the first dependent read
constexpr int SORT_THREADS = MSGS_SORT_THREADS;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int DIGIT_WARPS = 256 / 32;              // warps holding the digit threads
constexpr int SORT_ITEMS = MSGS_SORT_ITEMS;
constexpr int SORT_WINDOW = MSGS_SORT_WINDOW;
constexpr int HIST_BATCH = MSGS_SORT_HIST_BATCH;
#ifdef MSGS_SORT_PHASES
// tools/sort_microbench.cu: ns spent by thread 0 of every CTA in [load + rank | scan + look-back | scatter + store], and CTAs
__device__ unsigned long long g_sort_phase[6];
__device__ unsigned long long g_sort_trace[1024 * 4];     // pass 1 of the last sort: per tile {start, ranked, looked back, end}
__device__ __forceinline__ unsigned long long sort_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#endif
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;      // 8192 pairs per tile
constexpr uint32_t COUNT_MASK = (1u << 28) - 1u;          // look-back word: 4 flag bits (2 values per pass) | count

static_assert(SORT_MAX_PASSES == 7, "look-back flag encoding assumes at most 7 passes");

SortPlan make_sort_plan(int begin_bit, int end_bit) {
    SortPlan pl;
    pl.passes = 0;
    for (int b = begin_bit; b < end_bit && pl.passes < SORT_MAX_PASSES; b += RADIX_BITS) {
        const int nbits = min(RADIX_BITS, end_bit - b);
        pl.shift[pl.passes] = b;
        pl.mask[pl.passes] = (1u << nbits) - 1u;      // bits >= end_bit never influence the order
        ++pl.passes;
    }
    return pl;
}

static size_t sort_tiles(size_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }
constexpr int SORT_GROUP = 16;                             // tiles per look-back group
static size_t sort_groups(size_t n) { return sort_tiles(n) / SORT_GROUP + 1; }
// temp layout: [hist: SORT_MAX_PASSES x BINS u32][tickets: 8 u32][look-back: tiles x BINS u32][group look-back: groups x BINS u32]
size_t sort_temp_bytes(size_t n) {
    return (size_t)SORT_MAX_PASSES * BINS * 4 + 8 * 4 + (sort_tiles(n) + sort_groups(n) + 1) * BINS * 4 + 256;
}
uint32_t* sort_temp_hist(void* temp) { return reinterpret_cast<uint32_t*>(temp); }
size_t sort_temp_clear_bytes(size_t n) { return (size_t)SORT_MAX_PASSES * BINS * 4 + 8 * 4 + (sort_tiles(n) + sort_groups(n)) * BINS * 4; }
void sort_temp_clear(void* temp, size_t n, cudaStream_t st) { cudaMemsetAsync(temp, 0, sort_temp_clear_bytes(n), st); }

__device__ __forceinline__ void pdl_launch_dependents() {
#if MSGS_SORT_PDL
    asm volatile("griddepcontrol.launch_dependents;");
#endif
}
__device__ __forceinline__ void pdl_wait() {
#if MSGS_SORT_PDL
    asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}

// ---- up-front histogram of every pass's digit ----------------------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(MSGS_SORT_HIST_THREADS)
sort_histogram_kernel(const KeyT* __restrict__ keys, size_t n, SortPlan pl, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * BINS];
    pdl_launch_dependents();             // the first pass may set itself up meanwhile (it waits for this grid before reading)
    for (int i = threadIdx.x; i < pl.passes * BINS; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    // HIST_BATCH keys per thread are in flight before the first is counted (one memory round trip per batch, not per key)
    for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * HIST_BATCH) {
        KeyT kk[HIST_BATCH];
#pragma unroll
        for (int b = 0; b < HIST_BATCH; ++b) {
            const size_t i = i0 + (size_t)b * stride;
            kk[b] = (i < n) ? keys[i] : (KeyT)0;
        }
#pragma unroll
        for (int b = 0; b < HIST_BATCH; ++b) {
            const bool valid = i0 + (size_t)b * stride < n;
            if (!valid) break;
            const KeyT k = kk[b];
            // unrolled over the maximum: constant indices keep the plan in parameter space (a runtime index copies it to
            // local memory and reads shift / mask back from there for every key and pass)
#pragma unroll
            for (int p = 0; p < SORT_MAX_PASSES; ++p) {
                if (p >= pl.passes) break;
                const uint32_t d = (uint32_t)(k >> pl.shift[p]) & pl.mask[p];
                if (p + 1 < pl.passes) {
                    atomicAdd(&s_hist[p * BINS + d], 1u);          // low digits: spread over the bins
                } else {
                    // the top digit takes few distinct values (float exponents, tile ids): aggregate inside the warp
                    const unsigned peers = __match_any_sync(__activemask(), d);
                    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&s_hist[p * BINS + d], __popc(peers));
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pl.passes * BINS; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// exclusive scan of each pass's 256 bins (one CTA per pass)
__global__ void __launch_bounds__(BINS)
sort_scan_hist_kernel(uint32_t* hist) {
    __shared__ uint32_t s[BINS];
    uint32_t* h = hist + (size_t)blockIdx.x * BINS;
    const int t = threadIdx.x;
    const uint32_t v = h[t];
    s[t] = v;
    __syncthreads();
    for (int o = 1; o < BINS; o <<= 1) {
        const uint32_t a = (t >= o) ? s[t - o] : 0u;
        __syncthreads();
        s[t] += a;
        __syncthreads();
    }
    h[t] = s[t] - v;
}

// ---- one radix pass ---------------------------------------------------------------------------------------
// Lanes of the warp whose (valid) item has digit d.  match.any costs one round per DISTINCT value in the warp (measured:
// 5 us of a 15 us pass CTA on random low digits, tools/sort_microbench.cu); eight ballots cost the same whatever the data
// and are independent between items.  Invalid (padding) lanes match only themselves.
__device__ __forceinline__ unsigned match_digit(uint32_t d, bool valid, int lane) {
    unsigned m = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
    for (int b = 0; b < RADIX_BITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, bit);
        m &= bit ? bal : ~bal;
    }
    return valid ? m : (1u << lane);
}

template <typename KeyT>
struct SortSmem {
    KeyT     keys[SORT_TILE];               // 32 / 64 KB
    uint32_t vals[SORT_TILE];               // 32 KB
    uint32_t warp_hist[SORT_WARPS][BINS];   // 16 KB
    uint32_t bin_start[BINS];               // tile-local exclusive start of each digit
    int      bin_goff[BINS];                // global offset of the digit's run minus bin_start
    uint32_t wsum[DIGIT_WARPS];
    uint32_t hsum[DIGIT_WARPS];
    uint32_t tile;
};

template <typename KeyT, bool KEYS_OUT>
#ifdef MSGS_SORT_PHASES
__global__ void __launch_bounds__(SORT_THREADS, 2)      // the timers must not cost the second CTA per SM
#elif defined(MSGS_SORT_MIN_CTAS)
__global__ void __launch_bounds__(SORT_THREADS, MSGS_SORT_MIN_CTAS)
#else
__global__ void __launch_bounds__(SORT_THREADS)
#endif
sort_onesweep_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                     size_t n, int pass, int shift, uint32_t dmask, const uint32_t* __restrict__ hist,
                     uint32_t* tickets, volatile uint32_t* lookback) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<KeyT>& sm = *reinterpret_cast<SortSmem<KeyT>*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t FLAG_AGG = (uint32_t)(2 * pass + 1) << 28, FLAG_PRE = (uint32_t)(2 * pass + 2) << 28;

#ifdef MSGS_SORT_PHASES
    const unsigned long long t0 = sort_now();
#endif
    if (tid == 0) sm.tile = atomicAdd(&tickets[pass], 1u);
    for (int i = tid; i < SORT_WARPS * BINS; i += SORT_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    pdl_launch_dependents();
    pdl_wait();                          // everything above touched only this pass's ticket and shared memory
    __syncthreads();
#ifdef MSGS_SORT_PHASES
    const unsigned long long t0a = sort_now();
#endif
    const uint32_t tile = sm.tile;
    const bool dthread = tid < BINS;
    const uint32_t htotal = dthread ? hist[pass * BINS + tid] : 0u;      // digit tid: exclusive start (or, MSGS_SORT_FUSE_SCAN, its total)
    const size_t tile_base = (size_t)tile * SORT_TILE;
    const int tile_n = (int)min((size_t)SORT_TILE, n - tile_base);

    // 1. warp-striped load
    KeyT key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
    const int wbase = warp * 32 * SORT_ITEMS + lane;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int j = wbase + i * 32;
        if (j < tile_n) {
            key[i] = keys_in[tile_base + j];
            val[i] = vals_in ? vals_in[tile_base + j] : (uint32_t)(tile_base + j);      // no values: the pair's position
        } else {
            key[i] = (KeyT)~(KeyT)0;
            val[i] = 0;
        }
    }

    // 2. in-warp stable ranking
    uint32_t* wh = sm.warp_hist[warp];
#ifdef MSGS_SORT_PHASES
    unsigned long long t0b = 0;
#endif
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const bool valid = (wbase + i * 32) < tile_n;
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        // invalid (padding) lanes get a private pseudo-digit so they match nobody
#if MSGS_SORT_BALLOT
        const unsigned peers = match_digit(d, valid, lane);
#else
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, valid ? d : (uint32_t)(BINS + lane));
#endif
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (valid && lane == leader) {
            base = wh[d];
            wh[d] = base + __popc(peers);
        }
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        rank[i] = base + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
#ifdef MSGS_SORT_PHASES
        if (i == 0) t0b = sort_now();
#endif
    }
    __syncthreads();

#ifdef MSGS_SORT_PHASES
    const unsigned long long t1 = sort_now();
#endif
    // 3. digit d (= tid): exclusive scan over warps, tile count, then scan over digits
    uint32_t count = 0;
    if (dthread) {
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            const uint32_t c = sm.warp_hist[w][tid];
            sm.warp_hist[w][tid] = count;
            count += c;
        }
        // publish this tile's digit count as early as possible
        if (tile == 0 && !MSGS_SORT_TWO_LEVEL) lookback[tid] = FLAG_PRE | count;
        else                                   lookback[(size_t)tile * BINS + tid] = FLAG_AGG | count;
    }

    // block exclusive scan of `count` over the 256 digits
    uint32_t incl = count, hincl = htotal;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
        if (MSGS_SORT_FUSE_SCAN) {
            const uint32_t hv = __shfl_up_sync(0xFFFFFFFFu, hincl, o);
            if (lane >= o) hincl += hv;
        }
    }
    if (lane == 31 && dthread) { sm.wsum[warp] = incl; sm.hsum[warp] = hincl; }
    __syncthreads();
    uint32_t wex = 0, hwex = 0;
#pragma unroll
    for (int w = 0; w < DIGIT_WARPS; ++w) if (w < warp) { wex += sm.wsum[w]; hwex += sm.hsum[w]; }
    const uint32_t start = wex + incl - count;
    const uint32_t hstart = MSGS_SORT_FUSE_SCAN ? hwex + hincl - htotal : htotal;      // first output slot of digit tid
    if (dthread) sm.bin_start[tid] = start;

#if MSGS_SORT_LATE_LOOKBACK
    __syncthreads();
    // 4a. scatter to the tile-local sorted slot first: needs no global prefix, and the predecessors publish meanwhile
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        if ((wbase + i * 32) < tile_n) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
            const uint32_t pos = sm.bin_start[d] + sm.warp_hist[warp][d] + rank[i];
            sm.keys[pos] = key[i];
            sm.vals[pos] = val[i];
        }
    }
#endif
    // look back over earlier tiles for digit `tid`
    uint32_t prefix = 0;
#if MSGS_SORT_TWO_LEVEL
    // Two levels: the aggregates of the tiles before this one in its group of 16 (all loads in flight at once), plus the sum
    // of the earlier groups from a decoupled look-back over GROUP words, which the last tile of every group publishes.
    // A chain of tiles / 16 links instead of tiles; every wait is on a tile with an earlier ticket.
    if (dthread) {
        const int r = (int)tile & (SORT_GROUP - 1), g = (int)tile / SORT_GROUP;
        volatile uint32_t* grp = lookback + (size_t)gridDim.x * BINS;
        uint32_t insum = 0;
        {
            const volatile uint32_t* row = lookback + (size_t)(tile - r) * BINS + tid;
            unsigned pending = (1u << r) - 1u;
            uint32_t v[SORT_GROUP - 1];
            while (pending) {
#pragma unroll
                for (int k = 0; k < SORT_GROUP - 1; ++k) if ((pending >> k) & 1u) v[k] = row[(size_t)k * BINS];
#pragma unroll
                for (int k = 0; k < SORT_GROUP - 1; ++k)
                    if (((pending >> k) & 1u) && (v[k] & ~COUNT_MASK) == FLAG_AGG) { insum += v[k] & COUNT_MASK; pending &= ~(1u << k); }
            }
        }
        const bool last = r == SORT_GROUP - 1;
        if (last) grp[(size_t)g * BINS + tid] = (g == 0 ? FLAG_PRE : FLAG_AGG) | (insum + count);
        uint32_t gprefix = 0;
        if (g > 0) {
            int j = g - 1;
            bool done = false;
            while (!done) {
                uint32_t v[SORT_WINDOW];
#pragma unroll
                for (int k = 0; k < SORT_WINDOW; ++k) v[k] = (j - k >= 0) ? grp[(size_t)(j - k) * BINS + tid] : 0u;
                int used = 0;
#pragma unroll
                for (int k = 0; k < SORT_WINDOW; ++k) {
                    if (done || used != k) continue;
                    const uint32_t f = v[k] & ~COUNT_MASK;
                    if (f == FLAG_PRE) { gprefix += v[k] & COUNT_MASK; done = true; }
                    else if (f == FLAG_AGG) { gprefix += v[k] & COUNT_MASK; ++used; }
                }
                j -= used;
            }
            if (last) grp[(size_t)g * BINS + tid] = FLAG_PRE | (gprefix + insum + count);
        }
        prefix = gprefix + insum;
    }
#else
    if (tile > 0 && dthread) {
        int j = (int)tile - 1;
        // SORT_WINDOW predecessors per round trip (independent loads in flight), consumed nearest first.  Tile 0 always
        // publishes an inclusive prefix, so the walk ends at the latest there.
        bool done = false;
        while (!done) {
            uint32_t v[SORT_WINDOW];
#pragma unroll
            for (int k = 0; k < SORT_WINDOW; ++k) v[k] = (j - k >= 0) ? lookback[(size_t)(j - k) * BINS + tid] : 0u;
            int used = 0;
#pragma unroll
            for (int k = 0; k < SORT_WINDOW; ++k) {
                if (done || used != k) continue;               // stopped at an earlier word of this window
                const uint32_t f = v[k] & ~COUNT_MASK;
                if (f == FLAG_PRE) { prefix += v[k] & COUNT_MASK; done = true; }
                else if (f == FLAG_AGG) { prefix += v[k] & COUNT_MASK; ++used; }
                // any other flag value: that tile has not published for this pass yet -> read again from there
            }
            j -= used;
        }
        lookback[(size_t)tile * BINS + tid] = FLAG_PRE | (prefix + count);
    }
#endif
    if (dthread) sm.bin_goff[tid] = (int)(hstart + prefix) - (int)start;
    __syncthreads();

#ifdef MSGS_SORT_PHASES
    const unsigned long long t2 = sort_now();
#endif
#if !MSGS_SORT_LATE_LOOKBACK
    // 4. scatter to the tile-local sorted slot
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        if ((wbase + i * 32) < tile_n) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
            const uint32_t pos = sm.bin_start[d] + sm.warp_hist[warp][d] + rank[i];
            sm.keys[pos] = key[i];
            sm.vals[pos] = val[i];
        }
    }
    __syncthreads();
#endif
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const int j = tid + i * SORT_THREADS;
        if (j < tile_n) {
            const KeyT k = sm.keys[j];
            const uint32_t d = (uint32_t)(k >> shift) & dmask;
            const size_t g = (size_t)(sm.bin_goff[d] + j);
            if (KEYS_OUT) keys_out[g] = k;
            vals_out[g] = sm.vals[j];
        }
    }
#ifdef MSGS_SORT_PHASES
    if (tid == 0) {
        const unsigned long long t3 = sort_now();
        atomicAdd(&g_sort_phase[0], t1 - t0); atomicAdd(&g_sort_phase[1], t2 - t1); atomicAdd(&g_sort_phase[2], t3 - t2);
        atomicAdd(&g_sort_phase[3], 1ull);
        atomicAdd(&g_sort_phase[4], t0a - t0); atomicAdd(&g_sort_phase[5], t0b - t0a);
        if (pass == 1 && tile < 1024) { g_sort_trace[tile * 4] = t0; g_sort_trace[tile * 4 + 1] = t1; g_sort_trace[tile * 4 + 2] = t2; g_sort_trace[tile * 4 + 3] = t3; }
    }
#endif
}

template <typename KeyT>
void launch_sort_histogram(const KeyT* keys_in, size_t n, const SortPlan& pl, void* temp, cudaStream_t st) {
    if (n == 0 || pl.passes == 0) return;
    uint32_t* hist = sort_temp_hist(temp);
    // 2 CTAs per SM: every CTA ends with up to passes x 256 global atomics, so few, long-running CTAs
    const int hblocks = (int)min((size_t)148 * MSGS_SORT_HIST_CPS, (n + MSGS_SORT_HIST_THREADS - 1) / MSGS_SORT_HIST_THREADS);
    sort_histogram_kernel<KeyT><<<hblocks, MSGS_SORT_HIST_THREADS, 0, st>>>(keys_in, n, pl, hist);
    if (!MSGS_SORT_FUSE_SCAN) sort_scan_hist_kernel<<<pl.passes, BINS, 0, st>>>(hist);
}

template <typename KeyT>
void launch_sort_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out, size_t n,
                      const SortPlan& pl, int p, const uint32_t* hist, void* temp, cudaStream_t st) {
    if (n == 0) return;
    // With programmatic dependent launch the CTAs of the NEXT pass may become resident while this pass runs.  Measured
    // (tools/sort_microbench.cu): that pays when they follow the primaries SM by SM (one CTA per SM: 68.7 -> 58.0 us for
    // 1.2 M keys) and costs a factor 2 when they share the SMs with them (71 -> 130 us).  The shared-memory request is
    // therefore padded past half an SM's capacity: one pass CTA per SM, whatever the register count.
    const size_t smem = MSGS_SORT_PDL ? max(sizeof(SortSmem<KeyT>), (size_t)116 * 1024) : sizeof(SortSmem<KeyT>);
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaFuncSetAttribute(sort_onesweep_kernel<KeyT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(sort_onesweep_kernel<KeyT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set[dev] = true;
    }
    uint32_t* tickets = sort_temp_hist(temp) + SORT_MAX_PASSES * BINS;
    uint32_t* lookback = tickets + 8;
    const unsigned grid = (unsigned)sort_tiles(n);
#if MSGS_SORT_PDL
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(SORT_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    const int shift = pl.shift[p];
    const uint32_t dmask = pl.mask[p];
    KeyT* ko = keys_out;
    volatile uint32_t* lb = lookback;
    if (keys_out) cudaLaunchKernelEx(&cfg, sort_onesweep_kernel<KeyT, true>, keys_in, ko, vals_in, vals_out, n, p, shift, dmask, hist, tickets, lb);
    else          cudaLaunchKernelEx(&cfg, sort_onesweep_kernel<KeyT, false>, keys_in, ko, vals_in, vals_out, n, p, shift, dmask, hist, tickets, lb);
#else
    if (keys_out)
        sort_onesweep_kernel<KeyT, true><<<grid, SORT_THREADS, smem, st>>>(
            keys_in, keys_out, vals_in, vals_out, n, p, pl.shift[p], pl.mask[p], hist, tickets, lookback);
    else
        sort_onesweep_kernel<KeyT, false><<<grid, SORT_THREADS, smem, st>>>(
            keys_in, nullptr, vals_in, vals_out, n, p, pl.shift[p], pl.mask[p], hist, tickets, lookback);
#endif
}

template void launch_sort_histogram<uint32_t>(const uint32_t*, size_t, const SortPlan&, void*, cudaStream_t);
template void launch_sort_histogram<uint64_t>(const uint64_t*, size_t, const SortPlan&, void*, cudaStream_t);
template void launch_sort_pass<uint32_t>(const uint32_t*, uint32_t*, const uint32_t*, uint32_t*, size_t, const SortPlan&, int, const uint32_t*, void*, cudaStream_t);
template void launch_sort_pass<uint64_t>(const uint64_t*, uint64_t*, const uint32_t*, uint32_t*, size_t, const SortPlan&, int, const uint32_t*, void*, cudaStream_t);

}  // namespace msgs
