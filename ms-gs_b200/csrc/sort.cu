// Hand-written onesweep LSD radix sort of (key, u32 value) pairs — replaces cub::DeviceRadixSort::SortPairs
// (rasterizer_impl.cu:190-193, 327-332).  Stable, ascending.  Keys are u32 (binning stage) or u64 (public
// msgs_sort_pairs).  HBM-bound: one up-front histogram read of the keys plus, per pass, one read and one write of
// every pair; each pass is ONE kernel whose tiles chain their per-digit prefixes with a decoupled look-back (no
// separate scan/scatter launches, no per-tile histogram round trip through HBM).
//
// Pass kernel, per 256-thread CTA and tile of SORT_TILE pairs:
//   1. claim a tile through an atomic ticket (predecessors are then resident or finished);
//   2. load keys warp-striped (coalesced), rank them inside each warp with match.any (stable: item order, then
//      lane order), per-warp digit counters in shared memory;
//   3. scan counters across warps and digits; publish the tile's digit counts, look back for the digit prefixes of
//      earlier tiles (thread d owns digit d);
//   4. scatter keys and values to their tile-local sorted slot in shared memory, then stream them out so that each
//      digit's run is written with coalesced stores.  The last pass of the tile partition does not write keys.
#include "common.cuh"

namespace msgs {

constexpr int RADIX_BITS = 8;
constexpr int BINS = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;      // 4096 pairs per tile
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
// temp layout: [hist: SORT_MAX_PASSES x BINS u32][tickets: 8 u32][look-back: tiles x BINS u32]
size_t sort_temp_bytes(size_t n) {
    return (size_t)SORT_MAX_PASSES * BINS * 4 + 8 * 4 + (sort_tiles(n) + 1) * BINS * 4 + 256;
}
uint32_t* sort_temp_hist(void* temp) { return reinterpret_cast<uint32_t*>(temp); }
size_t sort_temp_clear_bytes(size_t n) { return (size_t)SORT_MAX_PASSES * BINS * 4 + 8 * 4 + sort_tiles(n) * BINS * 4; }
void sort_temp_clear(void* temp, size_t n, cudaStream_t st) { cudaMemsetAsync(temp, 0, sort_temp_clear_bytes(n), st); }

// ---- up-front histogram of every pass's digit ----------------------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const KeyT* __restrict__ keys, size_t n, SortPlan pl, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * BINS];
    for (int i = threadIdx.x; i < pl.passes * BINS; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const KeyT k = keys[i];
        for (int p = 0; p < pl.passes; ++p) {
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
template <typename KeyT>
struct SortSmem {
    KeyT     keys[SORT_TILE];               // 16 / 32 KB
    uint32_t vals[SORT_TILE];               // 16 KB
    uint32_t warp_hist[SORT_WARPS][BINS];   //  8 KB
    uint32_t bin_start[BINS];               // tile-local exclusive start of each digit
    int      bin_goff[BINS];                // global offset of the digit's run minus bin_start
    uint32_t wsum[SORT_WARPS];
    uint32_t tile;
};

template <typename KeyT, bool KEYS_OUT>
__global__ void __launch_bounds__(SORT_THREADS)
sort_onesweep_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                     size_t n, int pass, int shift, uint32_t dmask, const uint32_t* __restrict__ hist,
                     uint32_t* tickets, volatile uint32_t* lookback) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<KeyT>& sm = *reinterpret_cast<SortSmem<KeyT>*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t FLAG_AGG = (uint32_t)(2 * pass + 1) << 28, FLAG_PRE = (uint32_t)(2 * pass + 2) << 28;

    if (tid == 0) sm.tile = atomicAdd(&tickets[pass], 1u);
    for (int i = tid; i < SORT_WARPS * BINS; i += SORT_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = sm.tile;
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
            val[i] = vals_in[tile_base + j];
        } else {
            key[i] = (KeyT)~(KeyT)0;
            val[i] = 0;
        }
    }

    // 2. in-warp stable ranking
    uint32_t* wh = sm.warp_hist[warp];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const bool valid = (wbase + i * 32) < tile_n;
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        // invalid (padding) lanes get a private pseudo-digit so they match nobody
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, valid ? d : (uint32_t)(BINS + lane));
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (valid && lane == leader) {
            base = wh[d];
            wh[d] = base + __popc(peers);
        }
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        rank[i] = base + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();

    // 3. digit d (= tid): exclusive scan over warps, tile count, then scan over digits
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) {
        const uint32_t c = sm.warp_hist[w][tid];
        sm.warp_hist[w][tid] = count;
        count += c;
    }
    // publish this tile's digit count as early as possible
    if (tile == 0) lookback[tid] = FLAG_PRE | count;
    else           lookback[(size_t)tile * BINS + tid] = FLAG_AGG | count;

    // block exclusive scan of `count` over the 256 digits
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) sm.wsum[warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) if (w < warp) wex += sm.wsum[w];
    const uint32_t start = wex + incl - count;
    sm.bin_start[tid] = start;

    // look back over earlier tiles for digit `tid`
    uint32_t prefix = 0;
    if (tile > 0) {
        int j = (int)tile - 1;
        while (true) {
            const uint32_t v = lookback[(size_t)j * BINS + tid];
            const uint32_t f = v & ~COUNT_MASK;
            if (f == FLAG_PRE) { prefix += v & COUNT_MASK; break; }
            if (f == FLAG_AGG) { prefix += v & COUNT_MASK; --j; }
            // any other flag value: tile j has not published for this pass yet -> spin
        }
        lookback[(size_t)tile * BINS + tid] = FLAG_PRE | (prefix + count);
    }
    sm.bin_goff[tid] = (int)(hist[pass * BINS + tid] + prefix) - (int)start;
    __syncthreads();

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
}

template <typename KeyT>
void launch_sort_histogram(const KeyT* keys_in, size_t n, const SortPlan& pl, void* temp, cudaStream_t st) {
    if (n == 0 || pl.passes == 0) return;
    uint32_t* hist = sort_temp_hist(temp);
    // 2 CTAs per SM: every CTA ends with up to passes x 256 global atomics, so few, long-running CTAs
    const int hblocks = (int)min((size_t)148 * 2, (n + 255) / 256);
    sort_histogram_kernel<KeyT><<<hblocks, 256, 0, st>>>(keys_in, n, pl, hist);
    sort_scan_hist_kernel<<<pl.passes, BINS, 0, st>>>(hist);
}

template <typename KeyT>
void launch_sort_pass(const KeyT* keys_in, KeyT* keys_out, const uint32_t* vals_in, uint32_t* vals_out, size_t n,
                      const SortPlan& pl, int p, const uint32_t* hist, void* temp, cudaStream_t st) {
    if (n == 0) return;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(sort_onesweep_kernel<KeyT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem<KeyT>));
        cudaFuncSetAttribute(sort_onesweep_kernel<KeyT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem<KeyT>));
        attr_set = true;
    }
    uint32_t* tickets = sort_temp_hist(temp) + SORT_MAX_PASSES * BINS;
    uint32_t* lookback = tickets + 8;
    const unsigned grid = (unsigned)sort_tiles(n);
    if (keys_out)
        sort_onesweep_kernel<KeyT, true><<<grid, SORT_THREADS, sizeof(SortSmem<KeyT>), st>>>(
            keys_in, keys_out, vals_in, vals_out, n, p, pl.shift[p], pl.mask[p], hist, tickets, lookback);
    else
        sort_onesweep_kernel<KeyT, false><<<grid, SORT_THREADS, sizeof(SortSmem<KeyT>), st>>>(
            keys_in, nullptr, vals_in, vals_out, n, p, pl.shift[p], pl.mask[p], hist, tickets, lookback);
}

template void launch_sort_histogram<uint32_t>(const uint32_t*, size_t, const SortPlan&, void*, cudaStream_t);
template void launch_sort_histogram<uint64_t>(const uint64_t*, size_t, const SortPlan&, void*, cudaStream_t);
template void launch_sort_pass<uint32_t>(const uint32_t*, uint32_t*, const uint32_t*, uint32_t*, size_t, const SortPlan&, int, const uint32_t*, void*, cudaStream_t);
template void launch_sort_pass<uint64_t>(const uint64_t*, uint64_t*, const uint32_t*, uint32_t*, size_t, const SortPlan&, int, const uint32_t*, void*, cudaStream_t);

}  // namespace msgs
