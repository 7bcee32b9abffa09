// Binning stage: prefix sum of tiles_touched, (tile|depth) key emission, tile ranges.
// Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:301), duplicateWithKeys (:70-111),
// cudaMemset + identifyTileRanges (:116-138, :334-342).  All HBM-bound integer work.
#include "common.cuh"
#include "math.cuh"

namespace msgs {

// ------------------------------------------------------------------------------------------------
// state reset: look-back descriptors, ticket, depth-bit range.  One tiny launch per forward.
// ------------------------------------------------------------------------------------------------
__global__ void init_state_kernel(uint32_t* scan_state, size_t scan_words, uint32_t* meta) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < scan_words) scan_state[i] = 0u;
    if (i == 0) {
        meta[0] = 0u;            // num_rendered
        meta[1] = 0xFFFFFFFFu;   // min depth bits
        meta[2] = 0u;            // max depth bits
        meta[3] = 0u;
    }
}

// ------------------------------------------------------------------------------------------------
// Single-pass inclusive scan with decoupled look-back (one read + one write of 4 B per Gaussian).
// Tile descriptor: u64 = (status << 32) | value; status 1 = tile aggregate, 2 = inclusive prefix.
// Tiles are claimed through an atomic ticket so that a tile's predecessors are always resident or done.
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = (int)(SCAN_TILE / SCAN_THREADS);   // 8

__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int P,
            unsigned long long* state, uint32_t* ticket, uint32_t* meta) {
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const size_t base = (size_t)tile * SCAN_TILE + (size_t)tid * SCAN_ITEMS;

    uint32_t v[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= (size_t)P) {
        const uint4 a = *reinterpret_cast<const uint4*>(in + base);
        const uint4 b = *reinterpret_cast<const uint4*>(in + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < (size_t)P) ? in[base + k] : 0u;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { tsum += v[k]; v[k] = tsum; }   // thread-inclusive
    uint32_t wsum = tsum;                                                // warp-inclusive over thread sums
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, wsum, o);
        if (lane >= o) wsum += n;
    }
    if (lane == 31) s_warp[warp] = wsum;
    __syncthreads();
    uint32_t warp_excl = 0, block_sum = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const uint32_t s = s_warp[w];
        if (w < warp) warp_excl += s;
        block_sum += s;
    }
    if (tid == 0) {
        uint32_t prefix = 0;
        if (tile == 0) {
            atomicExch(&state[0], (2ull << 32) | block_sum);
        } else {
            atomicExch(&state[tile], (1ull << 32) | block_sum);
            int j = (int)tile - 1;
            while (true) {
                const unsigned long long d = *((volatile unsigned long long*)&state[j]);
                const uint32_t st = (uint32_t)(d >> 32);
                if (st == 0u) continue;
                prefix += (uint32_t)d;
                if (st == 2u) break;
                --j;
            }
            atomicExch(&state[tile], (2ull << 32) | (prefix + block_sum));
        }
        s_prefix = prefix;
        if ((size_t)(tile + 1) * SCAN_TILE >= (size_t)P) meta[0] = prefix + block_sum;   // num_rendered
    }
    __syncthreads();
    const uint32_t excl = s_prefix + warp_excl + (wsum - tsum);
    if (base + SCAN_ITEMS <= (size_t)P) {
        *reinterpret_cast<uint4*>(out + base) = make_uint4(excl + v[0], excl + v[1], excl + v[2], excl + v[3]);
        *reinterpret_cast<uint4*>(out + base + 4) = make_uint4(excl + v[4], excl + v[5], excl + v[6], excl + v[7]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k)
            if (base + k < (size_t)P) out[base + k] = excl + v[k];
    }
}

void launch_scan(GeomState g, int P, cudaStream_t st) {
    const int tiles = (int)((P + SCAN_TILE - 1) / SCAN_TILE);
    unsigned long long* state = reinterpret_cast<unsigned long long*>(g.scan_state);
    uint32_t* ticket = g.scan_state + 2 * (size_t)(tiles + 1);
    scan_kernel<<<tiles, SCAN_THREADS, 0, st>>>(g.tiles_touched, g.point_offsets, P, state, ticket, g.meta);
}

void launch_init_state(GeomState g, int P, cudaStream_t st) {
    const size_t words = 2 * ((P + SCAN_TILE - 1) / SCAN_TILE + 1) + 32;
    init_state_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(g.scan_state, words, g.meta);
}

// ------------------------------------------------------------------------------------------------
// Key emission.  key = (tile << depth_bits) | (float_bits(depth) - min_depth_bits): an order-preserving
// compaction of the reference key (tile << 32 | float_bits(depth), rasterizer_impl.cu:98-108); emission
// order is Gaussian-index ascending, row-major over the tile rectangle, as in the reference, so that the
// stable sort breaks (tile, depth) ties identically.
// Warp-cooperative: a Gaussian covering many tiles is emitted by all 32 lanes (coalesced 8 B / 4 B
// stores) instead of by one thread in a serial loop.
// ------------------------------------------------------------------------------------------------
constexpr int DUP_COOP_MIN = 16;   // rectangles of at least this many tiles are emitted by the whole warp

__global__ void __launch_bounds__(256)
duplicate_kernel(int P, const float4* __restrict__ rec, const uint32_t* __restrict__ offsets,
                 const int* __restrict__ radii, unsigned gx, unsigned gy, uint32_t min_depth_bits, int depth_bits,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t off = 0, count = 0, w = 0, dkey = 0;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    if (idx < P) {
        const int r = radii[idx];
        if (r > 0) {
            off = (idx == 0) ? 0u : offsets[idx - 1];
            const float4 a = rec[3 * (size_t)idx];
            get_rect(make_float2(a.x, a.y), r, rmin, rmax, gx, gy);
            w = rmax.x - rmin.x;
            count = w * (rmax.y - rmin.y);
            dkey = __float_as_uint(a.z) - min_depth_bits;
        }
    }
    // small rectangles: one thread, serial (most Gaussians touch 1-9 tiles)
    if (count > 0 && count < DUP_COOP_MIN) {
        uint32_t o = off;
        for (unsigned y = rmin.y; y < rmax.y; ++y)
            for (unsigned x = rmin.x; x < rmax.x; ++x) {
                const uint64_t tile = y * gx + x;
                keys[o] = (tile << depth_bits) | dkey;
                vals[o] = (uint32_t)idx;
                ++o;
            }
    }
    // large rectangles: the warp takes them one at a time
    unsigned todo = __ballot_sync(0xFFFFFFFFu, count >= DUP_COOP_MIN);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t c = __shfl_sync(0xFFFFFFFFu, count, src);
        const uint32_t o = __shfl_sync(0xFFFFFFFFu, off, src);
        const uint32_t ww = __shfl_sync(0xFFFFFFFFu, w, src);
        const uint32_t x0 = __shfl_sync(0xFFFFFFFFu, rmin.x, src);
        const uint32_t y0 = __shfl_sync(0xFFFFFFFFu, rmin.y, src);
        const uint32_t dk = __shfl_sync(0xFFFFFFFFu, dkey, src);
        const uint32_t id = (uint32_t)(idx - lane + src);
        for (uint32_t j = lane; j < c; j += 32) {
            const uint32_t yy = j / ww, xx = j - yy * ww;
            const uint64_t tile = (y0 + yy) * gx + (x0 + xx);
            keys[o + j] = (tile << depth_bits) | dk;
            vals[o + j] = id;
        }
    }
}

void launch_duplicate(GeomState g, int P, int gx, int gy, const int* radii, uint32_t min_depth_bits, int depth_bits,
                      uint64_t* keys, uint32_t* vals, cudaStream_t st) {
    duplicate_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, g.rec, g.point_offsets, radii, (unsigned)gx, (unsigned)gy,
                                                      min_depth_bits, depth_bits, keys, vals);
}

// ------------------------------------------------------------------------------------------------
// Tile ranges from the sorted keys (identifyTileRanges, rasterizer_impl.cu:116-138) with the zero-fill of
// empty tiles (cudaMemset, :334) folded in: ranges are zeroed by the same launch sequence's memset node.
// ------------------------------------------------------------------------------------------------
__global__ void tile_ranges_kernel(const uint64_t* __restrict__ keys, size_t L, int depth_bits, uint2* ranges) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t cur = (uint32_t)(keys[idx] >> depth_bits);
    if (idx == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = (uint32_t)(keys[idx - 1] >> depth_bits);
        if (cur != prev) {
            ranges[prev].y = (uint32_t)idx;
            ranges[cur].x = (uint32_t)idx;
        }
    }
    if (idx == L - 1) ranges[cur].y = (uint32_t)L;
}

void launch_tile_ranges(const uint64_t* keys_sorted, size_t R, int depth_bits, uint2* ranges, int tiles, cudaStream_t st) {
    cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), st);
    if (R > 0)
        tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(keys_sorted, R, depth_bits, ranges);
}

}  // namespace msgs
