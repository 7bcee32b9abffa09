// Binning stage (see common.cuh for the scheme): prefix sum of tiles_touched in depth order, per-tile counts /
// ranges / digit offsets from the corner grid, emission of (tile, Gaussian) instances in depth order.
// Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:301), duplicateWithKeys (:70-111),
// cudaMemset + identifyTileRanges (:116-138, :334-342).  All HBM-bound integer work.
#include "common.cuh"

namespace msgs {

// ------------------------------------------------------------------------------------------------
// state reset: look-back descriptors + ticket, corner grid, counters.  One small launch per forward.
// ------------------------------------------------------------------------------------------------
__global__ void init_state_kernel(uint32_t* scan_state, size_t scan_words, int* corners, size_t corner_words,
                                  uint32_t* meta) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < scan_words) scan_state[i] = 0u;
    if (i < corner_words) corners[i] = 0;
    if (i < 16) meta[i] = 0u;
}

void launch_init_state(GeomState g, ImageState img, int P, int gx, int gy, cudaStream_t st) {
    const size_t words = scan_state_words((size_t)P);
    const size_t cw = (size_t)(gx + 1) * (gy + 1);
    const size_t n = words > cw ? words : cw;
    init_state_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g.scan_state, words, img.corners, cw, g.meta);
}

// ------------------------------------------------------------------------------------------------
// Single-pass inclusive scan with decoupled look-back (one read + one write of 4 B per Gaussian) of
// tiles_touched gathered in depth order.  Tile descriptor: u64 = (status << 32) | value; status
// 1 = tile aggregate, 2 = inclusive prefix.  Tiles are claimed through an atomic ticket so that a tile's
// predecessors are always resident or done.
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = (int)(SCAN_TILE / SCAN_THREADS);   // 8

__global__ void __launch_bounds__(SCAN_THREADS)
scan_depth_order_kernel(const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ order,
                        uint32_t* __restrict__ out, int P, unsigned long long* state, uint32_t* ticket,
                        uint32_t* meta) {
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
        const uint4 a = *reinterpret_cast<const uint4*>(order + base);
        const uint4 b = *reinterpret_cast<const uint4*>(order + base + 4);
        const uint32_t id[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = __ldg(tiles_touched + id[k]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < (size_t)P) ? __ldg(tiles_touched + order[base + k]) : 0u;
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

void launch_scan_depth_order(GeomState g, int P, cudaStream_t st) {
    const int tiles = (int)((P + SCAN_TILE - 1) / SCAN_TILE);
    unsigned long long* state = reinterpret_cast<unsigned long long*>(g.scan_state);
    uint32_t* ticket = g.scan_state + 2 * (size_t)(tiles + 1);
    scan_depth_order_kernel<<<tiles, SCAN_THREADS, 0, st>>>(g.tiles_touched, g.dvals[0], g.point_offsets, P, state,
                                                            ticket, g.meta);
}

// ------------------------------------------------------------------------------------------------
// Per-tile instance counts WITHOUT touching the instances: every rendered Gaussian scattered +1/-1/-1/+1
// at the four corners of its tile rectangle (preprocess); the 2-D inclusive prefix sum of that grid is the
// number of Gaussians covering each tile.  From the counts: the tile ranges (identifyTileRanges,
// rasterizer_impl.cu:116-138; empty tiles stay {0,0} as after the reference's cudaMemset, :334) and the
// exclusive digit offsets of every pass of the tile partition (what a histogram pass over R keys would give).
// One CTA; T <= a few 10^4.
// ------------------------------------------------------------------------------------------------
constexpr int TS_THREADS = 1024;

__global__ void __launch_bounds__(TS_THREADS)
tile_scan_kernel(int* corners, int gx, int gy, uint2* __restrict__ ranges, SortPlan plan, uint32_t* __restrict__ part_hist) {
    __shared__ uint32_t s_chunk[TS_THREADS];
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * 256];
    const int tid = threadIdx.x;
    const int cw = gx + 1;
    for (int i = tid; i < plan.passes * 256; i += TS_THREADS) s_hist[i] = 0;
    // 2-D inclusive prefix, in place: along x for every row, then along y for every column
    for (int y = tid; y <= gy; y += TS_THREADS) {
        int run = 0;
        for (int x = 0; x <= gx; ++x) { run += corners[y * cw + x]; corners[y * cw + x] = run; }
    }
    __syncthreads();
    for (int x = tid; x <= gx; x += TS_THREADS) {
        int run = 0;
        for (int y = 0; y <= gy; ++y) { run += corners[y * cw + x]; corners[y * cw + x] = run; }
    }
    __syncthreads();
    // exclusive scan of the counts in tile-id order (row-major over the gx x gy tiles)
    const int T = gx * gy;
    const int per = (T + TS_THREADS - 1) / TS_THREADS;
    const int t0 = tid * per, t1 = min(T, t0 + per);
    uint32_t sum = 0;
    for (int t = t0; t < t1; ++t) sum += (uint32_t)corners[(t / gx) * cw + (t % gx)];
    s_chunk[tid] = sum;
    __syncthreads();
    for (int o = 1; o < TS_THREADS; o <<= 1) {
        const uint32_t a = (tid >= o) ? s_chunk[tid - o] : 0u;
        __syncthreads();
        s_chunk[tid] += a;
        __syncthreads();
    }
    uint32_t start = s_chunk[tid] - sum;
    for (int t = t0; t < t1; ++t) {
        const uint32_t c = (uint32_t)corners[(t / gx) * cw + (t % gx)];
        ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        start += c;
        if (c)
            for (int p = 0; p < plan.passes; ++p)
                atomicAdd(&s_hist[p * 256 + (((uint32_t)t >> plan.shift[p]) & plan.mask[p])], c);
    }
    __syncthreads();
    // exclusive scan of each pass's 256 digit counts (warp p handles pass p)
    const int warp = tid >> 5, lane = tid & 31;
    if (warp < plan.passes) {
        uint32_t carry = 0;
        for (int b = 0; b < 256; b += 32) {
            const uint32_t v = s_hist[warp * 256 + b + lane];
            uint32_t inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += n;
            }
            part_hist[warp * 256 + b + lane] = carry + inc - v;
            carry += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
    }
}

void launch_tile_scan(ImageState img, int gx, int gy, SortPlan plan, cudaStream_t st) {
    tile_scan_kernel<<<1, TS_THREADS, 0, st>>>(img.corners, gx, gy, img.ranges, plan, img.part_hist);
}

// ------------------------------------------------------------------------------------------------
// Instance emission in DEPTH order: thread k handles the k-th Gaussian of the depth-sorted list and writes
// (tile id, Gaussian index) for every tile of its rectangle, row-major as duplicateWithKeys does
// (rasterizer_impl.cu:98-108).  Depth order + a stable partition by tile id afterwards == the reference's
// stable sort on (tile << 32 | depth bits) of index-ordered instances.
// Warp-cooperative: a Gaussian covering many tiles is emitted by all 32 lanes (coalesced stores) instead of
// by one thread in a serial loop.
// ------------------------------------------------------------------------------------------------
constexpr int EMIT_COOP_MIN = 12;   // rectangles of at least this many tiles are emitted by the whole warp

__global__ void __launch_bounds__(256)
emit_kernel(int P, const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
            const ushort4* __restrict__ rect, unsigned gx, uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t off = 0, count = 0, id = 0, w = 0, x0 = 0, y0 = 0;
    if (k < P) {
        const uint32_t end = offsets[k];
        off = (k == 0) ? 0u : offsets[k - 1];
        count = end - off;
        if (count > 0) {
            id = order[k];
            const ushort4 r = rect[id];
            x0 = r.x; y0 = r.y; w = (uint32_t)r.z - r.x;
        }
    }
    if (count > 0 && count < EMIT_COOP_MIN) {
        uint32_t xx = 0, tile = y0 * gx + x0;
        for (uint32_t j = 0; j < count; ++j) {
            tkeys[off + j] = tile + xx;
            vals[off + j] = id;
            if (++xx == w) { xx = 0; tile += gx; }
        }
    }
    unsigned todo = __ballot_sync(0xFFFFFFFFu, count >= EMIT_COOP_MIN);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t c = __shfl_sync(0xFFFFFFFFu, count, src);
        const uint32_t o = __shfl_sync(0xFFFFFFFFu, off, src);
        const uint32_t ww = __shfl_sync(0xFFFFFFFFu, w, src);
        const uint32_t xs = __shfl_sync(0xFFFFFFFFu, x0, src);
        const uint32_t ys = __shfl_sync(0xFFFFFFFFu, y0, src);
        const uint32_t gid = __shfl_sync(0xFFFFFFFFu, id, src);
        for (uint32_t j = lane; j < c; j += 32) {
            const uint32_t yy = j / ww, xx = j - yy * ww;
            tkeys[o + j] = (ys + yy) * gx + (xs + xx);
            vals[o + j] = gid;
        }
    }
}

void launch_emit(GeomState g, int P, int gx, uint32_t* tkeys, uint32_t* vals, cudaStream_t st) {
    emit_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, g.dvals[0], g.point_offsets, g.rect, (unsigned)gx, tkeys, vals);
}

}  // namespace msgs
