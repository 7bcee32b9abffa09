// Binning stage: from depth-ordered Gaussians + their tile rectangles to the per-tile, depth-ordered lists
// (point_list) and their ranges.  Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:301),
// duplicateWithKeys (:70-111), cub::DeviceRadixSort::SortPairs on R 64-bit keys (:327-332), cudaMemset +
// identifyTileRanges (:116-138, :334-342).  The result is bit-identical to the reference's: a list ordered by
// (tile, depth bits, Gaussian index).
//
// Counting instead of sorting (see common.cuh).  With Gaussians already in depth order (dvals[0]):
//   chunk_count : chunk c (chunkG consecutive Gaussians) scatters +1/-1/-1/+1 at the 4 corners of every tile
//                 rectangle into a shared-memory grid; its 2-D prefix sum is, per tile, the number of the chunk's
//                 Gaussians covering it -> counts[c][tile].                         (P x 8 B read, NC x T x 4 B written)
//   chunk_scan  : exclusive scan of counts[.][tile] down the chunks, tile totals.  (NC x T x 8 B)
//   tile_scan   : exclusive scan of the totals -> tile_start, ranges, num_rendered.
//   place       : chunk c loads slot[tile] = tile_start + counts[c][tile] into shared memory and walks its
//                 Gaussians in depth order; every covered tile gets the index at slot[tile]++.  Warps own disjoint
//                 tile blocks, so the walk needs no atomics.                              (R x 4 B written)
// All integer work; HBM-bound only through the final 4-byte stores.
#include "common.cuh"

namespace msgs {

constexpr int BIN_THREADS = 256;
constexpr int BIN_WARPS = BIN_THREADS / 32;

__device__ __forceinline__ int odd_width(int gx) { return (gx + 1) | 1; }

// ------------------------------------------------------------------------------------------------
// chunk_count: grid (NC, nbands)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BIN_THREADS)
chunk_count_kernel(int P, const uint32_t* __restrict__ order, const ushort4* __restrict__ rect,
                   ushort4* __restrict__ rect_sorted, int gx, int gy, BinPlan bp, const uint32_t* __restrict__ meta,
                   uint32_t* __restrict__ counts) {
    extern __shared__ int s_grid[];                 // (rows + 1) x cw corner cells
    const int tid = threadIdx.x;
    const int c = blockIdx.x, band = blockIdx.y;
    const int y0 = band * bp.band_rows, y1 = min(gy, y0 + bp.band_rows), rows = y1 - y0;
    const int cw = odd_width(gx);
    const int cells = (rows + 1) * cw;
    for (int i = tid; i < cells; i += BIN_THREADS) s_grid[i] = 0;
    __syncthreads();
    const int V = (int)meta[1], chunkG = bin_chunk_size((uint32_t)V, bp.NC);
    const int k0 = min(V, c * chunkG), k1 = min(V, k0 + chunkG);
    for (int k = k0 + tid; k < k1; k += BIN_THREADS) {
        const ushort4 r = rect[order[k]];
        if (band == 0) rect_sorted[k] = r;          // depth-ordered copy for the placement pass (coalesced there)
        const int ya = max((int)r.y, y0) - y0, yb = min((int)r.w, y1) - y0;
        if (r.x < r.z && ya < yb) {
            atomicAdd(&s_grid[ya * cw + r.x], 1);
            atomicAdd(&s_grid[ya * cw + r.z], -1);
            atomicAdd(&s_grid[yb * cw + r.x], -1);
            atomicAdd(&s_grid[yb * cw + r.z], 1);
        }
    }
    __syncthreads();
    // 2-D inclusive prefix: along x (thread per row; odd row stride -> conflict-free), then along y
    for (int y = tid; y < rows; y += BIN_THREADS) {
        int run = 0;
        int* row = s_grid + y * cw;
        for (int x = 0; x < gx; ++x) { run += row[x]; row[x] = run; }
    }
    __syncthreads();
    for (int x = tid; x < gx; x += BIN_THREADS) {
        int run = 0;
        for (int y = 0; y < rows; ++y) { run += s_grid[y * cw + x]; s_grid[y * cw + x] = run; }
    }
    __syncthreads();
    uint32_t* out = counts + (size_t)c * gx * gy + (size_t)y0 * gx;
    const int n = rows * gx;
    for (int i = tid; i < n; i += BIN_THREADS) {
        const int y = i / gx, x = i - y * gx;
        out[i] = (uint32_t)s_grid[y * cw + x];
    }
}

void launch_chunk_count(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp, cudaStream_t st) {
    const size_t smem = (size_t)(bp.band_rows + 1) * ((gx + 1) | 1) * sizeof(int);
    dim3 grid(bp.NC, bp.nbands, 1);
    chunk_count_kernel<<<grid, BIN_THREADS, smem, st>>>(P, g.dvals[0], g.rect, g.rect_sorted, gx, gy, bp, g.meta, img.counts);
}

// ------------------------------------------------------------------------------------------------
// chunk_scan: exclusive scan of counts[c][t] over c (in place) + tile totals.  CTA = 32 tiles x 32 chunk groups;
// a warp row reads 32 consecutive tiles (128 B).  NC <= 1024 -> at most 32 chunks per group, held in registers.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
chunk_scan_kernel(uint32_t* __restrict__ counts, int tiles, int NC, uint32_t* __restrict__ tile_total) {
    __shared__ uint32_t s_sum[32][33];
    const int tx = threadIdx.x, g = threadIdx.y;
    const int t = blockIdx.x * 32 + tx;
    const int per = (NC + 31) / 32;
    const int c0 = g * per, c1 = min(NC, c0 + per);
    uint32_t v[32];
    uint32_t sum = 0;
    if (t < tiles) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int c = c0 + i;
            v[i] = (c < c1) ? counts[(size_t)c * tiles + t] : 0u;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sum += v[i];
    }
    s_sum[g][tx] = sum;
    __syncthreads();
    uint32_t base = 0, total = 0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
        const uint32_t s = s_sum[j][tx];
        if (j < g) base += s;
        total += s;
    }
    if (t < tiles) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int c = c0 + i;
            if (c < c1) counts[(size_t)c * tiles + t] = base;
            base += v[i];
        }
        if (g == 0) tile_total[t] = total;
    }
}

void launch_chunk_scan(const ImageState& img, int tiles, const BinPlan& bp, cudaStream_t st) {
    dim3 block(32, 32, 1);
    chunk_scan_kernel<<<(tiles + 31) / 32, block, 0, st>>>(img.counts, tiles, bp.NC, img.tile_total);
}

// ------------------------------------------------------------------------------------------------
// tile_scan: exclusive scan of the tile totals -> tile_start; ranges (identifyTileRanges,
// rasterizer_impl.cu:116-138; empty tiles stay {0,0} as after the reference's cudaMemset, :334);
// num_rendered -> meta[0].  One CTA; T <= a few 10^4.
// ------------------------------------------------------------------------------------------------
constexpr int TS_THREADS = 1024;

__global__ void __launch_bounds__(TS_THREADS)
tile_scan_kernel(const uint32_t* __restrict__ tile_total, int T, uint32_t* __restrict__ tile_start,
                 uint2* __restrict__ ranges, uint32_t* __restrict__ meta) {
    __shared__ uint32_t s_warp[TS_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (T + TS_THREADS - 1) / TS_THREADS;
    const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
    uint32_t sum = 0;
    for (int t = t0; t < t1; ++t) sum += tile_total[t];
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_warp[lane];
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (lane >= o) wi += n;
        }
        s_warp[lane] = wi - w;                       // exclusive over warps
        if (lane == 31) meta[0] = wi;                // num_rendered
    }
    __syncthreads();
    uint32_t start = s_warp[warp] + incl - sum;
    for (int t = t0; t < t1; ++t) {
        const uint32_t c = tile_total[t];
        tile_start[t] = start;
        ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        start += c;
    }
}

void launch_tile_scan(const ImageState& img, int tiles, uint32_t* meta, cudaStream_t st) {
    tile_scan_kernel<<<1, TS_THREADS, 0, st>>>(img.tile_total, tiles, img.tile_start, img.ranges, meta);
}

// ------------------------------------------------------------------------------------------------
// place: grid (NC chunks, super-blocks).  A CTA owns a super-block of 4bw x 2bh tiles, its 8 warps own one
// bw x bh block each (4 across, 2 down; BinPlan picks bw, bh).  The chunk is read 256 Gaussians at a time (thread =
// Gaussian, coalesced from the depth-ordered copies); the ones whose rectangle meets the super-block are compacted,
// IN ORDER, into a shared-memory queue (rectangle clipped to the super-block + index).  Every warp then walks the
// queue in order; for an entry that meets its block, lane = tile of the clipped rectangle:
// p = slot[tile]++; point_list[p] = index.  slot[] starts at tile_start + counts[chunk][tile].  Tiles of one
// rectangle are distinct and a warp's shared-memory accesses are ordered by __syncwarp, so no atomics are needed
// and the result is deterministic: depth order inside every tile.
// ------------------------------------------------------------------------------------------------
constexpr int PLACE_SLOTS = PLACE_BW_MAX * PLACE_BH_MAX;
// 4096 / w + 1 for w = 0..10: (j * kRecip12[w]) >> 12 == j / w for every j < 128
__constant__ int kRecip12[PLACE_BW_MAX + 1] = {0, 4097, 2049, 1366, 1025, 820, 683, 586, 513, 456, 410};

__global__ void __launch_bounds__(BIN_THREADS)
place_kernel(int P, const uint32_t* __restrict__ order, const ushort4* __restrict__ rect_sorted, int gx, int gy,
             BinPlan bp, int sb_x, const uint32_t* __restrict__ meta, const uint32_t* __restrict__ counts,
             const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ point_list, uint32_t capacity) {
    if (meta[0] > capacity) return;                 // speculatively sized list too small: the host relaunches (api.cu)
    __shared__ uint32_t s_slot[BIN_WARPS][PLACE_SLOTS];
    __shared__ uint32_t s_qx[2][BIN_THREADS], s_qy[2][BIN_THREADS], s_qid[2][BIN_THREADS];
    __shared__ uint32_t s_wcnt[2][BIN_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x;
    const int BW = bp.bw, BH = bp.bh;
    const int sbx0 = (blockIdx.y % sb_x) * 4 * BW, sby0 = (blockIdx.y / sb_x) * 2 * BH;
    const int sbx1 = min(gx, sbx0 + 4 * BW), sby1 = min(gy, sby0 + 2 * BH);
    // this warp's block, clipped to the grid (may be empty)
    const int bx0 = min(sbx1, sbx0 + (warp & 3) * BW), bx1 = min(sbx1, bx0 + BW);
    const int by0 = min(sby1, sby0 + (warp >> 2) * BH), by1 = min(sby1, by0 + BH);
    const int bwid = bx1 - bx0;
    const bool warp_has_tiles = bwid > 0 && by1 > by0;
    {
        const uint32_t* cnt = counts + (size_t)c * gx * gy;
        for (int i = lane; i < BW * BH; i += 32) {
            const int yy = i / BW, xx = i - yy * BW;
            const int x = bx0 + xx, y = by0 + yy;
            if (x < bx1 && y < by1) s_slot[warp][i] = tile_start[y * gx + x] + cnt[y * gx + x];
        }
    }
    const int V = (int)meta[1], chunkG = bin_chunk_size((uint32_t)V, bp.NC);
    const int k0 = min(V, c * chunkG), k1 = min(V, k0 + chunkG);
    if (k0 >= k1) return;                           // (uniform) chunk beyond the rendered Gaussians
    // register prefetch of the next batch
    ushort4 r_next = make_ushort4(0, 0, 0, 0);
    uint32_t id_next = 0;
    if (k0 + tid < k1) { r_next = rect_sorted[k0 + tid]; id_next = order[k0 + tid]; }
    int buf = 0;
    for (int kb = k0; kb < k1; kb += BIN_THREADS, buf ^= 1) {
        const ushort4 r = r_next;
        const uint32_t id = id_next;
        if (kb + BIN_THREADS + tid < k1) { r_next = rect_sorted[kb + BIN_THREADS + tid]; id_next = order[kb + BIN_THREADS + tid]; }
        else r_next = make_ushort4(0, 0, 0, 0);
        // clip to the super-block; ordered compaction of the survivors
        const int xa = max((int)r.x, sbx0), xb = min((int)r.z, sbx1);
        const int ya = max((int)r.y, sby0), yb = min((int)r.w, sby1);
        const bool hit = xa < xb && ya < yb;
        const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
        if (lane == 0) s_wcnt[buf][warp] = __popc(b);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < BIN_WARPS; ++w) {
            const int n = (int)s_wcnt[buf][w];
            if (w < warp) before += n;
            total += n;
        }
        if (hit) {
            const int pos = before + __popc(b & ((1u << lane) - 1u));
            s_qx[buf][pos] = (uint32_t)xa | ((uint32_t)xb << 16);
            s_qy[buf][pos] = (uint32_t)ya | ((uint32_t)yb << 16);
            s_qid[buf][pos] = id;
        }
        __syncthreads();
        if (!warp_has_tiles) continue;
        for (int q0 = 0; q0 < total; q0 += 32) {
            const int q = q0 + lane;
            // lane-parallel: clip entry q to this warp's block and pack everything the serial part needs into one
            // word: base slot (7 bits) | tile count (7) | reciprocal multiplier 4096/w + 1 (13) | width w (4)
            uint32_t pk = 0, eid = 0;
            bool mine = false;
            if (q < total) {
                const uint32_t ex = s_qx[buf][q], ey = s_qy[buf][q];
                eid = s_qid[buf][q];
                const int cxa = max((int)(ex & 0xFFFF), bx0), cxb = min((int)(ex >> 16), bx1);
                const int cya = max((int)(ey & 0xFFFF), by0), cyb = min((int)(ey >> 16), by1);
                mine = cxa < cxb && cya < cyb;
                const int w = cxb - cxa, h = cyb - cya;
                const int mul = kRecip12[mine ? w : 1];
                pk = (uint32_t)((cya - by0) * BW + (cxa - bx0)) | ((uint32_t)(w * h) << 7) | ((uint32_t)mul << 14) | ((uint32_t)w << 27);
            }
            unsigned todo = __ballot_sync(0xFFFFFFFFu, mine);
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const uint32_t sp = __shfl_sync(0xFFFFFFFFu, pk, src);
                const uint32_t gid = __shfl_sync(0xFFFFFFFFu, eid, src);
                const int base = sp & 127, cnt = (sp >> 7) & 127, mul = (sp >> 14) & 8191, w = sp >> 27;
                for (int j = lane; j < cnt; j += 32) {
                    const int yy = (j * mul) >> 12;                       // j / w, exact for j < 128, w <= 10
                    const int tl = base + yy * BW + (j - yy * w);
                    const uint32_t p = s_slot[warp][tl];
                    s_slot[warp][tl] = p + 1;
                    point_list[p] = gid;
                }
                __syncwarp();      // order this Gaussian's slot updates before the next one's reads (other lanes)
            }
        }
    }
}

void launch_place(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp,
                  uint32_t* point_list, uint32_t capacity, cudaStream_t st) {
    const int sb_x = (gx + 4 * bp.bw - 1) / (4 * bp.bw), sb_y = (gy + 2 * bp.bh - 1) / (2 * bp.bh);
    dim3 grid(bp.NC, sb_x * sb_y, 1);
    place_kernel<<<grid, BIN_THREADS, 0, st>>>(P, g.dvals[0], g.rect_sorted, gx, gy, bp, sb_x, g.meta, img.counts, img.tile_start, point_list, capacity);
}

}  // namespace msgs
