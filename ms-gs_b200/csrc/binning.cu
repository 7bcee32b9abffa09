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
//   l1_count / l1_scan / l1_emit : per (8 x 8-tile super-block, list chunk) runs of 8-byte entries {tile mask, index},
//                 one per (Gaussian, super-block it meets), in depth order (two-level placement, see below).
//   place2      : one CTA per (chunk group, super-block) walks ITS entries only; every tile's covering entries are found
//                 32 at a time with a ballot and leave as one run of consecutive 4-byte stores into the tile's list,
//                 starting at tile_start + counts[group][tile].  No atomics on global memory.   (R x 4 B written)
// All integer work.
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
                   ushort4* __restrict__ rect_sorted, int gx, int gy,
                   BinPlan bp, const uint32_t* __restrict__ meta, uint32_t* __restrict__ counts) {
    extern __shared__ int s_grid[];                 // (rows + 1) x cw corner cells
    const int tid = threadIdx.x;
    const int c = blockIdx.x, band = blockIdx.y;
    const int y0 = band * bp.band_rows, y1 = min(gy, y0 + bp.band_rows), rows = y1 - y0;
    const int cw = odd_width(gx);
    const int cells = (rows + 1) * cw;
    for (int i = tid; i < cells; i += BIN_THREADS) s_grid[i] = 0;
    __syncthreads();
    const int V = (int)meta[1], chunkG = bin_chunk_size((uint32_t)V, bp);
    int k0 = min(V, c * chunkG), k1 = min(V, k0 + chunkG);
    if (bp.split > 1) {                             // this CTA's part of the chunk (blockIdx.z of bp.split)
        const int part = (k1 - k0 + bp.split - 1) / bp.split;
        k0 = min(k1, k0 + (int)blockIdx.z * part);
        k1 = min(k1, k0 + part);
    }
    for (int k = k0 + tid; k < k1; k += BIN_THREADS) {
        const ushort4 r = rect[order[k]];
        if (band == 0) rect_sorted[k] = r;          // depth-ordered copy for the list kernels (coalesced there)
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
        const uint32_t v = (uint32_t)s_grid[y * cw + x];
        if (bp.split > 1) { if (v) atomicAdd(&out[i], v); }     // counts zeroed by the launcher
        else out[i] = v;
    }
}

void launch_chunk_count(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp,
                        cudaStream_t st) {
    const size_t smem = (size_t)(bp.band_rows + 1) * ((gx + 1) | 1) * sizeof(int);
    dim3 grid(bp.NC, bp.nbands, bp.split);
    if (bp.split > 1) cudaMemsetAsync(img.counts, 0, (size_t)bp.NC * gx * gy * sizeof(uint32_t), st);
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
                 uint2* __restrict__ ranges, uint32_t* __restrict__ meta, uint32_t* __restrict__ host_out) {
    // warp w owns the contiguous tiles [w * per, (w + 1) * per), per a multiple of 32: every load / store instruction of
    // a warp covers 32 consecutive tiles (coalesced), rows are scanned with shuffles and chained through a carry
    __shared__ uint32_t s_warp[TS_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = TS_THREADS / 32;
    const int per = ((T + NW - 1) / NW + 31) / 32 * 32;
    const int t0 = min(T, warp * per), t1 = min(T, t0 + per);
    uint32_t sum = 0;
    for (int t = t0 + lane; t < t1; t += 32) sum += tile_total[t];
    sum = __reduce_add_sync(0xFFFFFFFFu, sum);
    if (lane == 0) s_warp[warp] = sum;
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
        if (lane == 31) {
            meta[0] = wi;                            // num_rendered
            // ... and straight into the caller's pinned host word (mapped memory): the read-back needs no copy engine,
            // so it cannot queue behind a large device-to-host copy of another stream (a viewer's previous frame)
            if (host_out) { *reinterpret_cast<volatile uint32_t*>(host_out) = wi; __threadfence_system(); }
        }
    }
    __syncthreads();
    uint32_t carry = s_warp[warp];
    for (int tb = t0; tb < t1; tb += 32) {
        const int t = tb + lane;
        const uint32_t c = t < t1 ? tile_total[t] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += n;
        }
        const uint32_t start = carry + incl - c;
        if (t < t1) {
            tile_start[t] = start;
            ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        }
        carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
}

void launch_tile_scan(const ImageState& img, int tiles, uint32_t* meta, uint32_t* host_out, cudaStream_t st) {
    tile_scan_kernel<<<1, TS_THREADS, 0, st>>>(img.tile_total, tiles, img.tile_start, img.ranges, meta, host_out);
}

// ------------------------------------------------------------------------------------------------
// Two-level placement (common.cuh BinPlan).  Level 1: per (super-block, list chunk) entry lists.
//
// A level-1 CTA owns list chunk c = sub = 256 x ITEMS consecutive depth-ordered Gaussians; warp w, item i, lane l holds
// Gaussian c * sub + (w * ITEMS + i) * 32 + l, so (warp, item, lane) order is depth order.
//   l1_count : every Gaussian adds 1 to a shared-memory counter of each super-block its rectangle meets (usually 1-4)
//              -> l1off[s][c], and the per-segment sums the scan needs.
//   l1_scan  : exclusive scan of l1off in (super-block, chunk) order, in place.
//   l1_emit  : rank of a Gaussian's entry inside its (super-block, chunk) run = entries of the warps in front (counted
//              with shared-memory atomics, prefixed over the warps) + entries of the warp's earlier rows of 32 (a running
//              per-warp counter, bumped after each row) + the lower lanes of its own row that meet the same
//              super-block.  The last term comes from per-row ballot words — one per super-block COLUMN (which of the 32
//              rectangles reach it) and one per ROW, built with shared-memory atomicOr: lanes meeting (sx, sy) are
//              col[sx] & row[sy], so sbx + sby words describe all sbx * sby super-blocks.
// ------------------------------------------------------------------------------------------------
constexpr int L1_THREADS = 256;
constexpr int L1_WARPS = L1_THREADS / 32;

struct L1Row { int sx0, sx1, sy0, sy1; };       // super-block range of a rectangle; sx0 > sx1: meets nothing

__device__ __forceinline__ L1Row l1_range(const ushort4 r, int shift) {
    L1Row q;
    const bool ok = r.x < r.z && r.y < r.w;
    q.sx0 = ok ? ((int)r.x >> shift) : 1;
    q.sx1 = ok ? (((int)r.z - 1) >> shift) : 0;
    q.sy0 = ok ? ((int)r.y >> shift) : 1;
    q.sy1 = ok ? (((int)r.w - 1) >> shift) : 0;
    return q;
}

// l1_count: grid nl1.  l1off[s * nl1 + c] = entries of list chunk c in super-block s (zeros for chunks beyond V);
// segsum[(s * nl1 + c) / L1_SEG] += the same (zeroed by the launcher).
template <int ITEMS>
__global__ void __launch_bounds__(L1_THREADS)
l1_count_kernel(const ushort4* __restrict__ rect_sorted, BinPlan bp, const uint32_t* __restrict__ meta,
                uint32_t* __restrict__ l1off, uint32_t* __restrict__ segsum) {
    extern __shared__ uint32_t l1_smem[];
    uint32_t* cnt = l1_smem;                         // S
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int V = (int)meta[1], k0 = c * bp.sub;
    if (k0 >= V) {                                   // (uniform) no rendered Gaussian in this chunk
        for (int s = tid; s < bp.S; s += L1_THREADS) l1off[(size_t)s * bp.nl1 + c] = 0u;
        return;
    }
    for (int s = tid; s < bp.S; s += L1_THREADS) cnt[s] = 0u;
    ushort4 r[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int k = k0 + (warp * ITEMS + i) * 32 + lane;
        r[i] = make_ushort4(0, 0, 0, 0);
        if (k < V) r[i] = rect_sorted[k];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const L1Row q = l1_range(r[i], bp.sb_shift);
        for (int sy = q.sy0; sy <= q.sy1; ++sy)
            for (int sx = q.sx0; sx <= q.sx1; ++sx) atomicAdd(&cnt[sy * bp.sbx + sx], 1u);
    }
    __syncthreads();
    for (int s = tid; s < bp.S; s += L1_THREADS) {
        const uint32_t n = cnt[s];
        const size_t j = (size_t)s * bp.nl1 + c;
        l1off[j] = n;
        if (n) atomicAdd(&segsum[j / L1_SEG], n);
    }
}

// l1_scan: exclusive scan of l1off[0 .. n) in place, total -> l1off[n] and meta[2]: every segment adds the sums of the
// segments in front of it (l1_count accumulated them) to its local scan.  No spinning.
__global__ void __launch_bounds__(256)
l1_scan_kernel(uint32_t* __restrict__ data, int n, const uint32_t* __restrict__ segsum, uint32_t* __restrict__ meta) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg = blockIdx.x, base = seg * L1_SEG;
    // thread t scans the 16 consecutive elements [base + 16 t, + 16)
    constexpr int PER = L1_SEG / 256;
    uint32_t v[PER];
#pragma unroll
    for (int i = 0; i < PER; i += 4) {
        const int j = base + tid * PER + i;
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        if (j + 3 < n) x = *reinterpret_cast<const uint4*>(data + j);      // data is 128-byte aligned, j a multiple of 4
        else {
            if (j < n) x.x = data[j];
            if (j + 1 < n) x.y = data[j + 1];
            if (j + 2 < n) x.z = data[j + 2];
        }
        v[i] = x.x; v[i + 1] = x.y; v[i + 2] = x.z; v[i + 3] = x.w;
    }
    // sum of the segments in front of this one
    uint32_t pre = 0;
    for (int q = tid; q < seg; q += 256) pre += segsum[q];
    pre = __reduce_add_sync(0xFFFFFFFFu, pre);
    if (lane == 0) s_w[warp] = pre;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; ++w) t += s_w[w];
        s_base = t;
    }
    __syncthreads();
    const uint32_t seg_base = s_base;
    __syncthreads();
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) sum += v[i];
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t x = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += x;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) if (w < warp) wex += s_w[w];
    uint32_t run = seg_base + wex + incl - sum;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int j = base + tid * PER + i;
        if (j < n) data[j] = run;
        run += v[i];
    }
    // the thread holding the last element writes the total behind the array
    if (base + tid * PER < n && base + (tid + 1) * PER >= n) { data[n] = run; meta[2] = run; }
}

// l1_emit: grid nl1.  Entry of Gaussian k in super-block s: {tile columns of s it covers | tile rows << 16, index}.
template <int ITEMS>
__global__ void __launch_bounds__(L1_THREADS)
l1_emit_kernel(const ushort4* __restrict__ rect_sorted, const uint32_t* __restrict__ order, BinPlan bp,
               const uint32_t* __restrict__ meta, const uint32_t* __restrict__ l1off, uint2* __restrict__ l1,
               uint32_t capacity) {
    if (meta[0] > capacity) return;                  // speculatively sized buffer too small: the host relaunches
    extern __shared__ uint32_t l1_smem[];
    constexpr int WR = L1_WARPS * ITEMS;
    const int nb = bp.sbx + bp.sby;                  // ballot words per row of 32: columns, then rows
    uint32_t* ballots = l1_smem;                     // WR x nb
    uint32_t* wb = l1_smem + WR * nb;                // L1_WARPS x S: next entry slot of (super-block, warp)
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int V = (int)meta[1], k0 = c * bp.sub;
    if (k0 >= V) return;
    for (int i = tid; i < WR * nb + L1_WARPS * bp.S; i += L1_THREADS) l1_smem[i] = 0u;
    ushort4 r[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int k = k0 + (warp * ITEMS + i) * 32 + lane;
        r[i] = make_ushort4(0, 0, 0, 0);
        if (k < V) r[i] = rect_sorted[k];
    }
    __syncthreads();
    // ballot words of every row of 32 + entries per (warp, super-block)
    uint32_t* const wbw = wb + warp * bp.S;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const L1Row q = l1_range(r[i], bp.sb_shift);
        uint32_t* const brow = ballots + (warp * ITEMS + i) * nb;
        for (int sx = q.sx0; sx <= q.sx1; ++sx) atomicOr(&brow[sx], 1u << lane);
        for (int sy = q.sy0; sy <= q.sy1; ++sy) {
            atomicOr(&brow[bp.sbx + sy], 1u << lane);
            for (int sx = q.sx0; sx <= q.sx1; ++sx) atomicAdd(&wbw[sy * bp.sbx + sx], 1u);
        }
    }
    __syncthreads();
    // counts -> first slot: the (super-block, chunk) run's offset + the entries of the warps in front
    for (int s = tid; s < bp.S; s += L1_THREADS) {
        uint32_t run = l1off[(size_t)s * bp.nl1 + c];
#pragma unroll
        for (int w = 0; w < L1_WARPS; ++w) {
            const uint32_t n = wb[w * bp.S + s];
            wb[w * bp.S + s] = run;
            run += n;
        }
    }
    __syncthreads();
    const unsigned lt = (1u << lane) - 1u;
    const int sb = bp.sb;
#pragma unroll              // r[] stays in registers only with compile-time indices
    for (int i = 0; i < ITEMS; ++i) {
        const L1Row q = l1_range(r[i], bp.sb_shift);
        const bool ok = q.sx0 <= q.sx1;
        if (!__any_sync(0xFFFFFFFFu, ok)) continue;
        const uint32_t* const brow = ballots + (warp * ITEMS + i) * nb;
        const int k = k0 + (warp * ITEMS + i) * 32 + lane;
        const uint32_t id = ok ? order[k] : 0u;
        for (int sy = q.sy0; sy <= q.sy1; ++sy) {
            const uint32_t rowb = brow[bp.sbx + sy] & lt;
            const int oy = sy * sb;
            const int cy0 = max((int)r[i].y, oy) - oy, cy1 = min((int)r[i].w, oy + sb) - oy;
            const uint32_t rm = (((1u << (cy1 - cy0)) - 1u) << cy0) << 16;
            for (int sx = q.sx0; sx <= q.sx1; ++sx) {
                const uint32_t rank = wbw[sy * bp.sbx + sx] + __popc(brow[sx] & rowb);
                const int ox = sx * sb;
                const int cx0 = max((int)r[i].x, ox) - ox, cx1 = min((int)r[i].z, ox + sb) - ox;
                l1[rank] = make_uint2((((1u << (cx1 - cx0)) - 1u) << cx0) | rm, id);
            }
        }
        __syncwarp();        // every lane has read the row's first slots ...
        for (int sy = q.sy0; sy <= q.sy1; ++sy)
            for (int sx = q.sx0; sx <= q.sx1; ++sx) atomicAdd(&wbw[sy * bp.sbx + sx], 1u);
        __syncwarp();        // ... before they move on for the next row of 32
    }
}

// place2: grid (NC groups, S super-blocks).  Warp w owns tile rows w (+ 8) of the super-block.  It first compacts, in
// order, the entries that touch its row into a small queue; per 32 queued entries and per tile ONE ballot: the covering
// entries get consecutive slots and leave as one run of 4-byte stores.
template <int SB>
__global__ void __launch_bounds__(256)
place2_kernel(BinPlan bp, int gx, int gy, const uint32_t* __restrict__ meta, const uint32_t* __restrict__ counts,
              const uint32_t* __restrict__ tile_start, const uint32_t* __restrict__ l1off, const uint2* __restrict__ l1,
              uint32_t* __restrict__ point_list, uint32_t capacity) {
    if (meta[0] > capacity) return;
    constexpr int ROWS = SB / 8;
    __shared__ uint2 s_q[8][ROWS][64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = blockIdx.x, s = blockIdx.y;
    const int K = bin_group_chunks(meta[1], bp);
    const int c_lo = min(G * K, bp.nl1), c_hi = min(G * K + K, bp.nl1);
    const uint32_t e0 = l1off[(size_t)s * bp.nl1 + c_lo], e1 = l1off[(size_t)s * bp.nl1 + c_hi];
    if (e0 >= e1) return;
    const int sy = s / bp.sbx, sx = s - sy * bp.sbx;
    const int tiles = gx * gy;
    uint32_t slot[ROWS][SB];
    bool any_tile = false;
#pragma unroll
    for (int rr = 0; rr < ROWS; ++rr) {
        const int ty = sy * SB + warp + 8 * rr;
#pragma unroll
        for (int x = 0; x < SB; ++x) {
            const int tx = sx * SB + x;
            const bool ok = ty < gy && tx < gx;
            const int t = ok ? ty * gx + tx : 0;
            slot[rr][x] = ok ? tile_start[t] + counts[(size_t)G * tiles + t] : 0u;
            any_tile |= ok;
        }
    }
    if (!any_tile) return;                           // (warp-uniform) this warp's rows lie below the grid
    const unsigned lt = (1u << lane) - 1u;
    int qn[ROWS], qh[ROWS];
#pragma unroll
    for (int rr = 0; rr < ROWS; ++rr) { qn[rr] = 0; qh[rr] = 0; }

    // 32 queued entries of row rr (lane = entry; m = 0 for lanes beyond the queue's end): one ballot per tile
    auto drain = [&](int rr, uint32_t m, uint32_t id) {
#pragma unroll
        for (int x = 0; x < SB; ++x) {
            const bool hit = (m >> x) & 1u;
            const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
            if (hit) point_list[slot[rr][x] + __popc(b & lt)] = id;
            slot[rr][x] += __popc(b);
        }
    };

    uint2 nxt = make_uint2(0u, 0u);
    if (e0 + lane < e1) nxt = l1[e0 + lane];
    for (uint32_t b0 = e0; b0 < e1; b0 += 32) {
        const uint2 ent = nxt;
        nxt = make_uint2(0u, 0u);
        if (b0 + 32 + lane < e1) nxt = l1[b0 + 32 + lane];
#pragma unroll
        for (int rr = 0; rr < ROWS; ++rr) {
            const bool rh = (ent.x >> (16 + warp + 8 * rr)) & 1u;
            const unsigned b = __ballot_sync(0xFFFFFFFFu, rh);
            if (b == 0u) continue;
            if (rh) s_q[warp][rr][(qh[rr] + qn[rr] + __popc(b & lt)) & 63] = ent;
            qn[rr] += __popc(b);
            __syncwarp();
            if (qn[rr] >= 32) {
                const uint2 e = s_q[warp][rr][(qh[rr] + lane) & 63];
                __syncwarp();
                qh[rr] = (qh[rr] + 32) & 63;
                qn[rr] -= 32;
                drain(rr, e.x, e.y);
            }
        }
    }
#pragma unroll
    for (int rr = 0; rr < ROWS; ++rr) {
        if (qn[rr] > 0) {
            uint2 e = make_uint2(0u, 0u);
            if (lane < qn[rr]) e = s_q[warp][rr][(qh[rr] + lane) & 63];
            drain(rr, e.x, e.y);
        }
    }
}

static size_t l1_smem_bytes(const BinPlan& bp, bool emit) {
    if (!emit) return (size_t)bp.S * 4;
    return ((size_t)L1_WARPS * bp.items * (bp.sbx + bp.sby) + (size_t)L1_WARPS * bp.S) * 4;
}

void launch_l1_count_scan(const GeomState& g, const ImageState& img, int P, const BinPlan& bp, cudaStream_t st) {
    (void)P;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(l1_emit_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 << 10);
        cudaFuncSetAttribute(l1_emit_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 << 10);
        attr_set = true;
    }
    const size_t smem = l1_smem_bytes(bp, false);
    const int n = bp.S * bp.nl1;
    const int nseg = (n + L1_SEG - 1) / L1_SEG;
    cudaMemsetAsync(img.l1seg, 0, (size_t)nseg * sizeof(uint32_t), st);
    if (bp.items == 16) l1_count_kernel<16><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, bp, g.meta, img.l1off, img.l1seg);
    else                l1_count_kernel<4><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, bp, g.meta, img.l1off, img.l1seg);
    l1_scan_kernel<<<nseg, 256, 0, st>>>(img.l1off, n, img.l1seg, g.meta);
}

void launch_place2(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp,
                   const BinningState& bin, uint32_t capacity, cudaStream_t st) {
    (void)P;
    const size_t smem = l1_smem_bytes(bp, true);
    if (bp.items == 16)
        l1_emit_kernel<16><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, g.dvals[0], bp, g.meta, img.l1off, bin.l1, capacity);
    else
        l1_emit_kernel<4><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, g.dvals[0], bp, g.meta, img.l1off, bin.l1, capacity);
    dim3 grid(bp.NC, bp.S, 1);
    if (bp.sb == 16)
        place2_kernel<16><<<grid, 256, 0, st>>>(bp, gx, gy, g.meta, img.counts, img.tile_start, img.l1off, bin.l1, bin.point_list, capacity);
    else
        place2_kernel<8><<<grid, 256, 0, st>>>(bp, gx, gy, g.meta, img.counts, img.tile_start, img.l1off, bin.l1, bin.point_list, capacity);
}

}  // namespace msgs
