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
//   place       : chunk c starts from slot[tile] = tile_start + counts[c][tile] (kept in registers, lane = tile) and
//                 walks its Gaussians in depth order; every covered tile gets the index at slot[tile]++.  Warps own
//                 disjoint tile blocks, so the walk needs no atomics.                     (R x 4 B written)
// All integer work; HBM-bound only through the final 4-byte stores.
#include "common.cuh"

namespace msgs {

// How the placement pass walks its chunk (see place_kernel): 0 = rectangle + index of every Gaussian prefetched a
// round ahead; 1 = 4-byte super-block masks (written by the counting pass), rectangle + index fetched for the hits only.
// Masks pay when a CTA's super-block concerns few of the Gaussians it walks: many super-blocks (a 240 x 135 tile grid
// has 60: ~2 % hits), or a coarse unit of a view group (`sparse`: it walks the group's whole depth order and renders a
// small part of it).  With 8 super-blocks and every walked Gaussian rendered (C2) a sixth are hits and prefetching wins.
int place_walk_mode(const BinPlan& bp, int gx, int gy, bool sparse) {
    const int sb_x = (gx + 4 * bp.bw - 1) / (4 * bp.bw), sb_y = (gy + 2 * bp.bh - 1) / (2 * bp.bh);
    return (sb_x * sb_y >= 13 || sparse) ? 1 : 0;
}

constexpr int BIN_THREADS = 256;
constexpr int BIN_WARPS = BIN_THREADS / 32;

__device__ __forceinline__ int odd_width(int gx) { return (gx + 1) | 1; }

// ------------------------------------------------------------------------------------------------
// chunk_count: grid (NC, nbands)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BIN_THREADS)
chunk_count_kernel(int P, const uint32_t* __restrict__ order, const ushort4* __restrict__ rect,
                   ushort4* __restrict__ rect_sorted, uint32_t* __restrict__ sbmask, int want_mask, int gx, int gy,
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
        if (band == 0) {
            rect_sorted[k] = r;                     // depth-ordered copy for the placement pass (coalesced there)
            // which super-blocks of the placement pass (4bw x 2bh tiles each) the rectangle meets: column bits | row
            // bits << 16 — the 4 bytes a placement CTA reads per Gaussian to decide whether it concerns it at all
            uint32_t m = 0u;
            if (want_mask && r.x < r.z && r.y < r.w) {
                const int sbw = 4 * bp.bw, sbh = 2 * bp.bh;
                const int sx0 = min(15, (int)r.x / sbw), sx1 = min(15, ((int)r.z - 1) / sbw);
                const int sy0 = min(15, (int)r.y / sbh), sy1 = min(15, ((int)r.w - 1) / sbh);
                m = (((2u << sx1) - 1u) & ~((1u << sx0) - 1u)) | ((((2u << sy1) - 1u) & ~((1u << sy0) - 1u)) << 16);
            }
            if (want_mask) sbmask[k] = m;
        }
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

void launch_chunk_count(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp, int walk_mode,
                        cudaStream_t st) {
    const size_t smem = (size_t)(bp.band_rows + 1) * ((gx + 1) | 1) * sizeof(int);
    dim3 grid(bp.NC, bp.nbands, bp.split);
    if (bp.split > 1) cudaMemsetAsync(img.counts, 0, (size_t)bp.NC * gx * gy * sizeof(uint32_t), st);
    // the super-block masks go into the depth sort's spare key buffer (free once the sort has finished)
    chunk_count_kernel<<<grid, BIN_THREADS, smem, st>>>(P, g.dvals[0], g.rect, g.rect_sorted, g.dkeys[1], walk_mode, gx, gy, bp, g.meta, img.counts);
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
// place: grid (NC chunks, super-blocks).  A CTA owns a super-block of 4bw x 2bh tiles, its 8 warps own one
// bw x bh block each (4 across, 2 down; BinPlan picks bw, bh; at most 10 x 7 = 70 tiles).  Inside a warp, LANE = TILE:
// lane l holds, in registers, the next free list slot of tiles l, l + 32 and l + 64 of its block.  The chunk is read
// 1024 Gaussians at a time (4 per thread, coalesced from the depth-ordered copies); the ones whose rectangle meets
// the super-block are compacted, IN ORDER, into a shared-memory queue (rectangle clipped to the super-block + index).
// Every warp then walks the queue in order: the entries that meet its block are found 32 at a time (lane = entry) and
// compacted into the warp's queue as (column mask | row mask << 16, index) words; a lane whose tile lies in both masks stores
// the Gaussian index at its slot and bumps the slot.  No atomics, no shared-memory counters, deterministic: depth
// order inside every tile.  Slots start at tile_start + counts[chunk][tile].
// ------------------------------------------------------------------------------------------------
constexpr int PLACE_SLOTS = PLACE_BW_MAX * PLACE_BH_MAX;
constexpr int PLACE_K = (PLACE_SLOTS + 31) / 32;          // tiles per lane
constexpr int PLACE_ITEMS = 4;                             // Gaussians per thread and round
constexpr int PLACE_ROUND = BIN_THREADS * PLACE_ITEMS;
static_assert(PLACE_BW_MAX <= 16 && PLACE_BH_MAX <= 16, "column / row masks are 16 bits each");

template <bool MASK_WALK>
__global__ void __launch_bounds__(BIN_THREADS)
place_kernel(int P, const uint32_t* __restrict__ order, const ushort4* __restrict__ rect_sorted,
             const uint32_t* __restrict__ sbmask, int gx, int gy,
             BinPlan bp, int sb_x, const uint32_t* __restrict__ meta, const uint32_t* __restrict__ counts,
             const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ point_list, uint32_t capacity) {
    if (meta[0] > capacity) return;                 // speculatively sized list too small: the host relaunches (api.cu)
    __shared__ uint32_t s_qx[2][PLACE_ROUND], s_qy[2][PLACE_ROUND], s_qid[2][PLACE_ROUND];
    __shared__ uint32_t s_wcnt[2][BIN_WARPS];
    __shared__ uint2 s_wq[BIN_WARPS][32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x;
    const int BW = bp.bw, BH = bp.bh;
    const int sbx0 = (blockIdx.y % sb_x) * 4 * BW, sby0 = (blockIdx.y / sb_x) * 2 * BH;
    const int sbx1 = min(gx, sbx0 + 4 * BW), sby1 = min(gy, sby0 + 2 * BH);
    // this warp's block, clipped to the grid (may be empty)
    const int bx0 = min(sbx1, sbx0 + (warp & 3) * BW), bx1 = min(sbx1, bx0 + BW);
    const int by0 = min(sby1, sby0 + (warp >> 2) * BH), by1 = min(sby1, by0 + BH);
    const bool warp_has_tiles = bx1 > bx0 && by1 > by0;
    // my tiles: membership mask (column bit | row bit << 16; all ones = no tile) and next free slot
    uint32_t tmask[PLACE_K], slot[PLACE_K];
    {
        const uint32_t* cnt = counts + (size_t)c * gx * gy;
#pragma unroll
        for (int k = 0; k < PLACE_K; ++k) {
            const int i = lane + 32 * k;
            const int yy = i / BW, xx = i - yy * BW;
            const int x = bx0 + xx, y = by0 + yy;
            const bool ok = i < BW * BH && x < bx1 && y < by1;
            tmask[k] = ok ? ((1u << xx) | (1u << (16 + yy))) : 0xFFFFFFFFu;
            slot[k] = ok ? tile_start[y * gx + x] + cnt[y * gx + x] : 0u;
        }
    }
    const int V = (int)meta[1], chunkG = bin_chunk_size((uint32_t)V, bp);
    const int k0 = min(V, c * chunkG), k1 = min(V, k0 + chunkG);
    if (k0 >= k1) return;                           // (uniform) chunk beyond the rendered Gaussians
    // A round takes PLACE_ROUND consecutive Gaussians; warp w reads the PLACE_ITEMS x 32 of them starting at
    // w * 32 * PLACE_ITEMS (item i, lane l -> + 32 i + l), so (warp, item, lane) order is depth order.
    // Register prefetch of the next round.
    // The walk reads ONE 4-byte super-block mask per Gaussian (chunk_count wrote it: column bits | row bits << 16) and
    // tests it against this CTA's super-block with one AND + compare; rectangle and index are fetched for the hits only
    // (on a 240 x 135 tile grid a CTA's super-block concerns ~2 % of the Gaussians it walks).  Masks hold 16 columns and
    // 16 rows of super-blocks; beyond that (grids above 640 x 224 tiles) bit 15 stands for "15 or more" and the clip
    // below decides.
    const int my_sbx = blockIdx.y % sb_x, my_sby = blockIdx.y / sb_x;
    const uint32_t my_bits = (1u << min(15, my_sbx)) | (1u << (16 + min(15, my_sby)));
    // With few super-blocks (a 78 x 52 grid has 8) a sixth of the walked Gaussians are hits and fetching their
    // rectangles behind the mask test exposes that latency every round: there the rectangle + index of EVERY walked
    // Gaussian are prefetched a round ahead instead (MASK_WALK = false; the launcher picks by the super-block count).
    uint32_t m_next[PLACE_ITEMS];
    ushort4 r_next[PLACE_ITEMS];
    uint32_t id_next[PLACE_ITEMS];
    const int woff = warp * 32 * PLACE_ITEMS + lane;
#pragma unroll
    for (int i = 0; i < PLACE_ITEMS; ++i) {
        const int k = k0 + woff + 32 * i;
        if (MASK_WALK) {
            m_next[i] = k < k1 ? sbmask[k] : 0u;
        } else {
            r_next[i] = make_ushort4(0, 0, 0, 0); id_next[i] = 0;
            if (k < k1) { r_next[i] = rect_sorted[k]; id_next[i] = order[k]; }
        }
    }
    int buf = 0;
    for (int kb = k0; kb < k1; kb += PLACE_ROUND, buf ^= 1) {
        uint32_t qx[PLACE_ITEMS], qy[PLACE_ITEMS], qid[PLACE_ITEMS];
        unsigned bal[PLACE_ITEMS];
        bool cand[PLACE_ITEMS];
        ushort4 rr[PLACE_ITEMS];
        int wtotal = 0;
#pragma unroll
        for (int i = 0; i < PLACE_ITEMS; ++i) {
            const int kc = kb + woff + 32 * i;
            const int k = kc + PLACE_ROUND;
            if (MASK_WALK) {
                cand[i] = (m_next[i] & my_bits) == my_bits;
                rr[i] = make_ushort4(0, 0, 0, 0); qid[i] = 0;
                if (cand[i]) { rr[i] = rect_sorted[kc]; qid[i] = order[kc]; }
                m_next[i] = k < k1 ? sbmask[k] : 0u;
            } else {
                cand[i] = true;
                rr[i] = r_next[i]; qid[i] = id_next[i];
                r_next[i] = make_ushort4(0, 0, 0, 0);
                if (k < k1) { r_next[i] = rect_sorted[k]; id_next[i] = order[k]; }
            }
        }
#pragma unroll
        for (int i = 0; i < PLACE_ITEMS; ++i) {
            const ushort4 r = rr[i];
            // clip to the super-block
            const int xa = max((int)r.x, sbx0), xb = min((int)r.z, sbx1);
            const int ya = max((int)r.y, sby0), yb = min((int)r.w, sby1);
            const bool hit = cand[i] && xa < xb && ya < yb;
            qx[i] = (uint32_t)xa | ((uint32_t)xb << 16);
            qy[i] = (uint32_t)ya | ((uint32_t)yb << 16);
            bal[i] = __ballot_sync(0xFFFFFFFFu, hit);
            wtotal += __popc(bal[i]);
        }
        if (lane == 0) s_wcnt[buf][warp] = wtotal;
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < BIN_WARPS; ++w) {
            const int n = (int)s_wcnt[buf][w];
            if (w < warp) before += n;
            total += n;
        }
        // ordered compaction of the survivors into the queue
#pragma unroll
        for (int i = 0; i < PLACE_ITEMS; ++i) {
            if ((bal[i] >> lane) & 1u) {
                const int pos = before + __popc(bal[i] & ((1u << lane) - 1u));
                s_qx[buf][pos] = qx[i];
                s_qy[buf][pos] = qy[i];
                s_qid[buf][pos] = qid[i];
            }
            before += __popc(bal[i]);
        }
        __syncthreads();
        if (!warp_has_tiles) continue;
        for (int q0 = 0; q0 < total; q0 += 32) {
            const int q = q0 + lane;
            // lane-parallel: clip entry q to this warp's block -> column mask | row mask << 16 (block coordinates)
            uint32_t pk = 0, eid = 0;
            if (q < total) {
                const uint32_t ex = s_qx[buf][q], ey = s_qy[buf][q];
                eid = s_qid[buf][q];
                const int cxa = max((int)(ex & 0xFFFF), bx0), cxb = min((int)(ex >> 16), bx1);
                const int cya = max((int)(ey & 0xFFFF), by0), cyb = min((int)(ey >> 16), by1);
                if (cxa < cxb && cya < cyb)
                    pk = (((1u << (cxb - cxa)) - 1u) << (cxa - bx0)) | ((((1u << (cyb - cya)) - 1u) << (cya - by0)) << 16);
            }
            // the entries that meet this warp's block, compacted in order into the warp's own queue: the walk below then
            // reads one broadcast 8-byte word per entry instead of find-first-set + two shuffles
            const unsigned hitm = __ballot_sync(0xFFFFFFFFu, pk != 0u);
            if (pk != 0u) s_wq[warp][__popc(hitm & ((1u << lane) - 1u))] = make_uint2(pk, eid);
            __syncwarp();
            const int nh = __popc(hitm);
            for (int h = 0; h < nh; ++h) {
                const uint2 e = s_wq[warp][h];
#pragma unroll
                for (int k = 0; k < PLACE_K; ++k) {
                    if ((e.x & tmask[k]) == tmask[k]) { point_list[slot[k]] = e.y; ++slot[k]; }
                }
            }
            __syncwarp();
        }
    }
}

void launch_place(const GeomState& g, const ImageState& img, int P, int gx, int gy, const BinPlan& bp, int walk_mode,
                  uint32_t* point_list, uint32_t capacity, cudaStream_t st) {
    const int sb_x = (gx + 4 * bp.bw - 1) / (4 * bp.bw), sb_y = (gy + 2 * bp.bh - 1) / (2 * bp.bh);
    dim3 grid(bp.NC, sb_x * sb_y, 1);
    if (walk_mode)
        place_kernel<true><<<grid, BIN_THREADS, 0, st>>>(P, g.dvals[0], g.rect_sorted, g.dkeys[1], gx, gy, bp, sb_x, g.meta, img.counts, img.tile_start, point_list, capacity);
    else
        place_kernel<false><<<grid, BIN_THREADS, 0, st>>>(P, g.dvals[0], g.rect_sorted, g.dkeys[1], gx, gy, bp, sb_x, g.meta, img.counts, img.tile_start, point_list, capacity);
}


// ------------------------------------------------------------------------------------------------
// Two-level placement (common.cuh BinPlan).  Level 1: per (super-block, list chunk) entry lists.
//
// A level-1 CTA owns list chunk c = sub = 256 x ITEMS consecutive depth-ordered Gaussians; warp w, item i, lane l holds
// Gaussian c * sub + (w * ITEMS + i) * 32 + l, so (warp, item, lane) order is depth order.  For every such "warp row" of
// 32 Gaussians the CTA keeps, in shared memory, one ballot word per super-block COLUMN (which of the 32 rectangles reach
// it) and one per super-block ROW: the Gaussians of the row that meet super-block (sx, sy) are col[sx] & row[sy] —
// sbx + sby ballots describe all sbx * sby super-blocks.  Counting is popc over those words; the rank of a Gaussian's
// entry inside its (super-block, chunk) run is the popc over the earlier warp rows plus the lower lanes of its own.
// ------------------------------------------------------------------------------------------------
constexpr int L1_THREADS = 256;
constexpr int L1_WARPS = L1_THREADS / 32;

struct L1Row { int sx0, sx1, sy0, sy1; };       // super-block range of a rectangle; sx0 > sx1: meets nothing

template <int ITEMS>
__device__ __forceinline__ void l1_load_and_ballot(const ushort4* __restrict__ rect_sorted, int V, int k0, const BinPlan& bp,
                                                   uint32_t* colB, uint32_t* rowB, ushort4 (&r)[ITEMS], L1Row (&sr)[ITEMS]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int k = k0 + (warp * ITEMS + i) * 32 + lane;
        r[i] = make_ushort4(0, 0, 0, 0);
        if (k < V) r[i] = rect_sorted[k];
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int wr = warp * ITEMS + i;
        const bool ok = r[i].x < r[i].z && r[i].y < r[i].w;
        L1Row q;
        q.sx0 = ok ? ((int)r[i].x >> bp.sb_shift) : 1;
        q.sx1 = ok ? (((int)r[i].z - 1) >> bp.sb_shift) : 0;
        q.sy0 = ok ? ((int)r[i].y >> bp.sb_shift) : 1;
        q.sy1 = ok ? (((int)r[i].w - 1) >> bp.sb_shift) : 0;
        sr[i] = q;
        for (int sx = 0; sx < bp.sbx; ++sx) {
            const unsigned b = __ballot_sync(0xFFFFFFFFu, q.sx0 <= sx && sx <= q.sx1);
            if (lane == 0) colB[wr * bp.sbx + sx] = b;
        }
        for (int sy = 0; sy < bp.sby; ++sy) {
            const unsigned b = __ballot_sync(0xFFFFFFFFu, q.sy0 <= sy && sy <= q.sy1);
            if (lane == 0) rowB[wr * bp.sby + sy] = b;
        }
    }
}

// l1_count: grid nl1.  l1off[s * nl1 + c] = entries of list chunk c in super-block s (zeros for chunks beyond V).
template <int ITEMS>
__global__ void __launch_bounds__(L1_THREADS)
l1_count_kernel(const ushort4* __restrict__ rect_sorted, BinPlan bp, const uint32_t* __restrict__ meta,
                uint32_t* __restrict__ l1off) {
    extern __shared__ uint32_t l1_smem[];
    constexpr int WR = L1_WARPS * ITEMS;
    uint32_t* colB = l1_smem;
    uint32_t* rowB = l1_smem + WR * bp.sbx;
    const int c = blockIdx.x, tid = threadIdx.x;
    const int V = (int)meta[1], k0 = c * bp.sub;
    if (k0 >= V) {                                   // (uniform) no rendered Gaussian in this chunk
        for (int s = tid; s < bp.S; s += L1_THREADS) l1off[(size_t)s * bp.nl1 + c] = 0u;
        return;
    }
    ushort4 r[ITEMS];
    L1Row sr[ITEMS];
    l1_load_and_ballot<ITEMS>(rect_sorted, V, k0, bp, colB, rowB, r, sr);
    __syncthreads();
    for (int s = tid; s < bp.S; s += L1_THREADS) {
        const int sy = s / bp.sbx, sx = s - sy * bp.sbx;
        uint32_t n = 0;
        for (int wr = 0; wr < WR; ++wr) n += __popc(colB[wr * bp.sbx + sx] & rowB[wr * bp.sby + sy]);
        l1off[(size_t)s * bp.nl1 + c] = n;
    }
}

// l1_scan: exclusive scan of l1off[0 .. n) in place, total -> l1off[n] and meta[2].  Two launches, no spinning:
// per-segment sums, then every segment adds the sums in front of it to its local scan.
__global__ void __launch_bounds__(256)
l1_segsum_kernel(const uint32_t* __restrict__ data, int n, uint32_t* __restrict__ segsum) {
    __shared__ uint32_t s_w[8];
    const int base = blockIdx.x * L1_SEG;
    uint32_t sum = 0;
    for (int i = threadIdx.x; i < L1_SEG; i += 256) sum += (base + i < n) ? data[base + i] : 0u;
    sum = __reduce_add_sync(0xFFFFFFFFu, sum);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; ++w) t += s_w[w];
        segsum[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256)
l1_scan_kernel(uint32_t* __restrict__ data, int n, const uint32_t* __restrict__ segsum, uint32_t* __restrict__ meta) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg = blockIdx.x, base = seg * L1_SEG;
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
    // thread t scans the 16 consecutive elements [base + 16 t, + 16)
    constexpr int PER = L1_SEG / 256;
    uint32_t v[PER];
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int j = base + tid * PER + i;
        v[i] = j < n ? data[j] : 0u;
        sum += v[i];
    }
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
    uint32_t* colB = l1_smem;
    uint32_t* rowB = colB + WR * bp.sbx;
    uint32_t* base_s = rowB + WR * bp.sby;                               // S: first entry of (s, this chunk)
    uint16_t* warp_base = reinterpret_cast<uint16_t*>(base_s + bp.S);    // L1_WARPS x S: entries of the warps in front
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int V = (int)meta[1], k0 = c * bp.sub;
    if (k0 >= V) return;
    ushort4 r[ITEMS];
    L1Row sr[ITEMS];
    l1_load_and_ballot<ITEMS>(rect_sorted, V, k0, bp, colB, rowB, r, sr);
    __syncthreads();
    for (int s = tid; s < bp.S; s += L1_THREADS) {
        const int sy = s / bp.sbx, sx = s - sy * bp.sbx;
        uint32_t run = 0;
        for (int w = 0; w < L1_WARPS; ++w) {
            warp_base[w * bp.S + s] = (uint16_t)run;
            for (int i = 0; i < ITEMS; ++i) {
                const int wr = w * ITEMS + i;
                run += __popc(colB[wr * bp.sbx + sx] & rowB[wr * bp.sby + sy]);
            }
        }
        base_s[s] = l1off[(size_t)s * bp.nl1 + c];
    }
    __syncthreads();
    const unsigned lt = (1u << lane) - 1u;
    const int sb = bp.sb;
#pragma unroll              // r[] / sr[] stay in registers only with compile-time indices
    for (int i = 0; i < ITEMS; ++i) {
        const L1Row q = sr[i];
        const bool ok = q.sx0 <= q.sx1;
        if (!__any_sync(0xFFFFFFFFu, ok)) continue;
        const int k = k0 + (warp * ITEMS + i) * 32 + lane;
        const uint32_t id = ok ? order[k] : 0u;
        int sx = q.sx0, sy = q.sy0;
        bool active = ok;
        while (__any_sync(0xFFFFFFFFu, active)) {
            if (active) {
                const int s = sy * bp.sbx + sx;
                uint32_t rank = base_s[s] + warp_base[warp * bp.S + s];
                for (int ip = 0; ip < i; ++ip) {
                    const int wr = warp * ITEMS + ip;
                    rank += __popc(colB[wr * bp.sbx + sx] & rowB[wr * bp.sby + sy]);
                }
                const int wr = warp * ITEMS + i;
                rank += __popc(colB[wr * bp.sbx + sx] & rowB[wr * bp.sby + sy] & lt);
                const int ox = sx * sb, oy = sy * sb;
                const int cx0 = max((int)r[i].x, ox) - ox, cx1 = min((int)r[i].z, ox + sb) - ox;
                const int cy0 = max((int)r[i].y, oy) - oy, cy1 = min((int)r[i].w, oy + sb) - oy;
                const uint32_t cm = ((1u << (cx1 - cx0)) - 1u) << cx0;
                const uint32_t rm = ((1u << (cy1 - cy0)) - 1u) << cy0;
                l1[rank] = make_uint2(cm | (rm << 16), id);
                if (++sx > q.sx1) { sx = q.sx0; if (++sy > q.sy1) active = false; }
            }
        }
    }
}

// place2: grid (NC groups, S super-blocks).  Warp w owns tile rows w (+ 8) of the super-block; per 32 entries and per
// tile ONE ballot: the covering entries get consecutive slots and leave as one run of 4-byte stores.
template <int SB>
__global__ void __launch_bounds__(256)
place2_kernel(BinPlan bp, int gx, int gy, const uint32_t* __restrict__ meta, const uint32_t* __restrict__ counts,
              const uint32_t* __restrict__ tile_start, const uint32_t* __restrict__ l1off, const uint2* __restrict__ l1,
              uint32_t* __restrict__ point_list, uint32_t capacity) {
    if (meta[0] > capacity) return;
    constexpr int ROWS = SB / 8;
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
    uint2 nxt = make_uint2(0u, 0u);
    if (e0 + lane < e1) nxt = l1[e0 + lane];
    for (uint32_t b0 = e0; b0 < e1; b0 += 32) {
        const uint2 ent = nxt;
        nxt = make_uint2(0u, 0u);
        if (b0 + 32 + lane < e1) nxt = l1[b0 + 32 + lane];
#pragma unroll
        for (int rr = 0; rr < ROWS; ++rr) {
            const bool rh = (ent.x >> (16 + warp + 8 * rr)) & 1u;
            if (__ballot_sync(0xFFFFFFFFu, rh) == 0u) continue;
#pragma unroll
            for (int x = 0; x < SB; ++x) {
                const bool hit = rh && ((ent.x >> x) & 1u);
                const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
                if (hit) point_list[slot[rr][x] + __popc(b & lt)] = ent.y;
                slot[rr][x] += __popc(b);
            }
        }
    }
}

static size_t l1_smem_bytes(const BinPlan& bp, bool emit) {
    const size_t wr = (size_t)L1_WARPS * bp.items;
    size_t b = wr * (bp.sbx + bp.sby) * 4;
    if (emit) b += (size_t)bp.S * 4 + (size_t)L1_WARPS * bp.S * 2 + 16;
    return b;
}

void launch_l1_count_scan(const GeomState& g, const ImageState& img, int P, const BinPlan& bp, cudaStream_t st) {
    (void)P;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(l1_count_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 << 10);
        cudaFuncSetAttribute(l1_count_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 << 10);
        cudaFuncSetAttribute(l1_emit_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 << 10);
        cudaFuncSetAttribute(l1_emit_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 << 10);
        attr_set = true;
    }
    const size_t smem = l1_smem_bytes(bp, false);
    if (bp.items == 16) l1_count_kernel<16><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, bp, g.meta, img.l1off);
    else                l1_count_kernel<4><<<bp.nl1, L1_THREADS, smem, st>>>(g.rect_sorted, bp, g.meta, img.l1off);
    const int n = bp.S * bp.nl1;
    const int nseg = (n + L1_SEG - 1) / L1_SEG;
    l1_segsum_kernel<<<nseg, 256, 0, st>>>(img.l1off, n, img.l1seg);
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
