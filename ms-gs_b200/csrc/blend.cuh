// Shared pieces of the two blend kernels (forward: blend_fwd.cu, backward: blend_bwd.cu).
//
// Thread -> pixel map: a 16x16 tile is 8 warps, each owning an 8x4 pixel block (2 blocks across, 4 down).
// The map is free (results are per pixel); 8x4 blocks make the warp-level culling below much tighter than
// the reference's 16x2 strips (forward.cu:444-446).
//
// Staging: every round the CTA takes the next BATCH = 256 entries of its tile's list, one per thread.  The thread
// loads its splat record (64 B, four 16-byte loads issued a round ahead so the latency hides behind the previous
// round's blending), computes ONCE for the whole CTA which of the 8 warp blocks the splat can touch — its
// alpha >= 1/255 footprint (the ellipse power >= -ln(255*opacity), inflated for rounding; cull.cuh) against the
// 2 x 4 block grid, exactly — and writes record + 8-bit warp mask to shared memory.  Each warp then compacts the indices of
// the splats whose mask names it (ballot + popc) and runs a dense counted loop over them.
// The cull is conservative: a splat it rejects for a warp fails the reference's own per-pixel tests (power > 0 or
// alpha < 1/255, forward.cu:535,560 / backward.cu:632,669) at all 32 pixels of that warp, so skipping it changes
// no result.
#pragma once
#include "common.cuh"
#include "cull.cuh"

namespace msgs {

constexpr int BLEND_THREADS = 256;
constexpr int BATCH = 256;      // splats staged per round

// pixel owned by thread `tid` inside its tile
__device__ __forceinline__ void thread_pixel(int tid, int& lx, int& ly) {
    const int warp = tid >> 5, lane = tid & 31;
    lx = (warp & 1) * 8 + (lane & 7);
    ly = (warp >> 1) * 4 + (lane >> 3);
}

// One warp's compaction of the staged splats that name it: idx[] receives their slots (ascending = list order).
// Returns the number of hits.  `masks` are the CTA's 8-bit warp masks of this round, `count` the staged entries;
// slots below `first` are left out (backward: entries behind every pixel's last contributor).
// idx[] entries are slot << SHIFT (the blend loops use them as byte offsets into their staged arrays).
template <typename IdxT = uint8_t, int SHIFT = 0>
__device__ __forceinline__ int compact_hits(const uint8_t* masks, int count, int warp, int lane, IdxT* idx, int first = 0) {
    int n = 0;
    for (int g0 = first & ~31; g0 < count; g0 += 32) {
        const int mine = g0 + lane;
        const bool hit = mine >= first && mine < count && ((masks[mine] >> warp) & 1u);
        const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
        if (hit) idx[n + __popc(b & ((1u << lane) - 1u))] = (IdxT)(mine << SHIFT);
        n += __popc(b);
    }
    __syncwarp();
    return n;
}

}  // namespace msgs
