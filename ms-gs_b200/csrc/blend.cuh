// Shared pieces of the two blend kernels (forward: blend_fwd.cu, backward: blend_bwd.cu).
//
// Thread -> pixel map: a 16x16 tile is 8 warps, each owning an 8x4 pixel block (2 blocks across, 4 down).
// The map is free (results are per pixel); 8x4 blocks make the warp-level culling below much tighter than
// the reference's 16x2 strips (forward.cu:444-446).
//
// Staging: every round the CTA takes the next BATCH = 256 entries of its tile's list, one per thread.  The thread
// loads its splat record (48 B, three 16-byte loads issued a round ahead so the latency hides behind the previous
// round's blending), computes ONCE for the whole CTA which of the 8 warp blocks the splat can touch — the axis-aligned
// bounding box of its alpha >= 1/255 footprint (the ellipse power >= -ln(255*opacity), inflated for rounding) against
// the 2 x 4 block grid — and writes record + 8-bit warp mask to shared memory.  Each warp then compacts the indices of
// the splats whose mask names it (ballot + popc) and runs a dense counted loop over them.
// The cull is conservative: a splat it rejects for a warp fails the reference's own per-pixel tests (power > 0 or
// alpha < 1/255, forward.cu:535,560 / backward.cu:632,669) at all 32 pixels of that warp, so skipping it changes
// no result.
#pragma once
#include "common.cuh"

namespace msgs {

constexpr int BLEND_THREADS = 256;
constexpr int BATCH = 256;      // splats staged per round

// pixel owned by thread `tid` inside its tile
__device__ __forceinline__ void thread_pixel(int tid, int& lx, int& ly) {
    const int warp = tid >> 5, lane = tid & 31;
    lx = (warp & 1) * 8 + (lane & 7);
    ly = (warp >> 1) * 4 + (lane >> 3);
}

// Bit w of the result is clear only if the splat provably contributes to no pixel of warp w's 8x4 block
// (warp w: columns [8*(w&1), +7], rows [4*(w>>1), +3] of the tile whose first pixel is (tx0, ty0)).
// a = {x, y, depth, -}, b = {conic.x, conic.y, conic.z, opacity}.
__device__ __forceinline__ unsigned splat_warp_mask(const float4 a, const float4 b, float tx0, float ty0) {
    const float o = b.w;
    if (o < 0.0039f) return 0u;                    // alpha <= opacity*exp(power<=0) < 1/255 everywhere
    const float det = b.x * b.z - b.y * b.y;
    if (!(det > 0.0f && b.x > 0.0f && b.z > 0.0f)) return 0xFFu;    // not an ellipse (or NaN): no culling
    const float tau2 = 2.0f * (__logf(255.0f * o) + 0.02f);         // power >= -tau inside the footprint
    const float inv = tau2 / det;
    const float hx = sqrtf(inv * b.z) * 1.001f + 0.02f;
    const float hy = sqrtf(inv * b.x) * 1.001f + 0.02f;
    const float x0 = a.x - hx - tx0, x1 = a.x + hx - tx0;           // footprint box relative to the tile
    const float y0 = a.y - hy - ty0, y1 = a.y + hy - ty0;
    // NaN coordinates compare false everywhere: treat as "touches everything" by testing for the misses
    const unsigned cm = (!(x1 < 0.0f || x0 > 7.0f) ? 1u : 0u) | (!(x1 < 8.0f || x0 > 15.0f) ? 2u : 0u);
    unsigned m = 0u;
    if (!(y1 < 0.0f || y0 > 3.0f)) m |= cm;
    if (!(y1 < 4.0f || y0 > 7.0f)) m |= cm << 2;
    if (!(y1 < 8.0f || y0 > 11.0f)) m |= cm << 4;
    if (!(y1 < 12.0f || y0 > 15.0f)) m |= cm << 6;
    return m;
}

// One warp's compaction of the staged splats that name it: idx[] receives their slots (ascending = list order).
// Returns the number of hits.  `masks` are the CTA's 8-bit warp masks of this round, `count` the staged entries.
__device__ __forceinline__ int compact_hits(const uint8_t* masks, int count, int warp, int lane, uint8_t* idx) {
    int n = 0;
    for (int g0 = 0; g0 < count; g0 += 32) {
        const int mine = g0 + lane;
        const bool hit = mine < count && ((masks[mine] >> warp) & 1u);
        const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
        if (hit) idx[n + __popc(b & ((1u << lane) - 1u))] = (uint8_t)mine;
        n += __popc(b);
    }
    __syncwarp();
    return n;
}

}  // namespace msgs
