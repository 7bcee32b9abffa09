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

// Smallest value of the quadratic form q(d) = A dx^2 + 2 B dx dy + C dy^2 (d measured from the splat centre) over the
// axis-aligned box [x0, x1] x [y0, y1] (coordinates relative to the centre).  q is convex with its minimum 0 at the
// centre, so over a box that does not contain the centre the minimum sits on an edge that faces the centre: on the
// facing vertical edge (dx fixed) at dy* = clamp(-B dx / C), on the facing horizontal edge at dx* = clamp(-B dy / A).
__device__ __forceinline__ float min_quadratic_over_box(float A, float B, float C, float rA, float rC,
                                                        float x0, float x1, float y0, float y1) {
    const float dx = fminf(fmaxf(0.0f, x0), x1);        // nearest point of the box to the centre, per axis
    const float dy = fminf(fmaxf(0.0f, y0), y1);
    float q = 3.0e38f;
    if (dx != 0.0f) {
        const float t = fminf(fmaxf(-B * dx * rC, y0), y1);
        q = A * dx * dx + 2.0f * B * dx * t + C * t * t;
    }
    if (dy != 0.0f) {
        const float t = fminf(fmaxf(-B * dy * rA, x0), x1);
        q = fminf(q, A * t * t + 2.0f * B * t * dy + C * dy * dy);
    }
    return (dx == 0.0f && dy == 0.0f) ? 0.0f : q;
}

// Bit w of the result is clear only if the splat provably contributes to no pixel of warp w's 8x4 block
// (warp w: columns [8*(w&1), +7], rows [4*(w>>1), +3] of the tile whose first pixel is (tx0, ty0)): the minimum of
// -2 * power over the block exceeds 2 ln(255 * opacity) (with slack for rounding), i.e. alpha < 1/255 at all 32 pixels.
// a = {x, y, depth, -}, b = {conic.x, conic.y, conic.z, opacity}.
__device__ __forceinline__ unsigned splat_warp_mask(const float4 a, const float4 b, float tx0, float ty0) {
    const float o = b.w;
    if (o < 0.0039f) return 0u;                    // alpha <= opacity*exp(power<=0) < 1/255 everywhere
    const float A = b.x, B = b.y, C = b.z;
    const float det = A * C - B * B;
    if (!(det > 0.0f && A > 0.0f && C > 0.0f)) return 0xFFu;        // not an ellipse (or NaN): no culling
    const float thr = 2.0f * (__logf(255.0f * o) + 0.02f) * 1.002f + 0.01f;   // -2 power <= thr inside the footprint
    const float rA = 1.0f / A, rC = 1.0f / C;
    const float cx = a.x - tx0, cy = a.y - ty0;                     // centre relative to the tile
    if (!(cx == cx && cy == cy)) return 0xFFu;                      // NaN centre: no culling
    // bounding box of the footprint first (cheap reject of whole columns / rows of blocks)
    const float inv = thr / det;
    const float hx = sqrtf(inv * C) * 1.001f + 0.02f, hy = sqrtf(inv * A) * 1.001f + 0.02f;
    unsigned m = 0u;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const float bx0 = (float)((w & 1) * 8), by0 = (float)((w >> 1) * 4);
        const float x0 = bx0 - cx, x1 = bx0 + 7.0f - cx, y0 = by0 - cy, y1 = by0 + 3.0f - cy;
        if (x1 < -hx || x0 > hx || y1 < -hy || y0 > hy) continue;
        if (min_quadratic_over_box(A, B, C, rA, rC, x0, x1, y0, y1) <= thr) m |= 1u << w;
    }
    return m;
}

// One warp's compaction of the staged splats that name it: idx[] receives their slots (ascending = list order).
// Returns the number of hits.  `masks` are the CTA's 8-bit warp masks of this round, `count` the staged entries;
// slots below `first` are left out (backward: entries behind every pixel's last contributor).
__device__ __forceinline__ int compact_hits(const uint8_t* masks, int count, int warp, int lane, uint8_t* idx, int first = 0) {
    int n = 0;
    for (int g0 = first & ~31; g0 < count; g0 += 32) {
        const int mine = g0 + lane;
        const bool hit = mine >= first && mine < count && ((masks[mine] >> warp) & 1u);
        const unsigned b = __ballot_sync(0xFFFFFFFFu, hit);
        if (hit) idx[n + __popc(b & ((1u << lane) - 1u))] = (uint8_t)mine;
        n += __popc(b);
    }
    __syncwarp();
    return n;
}

}  // namespace msgs
