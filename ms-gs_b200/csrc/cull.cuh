// Conservative per-warp culling of splats in the blend kernels (blend.cuh) and the per-Gaussian footprint it uses
// (written by preprocess_fwd.cu into the 4th float4 of the splat record).
//
// A (pixel, splat) pair is skipped by the reference when power > 0 or alpha = opacity * exp(power) < 1/255
// (forward.cu:535,560 / backward.cu:632,669), i.e. unless  q(d) = -2 power = A dx^2 + 2 B dx dy + C dy^2  <=  2 ln(255 o).
// With slack for rounding this is q <= thr; the footprint's bounding box is |dx| <= hx, |dy| <= hy.  A warp owns an
// 8x4 pixel block: if the minimum of q over that block exceeds thr, all 32 of its pixels skip the splat and the warp
// need not look at it.  The test never rejects a splat the reference would blend.
#pragma once
#include <cuda_runtime.h>

namespace msgs {

// {hx, hy, thr, 0}.  thr < 0: contributes nowhere; thr = hx = hy = huge: no culling (not an ellipse / NaN).
__device__ __forceinline__ float4 splat_footprint(float A, float B, float C, float o) {
    if (o < 0.0039f) return make_float4(0.f, 0.f, -1.0f, 0.f);      // alpha <= o * exp(power <= 0) < 1/255 everywhere
    const float det = A * C - B * B;
    if (!(det > 0.0f && A > 0.0f && C > 0.0f)) return make_float4(3.0e38f, 3.0e38f, 3.0e38f, 0.f);
    const float thr = 2.0f * (__logf(255.0f * o) + 0.02f) * 1.002f + 0.01f;
    const float inv = thr / det;
    return make_float4(sqrtf(inv * C) * 1.001f + 0.02f, sqrtf(inv * A) * 1.001f + 0.02f, thr, 0.f);
}

// Smallest value of q over the box [x0, x1] x [y0, y1] (coordinates relative to the splat centre), or an upper
// bound of it that is exact whenever it matters.  q is convex with minimum 0 at the centre; over a box the minimum
// sits on an edge facing the centre: on the vertical edge nearest the centre at dy* = clamp(-B dx / C), or on the
// nearest horizontal edge at dx* = clamp(-B dy / A).  (If the centre lies inside the box's x-range, dx = 0 and the
// first candidate is q(0, clamp(0)) — a point of the box, so still a valid upper bound; likewise for y.)
__device__ __forceinline__ float min_quadratic_over_box(float A, float B, float C, float rA, float rC,
                                                        float x0, float x1, float y0, float y1) {
    const float dx = fminf(fmaxf(0.0f, x0), x1);        // nearest point of the box to the centre, per axis
    const float dy = fminf(fmaxf(0.0f, y0), y1);
    const float t1 = fminf(fmaxf(-B * dx * rC, y0), y1);
    const float q1 = A * dx * dx + t1 * (2.0f * B * dx + C * t1);
    const float t2 = fminf(fmaxf(-B * dy * rA, x0), x1);
    const float q2 = C * dy * dy + t2 * (2.0f * B * dy + A * t2);
    return fminf(q1, q2);
}

// Bit w of the result is clear only if the splat provably contributes to no pixel of warp w's 8x4 block
// (warp w: columns [8*(w&1), +7], rows [4*(w>>1), +3] of the tile whose first pixel is (tx0, ty0)).
// a = {x, y, depth, -}, b = {conic.x, conic.y, conic.z, opacity}, f = splat_footprint(...).
__device__ __forceinline__ unsigned splat_warp_mask(const float4 a, const float4 b, const float4 f, float tx0, float ty0) {
    const float thr = f.z;
    if (thr < 0.0f) return 0u;
    const float A = b.x, B = b.y, C = b.z, hx = f.x, hy = f.y;
    const float cx = a.x - tx0, cy = a.y - ty0;                     // centre relative to the tile
    if (!(cx == cx && cy == cy) || thr > 1.0e38f) return 0xFFu;     // NaN centre / not an ellipse: no culling
    const float rA = __fdividef(1.0f, A), rC = __fdividef(1.0f, C);
    // rows and columns of blocks the bounding box reaches
    unsigned cols = 0u, rows = 0u;
    if (!(7.0f - cx < -hx || 0.0f - cx > hx)) cols |= 1u;
    if (!(15.0f - cx < -hx || 8.0f - cx > hx)) cols |= 2u;
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (!((float)(4 * r + 3) - cy < -hy || (float)(4 * r) - cy > hy)) rows |= 1u << r;
    unsigned m = 0u;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        if (!((cols >> (w & 1)) & (rows >> (w >> 1)) & 1u)) continue;
        const float bx0 = (float)((w & 1) * 8), by0 = (float)((w >> 1) * 4);
        const float q = min_quadratic_over_box(A, B, C, rA, rC, bx0 - cx, bx0 + 7.0f - cx, by0 - cy, by0 + 3.0f - cy);
        if (!(q > thr)) m |= 1u << w;
    }
    return m;
}

}  // namespace msgs
