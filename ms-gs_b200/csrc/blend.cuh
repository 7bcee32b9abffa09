// Shared pieces of the two blend kernels (forward: blend_fwd.cu, backward: blend_bwd.cu).
//
// Thread -> pixel map: a 16x16 tile is 8 warps, each owning an 8x4 pixel block (2 blocks across, 4 down).
// The map is free (results are per pixel); 8x4 blocks make the warp-level culling below much tighter than
// the reference's 16x2 strips (forward.cu:444-446).
//
// Warp-level culling: for every staged splat one lane tests the splat's alpha >= 1/255 footprint (the
// axis-aligned bounding box of the ellipse power >= -ln(255*opacity), inflated for rounding) against the
// warp's 8x4 block; a ballot gives the warp the list of splats it must actually evaluate.  The test is
// conservative: a splat it rejects fails the reference's own per-pixel tests (power > 0 or alpha < 1/255,
// forward.cu:535,560 / backward.cu:632,669) at all 32 pixels, so skipping it changes no result.
#pragma once
#include "common.cuh"

namespace msgs {

constexpr int BLEND_THREADS = 256;
constexpr int BATCH = 256;      // splats staged per round

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// pixel owned by thread `tid` inside its tile
__device__ __forceinline__ void thread_pixel(int tid, int& lx, int& ly) {
    const int warp = tid >> 5, lane = tid & 31;
    lx = (warp & 1) * 8 + (lane & 7);
    ly = (warp >> 1) * 4 + (lane >> 3);
}

// true unless the splat provably contributes to no pixel of [x0,x1] x [y0,y1] (pixel-centre coordinates)
__device__ __forceinline__ bool splat_hits_rect(const float4 a, const float4 b, float x0, float y0, float x1, float y1) {
    const float o = b.w;
    if (o < 0.0039f) return false;                 // alpha <= opacity*exp(power<=0) < 1/255 everywhere
    const float det = b.x * b.z - b.y * b.y;
    if (!(det > 0.0f && b.x > 0.0f && b.z > 0.0f)) return true;     // not an ellipse (or NaN): no culling
    const float tau2 = 2.0f * (__logf(255.0f * o) + 0.02f);         // power >= -tau inside the footprint
    const float inv = tau2 / det;
    const float hx = sqrtf(inv * b.z) * 1.001f + 0.02f;
    const float hy = sqrtf(inv * b.x) * 1.001f + 0.02f;
    const bool out = (a.x + hx < x0) || (a.x - hx > x1) || (a.y + hy < y0) || (a.y - hy > y1);
    return !out;
}

}  // namespace msgs
