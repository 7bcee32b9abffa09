// Forward tile blend (replaces renderCUDA<3>, forward.cu:420-599).  FP32-pipe / issue bound.
// One CTA per 16x16 tile, one thread per pixel; splat records are gathered into shared memory with
// cp.async (16 B x 3 per splat), double-buffered so the gather of round i+1 overlaps the blending of
// round i; warp ballots (blend.cuh) skip splats that cannot touch a warp's 8x4 pixel block, and the tile
// stops as soon as every pixel's transmittance has saturated.
//
// Per pixel the arithmetic, its order, the skip rules and the termination rule are the reference's
// (forward.cu:524-582): acc_pixel_size / acc_depth accumulate BEFORE the T < 1e-4 test, colour after it
// (quirk Q7); n_contrib is the 1-based list position of the last splat that updated T.
#include "blend.cuh"

namespace msgs {

struct FwdStage {
    float4 a[BATCH];   // {x, y, depth, blend_px}
    float4 b[BATCH];   // {conic.x, conic.y, conic.z, opacity}
    float4 c[BATCH];   // {r, g, b, -}
};

__global__ void __launch_bounds__(BLEND_THREADS)
blend_fwd_kernel(int W, int H, int gx, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                 const float4* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                 uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ tile_max_contrib,
                 float* __restrict__ out_color, float* __restrict__ out_acc_px, float* __restrict__ out_depth) {
    __shared__ FwdStage stage[2];
    __shared__ uint32_t s_max_contrib;

    const int tid = threadIdx.x, lane = tid & 31;
    const int tile = blockIdx.y * gx + blockIdx.x;
    int lx, ly;
    thread_pixel(tid, lx, ly);
    const uint32_t px = blockIdx.x * TILE_X + lx, py = blockIdx.y * TILE_Y + ly;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float2 pixf = {(float)px, (float)py};
    // the warp's 8x4 block in pixel-centre coordinates
    const float bx0 = (float)(blockIdx.x * TILE_X + ((tid >> 5) & 1) * 8), by0 = (float)(blockIdx.y * TILE_Y + (tid >> 6) * 4);
    const float bx1 = bx0 + 7.0f, by1 = by0 + 3.0f;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;
    if (tid == 0) s_max_contrib = 0;

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float acc_pixel_size = 0.f, acc_depth = 0.f;

    // software pipeline: ids of round r+1 are fetched while round r's records are in flight
    uint32_t id_next = 0;
    if (tid < total) id_next = point_list[range.x + tid];
    if (rounds > 0) {
        if (tid < total) {
            const float4* src = rec + 3 * (size_t)id_next;
            cp_async16(&stage[0].a[tid], src);
            cp_async16(&stage[0].b[tid], src + 1);
            cp_async16(&stage[0].c[tid], src + 2);
        }
        cp_async_commit();
        if (BATCH + tid < total) id_next = point_list[range.x + BATCH + tid];
    }

    for (int r = 0; r < rounds; ++r) {
        const int cur = r & 1;
        const int count = min(BATCH, total - r * BATCH);
        if (r + 1 < rounds) {
            const int p = (r + 1) * BATCH + tid;
            if (p < total) {
                const float4* src = rec + 3 * (size_t)id_next;
                cp_async16(&stage[cur ^ 1].a[tid], src);
                cp_async16(&stage[cur ^ 1].b[tid], src + 1);
                cp_async16(&stage[cur ^ 1].c[tid], src + 2);
            }
            cp_async_commit();
            if (p + BATCH < total) id_next = point_list[range.x + p + BATCH];
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        // forward.cu:485-487 — stop when the whole tile is saturated (also publishes the staged round)
        if (__syncthreads_count(done) == BLEND_THREADS) break;

        const FwdStage& s = stage[cur];
        const bool warp_live = __any_sync(0xFFFFFFFFu, !done);
        if (warp_live) {
            for (int g0 = 0; g0 < count; g0 += 32) {
                const int mine = g0 + lane;
                bool hit = false;
                if (mine < count) hit = splat_hits_rect(s.a[mine], s.b[mine], bx0, by0, bx1, by1);
                unsigned todo = __ballot_sync(0xFFFFFFFFu, hit);
                while (todo) {
                    const int j = g0 + __ffs(todo) - 1;
                    todo &= todo - 1;
                    if (done) continue;
                    const float4 xy = s.a[j];
                    const float4 con_o = s.b[j];
                    const float2 d = {xy.x - pixf.x, xy.y - pixf.y};
                    const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                    if (power > 0.0f) continue;
                    const float alpha = min(0.99f, con_o.w * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;

                    acc_pixel_size += xy.w * T * alpha;
                    acc_depth += xy.z * T * alpha;

                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) {
                        done = true;
                        continue;
                    }
                    const float4 col = s.c[j];
                    C[0] += col.x * alpha * T;
                    C[1] += col.y * alpha * T;
                    C[2] += col.z * alpha * T;
                    T = test_T;
                    last_contributor = (uint32_t)(r * BATCH + j + 1);
                }
            }
        }
        __syncthreads();   // everyone is done reading stage[cur] before round r+2 overwrites it
    }
    cp_async_wait<0>();

    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        const size_t HW = (size_t)H * W;
        out_color[0 * HW + pix_id] = C[0] + T * bg[0];
        out_color[1 * HW + pix_id] = C[1] + T * bg[1];
        out_color[2 * HW + pix_id] = C[2] + T * bg[2];
        out_acc_px[pix_id] = acc_pixel_size;
        out_depth[pix_id] = acc_depth;
    }
    // tile-wide max of n_contrib: where the backward starts its back-to-front walk
    uint32_t m = __reduce_max_sync(0xFFFFFFFFu, inside ? last_contributor : 0u);
    __syncthreads();
    if (lane == 0) atomicMax(&s_max_contrib, m);
    __syncthreads();
    if (tid == 0) tile_max_contrib[tile] = s_max_contrib;
}

void launch_blend_fwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, float* final_T, uint32_t* n_contrib,
                      uint32_t* tile_max_contrib, float* out_color, float* out_acc_px, float* out_depth,
                      cudaStream_t st) {
    dim3 grid(gx, gy, 1);
    blend_fwd_kernel<<<grid, BLEND_THREADS, 0, st>>>(W, H, gx, ranges, point_list, rec, bg, final_T, n_contrib,
                                                      tile_max_contrib, out_color, out_acc_px, out_depth);
}

}  // namespace msgs
