// Forward tile blend (replaces renderCUDA<3>, forward.cu:420-599).  FP32-pipe / issue bound.
// One CTA per 16x16 tile, one thread per pixel.  Rounds of 256 list entries are staged through shared memory
// (blend.cuh): records are loaded one round ahead into registers, the CTA computes each splat's 8-bit warp mask once,
// every warp compacts its own hits and blends them in a dense loop; the tile stops as soon as every pixel's
// transmittance has saturated.  One block barrier per round.
//
// Per pixel the arithmetic, its order, the skip rules and the termination rule are the reference's
// (forward.cu:524-582): acc_pixel_size / acc_depth accumulate BEFORE the T < 1e-4 test, colour after it
// (quirk Q7); n_contrib is the 1-based list position of the last splat that updated T.
#include "blend.cuh"

namespace msgs {

struct FwdStage {
    float4 a[BATCH];      // {x, y, depth, blend_px}
    float4 b[BATCH];      // {conic.x, conic.y, conic.z, opacity}
    float4 c[BATCH];      // {r, g, b, -}
    uint8_t mask[BATCH];  // warps the splat can touch
};

__global__ void __launch_bounds__(BLEND_THREADS)
blend_fwd_kernel(int W, int H, int gx, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                 const float4* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                 uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ tile_max_contrib,
                 float* __restrict__ out_color, float* __restrict__ out_acc_px, float* __restrict__ out_depth,
                 const uint32_t* __restrict__ meta, uint32_t capacity) {
    if (meta[0] > capacity) return;                 // the list did not fit its speculative buffer: relaunched by the host
    __shared__ FwdStage stage[2];
    __shared__ uint8_t s_hits[BLEND_THREADS / 32][BATCH];
    __shared__ uint32_t s_max_contrib;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.y * gx + blockIdx.x;
    int lx, ly;
    thread_pixel(tid, lx, ly);
    const uint32_t px = blockIdx.x * TILE_X + lx, py = blockIdx.y * TILE_Y + ly;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float2 pixf = {(float)px, (float)py};
    const float tx0 = (float)(blockIdx.x * TILE_X), ty0 = (float)(blockIdx.y * TILE_Y);

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;
    if (tid == 0) s_max_contrib = 0;

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float acc_pixel_size = 0.f, acc_depth = 0.f;

    // software pipeline: records of round r+1 and ids of round r+2 are in flight while round r is blended
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 ra = zero4, rb = zero4, rc = zero4, rf = zero4;
    uint32_t id_next = 0;
    if (tid < total) {
        const float4* src = rec + REC_F4 * (size_t)point_list[range.x + tid];
        ra = src[0]; rb = src[1]; rc = src[2]; rf = src[3];
        stage[0].a[tid] = ra; stage[0].b[tid] = rb; stage[0].c[tid] = rc;
        stage[0].mask[tid] = (uint8_t)splat_warp_mask(ra, rb, rf, tx0, ty0);
    }
    if (BATCH + tid < total) id_next = point_list[range.x + BATCH + tid];

    for (int r = 0; r < rounds; ++r) {
        const int cur = r & 1;
        const int count = min(BATCH, total - r * BATCH);
        // forward.cu:485-487 — stop when the whole tile is saturated.  The barrier also publishes stage[cur] and
        // tells that everyone has finished reading stage[cur ^ 1] (round r - 1).
        if (__syncthreads_count(done) == BLEND_THREADS) break;
        const int pn = (r + 1) * BATCH + tid;         // my entry of the next round
        const bool have_next = pn < total;
        if (have_next) {
            const float4* src = rec + REC_F4 * (size_t)id_next;
            ra = __ldg(src); rb = __ldg(src + 1); rc = __ldg(src + 2); rf = __ldg(src + 3);
            if (pn + BATCH < total) id_next = point_list[range.x + pn + BATCH];
        }

        if (__any_sync(0xFFFFFFFFu, !done)) {
            const FwdStage& s = stage[cur];
            const int n = compact_hits(s.mask, count, warp, lane, s_hits[warp]);
            const uint32_t pos0 = (uint32_t)(r * BATCH + 1);
            for (int k = 0; k < n; ++k) {
                const int j = s_hits[warp][k];
                const float4 xy = s.a[j];
                const float4 con_o = s.b[j];
                const float2 d = {xy.x - pixf.x, xy.y - pixf.y};
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                if (done || power > 0.0f) continue;
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
                last_contributor = pos0 + (uint32_t)j;
            }
        }
        if (r + 1 < rounds) {
            FwdStage& nx = stage[cur ^ 1];
            if (have_next) {
                nx.a[tid] = ra; nx.b[tid] = rb; nx.c[tid] = rc;
                nx.mask[tid] = (uint8_t)splat_warp_mask(ra, rb, rf, tx0, ty0);
            }
        }
    }

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
                      const uint32_t* meta, uint32_t capacity, cudaStream_t st) {
    dim3 grid(gx, gy, 1);
    blend_fwd_kernel<<<grid, BLEND_THREADS, 0, st>>>(W, H, gx, ranges, point_list, rec, bg, final_T, n_contrib,
                                                      tile_max_contrib, out_color, out_acc_px, out_depth, meta, capacity);
}

}  // namespace msgs
