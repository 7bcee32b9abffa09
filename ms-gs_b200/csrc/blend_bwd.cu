// Backward tile blend (replaces renderCUDA<3> of backward.cu:472-734).
// Per pixel: the reference's back-to-front recurrence (backward.cu:619-731); its two divisions by (1 - alpha) share
// one correctly rounded reciprocal (gradients are toleranced, not bit-compared).  What changes is
// where the nine per-splat gradient sums go: the reference issues 9 global atomicAdd per (pixel, splat)
// pair; here a warp (8x4 pixels) first reduces its 32 partial sums with shuffles — the 9 values share
// one recursive-halving butterfly (12 shuffles instead of 45) — and then 9 lanes issue ONE
// red.global in all into the splat's 48-byte accumulator record (gacc, common.cuh).  Warps that the CTA-level
// footprint mask (blend.cuh) proves untouched by a splat skip it entirely, and the walk starts at the tile's
// max n_contrib instead of at the end of the tile's list.
#include "blend.cuh"

// A/B switch ("tried and rejected"): 1 stages the three record quarters the blend loop reads with
// cp.async (global -> shared without registers) and recomputes the cull mask from shared memory after the round; 0 (the
// product) is LDG -> registers -> mask -> STS with the loads issued a round ahead.
#ifndef MSGS_BLEND_BWD_CPASYNC
#define MSGS_BLEND_BWD_CPASYNC 0
#endif

namespace msgs {

__device__ __forceinline__ void cp_async16_cg(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}

__device__ __forceinline__ float exp2f_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// one staged list entry: 64 bytes, so that the blend loop addresses it as base + slot * 64 + constant
struct __align__(16) BwdSplat {
    float4 a;             // {x, y, depth, blend_px}
    float4 b;             // {conic.x, conic.y, conic.z, opacity}
    float4 c;             // {r, g, b, filter bits}
    uint32_t id;          // Gaussian index
    uint32_t pad[3];
};
struct BwdStage {
    BwdSplat e[BATCH];
    uint8_t mask[BATCH];  // warps the splat can touch (0 if the backward's filter re-check skips it, quirk Q2)
};

// Sum v[0..7] and w over the 32 lanes with recursive halving (12 shuffles for 9 values).  Afterwards a lane with
// (lane & 2) == 0 returns the full sum of value number (lane >> 2), a lane with (lane & 2) != 0 the full sum of w.
__device__ __forceinline__ float butterfly9(float v[8], float w, int lane) {
    const bool u16 = lane & 16, u8 = lane & 8, u4 = lane & 4, u2 = lane & 2;
    float x4[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = u16 ? v[k] : v[k + 4];
        const float keep = u16 ? v[k + 4] : v[k];
        x4[k] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 16);
    }
    w += __shfl_xor_sync(0xFFFFFFFFu, w, 16);
    float x2[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float send = u8 ? x4[k] : x4[k + 2];
        const float keep = u8 ? x4[k + 2] : x4[k];
        x2[k] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 8);
    }
    w += __shfl_xor_sync(0xFFFFFFFFu, w, 8);
    float y;
    {
        const float send = u4 ? x2[0] : x2[1];
        const float keep = u4 ? x2[1] : x2[0];
        y = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 4);
    }
    w += __shfl_xor_sync(0xFFFFFFFFu, w, 4);
    // y: value (lane >> 2) summed over the lanes of my class (lane & 3); w: summed over my class.  Fold the classes:
    const float send = u2 ? y : w;
    const float keep = u2 ? w : y;
    float z = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 2);
    z += __shfl_xor_sync(0xFFFFFFFFu, z, 1);
    return z;
}

__global__ void __launch_bounds__(BLEND_THREADS)
blend_bwd_kernel(int W, int H, int gx, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                 const float4* __restrict__ rec, const float* __restrict__ bg, const float* __restrict__ final_Ts,
                 const uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ tile_max_contrib,
                 const float* __restrict__ dL_dpixels, float* __restrict__ gacc, int skip_bits) {
    __shared__ BwdStage stage[2];
    __shared__ uint16_t s_hits[BLEND_THREADS / 32][BATCH];      // byte offsets of this warp's hits inside stage[].e

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
    // Entries at list positions >= max n_contrib are skipped by every pixel (backward.cu:624): start there.
    const int total = (int)min(range.y - range.x, tile_max_contrib[tile]);
    const int rounds = (total + BATCH - 1) / BATCH;
    if (rounds == 0) return;
    const uint32_t last_idx = range.x + total - 1;    // list index of the first entry walked (the backmost)

    const float T_final = inside ? final_Ts[pix_id] : 0;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
    // entries at list positions >= the warp's max n_contrib are skipped by all 32 pixels of this warp
    const int warp_last = __reduce_max_sync(0xFFFFFFFFu, last_contributor);
    float accum_rec[3] = {0.f, 0.f, 0.f};
    float dL_dpixel[3] = {0.f, 0.f, 0.f};
    if (inside) {
        const size_t HW = (size_t)H * W;
        dL_dpixel[0] = dL_dpixels[0 * HW + pix_id];
        dL_dpixel[1] = dL_dpixels[1 * HW + pix_id];
        dL_dpixel[2] = dL_dpixels[2 * HW + pix_id];
    }
    float last_alpha = 0;
    float last_color[3] = {0.f, 0.f, 0.f};
    // backward.cu:551-552 — double literal 0.5 (quirk Q4)
    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;
    // backward.cu:700-702 — constant per pixel, hoisted (with T_final: -T_final * bg . dL_dpixel)
    float bg_dot_dpixel = 0;
    for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];
    const float neg_Tf_bg = -T_final * bg_dot_dpixel;

    // slot k of round r holds list entry last_idx - (r*BATCH + k): back to front.  Records of round r+1 and ids of
    // round r+2 are in flight while round r is processed.
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 ra = zero4, rb = zero4, rc = zero4, rf = zero4;
    uint32_t id_cur = 0, id_next = 0;
    if (tid < total) {
        id_cur = point_list[last_idx - tid];
        const float4* src = rec + REC_F4 * (size_t)id_cur;
        ra = src[0]; rb = src[1]; rc = src[2]; rf = src[3];
        BwdSplat& e = stage[0].e[tid];
        e.a = ra; e.b = rb; e.c = rc; e.id = id_cur;
        stage[0].mask[tid] = (((int)rc.w & skip_bits) == 0) ? (uint8_t)splat_warp_mask(ra, rb, rf, tx0, ty0) : (uint8_t)0;
    }
    // the 9 sums of a splat leave the warp through ONE red instruction: after butterfly9 lane 4v holds sum number
    // v (v < 8) and lane 2 the opacity sum
    const bool red_on = (lane & 3) == 0 || lane == 2;
    const int red_off = (lane & 3) == 0 ? (lane >> 2) : 8;
    if (BATCH + tid < total) id_next = point_list[last_idx - BATCH - tid];

    for (int r = 0; r < rounds; ++r) {
        const int cur = r & 1;
        const int count = min(BATCH, total - r * BATCH);
        __syncthreads();      // publishes stage[cur]; everyone has finished reading stage[cur ^ 1] (round r - 1)
        const int pn = (r + 1) * BATCH + tid;
        const bool have_next = pn < total;
        if (have_next) {
            id_cur = id_next;
            const float4* src = rec + REC_F4 * (size_t)id_cur;
#if MSGS_BLEND_BWD_CPASYNC
            BwdSplat& e = stage[cur ^ 1].e[tid];      // nobody reads stage[cur ^ 1] during round r
            cp_async16_cg(&e.a, src + 0);
            cp_async16_cg(&e.b, src + 1);
            cp_async16_cg(&e.c, src + 2);
            rf = src[3];
#else
            ra = src[0]; rb = src[1]; rc = src[2]; rf = src[3];
#endif
            if (pn + BATCH < total) id_next = point_list[last_idx - pn - BATCH];
        }
#if MSGS_BLEND_BWD_CPASYNC
        asm volatile("cp.async.commit_group;\n" ::);
#endif

        const BwdStage& s = stage[cur];
        // 0-based list position of slot j is top - j; the reference's `contributor` after its decrement
        const int top = total - 1 - r * BATCH;
        // slot j matters to this warp only if top - j < warp_last  <=>  j >= top - warp_last + 1
        const int n = compact_hits<uint16_t, 6>(s.mask, count, warp, lane, s_hits[warp], max(0, top - warp_last + 1));
        const unsigned char* sbase = reinterpret_cast<const unsigned char*>(s.e);
        for (int k = 0; k < n; ++k) {
            const int off = s_hits[warp][k];
            const BwdSplat& e = *reinterpret_cast<const BwdSplat*>(sbase + off);
            const int contributor = top - (off >> 6);

            float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            float g_op = 0.f;
            bool active = false;
            if (contributor < last_contributor) {
                const float4 xy = e.a;
                const float4 con_o = e.b;
                const float2 d = {xy.x - pixf.x, xy.y - pixf.y};
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                if (!(power > 0.0f)) {
                    // ex2.approx of power * log2(e) (2 instructions) instead of expf's range-reduced evaluation (10): power is
                    // in [-5.6, 0] for every pair that survives the alpha test, where the two differ by <= 4e-7 relative —
                    // three orders below the gradient bar (the forward keeps expf: n_contrib and final_T are bit-compared)
                    const float G = exp2f_approx(power * 1.4426950408889634f);
                    const float alpha = min(0.99f, con_o.w * G);
                    if (!(alpha < 1.0f / 255.0f)) {
                        active = true;
                        const float4 col = e.c;
                        // one reciprocal serves T / (1 - alpha) and T_final / (1 - alpha) (backward.cu:672, 705)
                        // (1 - alpha in [0.01, 1]: approximate reciprocal + one Newton step, no special cases)
                        const float oma = 1.f - alpha;
                        float inv1ma;
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv1ma) : "f"(oma));
                        inv1ma = fmaf(inv1ma, fmaf(-oma, inv1ma, 1.0f), inv1ma);
                        T = T * inv1ma;
                        const float dchannel_dcolor = alpha * T;
                        float dL_dalpha = 0.0f;
                        const float cc[3] = {col.x, col.y, col.z};
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            const float c = cc[ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = c;
                            const float dL_dchannel = dL_dpixel[ch];
                            dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
                            g[ch] = dchannel_dcolor * dL_dchannel;
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += neg_Tf_bg * inv1ma;

                        const float dL_dG = con_o.w * dL_dalpha;
                        const float gdx = G * d.x;
                        const float gdy = G * d.y;
                        const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
                        const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
                        g[3] = dL_dG * dG_ddelx * ddelx_dx;
                        g[4] = dL_dG * dG_ddely * ddely_dy;
                        g[5] = -0.5f * gdx * d.x * dL_dG;
                        g[6] = -0.5f * gdx * d.y * dL_dG;
                        g[7] = -0.5f * gdy * d.y * dL_dG;
                        g_op = G * dL_dalpha;
                    }
                }
            }
            if (!__any_sync(0xFFFFFFFFu, active)) continue;
            const float sum = butterfly9(g, g_op, lane);
            if (red_on) atomicAdd(gacc + 12 * (size_t)e.id + red_off, sum);
        }
#if MSGS_BLEND_BWD_CPASYNC
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        if (r + 1 < rounds && have_next) {
            BwdStage& nx = stage[cur ^ 1];
            BwdSplat& e = nx.e[tid];
            const float4 sa = e.a, sb = e.b;          // this thread's own copies: visible to it after the wait
            const float sw = e.c.w;
            e.id = id_cur;
            nx.mask[tid] = (((int)sw & skip_bits) == 0) ? (uint8_t)splat_warp_mask(sa, sb, rf, tx0, ty0) : (uint8_t)0;
        }
#else
        if (r + 1 < rounds && have_next) {
            BwdStage& nx = stage[cur ^ 1];
            BwdSplat& e = nx.e[tid];
            e.a = ra; e.b = rb; e.c = rc; e.id = id_cur;
            nx.mask[tid] = (((int)rc.w & skip_bits) == 0) ? (uint8_t)splat_warp_mask(ra, rb, rf, tx0, ty0) : (uint8_t)0;
        }
#endif
    }
}

void launch_blend_bwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                      const float4* rec, const float* bg, const float* final_T, const uint32_t* n_contrib,
                      const uint32_t* tile_max_contrib, const float* dL_dpix, float* gacc, int filter_small,
                      int filter_large, cudaStream_t st) {
    dim3 grid(gx, gy, 1);
    // quirk Q2: the backward skips a pair when the Gaussian fails filter_small OR filter_large, tested independently
    // (backward.cu:649-660); the per-Gaussian outcomes are bits 0 / 1 of the record's filter field (preprocess_fwd.cu)
    const int skip_bits = (filter_small ? 1 : 0) | (filter_large ? 2 : 0);
    blend_bwd_kernel<<<grid, BLEND_THREADS, 0, st>>>(W, H, gx, ranges, point_list, rec, bg, final_T, n_contrib,
                                                      tile_max_contrib, dL_dpix, gacc, skip_bits);
}

}  // namespace msgs
