// Per-thread body of the fused statistics update (stats.cu; include/msgs_stats.h).  __host__ __device__: the same code
// runs on the host in tests/test_zz_stats.py against golden vectors produced by the reference's own GaussianModel methods.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace msgs {

__host__ __device__ __forceinline__ void
update_stats_thread(long long i, int P, int L, int lvl, int what, const int32_t* __restrict__ radii,
                    const float* __restrict__ px, const int64_t* __restrict__ target, const float* __restrict__ g2d,
                    float* __restrict__ max_px, float* __restrict__ min_px, float* __restrict__ max_r,
                    float* __restrict__ accum, float* __restrict__ denom, uint8_t* __restrict__ base) {
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;                                  // visibility_filter = radii > 0: every update is masked by it
    if ((what & 1) && target[i] == (int64_t)lvl) {       // update_pixel_sizes (gaussian_model.py:662-686)
        const float p = px[i];
        if (lvl > 0) max_px[i] = fmaxf(max_px[i] * 0.95f, p);
        if (lvl < L - 1) {
            const float m = fmaxf(min_px[i] * 1.05f, -1.0f);          // torch.clip(x * 1.05, -1)
            min_px[i] = m < 0.0f ? (p > 0.0f ? p : m) : (p > 0.0f ? fminf(m, p) : m);
        }
    }
    if (what & 2) {                                      // train.py:249 + add_densification_stats (gaussian_model.py:698-701)
        max_r[i] = fmaxf(max_r[i], (float)r);
        const float gx = g2d[i * 3], gy = g2d[i * 3 + 1];
        accum[i * L + lvl] += sqrtf(gx * gx + gy * gy);
        denom[i * L + lvl] += 1.0f;
    }
    if (what & 4) base[i] = 1;                           // update_base_gaussian_mask (gaussian_model.py:703-705)
}

// ---- all units of a step in one launch, and the sharded step's delta / apply pair (include/msgs_stats.h) ----
constexpr int STATS_MAX_UNITS = 8;
struct StatsUnits {
    int n;
    int lvl[STATS_MAX_UNITS], what[STATS_MAX_UNITS];
    const int32_t* radii[STATS_MAX_UNITS];
    const float* px[STATS_MAX_UNITS];
    const float* g2d[STATS_MAX_UNITS];
};

// the units in order, exactly as n consecutive single-unit launches would apply them
__host__ __device__ __forceinline__ void
update_stats_units_thread(long long i, int P, int L, const StatsUnits& U, const int64_t* __restrict__ target,
                          float* __restrict__ max_px, float* __restrict__ min_px, float* __restrict__ max_r,
                          float* __restrict__ accum, float* __restrict__ denom, uint8_t* __restrict__ base) {
    for (int u = 0; u < U.n; ++u)
        update_stats_thread(i, P, L, U.lvl[u], U.what[u], U.radii[u], U.px[u], target, U.g2d[u], max_px, min_px, max_r, accum,
                            denom, base);
}

// Delta lanes of one Gaussian (what a rank's units observed this step; see msgs_stats.h for the layout):
//   sum  dacc[i*L + lvl], ddenom[i*L + lvl]
//   max  radius[i] (float, 0 = never visible), obs_hi[i] (largest pixel size observed by a unit of the Gaussian's target
//        level; 0 = no observation; an observation of size <= 0 is stored as the smallest denormal), base_seen[i] (0/1.0f)
//   min  obs_lo[i] (smallest observed pixel size > 0; +inf = none)
// Non-negative floats order like their bit patterns, which is what the switch's u32 max / min reduce.
#define MSGS_OBS_ZERO 1.401298464e-45f
__host__ __device__ __forceinline__ void
stats_delta_thread(long long i, int P, int L, const StatsUnits& U, const int64_t* __restrict__ target,
                   float* __restrict__ dacc, float* __restrict__ ddenom, float* __restrict__ radius,
                   float* __restrict__ obs_hi, float* __restrict__ obs_lo, float* __restrict__ base_seen) {
    if (i >= P) return;
    float rad = 0.f, hi = 0.f, lo = INFINITY, seen = 0.f;
    bool any = false;
    for (int u = 0; u < U.n; ++u) {
        const int r = U.radii[u][i];
        if (r <= 0) continue;
        any = true;
        const int lvl = U.lvl[u], what = U.what[u];
        if ((what & 1) && target[i] == (int64_t)lvl) {
            const float p = U.px[u][i];
            hi = fmaxf(hi, p > 0.f ? p : MSGS_OBS_ZERO);
            if (p > 0.f) lo = fminf(lo, p);
        }
        if (what & 2) {
            rad = fmaxf(rad, (float)r);
            const float gx = U.g2d[u][i * 3], gy = U.g2d[u][i * 3 + 1];
            dacc[i * L + lvl] += sqrtf(gx * gx + gy * gy);
            ddenom[i * L + lvl] += 1.0f;
        }
        if (what & 4) seen = 1.0f;
    }
    if (!any) return;
    if (rad > 0.f) radius[i] = fmaxf(radius[i], rad);
    if (hi > 0.f) obs_hi[i] = fmaxf(obs_hi[i], hi);
    if (lo < INFINITY) obs_lo[i] = fminf(obs_lo[i], lo);
    if (seen > 0.f) base_seen[i] = 1.0f;
}

// Apply the (reduced) deltas to the state: for a Gaussian observed by ONE unit of its target level this is
// update_stats_thread exactly; several observations of one step (other ranks' views) act as one with the largest / smallest
// observed size, i.e. the 0.95 / 1.05 ageing (gaussian_model.py:669-672) is applied once per step.
__host__ __device__ __forceinline__ void
stats_apply_thread(long long i, int P, int L, const int64_t* __restrict__ target, const float* __restrict__ dacc,
                   const float* __restrict__ ddenom, const float* __restrict__ radius, const float* __restrict__ obs_hi,
                   const float* __restrict__ obs_lo, const float* __restrict__ base_seen, float* __restrict__ max_px,
                   float* __restrict__ min_px, float* __restrict__ max_r, float* __restrict__ accum,
                   float* __restrict__ denom, uint8_t* __restrict__ base) {
    if (i >= P) return;
    const float hi = obs_hi[i];
    if (hi > 0.f) {
        const int lvl = (int)target[i];
        const float p = hi == MSGS_OBS_ZERO ? 0.f : hi;
        if (lvl > 0) max_px[i] = fmaxf(max_px[i] * 0.95f, p);
        if (lvl < L - 1) {
            const float lo = obs_lo[i];
            const bool valid = lo < INFINITY;
            const float m = fmaxf(min_px[i] * 1.05f, -1.0f);
            min_px[i] = m < 0.0f ? (valid ? lo : m) : (valid ? fminf(m, lo) : m);
        }
    }
    const float rad = radius[i];
    if (rad > 0.f) {
        max_r[i] = fmaxf(max_r[i], rad);
        for (int l = 0; l < L; ++l) {
            const float dn = ddenom[i * L + l];
            if (dn != 0.f) { accum[i * L + l] += dacc[i * L + l]; denom[i * L + l] += dn; }
        }
    }
    if (base && base_seen[i] > 0.f) base[i] = 1;
}

}  // namespace msgs
