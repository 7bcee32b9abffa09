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

}  // namespace msgs
