// Test-only: the per-thread body of ms-gs_b200/csrc/stats_body.cuh run on the HOST for every thread index of the launch
// grid (256-thread blocks, overshoot included).  Built by tests/test_zz_stats.py.
#include "../ms-gs_b200/csrc/stats_body.cuh"

extern "C" void host_update_stats(int P, int L, int lvl, int what, const int32_t* radii, const float* px,
                                  const int64_t* target, const float* g2d, float* max_px, float* min_px, float* max_r,
                                  float* accum, float* denom, uint8_t* base) {
    const long long threads = ((long long)P + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i)
        msgs::update_stats_thread(i, P, L, lvl, what, radii, px, target, g2d, max_px, min_px, max_r, accum, denom, base);
}
