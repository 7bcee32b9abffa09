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

using namespace msgs;
static StatsUnits pack(int n, const int* lvl, const int* what, const int32_t* const* radii, const float* const* px,
                       const float* const* g2d) {
    StatsUnits U;
    U.n = n;
    for (int u = 0; u < n; ++u) { U.lvl[u] = lvl[u]; U.what[u] = what[u]; U.radii[u] = radii[u]; U.px[u] = px[u]; U.g2d[u] = g2d[u]; }
    return U;
}

extern "C" void host_update_stats_units(int P, int L, int n, const int* lvl, const int* what, const int32_t* const* radii,
                                        const float* const* px, const float* const* g2d, const int64_t* target, float* max_px,
                                        float* min_px, float* max_r, float* accum, float* denom, uint8_t* base) {
    const StatsUnits U = pack(n, lvl, what, radii, px, g2d);
    const long long threads = ((long long)P + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i) update_stats_units_thread(i, P, L, U, target, max_px, min_px, max_r, accum, denom, base);
}

// delta lanes as separate arrays (the device code derives them from one buffer; the bodies take pointers)
extern "C" void host_stats_delta_units(int P, int L, int n, const int* lvl, const int* what, const int32_t* const* radii,
                                       const float* const* px, const float* const* g2d, const int64_t* target, float* dacc,
                                       float* ddenom, float* radius, float* obs_hi, float* obs_lo, float* base_seen) {
    const StatsUnits U = pack(n, lvl, what, radii, px, g2d);
    const long long threads = ((long long)P + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i) stats_delta_thread(i, P, L, U, target, dacc, ddenom, radius, obs_hi, obs_lo, base_seen);
}

extern "C" void host_stats_delta_apply(int P, int L, const int64_t* target, const float* dacc, const float* ddenom,
                                       const float* radius, const float* obs_hi, const float* obs_lo, const float* base_seen,
                                       float* max_px, float* min_px, float* max_r, float* accum, float* denom, uint8_t* base) {
    const long long threads = ((long long)P + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i)
        stats_apply_thread(i, P, L, target, dacc, ddenom, radius, obs_hi, obs_lo, base_seen, max_px, min_px, max_r, accum, denom, base);
}
