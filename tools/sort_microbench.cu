// Micro-benchmark of the depth sort (ms-gs_b200/csrc/sort.cu, u32 keys + u32 values, 4 passes) on its own: per-launch times
// with CUDA events, the whole sort back to back, and (-DMSGS_SORT_PHASES) where a pass CTA spends its time.  Built as ONE
// translation unit around sort.cu so that the tunables (-DMSGS_SORT_ITEMS / _WINDOW / _HIST_BATCH / _FUSE_SCAN) can be
// compared without touching the product library:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DMSGS_SORT_WINDOW=8 tools/sort_microbench.cu -o build/sortbench_w8
//   build/sortbench_w8 [n=1200000] [reps=200]
// Keys: float bits of a view depth (log-uniform in [0.5, 40]) with 15 % not rendered (0xFFFFFFFF), values = index, exactly
// what preprocess_fwd hands the sort.  The result is checked against std::stable_sort.
#include "../ms-gs_b200/csrc/sort.cu"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <random>
#include <vector>

using namespace msgs;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main(int argc, char** argv) {
    const size_t n = argc > 1 ? (size_t)atoll(argv[1]) : 1200000;
    const int reps = argc > 2 ? atoi(argv[2]) : 200;
    std::mt19937 rng(7);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::vector<uint32_t> hk(n), hv(n);
    for (size_t i = 0; i < n; ++i) {
        const float z = std::exp(std::log(0.5f) + u(rng) * (std::log(40.f) - std::log(0.5f)));
        uint32_t bits; memcpy(&bits, &z, 4);
        hk[i] = u(rng) < 0.15f ? 0xFFFFFFFFu : bits;
        hv[i] = (uint32_t)i;
    }
    uint32_t *k0, *k[2], *v0, *v[2]; void* temp;
    CK(cudaMalloc(&k0, n * 4)); CK(cudaMalloc(&v0, n * 4));
    for (int i = 0; i < 2; ++i) { CK(cudaMalloc(&k[i], n * 4)); CK(cudaMalloc(&v[i], n * 4)); }
    CK(cudaMalloc(&temp, sort_temp_bytes(n)));
    CK(cudaMemcpy(k0, hk.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(v0, hv.data(), n * 4, cudaMemcpyHostToDevice));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    const SortPlan pl = make_sort_plan(0, 32);
    const int NE = 3 + pl.passes;                    // memset | histogram (+ scan) | pass 0..3
    std::vector<cudaEvent_t> ev(NE + 1);
    for (auto& e : ev) CK(cudaEventCreate(&e));

    auto sort_once = [&](bool events) {
        // the forward sorts dkeys[0] in place of a copy; here the inputs are restored first (not timed with events)
        cudaMemcpyAsync(k[0], k0, n * 4, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(v[0], v0, n * 4, cudaMemcpyDeviceToDevice, st);
        int e = 0;
        if (events) cudaEventRecord(ev[e++], st);
        sort_temp_clear(temp, n, st);
        if (events) cudaEventRecord(ev[e++], st);
        launch_sort_histogram<uint32_t>(k[0], n, pl, temp, st);
        if (events) cudaEventRecord(ev[e++], st);
        int cur = 0;
        for (int p = 0; p < pl.passes; ++p) {
            launch_sort_pass<uint32_t>(k[cur], k[cur ^ 1], v[cur], v[cur ^ 1], n, pl, p, sort_temp_hist(temp), temp, st);
            cur ^= 1;
            if (events) cudaEventRecord(ev[e++], st);
        }
        return cur;
    };
    int cur = 0;
    for (int i = 0; i < 5; ++i) cur = sort_once(false);
    CK(cudaStreamSynchronize(st));
    // correctness
    std::vector<uint32_t> gk(n), gv(n), order(n);
    CK(cudaMemcpy(gk.data(), k[cur], n * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(gv.data(), v[cur], n * 4, cudaMemcpyDeviceToHost));
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return hk[a] < hk[b]; });
    size_t bad = 0;
    for (size_t i = 0; i < n; ++i) bad += (gv[i] != order[i]) || (gk[i] != hk[order[i]]);
    printf("items=%d window=%d hist_batch=%d fuse_scan=%d threads=%d hist_cps=%d ballot=%d late_lb=%d two_level=%d hist_threads=%d pdl=%d n=%zu tiles=%zu  check: %s (%zu wrong)\n", SORT_ITEMS,
           SORT_WINDOW, HIST_BATCH, MSGS_SORT_FUSE_SCAN, SORT_THREADS, MSGS_SORT_HIST_CPS, MSGS_SORT_BALLOT, MSGS_SORT_LATE_LOOKBACK, MSGS_SORT_TWO_LEVEL, MSGS_SORT_HIST_THREADS, MSGS_SORT_PDL, n, sort_tiles(n), bad ? "FAIL" : "ok", bad);

    // per-launch times
    std::vector<double> acc(NE, 0.0);
    for (int r = 0; r < reps; ++r) {
        sort_once(true);
        CK(cudaStreamSynchronize(st));
        for (int i = 0; i < NE - 1; ++i) { float ms; cudaEventElapsedTime(&ms, ev[i], ev[i + 1]); acc[i] += ms; }
    }
    printf("  per launch (us, events between launches): memset %.2f | histogram%s %.2f |", acc[0] * 1e3 / reps,
           MSGS_SORT_FUSE_SCAN ? "" : "+scan", acc[1] * 1e3 / reps);
    double tot = acc[0] + acc[1];
    for (int p = 0; p < pl.passes; ++p) { printf(" pass%d %.2f", p, acc[2 + p] * 1e3 / reps); tot += acc[2 + p]; }
    printf(" | sum %.2f\n", tot * 1e3 / reps);

    // the whole sort back to back (the restore copies timed separately and subtracted)
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float ms_all, ms_copy;
    CK(cudaEventRecord(a, st));
    for (int r = 0; r < reps; ++r) sort_once(false);
    CK(cudaEventRecord(b, st)); CK(cudaStreamSynchronize(st)); cudaEventElapsedTime(&ms_all, a, b);
    CK(cudaEventRecord(a, st));
    for (int r = 0; r < reps; ++r) {
        cudaMemcpyAsync(k[0], k0, n * 4, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(v[0], v0, n * 4, cudaMemcpyDeviceToDevice, st);
    }
    CK(cudaEventRecord(b, st)); CK(cudaStreamSynchronize(st)); cudaEventElapsedTime(&ms_copy, a, b);
    printf("  whole sort back to back: %.2f us (restore copies %.2f us subtracted)\n", (ms_all - ms_copy) * 1e3 / reps, ms_copy * 1e3 / reps);
#ifdef MSGS_SORT_PHASES
    unsigned long long ph[6] = {0, 0, 0, 0, 0, 0};
    CK(cudaMemcpyToSymbol(g_sort_phase, ph, sizeof(ph)));
    for (int r = 0; r < 20; ++r) sort_once(false);
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpyFromSymbol(ph, g_sort_phase, sizeof(ph)));
    printf("  pass CTA phases (us, mean over %llu CTAs): load+rank %.2f (ticket %.2f, first key ranked %.2f) | scan+look-back %.2f | "
           "scatter+store %.2f\n", ph[3], ph[0] * 1e-3 / ph[3], ph[4] * 1e-3 / ph[3], ph[5] * 1e-3 / ph[3], ph[1] * 1e-3 / ph[3],
           ph[2] * 1e-3 / ph[3]);
    {
        const size_t T = std::min(sort_tiles(n), (size_t)1024);
        std::vector<unsigned long long> tr(T * 4);
        CK(cudaMemcpyFromSymbol(tr.data(), g_sort_trace, T * 4 * sizeof(unsigned long long)));
        unsigned long long lo = ~0ull, hi = 0;
        for (size_t t = 0; t < T; ++t) { lo = std::min(lo, tr[t * 4]); hi = std::max(hi, tr[t * 4 + 3]); }
        printf("  pass 1 of the last sort: first CTA start -> last CTA end %.2f us; per tile (us after the first start) start / ranked / looked back / end\n", (hi - lo) * 1e-3);
        for (size_t t = 0; t < T; t += (T + 11) / 12)
            printf("    tile %3zu: %.2f %.2f %.2f %.2f\n", t, (tr[t * 4] - lo) * 1e-3, (tr[t * 4 + 1] - lo) * 1e-3, (tr[t * 4 + 2] - lo) * 1e-3, (tr[t * 4 + 3] - lo) * 1e-3);
        printf("    tile %3zu: %.2f %.2f %.2f %.2f\n", T - 1, (tr[(T - 1) * 4] - lo) * 1e-3, (tr[(T - 1) * 4 + 1] - lo) * 1e-3, (tr[(T - 1) * 4 + 2] - lo) * 1e-3, (tr[(T - 1) * 4 + 3] - lo) * 1e-3);
    }
#endif
    return bad ? 2 : 0;
}
