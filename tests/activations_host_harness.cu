// Test-only: runs the per-thread bodies of ms-gs_b200/csrc/activations_body.cuh on the HOST, one call per thread index of
// the grid the library would launch (256-thread blocks, overshoot included), so that the index arithmetic of the product
// kernels is checked against the golden vectors in a container without a GPU.  Built by tests/test_zz_activations.py.
#include "../ms-gs_b200/csrc/activations_body.cuh"

extern "C" void host_activate(int P, int D, const float* fdc, const float* frest, const float* sraw, const float* rraw,
                              const float* oraw, float* shs, float* scales, float* rots, float* opac) {
    const long long threads = (msgs::activate_threads(P, D) + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i)
        msgs::activate_thread(i, P, D, fdc, frest, sraw, rraw, oraw, shs, scales, rots, opac);
}

extern "C" void host_activate_backward(int P, int D, const float* scales, const float* rraw, const float* opac,
                                       const float* g_shs, const float* g_scales, const float* g_rots, const float* g_opac,
                                       float* g_fdc, float* g_frest, float* g_sraw, float* g_rraw, float* g_oraw) {
    const long long threads = (msgs::activate_threads(P, D) + 255) / 256 * 256;
    for (long long i = 0; i < threads; ++i)
        msgs::activate_backward_thread(i, P, D, scales, rraw, opac, g_shs, g_scales, g_rots, g_opac, g_fdc, g_frest, g_sraw,
                                       g_rraw, g_oraw);
}
