// Per-thread bodies of the fused activation kernels (activations.cu).  __host__ __device__ so that the SAME code runs
// on the host in tests/test_zz_activations.py::test_thread_bodies_on_host (a loop over the thread index, checked against
// the golden vectors without a GPU): the index arithmetic of the SH rows is the part a GPU-less container can verify.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "act.cuh"

namespace msgs {

#ifdef __CUDA_ARCH__
#define MSGS_LDG(p) __ldg(p)
#else
#define MSGS_LDG(p) (*(p))
#endif

__host__ __device__ __forceinline__ void
activate_thread(long long i, int P, int D, const float* __restrict__ fdc, const float* __restrict__ frest,
                const float* __restrict__ sraw, const float* __restrict__ rraw, const float* __restrict__ oraw,
                float* __restrict__ shs, float* __restrict__ scales, float* __restrict__ rots, float* __restrict__ opac) {
    const int row = 3 * (1 + D), rest = 3 * D;
    const long long total = (long long)P * row, e0 = i * 4;
    if (e0 < total) {
        long long p = e0 / row;
        int f = (int)(e0 - p * row);
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = e0 + k >= total ? 0.f : (f < 3 ? MSGS_LDG(fdc + p * 3 + f) : MSGS_LDG(frest + p * rest + (f - 3)));
            if (++f == row) { f = 0; ++p; }
        }
        if (e0 + 4 <= total) reinterpret_cast<float4*>(shs)[i] = make_float4(v[0], v[1], v[2], v[3]);
        else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (e0 + k < total) shs[e0 + k] = v[k];
        }
    }
    if (i < P) {
#pragma unroll
        for (int k = 0; k < 3; ++k) scales[i * 3 + k] = act_exp(MSGS_LDG(sraw + i * 3 + k));
        reinterpret_cast<float4*>(rots)[i] = act_normalize(MSGS_LDG(reinterpret_cast<const float4*>(rraw) + i));
        opac[i] = act_sigmoid(MSGS_LDG(oraw + i));
    }
}

__host__ __device__ __forceinline__ void
activate_backward_thread(long long i, int P, int D, const float* __restrict__ scales, const float* __restrict__ rraw,
                         const float* __restrict__ opac, const float* __restrict__ g_shs,
                         const float* __restrict__ g_scales, const float* __restrict__ g_rots,
                         const float* __restrict__ g_opac, float* __restrict__ g_fdc, float* __restrict__ g_frest,
                         float* __restrict__ g_sraw, float* __restrict__ g_rraw, float* __restrict__ g_oraw) {
    const int row = 3 * (1 + D), rest = 3 * D;
    const long long total = (long long)P * row, e0 = i * 4;
    if (e0 < total && (g_fdc || g_frest)) {
        long long p = e0 / row;
        int f = (int)(e0 - p * row);
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        if (g_shs) {
            if (e0 + 4 <= total) {
                const float4 g = MSGS_LDG(reinterpret_cast<const float4*>(g_shs) + i);
                v[0] = g.x; v[1] = g.y; v[2] = g.z; v[3] = g.w;
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (e0 + k < total) v[k] = MSGS_LDG(g_shs + e0 + k);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (e0 + k < total) {
                if (f < 3) { if (g_fdc) g_fdc[p * 3 + f] = v[k]; }
                else if (g_frest) g_frest[p * rest + (f - 3)] = v[k];
            }
            if (++f == row) { f = 0; ++p; }
        }
    }
    if (i < P) {
        if (g_sraw) {
#pragma unroll
            for (int k = 0; k < 3; ++k)          // d exp(x) = exp(x) = the forward output
                g_sraw[i * 3 + k] = g_scales ? MSGS_LDG(g_scales + i * 3 + k) * MSGS_LDG(scales + i * 3 + k) : 0.f;
        }
        if (g_rraw) {
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g_rots) {
                o = act_normalize_bwd(MSGS_LDG(reinterpret_cast<const float4*>(rraw) + i),
                                      MSGS_LDG(reinterpret_cast<const float4*>(g_rots) + i));
            }
            reinterpret_cast<float4*>(g_rraw)[i] = o;
        }
        if (g_oraw) {
            g_oraw[i] = g_opac ? act_sigmoid_bwd(MSGS_LDG(opac + i), MSGS_LDG(g_opac + i)) : 0.f;
        }
    }
}

// threads one launch needs: one per 16 bytes of the SH rows, and at least one per Gaussian
__host__ __device__ inline long long activate_threads(int P, int D) {
    long long threads = ((long long)P * 3 * (1 + D) + 3) / 4;
    return threads < P ? (long long)P : threads;
}

}  // namespace msgs
