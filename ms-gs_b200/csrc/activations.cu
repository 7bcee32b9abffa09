// Fused parameter activations (include/msgs_activations.h): the five element-wise passes and the
// (P,16,3) concat that gaussian_renderer.render() runs per call (gaussian_renderer/__init__.py:57-89 through the
// GaussianModel getters, scene/gaussian_model.py:127-155) as ONE launch, and their backward as ONE launch.
// HBM-bound: 448 B per Gaussian forward (224 read + 224 written at D = 15), the same backward.  Layout: the SH rows are
// moved by one thread per 16 bytes of the flat (P, 1 + D, 3) array (the concat shifts the source by 3 floats, so the reads are scalar but
// contiguous across the warp; the stores are 16-byte); the first P threads also do the per-Gaussian scale / rotation /
// opacity work (float4 rotation rows, contiguous 3-float scale rows).
#include "common.cuh"
#include "../../include/msgs_raster.h"
#include "../../include/msgs_activations.h"

namespace msgs {

__global__ void __launch_bounds__(256)
activate_kernel(int P, int D, const float* __restrict__ fdc, const float* __restrict__ frest,
                const float* __restrict__ sraw, const float* __restrict__ rraw, const float* __restrict__ oraw,
                float* __restrict__ shs, float* __restrict__ scales, float* __restrict__ rots, float* __restrict__ opac) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = 3 * (1 + D), rest = 3 * D;
    const long long total = (long long)P * row, e0 = i * 4;
    if (e0 < total) {
        long long p = e0 / row;
        int f = (int)(e0 - p * row);
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = e0 + k >= total ? 0.f : (f < 3 ? __ldg(fdc + p * 3 + f) : __ldg(frest + p * rest + (f - 3)));
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
        for (int k = 0; k < 3; ++k) scales[i * 3 + k] = expf(__ldg(sraw + i * 3 + k));
        const float4 q = __ldg(reinterpret_cast<const float4*>(rraw) + i);
        // torch.nn.functional.normalize: x / max(||x||_2, eps), eps = 1e-12
        const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
        reinterpret_cast<float4*>(rots)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
        opac[i] = 1.0f / (1.0f + expf(-__ldg(oraw + i)));
    }
}

__global__ void __launch_bounds__(256)
activate_backward_kernel(int P, int D, const float* __restrict__ scales, const float* __restrict__ rraw,
                         const float* __restrict__ opac, const float* __restrict__ g_shs,
                         const float* __restrict__ g_scales, const float* __restrict__ g_rots,
                         const float* __restrict__ g_opac, float* __restrict__ g_fdc, float* __restrict__ g_frest,
                         float* __restrict__ g_sraw, float* __restrict__ g_rraw, float* __restrict__ g_oraw) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = 3 * (1 + D), rest = 3 * D;
    const long long total = (long long)P * row, e0 = i * 4;
    if (e0 < total && (g_fdc || g_frest)) {
        long long p = e0 / row;
        int f = (int)(e0 - p * row);
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        if (g_shs) {
            if (e0 + 4 <= total) {
                const float4 g = __ldg(reinterpret_cast<const float4*>(g_shs) + i);
                v[0] = g.x; v[1] = g.y; v[2] = g.z; v[3] = g.w;
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (e0 + k < total) v[k] = __ldg(g_shs + e0 + k);
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
                g_sraw[i * 3 + k] = g_scales ? __ldg(g_scales + i * 3 + k) * __ldg(scales + i * 3 + k) : 0.f;
        }
        if (g_rraw) {
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g_rots) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(rraw) + i);
                const float4 g = __ldg(reinterpret_cast<const float4*>(g_rots) + i);
                const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
                if (n > 1e-12f) {
                    // y = x / n:  dx = (g - y (y . g)) / n
                    const float inv = 1.0f / n;
                    const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
                    const float d = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
                    o = make_float4((g.x - y.x * d) * inv, (g.y - y.y * d) * inv, (g.z - y.z * d) * inv, (g.w - y.w * d) * inv);
                } else {
                    // clamped denominator (constant 1e-12): y = x / eps
                    o = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
                }
            }
            reinterpret_cast<float4*>(g_rraw)[i] = o;
        }
        if (g_oraw) {
            const float s = __ldg(opac + i);
            g_oraw[i] = g_opac ? __ldg(g_opac + i) * s * (1.0f - s) : 0.f;
        }
    }
}

static inline bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace msgs

extern "C" int msgs_activate(int P, int D, const float* features_dc, const float* features_rest, const float* scaling_raw,
                             const float* rotation_raw, const float* opacity_raw, float* shs, float* scales,
                             float* rotations, float* opacities, void* stream) {
    using namespace msgs;
    if (P < 0 || D < 0) return MSGS_ERR_ARG;
    if (P == 0) return MSGS_OK;
    if (!features_dc || (D > 0 && !features_rest) || !scaling_raw || !rotation_raw || !opacity_raw || !shs || !scales ||
        !rotations || !opacities || !aligned16(shs) || !aligned16(rotation_raw) || !aligned16(rotations))
        return MSGS_ERR_ARG;
    long long threads = ((long long)P * 3 * (1 + D) + 3) / 4;      // one per 16 bytes of the SH rows,
    if (threads < P) threads = P;                                   // and at least one per Gaussian
    const unsigned blocks = (unsigned)((threads + 255) / 256);
    activate_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(P, D, features_dc, features_rest, scaling_raw, rotation_raw,
                                                              opacity_raw, shs, scales, rotations, opacities);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

extern "C" int msgs_activate_backward(int P, int D, const float* scales, const float* rotation_raw, const float* opacities,
                                      const float* dL_dshs, const float* dL_dscales, const float* dL_drotations,
                                      const float* dL_dopacities, float* dL_dfeatures_dc, float* dL_dfeatures_rest,
                                      float* dL_dscaling_raw, float* dL_drotation_raw, float* dL_dopacity_raw, void* stream) {
    using namespace msgs;
    if (P < 0 || D < 0) return MSGS_ERR_ARG;
    if (P == 0) return MSGS_OK;
    if ((dL_dscaling_raw && dL_dscales && !scales) || (dL_drotation_raw && dL_drotations && !rotation_raw) ||
        (dL_dopacity_raw && dL_dopacities && !opacities) || !aligned16(dL_dshs) || !aligned16(rotation_raw) ||
        !aligned16(dL_drotations) || !aligned16(dL_drotation_raw))
        return MSGS_ERR_ARG;
    long long threads = ((long long)P * 3 * (1 + D) + 3) / 4;      // one per 16 bytes of the SH rows,
    if (threads < P) threads = P;                                   // and at least one per Gaussian
    const unsigned blocks = (unsigned)((threads + 255) / 256);
    activate_backward_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
        P, D, scales, rotation_raw, opacities, dL_dshs, dL_dscales, dL_drotations, dL_dopacities, dL_dfeatures_dc,
        dL_dfeatures_rest, dL_dscaling_raw, dL_drotation_raw, dL_dopacity_raw);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
