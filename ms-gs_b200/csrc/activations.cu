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
#include "activations_body.cuh"

namespace msgs {

__global__ void __launch_bounds__(256)
activate_kernel(int P, int D, const float* __restrict__ fdc, const float* __restrict__ frest,
                const float* __restrict__ sraw, const float* __restrict__ rraw, const float* __restrict__ oraw,
                float* __restrict__ shs, float* __restrict__ scales, float* __restrict__ rots, float* __restrict__ opac) {
    activate_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, D, fdc, frest, sraw, rraw, oraw, shs, scales, rots, opac);
}

__global__ void __launch_bounds__(256)
activate_backward_kernel(int P, int D, const float* __restrict__ scales, const float* __restrict__ rraw,
                         const float* __restrict__ opac, const float* __restrict__ g_shs,
                         const float* __restrict__ g_scales, const float* __restrict__ g_rots,
                         const float* __restrict__ g_opac, float* __restrict__ g_fdc, float* __restrict__ g_frest,
                         float* __restrict__ g_sraw, float* __restrict__ g_rraw, float* __restrict__ g_oraw) {
    activate_backward_thread((long long)blockIdx.x * blockDim.x + threadIdx.x, P, D, scales, rraw, opac, g_shs, g_scales,
                             g_rots, g_opac, g_fdc, g_frest, g_sraw, g_rraw, g_oraw);
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
    const long long threads = activate_threads(P, D);
    const unsigned blocks = (unsigned)((threads + 255) / 256);
    note_launches(1);
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
    const long long threads = activate_threads(P, D);
    const unsigned blocks = (unsigned)((threads + 255) / 256);
    note_launches(1);
    activate_backward_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
        P, D, scales, rotation_raw, opacities, dL_dshs, dL_dscales, dL_drotations, dL_dopacities, dL_dfeatures_dc,
        dL_dfeatures_rest, dL_dscaling_raw, dL_drotation_raw, dL_dopacity_raw);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
