// Parameter activations of GaussianModel (scene/gaussian_model.py:37-47, 127-155), shared by the stand-alone activation
// kernels (activations.cu) and the raw-parameter mode of the preprocess kernels (msgs_*_raw entry points: the activations
// run in registers inside preprocess_fwd and their backward inside preprocess_bwd), so that both paths give identical bits.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace msgs {

__host__ __device__ __forceinline__ float act_exp(float x) { return expf(x); }                           // scaling_activation
__host__ __device__ __forceinline__ float act_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }     // opacity_activation
// torch.nn.functional.normalize: x / max(||x||_2, eps), eps = 1e-12
__host__ __device__ __forceinline__ float act_norm4(const float4 q) {
    return sqrtf(fmaf(q.w, q.w, fmaf(q.z, q.z, fmaf(q.y, q.y, q.x * q.x))));
}
__host__ __device__ __forceinline__ float4 act_normalize(const float4 q) {
    const float n = fmaxf(act_norm4(q), 1e-12f);
    return make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
}
// backward of y = normalize(x) for the upstream gradient g:  dx = (g - y (y . g)) / n;  clamped denominator: g / eps
__host__ __device__ __forceinline__ float4 act_normalize_bwd(const float4 q, const float4 g) {
    const float n = act_norm4(q);
    if (n > 1e-12f) {
        const float inv = 1.0f / n;
        const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
        const float d = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
        return make_float4((g.x - y.x * d) * inv, (g.y - y.y * d) * inv, (g.z - y.z * d) * inv, (g.w - y.w * d) * inv);
    }
    return make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
}
__host__ __device__ __forceinline__ float act_sigmoid_bwd(float s, float g) { return g * s * (1.0f - s); }   // s = sigmoid(x)

}  // namespace msgs
