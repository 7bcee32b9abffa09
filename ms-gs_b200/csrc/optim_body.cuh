// Per-element body of the fused Adam step (optim.cu; include/msgs_optim.h).  __host__ __device__: the same code runs on
// the host in tests/test_optim.py against torch.optim.Adam without a GPU.
// Operation order of torch/optim/adam.py (_multi_tensor_adam / _single_tensor_adam, fp32 tensors, capturable=False):
//   exp_avg.lerp_(grad, 1 - beta1)                      -> m + w1 (g - m)          (|w1| < 0.5 branch of lerp)
//   exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2) -> (b2 v) + (w2 g) g
//   denom = exp_avg_sq.sqrt() / sqrt(bc2) + eps
//   param.addcdiv_(exp_avg, denom, value=-(lr / bc1))   -> p + step (m / denom)
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace msgs {

struct AdamConsts { float beta2, w1, w2, eps, bc2_sqrt; };
struct AdamFieldDev { long long begin4, end4; float step, step_alt; int period, alt_count; };   // float4 units; step = -(lr / bc1)
struct AdamFields { int n; AdamFieldDev f[8]; };

__host__ __device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float step, const AdamConsts& c) {
    m = fmaf(c.w1, g - m, m);
    v = fmaf(c.w2 * g, g, c.beta2 * v);
    const float denom = sqrtf(v) / c.bc2_sqrt + c.eps;
    p = fmaf(step, m / denom, p);
}

// the four step sizes of float4 chunk q, or false if q lies in no field (padding between fields: left untouched)
__host__ __device__ __forceinline__ bool adam_steps(long long q, const AdamFields& F, float st[4]) {
    for (int i = 0; i < F.n; ++i) {
        const AdamFieldDev& f = F.f[i];
        if (q >= f.begin4 && q < f.end4) {
            if (f.period <= 0) { st[0] = st[1] = st[2] = st[3] = f.step; return true; }
            int r = (int)(((q - f.begin4) * 4) % f.period);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                st[k] = r < f.alt_count ? f.step_alt : f.step;
                if (++r == f.period) r = 0;
            }
            return true;
        }
    }
    return false;
}

}  // namespace msgs
