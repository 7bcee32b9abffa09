// Per-element body of the fused Adam step (optim.cu; include/msgs_optim.h).  __host__ __device__: the same code runs on
// the host in tests/test_optim.py against torch.optim.Adam without a GPU.
// Operation order of torch/optim/adam.py (_multi_tensor_adam / _single_tensor_adam, fp32 tensors, capturable=False):
//   exp_avg.lerp_(grad, 1 - beta1)                      -> m + w1 (g - m)          (|w1| < 0.5 branch of lerp)
//   exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2) -> (b2 v) + w2 (g g)   (the foreach functor CUDA tensors take:
//                                                          input + scalar * (t1 * t2), ForeachPointwiseOp.cu)
//   denom = exp_avg_sq.sqrt() / sqrt(bc2) + eps
//   param.addcdiv_(exp_avg, denom, value=-(lr / bc1))   -> p + step (m / denom)
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <cmath>
#include "../../include/msgs_optim.h"

namespace msgs {

struct AdamConsts { float beta2, w1, w2, eps, bc2_sqrt; };
struct AdamFieldDev { long long begin4, end4; float step, step_alt; int period, alt_count; };   // float4 units; step = -(lr / bc1)
struct AdamFields { int n; AdamFieldDev f[8]; };

__host__ __device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float step, const AdamConsts& c) {
    m = fmaf(c.w1, g - m, m);
    v = fmaf(c.w2, g * g, c.beta2 * v);
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

// host: doubles -> the fp32 constants the kernels see (torch forms bias corrections and step sizes as Python floats)
inline int adam_build(const msgs_adam_field* fields, int n_fields, long long n_floats, double beta1, double beta2, double eps,
                        int step, AdamFields& F, AdamConsts& C) {
    if (!fields || n_fields < 1 || n_fields > MSGS_ADAM_MAX_FIELDS || step < 1 || n_floats < 0 || (n_floats & 3)) return -1;
    // torch/optim/adam.py: bias corrections and the step size are Python floats (doubles); the kernels see them as fp32
    const double bc1 = 1.0 - std::pow(beta1, (double)step);
    const double bc2 = 1.0 - std::pow(beta2, (double)step);
    C.beta2 = (float)beta2;
    C.w1 = (float)(1.0 - beta1);
    C.w2 = (float)(1.0 - beta2);
    C.eps = (float)eps;
    C.bc2_sqrt = (float)std::sqrt(bc2);
    F.n = n_fields;
    for (int i = 0; i < n_fields; ++i) {
        const msgs_adam_field& f = fields[i];
        if (f.begin < 0 || f.end < f.begin || f.end > n_floats || (f.begin & 3) || (f.end & 3) || f.period < 0 ||
            (f.period > 0 && (f.alt_count < 0 || f.alt_count > f.period)))
            return -1;
        F.f[i].begin4 = f.begin / 4; F.f[i].end4 = f.end / 4;
        F.f[i].step = (float)(-(f.lr / bc1));
        F.f[i].step_alt = (float)(-(f.lr_alt / bc1));
        F.f[i].period = f.period; F.f[i].alt_count = f.alt_count;
    }
    return 0;
}


}  // namespace msgs
