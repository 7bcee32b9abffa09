// Fused Adam step over the flat parameter buffer, and the sharded step's reduce-scatter + Adam + all-gather as ONE kernel
// over NVSwitch multicast memory (include/msgs_optim.h; reference: scene/gaussian_model.py:229-252, train.py:416-418).
// HBM-bound streaming kernels: 28 B per element (g, p, m, v read; p, m, v written), 16-byte accesses, every element once.
#include "common.cuh"
#include "optim_body.cuh"
#include "../../include/msgs_raster.h"
#include "../../include/msgs_optim.h"
#include <cmath>

namespace msgs {

__device__ __forceinline__ float4 mc_ld_reduce_add(const float4* addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float4* addr, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVLS = false: params / grads are this GPU's buffers.  NVLS = true: grads and params_out are multicast addresses (the
// gradient load is the switch's sum over the ranks, the parameter store lands in every replica), params_in is the local
// replica; [begin4, end4) is this rank's slice and m / v hold only that slice (state_base4 = begin4).
template <bool NVLS>
__global__ void __launch_bounds__(512)
adam_kernel(float4* __restrict__ params_out, const float4* __restrict__ grads, const float4* __restrict__ params_in,
            float4* __restrict__ exp_avg, float4* __restrict__ exp_avg_sq, long long begin4, long long end4,
            long long state_base4, const __grid_constant__ AdamFields F, AdamConsts C) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long q = begin4 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (NVLS) {
        // The sharded kernel is bound by the switch round trip of multimem.ld_reduce (a few microseconds): 4 independent
        // reductions in flight per thread, 512-thread CTAs.  Measured on 2 and 8 B200 (profiles/r02o_*): 1 in flight
        // 0.73 ms -> 4 in flight 0.60 ms per 283 MB step at N = 8; 128-thread CTAs with 16 in flight each (the same number
        // of requests, a quarter of the warps — tried so that the kernel could run under the preprocess backward without
        // taking its thread slots) are 2x SLOWER: the switch path wants many warps, not many requests per warp.
        constexpr int FLY = 4;
        for (; q + (FLY - 1) * stride < end4; q += FLY * stride) {
            float4 g[FLY];
#pragma unroll
            for (int k = 0; k < FLY; ++k) g[k] = mc_ld_reduce_add(grads + q + k * stride);
#pragma unroll
            for (int k = 0; k < FLY; ++k) {
                const long long e = q + k * stride;
                float st[4];
                if (!adam_steps(e, F, st)) continue;
                float4 p = params_in[e];
                float4 m = exp_avg[e - state_base4], v = exp_avg_sq[e - state_base4];
                adam_update(p.x, g[k].x, m.x, v.x, st[0], C);
                adam_update(p.y, g[k].y, m.y, v.y, st[1], C);
                adam_update(p.z, g[k].z, m.z, v.z, st[2], C);
                adam_update(p.w, g[k].w, m.w, v.w, st[3], C);
                exp_avg[e - state_base4] = m;
                exp_avg_sq[e - state_base4] = v;
                mc_st(params_out + e, p);
            }
        }
    }
    for (; q < end4; q += stride) {
        float st[4];
        if (!adam_steps(q, F, st)) continue;
        const float4 g = NVLS ? mc_ld_reduce_add(grads + q) : grads[q];
        float4 p = params_in[q];
        float4 m = exp_avg[q - state_base4], v = exp_avg_sq[q - state_base4];
        adam_update(p.x, g.x, m.x, v.x, st[0], C);
        adam_update(p.y, g.y, m.y, v.y, st[1], C);
        adam_update(p.z, g.z, m.z, v.z, st[2], C);
        adam_update(p.w, g.w, m.w, v.w, st[3], C);
        exp_avg[q - state_base4] = m;
        exp_avg_sq[q - state_base4] = v;
        if (NVLS) mc_st(params_out + q, p);
        else params_out[q] = p;
    }
    if (NVLS) __threadfence_system();
}

// two-shot all-reduce of typed lanes: [0, a4) f32 add (16-byte operations), [a4, b4) u32 max, [b4, c4) u32 min (word by word:
// multimem.ld_reduce takes no vector qualifier for integer types); rank r owns the r-th slice of the float4 index space
__global__ void __launch_bounds__(512)
stats_allreduce_kernel(float4* __restrict__ mc, long long begin4, long long end4, long long a4, long long b4) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q = begin4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; q < end4; q += stride) {
        if (q < a4) {
            mc_st(mc + q, mc_ld_reduce_add(mc + q));
        } else {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(mc + q);
            uint32_t r[4];
            if (q < b4) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    asm volatile("multimem.ld_reduce.relaxed.sys.global.max.u32 %0, [%1];" : "=r"(r[k]) : "l"(w + k) : "memory");
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    asm volatile("multimem.ld_reduce.relaxed.sys.global.min.u32 %0, [%1];" : "=r"(r[k]) : "l"(w + k) : "memory");
            }
            mc_st(mc + q, make_float4(__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])));
        }
    }
    __threadfence_system();
}

}  // namespace msgs

using namespace msgs;

extern "C" {

int msgs_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n_floats,
                   const msgs_adam_field* fields, int n_fields, double beta1, double beta2, double eps, int step, void* stream) {
    AdamFields F; AdamConsts C;
    const int rc = adam_build(fields, n_fields, n_floats, beta1, beta2, eps, step, F, C);
    if (rc != MSGS_OK) return rc;
    if (n_floats == 0) return MSGS_OK;
    if (!params || !grads || !exp_avg || !exp_avg_sq) return MSGS_ERR_ARG;
    if ((((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15)) return MSGS_ERR_ARG;
    const long long n4 = n_floats / 4;
    long long blocks = (n4 + 511) / 512;
    if (blocks > 148 * 8) blocks = 148 * 8;           // grid-stride: a few resident CTAs per SM keep HBM saturated
    note_launches(1);
    adam_kernel<false><<<(unsigned)blocks, 512, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(params), reinterpret_cast<const float4*>(grads), reinterpret_cast<const float4*>(params),
        reinterpret_cast<float4*>(exp_avg), reinterpret_cast<float4*>(exp_avg_sq), 0, n4, 0, F, C);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

long long msgs_adam_shard_floats(long long n_floats, int world_size) {
    if (world_size < 1 || n_floats < 0) return 0;
    const long long n4 = n_floats / 4;
    return 4 * ((n4 + world_size - 1) / world_size);
}
long long msgs_adam_shard_begin(long long n_floats, int rank, int world_size) {
    const long long per = msgs_adam_shard_floats(n_floats, world_size);
    const long long b = per * rank;
    return b < n_floats ? b : n_floats;
}

int msgs_adam_step_multimem(void* params_mc, const void* grads_mc, const float* params_local, float* exp_avg,
                            float* exp_avg_sq, long long n_floats, const msgs_adam_field* fields, int n_fields,
                            double beta1, double beta2, double eps, int step, int rank, int world_size, int ctas, void* stream) {
    AdamFields F; AdamConsts C;
    const int rc = adam_build(fields, n_fields, n_floats, beta1, beta2, eps, step, F, C);
    if (rc != MSGS_OK) return rc;
    if (world_size < 1 || rank < 0 || rank >= world_size) return MSGS_ERR_ARG;
    if (!params_mc || !grads_mc || !params_local || !exp_avg || !exp_avg_sq) return MSGS_ERR_ARG;
    if ((((uintptr_t)params_mc | (uintptr_t)grads_mc | (uintptr_t)params_local | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15))
        return MSGS_ERR_ARG;
    const long long begin = msgs_adam_shard_begin(n_floats, rank, world_size);
    long long end = begin + msgs_adam_shard_floats(n_floats, world_size);
    if (end > n_floats) end = n_floats;
    if (begin >= end) return MSGS_OK;
    if (ctas <= 0) ctas = 148;          // NVLink-bound, like msgs_allreduce_multimem
    note_launches(1);
    adam_kernel<true><<<ctas, 512, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(params_mc), reinterpret_cast<const float4*>(grads_mc),
        reinterpret_cast<const float4*>(params_local), reinterpret_cast<float4*>(exp_avg),
        reinterpret_cast<float4*>(exp_avg_sq), begin / 4, end / 4, begin / 4, F, C);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

/* the same over ONE range [range_begin, range_end) of the flat buffer (multiples of 4): rank r owns the r-th of world_size
 * equal parts of the RANGE; exp_avg / exp_avg_sq point at this rank's state of that part (element 0 = its first element) */
int msgs_adam_step_multimem_range(void* params_mc, const void* grads_mc, const float* params_local, float* exp_avg,
                                  float* exp_avg_sq, long long n_floats, const msgs_adam_field* fields, int n_fields,
                                  double beta1, double beta2, double eps, int step, long long range_begin, long long range_end,
                                  int rank, int world_size, int ctas, void* stream) {
    AdamFields F; AdamConsts C;
    const int rc = adam_build(fields, n_fields, n_floats, beta1, beta2, eps, step, F, C);
    if (rc != MSGS_OK) return rc;
    if (world_size < 1 || rank < 0 || rank >= world_size) return MSGS_ERR_ARG;
    if (range_begin < 0 || range_end < range_begin || range_end > n_floats || ((range_begin | range_end) & 3)) return MSGS_ERR_ARG;
    if (!params_mc || !grads_mc || !params_local || !exp_avg || !exp_avg_sq) return MSGS_ERR_ARG;
    if ((((uintptr_t)params_mc | (uintptr_t)grads_mc | (uintptr_t)params_local | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15))
        return MSGS_ERR_ARG;
    const long long len = range_end - range_begin;
    const long long begin = range_begin + msgs_adam_shard_begin(len, rank, world_size);
    long long end = begin + msgs_adam_shard_floats(len, world_size);
    if (end > range_end) end = range_end;
    if (begin >= end) return MSGS_OK;
    if (ctas <= 0) ctas = 148;
    note_launches(1);
    adam_kernel<true><<<ctas, 512, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(params_mc), reinterpret_cast<const float4*>(grads_mc),
        reinterpret_cast<const float4*>(params_local), reinterpret_cast<float4*>(exp_avg),
        reinterpret_cast<float4*>(exp_avg_sq), begin / 4, end / 4, begin / 4, F, C);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

int msgs_stats_allreduce_multimem(void* stats_mc, long long n_sum, long long n_max, long long n_min, int rank,
                                  int world_size, int ctas, void* stream) {
    if (!stats_mc || ((uintptr_t)stats_mc & 15) || n_sum < 0 || n_max < 0 || n_min < 0 || ((n_sum | n_max | n_min) & 3) ||
        world_size < 1 || rank < 0 || rank >= world_size)
        return MSGS_ERR_ARG;
    const long long a4 = n_sum / 4, b4 = a4 + n_max / 4, c4 = b4 + n_min / 4;
    const long long per = (c4 + world_size - 1) / world_size;
    const long long begin = per * rank, end = begin + per < c4 ? begin + per : c4;
    if (begin >= end) return MSGS_OK;
    if (ctas <= 0) ctas = 148;
    note_launches(1);
    stats_allreduce_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(stats_mc), begin, end, a4, b4);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}

}  // extern "C"
