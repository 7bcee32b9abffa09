// Gradient exchange over NVSwitch multicast (NVLS), hand-written: a two-shot all-reduce of a SYMMETRIC buffer (same
// size and offset on every rank, mapped through one multicast object; msgs_sharding.SymmetricGradientExchange owns it).
// Rank r owns the r-th slice of the buffer: for every 16 bytes of its slice it issues ONE multimem.ld_reduce — the
// switch fetches that address from all ranks' copies and returns their fp32 sum — and ONE multimem.st, which the switch
// writes into all ranks' copies.  Per GPU and direction that is (1 + 1/N) x the buffer over NVLink, against 2 (N-1)/N x
// for a ring; no staging buffers, no copy through shared memory.
// Ordering: the caller brackets the launch with stream-ordered barriers across the ranks (before: every rank's gradients
// are complete in its copy; after: every rank's stores have landed) — torch.distributed._symmetric_memory handle.barrier.
// Replaces the NCCL all-reduce of the per-Gaussian gradients that the sharded training step runs before the optimiser
// (SURVEY.md §8e); the reference itself has no multi-GPU path.
#include "common.cuh"
#include "../../include/msgs_raster.h"

namespace msgs {

__global__ void __launch_bounds__(512)
allreduce_multimem_kernel(float4* __restrict__ mc, size_t begin, size_t end) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // 4 independent 16-byte reductions in flight per thread
    for (; i + 3 * stride < end; i += 4 * stride) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v[k].x), "=f"(v[k].y), "=f"(v[k].z), "=f"(v[k].w) : "l"(mc + i + k * stride) : "memory");
#pragma unroll
        for (int k = 0; k < 4; ++k)
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                         ::"l"(mc + i + k * stride), "f"(v[k].x), "f"(v[k].y), "f"(v[k].z), "f"(v[k].w) : "memory");
    }
    for (; i < end; i += stride) {
        float4 v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc + i) : "memory");
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(mc + i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
    __threadfence_system();
}

}  // namespace msgs

extern "C" int msgs_allreduce_multimem(void* multicast_ptr, size_t n_floats, int rank, int world_size, int ctas, void* stream) {
    using namespace msgs;
    if (!multicast_ptr || ((uintptr_t)multicast_ptr & 15) || (n_floats & 3) || world_size < 1 || rank < 0 || rank >= world_size)
        return MSGS_ERR_ARG;
    const size_t n4 = n_floats / 4;
    const size_t per = (n4 + world_size - 1) / world_size;
    const size_t begin = (size_t)rank * per, end = begin + per < n4 ? begin + per : n4;
    if (begin >= end) return MSGS_OK;
    if (ctas <= 0) ctas = 148;          // one CTA per SM: the kernel is NVLink-bound, not SM-bound
    note_launches(1);
    allreduce_multimem_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>((float4*)multicast_ptr, begin, end);
    return cudaPeekAtLastError() == cudaSuccess ? MSGS_OK : MSGS_ERR_CUDA;
}
