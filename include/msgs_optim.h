/* msgs_optim.h — C ABI of the fused optimiser step of the MS-GS training loop (device half) and of
 * the sharded step's exchange that is fused with it (SURVEY.md §8 row e).
 *
 * Reference semantics: scene/gaussian_model.py:229-252 builds `torch.optim.Adam(l, lr=0.0, eps=1e-15)` with one param
 * group per tensor (xyz, f_dc, f_rest, opacity, scaling, rotation; occ_multiplier / dc_delta have lr 0) and train.py:416-418
 * calls `optimizer.step(); optimizer.zero_grad(set_to_none=True)` once per iteration: default betas (0.9, 0.999), no weight
 * decay, no amsgrad, i.e. per element (torch/optim/adam.py, _multi_tensor_adam order of operations)
 *     m = m + (1 - b1) (g - m);   v = b2 v + (1 - b2) g g;
 *     p = p - (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
 * In the reference that is ~9 foreach launches over 8 tensors and 28 B/element of traffic spread over them; here ONE launch
 * over ONE flat buffer that holds every trainable tensor field by field (the layout of the flat gradient buffer the
 * rasterizer's backward writes: ms-gs_b200/msgs_optim.py FlatLayout), each element read and written once.
 *
 * Sharded step (one process per GPU, Gaussians replicated): msgs_adam_step_multimem is reduce-scatter + Adam + all-gather
 * as ONE kernel over NVSwitch multicast memory — rank r, for every 16 bytes of ITS 1/N slice: one multimem.ld_reduce of
 * the gradient (the switch sums the N ranks' copies), the Adam update with the rank's shard of the optimiser state (m, v
 * exist only for the owned slice), one multimem.st of the updated parameters into ALL ranks' replicas.  Per GPU the
 * optimiser traffic drops to 1/N and no separate collective runs.  The caller brackets the launch with stream-ordered
 * barriers across the ranks (before: every rank's gradients are complete; after: every rank's parameter stores landed).
 *
 * Device pointers, fp32, `stream` is a cudaStream_t.  Exported by libmsgs_raster.so.  MSGS_OK or MSGS_ERR_* (msgs_raster.h).
 * No CPU fallback.
 */
#ifndef MSGS_OPTIM_H
#define MSGS_OPTIM_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One param group inside the flat buffer: elements [begin, end) (multiples of 4 floats, fields do not overlap).
 * period > 0: the field is made of rows of `period` floats whose first `alt_count` floats use lr_alt instead of lr
 * (the (P,16,3) SH tensor: period 48, alt_count 3 -> f_dc with feature_lr, f_rest with feature_lr / 20).  period == 0: lr only. */
typedef struct {
    long long begin, end;
    double lr, lr_alt;          /* doubles, as torch keeps them: lr / (1 - b1^t) is formed in double and then narrowed */
    int period, alt_count;
} msgs_adam_field;

#define MSGS_ADAM_MAX_FIELDS 8

/* Single-GPU step.  params / grads / exp_avg / exp_avg_sq: n_floats each (flat layout); `step` = t >= 1 (this call's step
 * count, as torch's state['step'] AFTER its increment).  Elements outside every field are left untouched. */
int msgs_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n_floats,
                   const msgs_adam_field* fields, int n_fields, double beta1, double beta2, double eps, int step, void* stream);

/* Sharded step.  params_mc / grads_mc: the flat parameter / gradient buffers' addresses INSIDE their multicast mappings
 * (symmetric buffers, same layout on every rank); params_local: this rank's replica (plain address); exp_avg / exp_avg_sq:
 * this rank's SHARD of the state: msgs_adam_shard_floats(n_floats, world_size) floats each, element i of the shard is flat
 * element msgs_adam_shard_begin(n_floats, rank, world_size) + i.  ctas <= 0: one CTA per SM. */
int msgs_adam_step_multimem(void* params_mc, const void* grads_mc, const float* params_local, float* exp_avg,
                            float* exp_avg_sq, long long n_floats, const msgs_adam_field* fields, int n_fields,
                            double beta1, double beta2, double eps, int step, int rank, int world_size, int ctas, void* stream);
/* The same over ONE range [range_begin, range_end) of the flat buffer (multiples of 4 floats): rank r owns the r-th of
 * world_size equal parts of the RANGE (msgs_adam_shard_floats / _begin applied to the range's length); exp_avg / exp_avg_sq
 * point at this rank's state of that part.  A step issued as several ranges (the SH rows of successive slices of the
 * Gaussians, then the rest) lets the exchange of one slice run while the backward still computes the next one. */
int msgs_adam_step_multimem_range(void* params_mc, const void* grads_mc, const float* params_local, float* exp_avg,
                                  float* exp_avg_sq, long long n_floats, const msgs_adam_field* fields, int n_fields,
                                  double beta1, double beta2, double eps, int step, long long range_begin, long long range_end,
                                  int rank, int world_size, int ctas, void* stream);
long long msgs_adam_shard_floats(long long n_floats, int world_size);
long long msgs_adam_shard_begin(long long n_floats, int rank, int world_size);

/* Typed-lane all-reduce of the per-step statistics deltas (sums for the gradient-norm accumulators and
 * visibility counts, max for max_radii2D and the observed pixel sizes, min for the observed pixel sizes) over the
 * multicast mapping of a symmetric buffer, two-shot like msgs_allreduce_multimem, ONE launch for all three lane groups:
 * words [0, n_sum) fp32 add, [n_sum, n_sum + n_max) u32 max, [.., + n_min) u32 min (non-negative floats order like their
 * bit patterns).  Every count is a multiple of 4 words. */
int msgs_stats_allreduce_multimem(void* stats_mc, long long n_sum, long long n_max, long long n_min, int rank,
                                  int world_size, int ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif
