// Test-only: the per-element body of ms-gs_b200/csrc/optim_body.cuh (the fused Adam step) run on the HOST over every
// float4 chunk of a flat buffer.  Built by tests/test_optim.py.
#include "../ms-gs_b200/csrc/optim_body.cuh"

extern "C" int host_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n_floats,
                              const msgs_adam_field* fields, int n_fields, double beta1, double beta2, double eps, int step) {
    msgs::AdamFields F; msgs::AdamConsts C;
    const int rc = msgs::adam_build(fields, n_fields, n_floats, beta1, beta2, eps, step, F, C);
    if (rc != 0) return rc;
    for (long long q = 0; q < n_floats / 4; ++q) {
        float st[4];
        if (!msgs::adam_steps(q, F, st)) continue;
        for (int k = 0; k < 4; ++k)
            msgs::adam_update(params[4 * q + k], grads[4 * q + k], exp_avg[4 * q + k], exp_avg_sq[4 * q + k], st[k], C);
    }
    return 0;
}
