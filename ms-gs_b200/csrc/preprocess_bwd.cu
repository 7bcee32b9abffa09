// Backward per-Gaussian preprocess: ONE kernel for what the reference runs as computeCov2DCUDA
// (backward.cu:144-277) followed by preprocessCUDA<3> (backward.cu:349-469), and it also replaces the
// binding's 11 torch::zeros fills (rasterize_points.cu:184-194, 364 B per Gaussian): every gradient element
// is written exactly once here, zeros included.  HBM-bound (dL_dsh: 192 B per Gaussian dominates).
//
// STAGED variant (SH path with M = 16, the training configuration): the 192-byte SH rows of a warp's 32 Gaussians are
// copied global -> shared with coalesced 16-byte cp.async chunks while the covariance backward runs; every lane then
// turns its row IN PLACE into its dL_dsh row (zeros for Gaussians that get no SH gradient), and the warp streams the
// 32 rows out with coalesced 16-byte stores.  The generic variant (precomputed colours, other M) keeps per-thread rows.
#include "common.cuh"
#include "math.cuh"
#include "act.cuh"

namespace msgs {

constexpr int PB_THREADS = 128;

// red.global.add.v4.f32: one 16-byte reduction into L2 (no read-back).  The accumulating backward uses reductions, not
// load-add-store, so that the backward passes of several (view, scale) units may run CONCURRENTLY on different streams
// into the same accumulators (msgs_sharding.UnitStreams).
__device__ __forceinline__ void red_add4(float4* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ACC = 1 (msgs_backward_accumulate): the five gradients that training sums over the (view, scale) units of a step —
// dL_dmean3D, dL_dsh, dL_dopacity, dL_dscale, dL_drot — are ADDED to what the output arrays hold, and only for the
// Gaussians this unit rendered (the others contribute zero, so they cost no traffic: at the coarse scales of a
// multi-scale step that is almost all of them).  dL_dmean2D (per-view densification statistic) is still written for every
// Gaussian; the remaining outputs are optional (NULL = skipped) and written, not accumulated.
template <bool STAGED, int ACC, bool RAW>
__global__ void __launch_bounds__(PB_THREADS, STAGED ? 8 : 4)
preprocess_bwd_kernel(BwdParams prm, const float* __restrict__ means3D, const float* __restrict__ shs,
                      const float* __restrict__ scales, const float* __restrict__ rotations,
                      const float* __restrict__ cov3D_precomp, const int* __restrict__ radii,
                      const float* __restrict__ pixel_sizes, const float* __restrict__ max_px,
                      const float* __restrict__ min_px, const uint8_t* __restrict__ base_mask, GeomState g,
                      float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic_out,
                      float* __restrict__ dL_dopacity, float* __restrict__ dL_docc, float* __restrict__ dL_ddc,
                      float* __restrict__ dL_dcolor, float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D,
                      float* __restrict__ dL_dsh, float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
    __shared__ __align__(16) float s_rows[STAGED ? (PB_THREADS / 32) * 32 * SH_ROW_STRIDE : 4];
    const int idx = prm.first + blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = idx < prm.P;
    if (!STAGED && !valid) return;
    const int M = prm.M;
    const bool rendered = valid && radii[idx] > 0;
    float* const wrows = s_rows + (STAGED ? (threadIdx.x >> 5) * 32 * SH_ROW_STRIDE : 0);   // this warp's 32 rows
    float* const myrow = wrows + (STAGED ? lane * SH_ROW_STRIDE : 0);
    const size_t warp_g0 = (size_t)(idx - lane);                                            // first Gaussian of the warp
    if (STAGED) {
        // start the SH copy of the rendered Gaussians' rows; it lands while the covariance backward runs
        const unsigned rmask = __ballot_sync(0xFFFFFFFFu, rendered);
        stage_sh_rows(wrows, shs + warp_g0 * SH_ROW_FLOATS, rmask, lane);
    }

    float o_mean2D[2] = {0.f, 0.f};
    float o_conic[3] = {0.f, 0.f, 0.f};
    float o_opacity = 0.f;
    float o_color[3] = {0.f, 0.f, 0.f};
    float o_mean[3] = {0.f, 0.f, 0.f};
    float o_cov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float o_scale[3] = {0.f, 0.f, 0.f};
    float o_rot[4] = {0.f, 0.f, 0.f, 0.f};
    bool sh_written = false;
    bool do_sh = false;            // rendered and not filtered by the backward's own test (quirk Q2)
    float3 mean = {0.f, 0.f, 0.f};

    if (rendered) {
        // sums accumulated by the blend backward
        const float4* ga = reinterpret_cast<const float4*>(g.gacc) + 3 * (size_t)idx;
        const float4 g0 = ga[0], g1 = ga[1], g2 = ga[2];
        o_color[0] = g0.x; o_color[1] = g0.y; o_color[2] = g0.z;
        o_mean2D[0] = g0.w; o_mean2D[1] = g1.x;
        o_conic[0] = g1.y; o_conic[1] = g1.z; o_conic[2] = g1.w;
        o_opacity = g2.x;

        const float* view_matrix = prm.view;
        mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        const float* cov3D = (cov3D_precomp != nullptr ? cov3D_precomp : g.cov3D) + 6 * (size_t)idx;

        // ---------------- computeCov2DCUDA (backward.cu:155-276) ----------------
        {
            const float h_x = prm.focal_x, h_y = prm.focal_y;
            const float3 dL_dconic = {o_conic[0], o_conic[1], o_conic[2]};
            float3 t = xform4x3(mean, view_matrix);
            const float limx = 1.3f * prm.tan_fovx;
            const float limy = 1.3f * prm.tan_fovy;
            const float txtz = t.x / t.z;
            const float tytz = t.y / t.z;
            t.x = min(limx, max(-limx, txtz)) * t.z;
            t.y = min(limy, max(-limy, tytz)) * t.z;
            const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
            const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;

            const m3 J = make_m3(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z),
                                 0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z),
                                 0, 0, 0);
            const m3 W = make_m3(view_matrix[0], view_matrix[4], view_matrix[8],
                                 view_matrix[1], view_matrix[5], view_matrix[9],
                                 view_matrix[2], view_matrix[6], view_matrix[10]);
            const m3 Vrk = make_m3(cov3D[0], cov3D[1], cov3D[2],
                                   cov3D[1], cov3D[3], cov3D[4],
                                   cov3D[2], cov3D[4], cov3D[5]);
            const m3 T = W * J;
            m3 cov2D = transpose(T) * transpose(Vrk) * T;
            const float a = cov2D.c[0][0] += 0.3f;
            const float b = cov2D.c[0][1];
            const float c = cov2D.c[1][1] += 0.3f;

            const float denom = a * c - b * b;
            float dL_da = 0, dL_db = 0, dL_dc = 0;
            const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
            if (denom2inv != 0) {
                dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
                dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
                dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
                o_cov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
                o_cov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
                o_cov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
                o_cov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
                o_cov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
                o_cov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
            }
            const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                                  (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
            const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                                  (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
            const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                                  (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
            const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                                  (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
            const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                                  (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
            const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                                  (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

            const float dL_dJ00 = W.c[0][0] * dL_dT00 + W.c[0][1] * dL_dT01 + W.c[0][2] * dL_dT02;
            const float dL_dJ02 = W.c[2][0] * dL_dT00 + W.c[2][1] * dL_dT01 + W.c[2][2] * dL_dT02;
            const float dL_dJ11 = W.c[1][0] * dL_dT10 + W.c[1][1] * dL_dT11 + W.c[1][2] * dL_dT12;
            const float dL_dJ12 = W.c[2][0] * dL_dT10 + W.c[2][1] * dL_dT11 + W.c[2][2] * dL_dT12;

            const float tz = 1.f / t.z;
            const float tz2 = tz * tz;
            const float tz3 = tz2 * tz;
            const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
            const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
            const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
            const float3 dm = xformVec4x3T({dL_dtx, dL_dty, dL_dtz}, view_matrix);
            o_mean[0] = dm.x; o_mean[1] = dm.y; o_mean[2] = dm.z;
        }

        // ---------------- preprocessCUDA backward (backward.cu:378-468) ----------------
        // independent small / large re-check (quirk Q2)
        const float pixel_size = pixel_sizes[idx];
        float rel_min_pixel_size = 1.0f;
        const float mn = min_px[idx];
        if (mn > 0) rel_min_pixel_size = pixel_size / mn;
        float rel_max_pixel_size = 1.0f;
        const float mx = max_px[idx];
        if (mx > 0) rel_max_pixel_size = pixel_size / mx;
        const bool base = base_mask[idx] != 0;
        bool filtered = false;
        if (prm.filter_small && rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) filtered = true;
        if (prm.filter_large && rel_max_pixel_size > 2.0f && !base) filtered = true;

        if (!filtered) {
            const float* proj = prm.proj;
            const float3 m = mean;
            const float4 m_hom = xform4x4(m, proj);
            const float m_w = 1.0f / (m_hom.w + 0.0000001f);
            const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
            const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
            v3 dL_dmean;
            dL_dmean.x = (proj[0] * m_w - proj[3] * mul1) * o_mean2D[0] + (proj[1] * m_w - proj[3] * mul2) * o_mean2D[1];
            dL_dmean.y = (proj[4] * m_w - proj[7] * mul1) * o_mean2D[0] + (proj[5] * m_w - proj[7] * mul2) * o_mean2D[1];
            dL_dmean.z = (proj[8] * m_w - proj[11] * mul1) * o_mean2D[0] + (proj[9] * m_w - proj[11] * mul2) * o_mean2D[1];
            o_mean[0] += dL_dmean.x; o_mean[1] += dL_dmean.y; o_mean[2] += dL_dmean.z;

            do_sh = true;
            if (!STAGED && shs != nullptr) {
                // SH backward (backward.cu:20-139)
                const v3 pos{mean.x, mean.y, mean.z};
                const v3 campos{prm.campos[0], prm.campos[1], prm.campos[2]};
                const v3 dir_orig = pos - campos;
                const v3 dir = dir_orig / length(dir_orig);
                const float* shf = shs + (size_t)idx * M * 3;
                auto SH = [&](int k) { return v3{shf[3 * k], shf[3 * k + 1], shf[3 * k + 2]}; };
                const uint8_t cl = g.clamped[idx];
                v3 dL_dRGB{o_color[0], o_color[1], o_color[2]};
                dL_dRGB.x *= (cl & 1) ? 0 : 1;
                dL_dRGB.y *= (cl & 2) ? 0 : 1;
                dL_dRGB.z *= (cl & 4) ? 0 : 1;
                v3 dRGBdx{0, 0, 0}, dRGBdy{0, 0, 0}, dRGBdz{0, 0, 0};
                const float x = dir.x, y = dir.y, z = dir.z;
                float* out = dL_dsh + (size_t)idx * M * 3;
                int written = 0;
                auto PUT = [&](int k, float w) {
                    const v3 r = w * dL_dRGB;
                    if (ACC) { atomicAdd(out + 3 * k, r.x); atomicAdd(out + 3 * k + 1, r.y); atomicAdd(out + 3 * k + 2, r.z); }
                    else { out[3 * k] = r.x; out[3 * k + 1] = r.y; out[3 * k + 2] = r.z; }
                    written = k + 1;
                };
                PUT(0, kSH0);
                if (prm.D > 0) {
                    PUT(1, -kSH1 * y);
                    PUT(2, kSH1 * z);
                    PUT(3, -kSH1 * x);
                    dRGBdx = -kSH1 * SH(3);
                    dRGBdy = -kSH1 * SH(1);
                    dRGBdz = kSH1 * SH(2);
                    if (prm.D > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z;
                        const float xy = x * y, yz = y * z, xz = x * z;
                        PUT(4, kSH2[0] * xy);
                        PUT(5, kSH2[1] * yz);
                        PUT(6, kSH2[2] * (2.f * zz - xx - yy));
                        PUT(7, kSH2[3] * xz);
                        PUT(8, kSH2[4] * (xx - yy));
                        dRGBdx += kSH2[0] * y * SH(4) + kSH2[2] * 2.f * -x * SH(6) + kSH2[3] * z * SH(7) + kSH2[4] * 2.f * x * SH(8);
                        dRGBdy += kSH2[0] * x * SH(4) + kSH2[1] * z * SH(5) + kSH2[2] * 2.f * -y * SH(6) + kSH2[4] * 2.f * -y * SH(8);
                        dRGBdz += kSH2[1] * y * SH(5) + kSH2[2] * 2.f * 2.f * z * SH(6) + kSH2[3] * x * SH(7);
                        if (prm.D > 2) {
                            PUT(9, kSH3[0] * y * (3.f * xx - yy));
                            PUT(10, kSH3[1] * xy * z);
                            PUT(11, kSH3[2] * y * (4.f * zz - xx - yy));
                            PUT(12, kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                            PUT(13, kSH3[4] * x * (4.f * zz - xx - yy));
                            PUT(14, kSH3[5] * z * (xx - yy));
                            PUT(15, kSH3[6] * x * (xx - 3.f * yy));
                            dRGBdx += (kSH3[0] * SH(9) * 3.f * 2.f * xy + kSH3[1] * SH(10) * yz + kSH3[2] * SH(11) * -2.f * xy +
                                       kSH3[3] * SH(12) * -3.f * 2.f * xz + kSH3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) +
                                       kSH3[5] * SH(14) * 2.f * xz + kSH3[6] * SH(15) * 3.f * (xx - yy));
                            dRGBdy += (kSH3[0] * SH(9) * 3.f * (xx - yy) + kSH3[1] * SH(10) * xz +
                                       kSH3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + kSH3[3] * SH(12) * -3.f * 2.f * yz +
                                       kSH3[4] * SH(13) * -2.f * xy + kSH3[5] * SH(14) * -2.f * yz +
                                       kSH3[6] * SH(15) * -3.f * 2.f * xy);
                            dRGBdz += (kSH3[1] * SH(10) * xy + kSH3[2] * SH(11) * 4.f * 2.f * yz +
                                       kSH3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + kSH3[4] * SH(13) * 4.f * 2.f * xz +
                                       kSH3[5] * SH(14) * (xx - yy));
                        }
                    }
                }
                if (!ACC) for (int k = written; k < M; ++k) { out[3 * k] = 0.f; out[3 * k + 1] = 0.f; out[3 * k + 2] = 0.f; }
                sh_written = true;

                const v3 dL_ddir{dot(dRGBdx, dL_dRGB), dot(dRGBdy, dL_dRGB), dot(dRGBdz, dL_dRGB)};
                // dnormvdv (auxiliary.h:107-117)
                const float3 v = {dir_orig.x, dir_orig.y, dir_orig.z};
                const float3 dv = {dL_ddir.x, dL_ddir.y, dL_ddir.z};
                const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
                const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
                float3 dn;
                dn.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
                dn.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
                dn.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
                o_mean[0] += dn.x; o_mean[1] += dn.y; o_mean[2] += dn.z;
            }

            if (scales != nullptr) {
                // cov3D backward (backward.cu:281-344); no quaternion-normalisation backward (quirk Q3)
                const float4 q_raw = reinterpret_cast<const float4*>(rotations)[idx];
                v3 s_act{scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
                if (RAW) s_act = v3{act_exp(s_act.x), act_exp(s_act.y), act_exp(s_act.z)};
                const float4 q = RAW ? act_normalize(q_raw) : q_raw;
                const float r = q.x, x = q.y, y = q.z, z = q.w;
                const m3 R = make_m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                                     2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                                     2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
                m3 S = make_m3(1.0f, 0.f, 0.f, 0.f, 1.0f, 0.f, 0.f, 0.f, 1.0f);
                const v3 s = prm.scale_modifier * s_act;
                S.c[0][0] = s.x;
                S.c[1][1] = s.y;
                S.c[2][2] = s.z;
                const m3 Mm = S * R;
                const m3 dL_dSigma = make_m3(o_cov[0], 0.5f * o_cov[1], 0.5f * o_cov[2],
                                             0.5f * o_cov[1], o_cov[3], 0.5f * o_cov[4],
                                             0.5f * o_cov[2], 0.5f * o_cov[4], o_cov[5]);
                const m3 dL_dM = 2.0f * Mm * dL_dSigma;
                const m3 Rt = transpose(R);
                m3 dL_dMt = transpose(dL_dM);
                o_scale[0] = dot(col(Rt, 0), col(dL_dMt, 0));
                o_scale[1] = dot(col(Rt, 1), col(dL_dMt, 1));
                o_scale[2] = dot(col(Rt, 2), col(dL_dMt, 2));
#pragma unroll
                for (int k = 0; k < 3; ++k) { dL_dMt.c[0][k] *= s.x; dL_dMt.c[1][k] *= s.y; dL_dMt.c[2][k] *= s.z; }
                o_rot[0] = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
                o_rot[1] = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
                o_rot[2] = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
                o_rot[3] = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
                if (RAW) {       // chain through exp / normalize: gradients of the RAW scaling / rotation
                    o_scale[0] *= s_act.x; o_scale[1] *= s_act.y; o_scale[2] *= s_act.z;
                    const float4 gq = act_normalize_bwd(q_raw, make_float4(o_rot[0], o_rot[1], o_rot[2], o_rot[3]));
                    o_rot[0] = gq.x; o_rot[1] = gq.y; o_rot[2] = gq.z; o_rot[3] = gq.w;
                }
            }
        }
        if (RAW) o_opacity = act_sigmoid_bwd(act_sigmoid(prm.opacity_raw[idx]), o_opacity);
    }

    if (STAGED) {
        stage_wait_all();        // uniform: every lane's copies have landed before any lane reads a row
        __syncwarp();
    }
    if (STAGED && do_sh) {
        // SH backward (backward.cu:20-139), coefficient by coefficient on the staged row, in place
        // SH backward (backward.cu:20-139), coefficient by coefficient on the staged row, in place
        const v3 pos{mean.x, mean.y, mean.z};
        const v3 campos{prm.campos[0], prm.campos[1], prm.campos[2]};
        const v3 dir_orig = pos - campos;
        const ShDir sd = make_shdir(dir_orig / length(dir_orig));
        const uint8_t cl = g.clamped[idx];
        v3 dL_dRGB{o_color[0], o_color[1], o_color[2]};
        dL_dRGB.x *= (cl & 1) ? 0 : 1;
        dL_dRGB.y *= (cl & 2) ? 0 : 1;
        dL_dRGB.z *= (cl & 4) ? 0 : 1;
        v3 dRGBdx{0, 0, 0}, dRGBdy{0, 0, 0}, dRGBdz{0, 0, 0};
        const int ncoef = (prm.D + 1) * (prm.D + 1);
        float4* row4 = reinterpret_cast<float4*>(myrow);
#pragma unroll
        for (int grp = 0; grp < 4; ++grp) {
            float4 q[3] = {row4[3 * grp], row4[3 * grp + 1], row4[3 * grp + 2]};
            float* f = reinterpret_cast<float*>(q);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int k = 4 * grp + kk;
                const v3 shk{f[3 * kk], f[3 * kk + 1], f[3 * kk + 2]};
                v3 outk{0.f, 0.f, 0.f};
                if (k < ncoef) {
                    outk = sh_weight(k, sd) * dL_dRGB;
                    const v3 gw = sh_weight_grad(k, sd);
                    dRGBdx += gw.x * shk;
                    dRGBdy += gw.y * shk;
                    dRGBdz += gw.z * shk;
                }
                f[3 * kk] = outk.x; f[3 * kk + 1] = outk.y; f[3 * kk + 2] = outk.z;
            }
            row4[3 * grp] = q[0]; row4[3 * grp + 1] = q[1]; row4[3 * grp + 2] = q[2];
        }
        sh_written = true;
        const v3 dL_ddir{dot(dRGBdx, dL_dRGB), dot(dRGBdy, dL_dRGB), dot(dRGBdz, dL_dRGB)};
        // dnormvdv (auxiliary.h:107-117)
        const float3 v = {dir_orig.x, dir_orig.y, dir_orig.z};
        const float3 dv = {dL_ddir.x, dL_ddir.y, dL_ddir.z};
        const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        float3 dn;
        dn.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
        dn.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
        dn.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
        o_mean[0] += dn.x; o_mean[1] += dn.y; o_mean[2] += dn.z;
    }

    // ---- write every output element exactly once (ACC: add the training gradients of the rendered Gaussians) ----
    if (STAGED) {
        const unsigned wmask = __ballot_sync(0xFFFFFFFFu, sh_written);
        // rows without an SH gradient become zeros; then the warp streams its 32 rows out, coalesced
        if (!ACC && !sh_written) {
            float4* row4 = reinterpret_cast<float4*>(myrow);
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 12; ++k) row4[k] = z;
        }
        __syncwarp();
        const size_t rows_valid = (size_t)prm.P > warp_g0 ? (size_t)prm.P - warp_g0 : 0;     // rows of this warp inside P
        float4* gout = reinterpret_cast<float4*>(dL_dsh + warp_g0 * SH_ROW_FLOATS);
        if (!ACC || wmask) {
#pragma unroll
            for (int it = 0; it < 12; ++it) {
                const int c = it * 32 + lane;
                const int row = (c * 2731) >> 15;        // c / 12
                const int col = c - row * 12;
                if ((size_t)row < rows_valid) {
                    const float4 v = *reinterpret_cast<const float4*>(wrows + row * SH_ROW_STRIDE + col * 4);
                    if (!ACC) gout[c] = v;
                    else if ((wmask >> row) & 1u) {
                        red_add4(gout + c, v);
                    }
                }
            }
        }
        if (dL_ddc) {   // (P, 12) zeros, coalesced
            float4* o4 = reinterpret_cast<float4*>(dL_ddc) + 3 * warp_g0;
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int it = 0; it < 3; ++it) {
                const int c = it * 32 + lane;
                if ((size_t)((c * 21846) >> 16) < rows_valid) o4[c] = z;      // c / 3 for c < 96
            }
        }
        if (!valid) return;
    }
    if (dL_dmean2D) {
        dL_dmean2D[3 * (size_t)idx + 0] = o_mean2D[0];
        dL_dmean2D[3 * (size_t)idx + 1] = o_mean2D[1];
        dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
    }
    if (dL_dconic_out) {
        reinterpret_cast<float4*>(dL_dconic_out)[idx] = make_float4(o_conic[0], o_conic[1], 0.f, o_conic[2]);
    }
    if (dL_dcolor) {
        dL_dcolor[3 * (size_t)idx + 0] = o_color[0];
        dL_dcolor[3 * (size_t)idx + 1] = o_color[1];
        dL_dcolor[3 * (size_t)idx + 2] = o_color[2];
    }
    if (dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; ++k) dL_dcov3D[6 * (size_t)idx + k] = o_cov[k];
    }
    if (ACC) {
        if (rendered) {
            atomicAdd(dL_dopacity + idx, o_opacity);
#pragma unroll
            for (int k = 0; k < 3; ++k) atomicAdd(dL_dmean3D + 3 * (size_t)idx + k, o_mean[k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) atomicAdd(dL_dscale + 3 * (size_t)idx + k, o_scale[k]);
            red_add4(reinterpret_cast<float4*>(dL_drot) + idx, make_float4(o_rot[0], o_rot[1], o_rot[2], o_rot[3]));
        }
    } else {
        dL_dopacity[idx] = o_opacity;
        dL_dmean3D[3 * (size_t)idx + 0] = o_mean[0];
        dL_dmean3D[3 * (size_t)idx + 1] = o_mean[1];
        dL_dmean3D[3 * (size_t)idx + 2] = o_mean[2];
#pragma unroll
        for (int k = 0; k < 3; ++k) dL_dscale[3 * (size_t)idx + k] = o_scale[k];
        reinterpret_cast<float4*>(dL_drot)[idx] = make_float4(o_rot[0], o_rot[1], o_rot[2], o_rot[3]);
    }
    if (!ACC && !STAGED && !sh_written && M > 0 && dL_dsh != nullptr) {
        float* out = dL_dsh + (size_t)idx * M * 3;
        if (M == 16) {
            float4* o4 = reinterpret_cast<float4*>(out);
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 12; ++k) o4[k] = z;
        } else {
            for (int k = 0; k < 3 * M; ++k) out[k] = 0.f;
        }
    }
    if (rendered) {
        // accumulators consumed: back to zero, so that a second backward over the same state starts clean (common.cuh).
        // Last, behind every load of the thread: stores into the state buffer may alias the kernel's other inputs as far as
        // the compiler knows, and issued early they pin all later loads behind them (measured: 0.35 ms instead of 0.15 ms).
        float4* gz = reinterpret_cast<float4*>(g.gacc) + 3 * (size_t)idx;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        gz[0] = z4; gz[1] = z4; gz[2] = z4;
    }
    if (dL_docc) reinterpret_cast<float4*>(dL_docc)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!STAGED && dL_ddc) {
        float4* o4 = reinterpret_cast<float4*>(dL_ddc) + 3 * (size_t)idx;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        o4[0] = z; o4[1] = z; o4[2] = z;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// View group backward (msgs_backward_group_accumulate): the preprocess backward of the S units of ONE view (same camera,
// different image sizes) in one pass over the Gaussians.  Per unit only the EWA-covariance backward depends on the
// resolution (focal length); everything behind it — projection backward, SH backward (192-byte row read + 192-byte
// gradient row), cov3D backward — is LINEAR in the per-unit sums the blend backward accumulated, so those sums are added
// over the units first and the linear maps run once: one SH row read and one gradient row reduction per Gaussian and
// view instead of one per unit.  Equals the sum of the units' msgs_backward_accumulate results up to fp32 summation
// order (which the per-unit reductions do not fix either).  Per-unit outputs stay per unit: dL_dmean2D (the per-view,
// per-level densification statistic) is written for every Gaussian of every unit, and every unit's blend-backward
// accumulators are re-zeroed.  SH path with M = 16 and scale/rotation or precomputed covariance inputs.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cov2d_bwd(const float3 mean, const float* __restrict__ cov3D, float h_x, float h_y,
                                          float tan_fovx, float tan_fovy, const float* __restrict__ view_matrix,
                                          const float3 dL_dconic, float* o_cov, float* o_mean) {
    // computeCov2DCUDA (backward.cu:155-276), as in preprocess_bwd_kernel
    float3 t = xform4x3(mean, view_matrix);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;
    const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
    const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
    const m3 J = make_m3(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z),
                         0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z),
                         0, 0, 0);
    const m3 W = make_m3(view_matrix[0], view_matrix[4], view_matrix[8],
                         view_matrix[1], view_matrix[5], view_matrix[9],
                         view_matrix[2], view_matrix[6], view_matrix[10]);
    const m3 Vrk = make_m3(cov3D[0], cov3D[1], cov3D[2],
                           cov3D[1], cov3D[3], cov3D[4],
                           cov3D[2], cov3D[4], cov3D[5]);
    const m3 T = W * J;
    m3 cov2D = transpose(T) * transpose(Vrk) * T;
    const float a = cov2D.c[0][0] += 0.3f;
    const float b = cov2D.c[0][1];
    const float c = cov2D.c[1][1] += 0.3f;
    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
#pragma unroll
    for (int k = 0; k < 6; ++k) o_cov[k] = 0.f;
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
        dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
        dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
        o_cov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        o_cov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        o_cov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        o_cov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        o_cov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        o_cov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
    }
    const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dJ00 = W.c[0][0] * dL_dT00 + W.c[0][1] * dL_dT01 + W.c[0][2] * dL_dT02;
    const float dL_dJ02 = W.c[2][0] * dL_dT00 + W.c[2][1] * dL_dT01 + W.c[2][2] * dL_dT02;
    const float dL_dJ11 = W.c[1][0] * dL_dT10 + W.c[1][1] * dL_dT11 + W.c[1][2] * dL_dT12;
    const float dL_dJ12 = W.c[2][0] * dL_dT10 + W.c[2][1] * dL_dT11 + W.c[2][2] * dL_dT12;
    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    const float3 dm = xformVec4x3T({dL_dtx, dL_dty, dL_dtz}, view_matrix);
    o_mean[0] = dm.x; o_mean[1] = dm.y; o_mean[2] = dm.z;
}

__global__ void __launch_bounds__(PB_THREADS, 4)
preprocess_bwd_group_kernel(BwdParams prm, const __grid_constant__ GroupBwdParams gp, const float* __restrict__ means3D,
                            const float* __restrict__ shs, const float* __restrict__ scales,
                            const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                            const float* __restrict__ max_px, const float* __restrict__ min_px,
                            const uint8_t* __restrict__ base_mask, float* __restrict__ dL_dopacity,
                            float* __restrict__ dL_dmean3D, float* __restrict__ dL_dsh, float* __restrict__ dL_dscale,
                            float* __restrict__ dL_drot) {
    __shared__ __align__(16) float s_rows[(PB_THREADS / 32) * 32 * SH_ROW_STRIDE];
    const int idx = prm.first + blockIdx.x * blockDim.x + threadIdx.x;      // Gaussians [first, P): a slice of the pass
    const int lane = threadIdx.x & 31;
    const bool valid = idx < prm.P;
    const int S = gp.S;
    uint32_t rendered_at = 0;
    if (valid) {
#pragma unroll
        for (int s = 0; s < MAX_GROUP; ++s)
            if (s < S && gp.sc[s].radii[idx] > 0) rendered_at |= 1u << s;
    }
    const bool rendered = rendered_at != 0u;
    float* const wrows = s_rows + (threadIdx.x >> 5) * 32 * SH_ROW_STRIDE;
    float* const myrow = wrows + lane * SH_ROW_STRIDE;
    const size_t warp_g0 = (size_t)(idx - lane);
    const unsigned rmask = __ballot_sync(0xFFFFFFFFu, rendered);
    stage_sh_rows(wrows, shs + warp_g0 * SH_ROW_FLOATS, rmask, lane);

    float m2x[MAX_GROUP], m2y[MAX_GROUP];        // per-unit dL_dmean2D (static indexing only: stays in registers)
#pragma unroll
    for (int s = 0; s < MAX_GROUP; ++s) { m2x[s] = 0.f; m2y[s] = 0.f; }
    float sum_color[3] = {0.f, 0.f, 0.f}, sum_m2d[2] = {0.f, 0.f}, sum_cov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float o_mean[3] = {0.f, 0.f, 0.f}, o_scale[3] = {0.f, 0.f, 0.f}, o_rot[4] = {0.f, 0.f, 0.f, 0.f};
    float o_opacity = 0.f;
    bool do_sh = false;                          // some unit rendered the Gaussian and did not filter it (quirk Q2)
    float3 mean = {0.f, 0.f, 0.f};
    uint8_t cl = 0;

    if (rendered) {
        // the view matrix in registers (the loop's stores... there are none, but loads through gp pointers are not
        // restrict-qualified either; a local copy keeps the 12 values out of the per-scale dependency chains)
        float vm[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) vm[k] = prm.view[k];
        const int s0 = __ffs(rendered_at) - 1;   // every rendering unit holds the same cov3D / clamp bits
        mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        float c3[6];
        {
            const float* src = (cov3D_precomp != nullptr ? cov3D_precomp : gp.sc[s0].cov3D) + 6 * (size_t)idx;
#pragma unroll
            for (int k = 0; k < 6; ++k) c3[k] = src[k];
        }
        cl = gp.sc[s0].clamped[idx];
        const float mn = min_px[idx], mx = max_px[idx];
        const bool base = base_mask[idx] != 0;
        // software pipeline over the rendering units: the next unit's sums (48 B) and pixel size are in flight while
        // this unit's covariance backward runs
        uint32_t todo = rendered_at;
        float4 n0, n1, n2;
        float npx;
        {
            const GroupBwdScale& u = gp.sc[s0];
            const float4* ga = reinterpret_cast<const float4*>(u.gacc) + 3 * (size_t)idx;
            n0 = ga[0]; n1 = ga[1]; n2 = ga[2];
            npx = u.pixel_sizes[idx];
        }
#pragma unroll 1
        while (todo) {
            const int s = __ffs(todo) - 1;
            todo &= todo - 1;
            const GroupBwdScale& u = gp.sc[s];
            const float4 g0 = n0, g1 = n1, g2 = n2;
            const float pixel_size = npx;
            if (todo) {
                const GroupBwdScale& un = gp.sc[__ffs(todo) - 1];
                const float4* ga = reinterpret_cast<const float4*>(un.gacc) + 3 * (size_t)idx;
                n0 = ga[0]; n1 = ga[1]; n2 = ga[2];
                npx = un.pixel_sizes[idx];
            }
            float oc[6], om[3];
            cov2d_bwd(mean, c3, u.focal_x, u.focal_y, prm.tan_fovx, prm.tan_fovy, vm, make_float3(g1.y, g1.z, g1.w), oc, om);
            o_mean[0] += om[0]; o_mean[1] += om[1]; o_mean[2] += om[2];
            o_opacity += g2.x;
#pragma unroll
            for (int k = 0; k < MAX_GROUP; ++k) if (k == s) { m2x[k] = g0.w; m2y[k] = g1.x; }
            // independent small / large re-check (quirk Q2, backward.cu:399-412)
            float rel_min_pixel_size = 1.0f;
            if (mn > 0) rel_min_pixel_size = pixel_size / mn;
            float rel_max_pixel_size = 1.0f;
            if (mx > 0) rel_max_pixel_size = pixel_size / mx;
            bool filtered = false;
            if (prm.filter_small && rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) filtered = true;
            if (prm.filter_large && rel_max_pixel_size > 2.0f && !base) filtered = true;
            if (!filtered) {
                do_sh = true;
                sum_color[0] += g0.x; sum_color[1] += g0.y; sum_color[2] += g0.z;
                sum_m2d[0] += g0.w; sum_m2d[1] += g1.x;
#pragma unroll
                for (int k = 0; k < 6; ++k) sum_cov[k] += oc[k];
            }
        }
        if (do_sh) {
            // projection backward (backward.cu:435-452), on the summed screen-space gradient
            const float* proj = prm.proj;
            const float3 m = mean;
            const float4 m_hom = xform4x4(m, proj);
            const float m_w = 1.0f / (m_hom.w + 0.0000001f);
            const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
            const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
            o_mean[0] += (proj[0] * m_w - proj[3] * mul1) * sum_m2d[0] + (proj[1] * m_w - proj[3] * mul2) * sum_m2d[1];
            o_mean[1] += (proj[4] * m_w - proj[7] * mul1) * sum_m2d[0] + (proj[5] * m_w - proj[7] * mul2) * sum_m2d[1];
            o_mean[2] += (proj[8] * m_w - proj[11] * mul1) * sum_m2d[0] + (proj[9] * m_w - proj[11] * mul2) * sum_m2d[1];
            if (scales != nullptr) {
                // cov3D backward (backward.cu:281-344) on the summed dL_dcov3D; no quaternion-normalisation backward (Q3)
                const float4 q_raw = reinterpret_cast<const float4*>(rotations)[idx];
                v3 s_act{scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
                if (prm.raw) s_act = v3{act_exp(s_act.x), act_exp(s_act.y), act_exp(s_act.z)};
                const float4 q = prm.raw ? act_normalize(q_raw) : q_raw;
                const float r = q.x, x = q.y, y = q.z, z = q.w;
                const m3 R = make_m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                                     2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                                     2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
                m3 Sm = make_m3(1.0f, 0.f, 0.f, 0.f, 1.0f, 0.f, 0.f, 0.f, 1.0f);
                const v3 sv = prm.scale_modifier * s_act;
                Sm.c[0][0] = sv.x;
                Sm.c[1][1] = sv.y;
                Sm.c[2][2] = sv.z;
                const m3 Mm = Sm * R;
                const m3 dL_dSigma = make_m3(sum_cov[0], 0.5f * sum_cov[1], 0.5f * sum_cov[2],
                                             0.5f * sum_cov[1], sum_cov[3], 0.5f * sum_cov[4],
                                             0.5f * sum_cov[2], 0.5f * sum_cov[4], sum_cov[5]);
                const m3 dL_dM = 2.0f * Mm * dL_dSigma;
                const m3 Rt = transpose(R);
                m3 dL_dMt = transpose(dL_dM);
                o_scale[0] = dot(col(Rt, 0), col(dL_dMt, 0));
                o_scale[1] = dot(col(Rt, 1), col(dL_dMt, 1));
                o_scale[2] = dot(col(Rt, 2), col(dL_dMt, 2));
#pragma unroll
                for (int k = 0; k < 3; ++k) { dL_dMt.c[0][k] *= sv.x; dL_dMt.c[1][k] *= sv.y; dL_dMt.c[2][k] *= sv.z; }
                o_rot[0] = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
                o_rot[1] = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
                o_rot[2] = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
                o_rot[3] = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
                if (prm.raw) {       // chain through exp / normalize: gradients of the RAW scaling / rotation
                    o_scale[0] *= s_act.x; o_scale[1] *= s_act.y; o_scale[2] *= s_act.z;
                    const float4 gq = act_normalize_bwd(q_raw, make_float4(o_rot[0], o_rot[1], o_rot[2], o_rot[3]));
                    o_rot[0] = gq.x; o_rot[1] = gq.y; o_rot[2] = gq.z; o_rot[3] = gq.w;
                }
            }
        }
        if (prm.raw) o_opacity = act_sigmoid_bwd(act_sigmoid(prm.opacity_raw[idx]), o_opacity);
    }

    stage_wait_all();
    __syncwarp();
    if (do_sh) {
        // SH backward (backward.cu:20-139) on the summed colour gradient, in place on the staged row
        const v3 pos{mean.x, mean.y, mean.z};
        const v3 campos{prm.campos[0], prm.campos[1], prm.campos[2]};
        const v3 dir_orig = pos - campos;
        const ShDir sd = make_shdir(dir_orig / length(dir_orig));
        v3 dL_dRGB{sum_color[0], sum_color[1], sum_color[2]};
        dL_dRGB.x *= (cl & 1) ? 0 : 1;
        dL_dRGB.y *= (cl & 2) ? 0 : 1;
        dL_dRGB.z *= (cl & 4) ? 0 : 1;
        v3 dRGBdx{0, 0, 0}, dRGBdy{0, 0, 0}, dRGBdz{0, 0, 0};
        const int ncoef = (prm.D + 1) * (prm.D + 1);
        float4* row4 = reinterpret_cast<float4*>(myrow);
#pragma unroll
        for (int grp = 0; grp < 4; ++grp) {
            float4 q[3] = {row4[3 * grp], row4[3 * grp + 1], row4[3 * grp + 2]};
            float* f = reinterpret_cast<float*>(q);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int k = 4 * grp + kk;
                const v3 shk{f[3 * kk], f[3 * kk + 1], f[3 * kk + 2]};
                v3 outk{0.f, 0.f, 0.f};
                if (k < ncoef) {
                    outk = sh_weight(k, sd) * dL_dRGB;
                    const v3 gw = sh_weight_grad(k, sd);
                    dRGBdx += gw.x * shk;
                    dRGBdy += gw.y * shk;
                    dRGBdz += gw.z * shk;
                }
                f[3 * kk] = outk.x; f[3 * kk + 1] = outk.y; f[3 * kk + 2] = outk.z;
            }
            row4[3 * grp] = q[0]; row4[3 * grp + 1] = q[1]; row4[3 * grp + 2] = q[2];
        }
        const v3 dL_ddir{dot(dRGBdx, dL_dRGB), dot(dRGBdy, dL_dRGB), dot(dRGBdz, dL_dRGB)};
        // dnormvdv (auxiliary.h:107-117)
        const float3 v = {dir_orig.x, dir_orig.y, dir_orig.z};
        const float3 dv = {dL_ddir.x, dL_ddir.y, dL_ddir.z};
        const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        o_mean[0] += ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
        o_mean[1] += (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
        o_mean[2] += (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    }
    const unsigned wmask = __ballot_sync(0xFFFFFFFFu, do_sh);
    __syncwarp();
    if (wmask) {
        const size_t rows_valid = (size_t)prm.P > warp_g0 ? (size_t)prm.P - warp_g0 : 0;
        float4* gout = reinterpret_cast<float4*>(dL_dsh + warp_g0 * SH_ROW_FLOATS);
#pragma unroll
        for (int it = 0; it < 12; ++it) {
            const int c = it * 32 + lane;
            const int row = (c * 2731) >> 15;        // c / 12
            const int col = c - row * 12;
            if ((size_t)row < rows_valid && ((wmask >> row) & 1u))
                red_add4(gout + c, *reinterpret_cast<const float4*>(wrows + row * SH_ROW_STRIDE + col * 4));
        }
    }
    if (!valid) return;
    if (rendered) {
        atomicAdd(dL_dopacity + idx, o_opacity);
#pragma unroll
        for (int k = 0; k < 3; ++k) atomicAdd(dL_dmean3D + 3 * (size_t)idx + k, o_mean[k]);
#pragma unroll
        for (int k = 0; k < 3; ++k) atomicAdd(dL_dscale + 3 * (size_t)idx + k, o_scale[k]);
        red_add4(reinterpret_cast<float4*>(dL_drot) + idx, make_float4(o_rot[0], o_rot[1], o_rot[2], o_rot[3]));
    }
    // per unit: dL_dmean2D of every Gaussian; accumulators consumed -> back to zero (last: see preprocess_bwd_kernel)
#pragma unroll
    for (int s = 0; s < MAX_GROUP; ++s) {
        if (s >= S) break;
        float* d2 = gp.sc[s].dL_dmean2D;
        if (d2) { d2[3 * (size_t)idx] = m2x[s]; d2[3 * (size_t)idx + 1] = m2y[s]; d2[3 * (size_t)idx + 2] = 0.f; }
        if ((rendered_at >> s) & 1u) {
            float4* gz = reinterpret_cast<float4*>(gp.sc[s].gacc) + 3 * (size_t)idx;
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
            gz[0] = z4; gz[1] = z4; gz[2] = z4;
        }
    }
}

void launch_preprocess_bwd_group(const BwdParams& prm, const GroupBwdParams& gp, const float* means3D, const float* shs,
                                 const float* scales, const float* rotations, const float* cov3D_precomp,
                                 const float* max_px, const float* min_px, const uint8_t* base_mask, float* dL_dopacity,
                                 float* dL_dmean3D, float* dL_dsh, float* dL_dscale, float* dL_drot, cudaStream_t st) {
    if (prm.P <= prm.first) return;
    const unsigned grid = (unsigned)((prm.P - prm.first + PB_THREADS - 1) / PB_THREADS);
    preprocess_bwd_group_kernel<<<grid, PB_THREADS, 0, st>>>(prm, gp, means3D, shs, scales, rotations, cov3D_precomp, max_px,
                                                             min_px, base_mask, dL_dopacity, dL_dmean3D, dL_dsh, dL_dscale,
                                                             dL_drot);
}

void launch_preprocess_bwd(const BwdParams& prm, const float* means3D, const float* shs, const float* scales,
                           const float* rotations, const float* cov3D_precomp, const int* radii,
                           const float* pixel_sizes, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, GeomState g,
                           float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_docc, float* dL_ddc,
                           float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                           float* dL_dscale, float* dL_drot, int accumulate, cudaStream_t st) {
    if (prm.P <= prm.first) return;
    const unsigned grid = (unsigned)((prm.P - prm.first + PB_THREADS - 1) / PB_THREADS);
    const bool staged = shs != nullptr && prm.M == 16 && dL_dsh != nullptr;
#define MSGS_PB_LAUNCH(S, A) (prm.raw ? preprocess_bwd_kernel<S, A, true> : preprocess_bwd_kernel<S, A, false>)<<<grid, PB_THREADS, 0, st>>>( \
        prm, means3D, shs, scales, rotations, cov3D_precomp, radii, pixel_sizes, max_px, min_px, base_mask, g, \
        dL_dmean2D, dL_dconic, dL_dopacity, dL_docc, dL_ddc, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot)
    if (staged) {
        if (accumulate) MSGS_PB_LAUNCH(true, 1); else MSGS_PB_LAUNCH(true, 0);
    } else {
        if (accumulate) MSGS_PB_LAUNCH(false, 1); else MSGS_PB_LAUNCH(false, 0);
    }
#undef MSGS_PB_LAUNCH
}

}  // namespace msgs
