// Forward per-Gaussian preprocess (replaces preprocessCUDA<3>, forward.cu:159-390, plus checkFrustum,
// rasterizer_impl.cu:54-66).  HBM-bound: ~53 B read per Gaussian + 192 B of SH per survivor, 28 B + 64 B record
// written.  Everything upstream of radii / tiles_touched / depth / pixel_sizes keeps the reference's expression
// trees (math.cuh) so the integers derived from them are bit-exact with the reference CUDA build.
#include "common.cuh"
#include "math.cuh"
#include "cull.cuh"
#include "act.cuh"
#include <cstdio>

namespace msgs {

// forward.cu:122-156 — quaternion is NOT normalised (quirk Q3)
__device__ __forceinline__ void cov3d_from_scale_rot(const v3 scale, float mod, const float4 rot, float* cov3D) {
    m3 S = make_m3(1.0f, 0.f, 0.f, 0.f, 1.0f, 0.f, 0.f, 0.f, 1.0f);
    S.c[0][0] = mod * scale.x;
    S.c[1][1] = mod * scale.y;
    S.c[2][2] = mod * scale.z;
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    const m3 R = make_m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                         2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                         2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    const m3 M = S * R;
    const m3 Sigma = transpose(M) * M;
    cov3D[0] = Sigma.c[0][0];
    cov3D[1] = Sigma.c[0][1];
    cov3D[2] = Sigma.c[0][2];
    cov3D[3] = Sigma.c[1][1];
    cov3D[4] = Sigma.c[1][2];
    cov3D[5] = Sigma.c[2][2];
}

// forward.cu:78-117 — EWA projection + 0.3 px low-pass
__device__ __forceinline__ float3 cov2d_fwd(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                            float tan_fovy, const float* cov3D, const float* viewmatrix) {
    float3 t = xform4x3(mean, viewmatrix);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;
    const m3 J = make_m3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
                         0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
                         0, 0, 0);
    const m3 W = make_m3(viewmatrix[0], viewmatrix[4], viewmatrix[8],
                         viewmatrix[1], viewmatrix[5], viewmatrix[9],
                         viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    const m3 T = W * J;
    const m3 Vrk = make_m3(cov3D[0], cov3D[1], cov3D[2],
                           cov3D[1], cov3D[3], cov3D[4],
                           cov3D[2], cov3D[4], cov3D[5]);
    m3 cov = transpose(T) * transpose(Vrk) * T;
    cov.c[0][0] += 0.3f;
    cov.c[1][1] += 0.3f;
    return {float(cov.c[0][0]), float(cov.c[0][1]), float(cov.c[1][1])};
}

// forward.cu:21-75 — SH (deg <= 3) -> RGB, +0.5, clamp at 0 with flags.  sh = 16 (M) x 3 floats of this Gaussian.
__device__ __forceinline__ v3 sh_to_rgb(int deg, const v3* sh, v3 pos, v3 campos, uint8_t& clamped_bits) {
    v3 dir = pos - campos;
    dir = dir / length(dir);
    const v3 zero{0.0f, 0.0f, 0.0f};                   // dc_delta_interp == 0 (quirk Q9, forward.cu:343-348)
    v3 result = kSH0 * (sh[0] + zero);
    if (deg > 0) {
        const float x = dir.x, y = dir.y, z = dir.z;
        result = result - kSH1 * y * sh[1] + kSH1 * z * sh[2] - kSH1 * x * sh[3];
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            result = result + kSH2[0] * xy * sh[4] + kSH2[1] * yz * sh[5] + kSH2[2] * (2.0f * zz - xx - yy) * sh[6] +
                     kSH2[3] * xz * sh[7] + kSH2[4] * (xx - yy) * sh[8];
            if (deg > 2) {
                result = result + kSH3[0] * y * (3.0f * xx - yy) * sh[9] + kSH3[1] * xy * z * sh[10] +
                         kSH3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
                         kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                         kSH3[4] * x * (4.0f * zz - xx - yy) * sh[13] + kSH3[5] * z * (xx - yy) * sh[14] +
                         kSH3[6] * x * (xx - 3.0f * yy) * sh[15];
            }
        }
    }
    result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
    clamped_bits = (uint8_t)((result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0));
    return {fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
}

// The MS-GS screen-footprint measure: half-extent (px) of the alpha = 1/255 level set along the tighter
// axis-aligned direction of a conic (forward.cu:249-257 and again :499-505).
__device__ __forceinline__ float level_set_px(float occ, float cx, float cz) {
    float level_set = -2 * logf(1 / (255.0f * occ));
    level_set = max(0.0f, level_set);
    const float dx = sqrtf(level_set / cx);
    const float dy = sqrtf(level_set / cz);
    return min(dx, dy);
}

constexpr int PF_THREADS = 128;

// STAGED (SH path, M = 16): the 192-byte SH rows of the Gaussians that survive every cull are copied global -> shared
// by the whole warp in coalesced 16-byte chunks (math.cuh) and each lane evaluates its colour from shared memory —
// no 48-register row, no 32-lines-per-instruction gathers, and culled Gaussians cost no SH traffic at all.
template <bool STAGED>
__global__ void __launch_bounds__(PF_THREADS, STAGED ? 8 : 4)
preprocess_fwd_kernel(FwdParams prm, const float* __restrict__ means3D, const float* __restrict__ scales,
                      const float* __restrict__ rotations, const float* __restrict__ opacities,
                      const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                      const float* __restrict__ colors_precomp, const float* __restrict__ max_px,
                      const float* __restrict__ min_px, const uint8_t* __restrict__ base_mask,
                      int* __restrict__ radii, float* __restrict__ pixel_sizes, GeomState g) {
    __shared__ __align__(16) float s_rows[STAGED ? (PF_THREADS / 32) * 32 * SH_ROW_STRIDE : 4];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = idx < prm.P;
    bool rendered = false;
    uint32_t depth_key = 0xFFFFFFFFu;     // not rendered: sorts behind every rendered Gaussian
    // forward.cu:203-205 — outputs default to "not rendered"
    int out_radius = 0;
    float out_px = 0.0f;
    uint32_t out_tiles = 0;
    ushort4 out_rect = make_ushort4(0, 0, 0, 0);        // zero area: not rendered
    float3 p_orig = {0.f, 0.f, 0.f};
    float4 rec0 = make_float4(0.f, 0.f, 0.f, 0.f), rec1 = rec0;
    float filter_bits = 0.0f;
    if (valid) {
        do {
            const float* viewmatrix = prm.view;
            const float* projmatrix = prm.proj;
            p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
            // every per-Gaussian input is requested up front: one memory round trip instead of a chain of five dependent
            // ones (position -> scale / rotation -> opacity -> min -> max pixel size); culled Gaussians (14 % at C2) pay
            // 41 bytes they would not have needed
            float occ = opacities[idx];
            const float mn = min_px[idx];
            const float mx = max_px[idx];
            const bool base = base_mask[idx] != 0;
            v3 sc_in{0.f, 0.f, 0.f};
            float4 q_in = make_float4(0.f, 0.f, 0.f, 0.f);
            if (cov3D_precomp == nullptr) {
                sc_in = v3{scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
                q_in = reinterpret_cast<const float4*>(rotations)[idx];
            }
            if (prm.raw) {       // gaussian_renderer/__init__.py:57-89: get_opacity / get_scaling / get_rotation, in registers
                occ = act_sigmoid(occ);
                sc_in = v3{act_exp(sc_in.x), act_exp(sc_in.y), act_exp(sc_in.z)};
                q_in = act_normalize(q_in);
            }
            // auxiliary.h:144-169 in_frustum: near plane only (quirk Q11)
            const float3 p_view = xform4x3(p_orig, viewmatrix);
            if (p_view.z <= 0.2f) {
                if (prm.prefiltered) {
                    printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                    __trap();
                }
                break;
            }
            const float4 p_hom = xform4x4(p_orig, projmatrix);
            const float p_w = 1.0f / (p_hom.w + 0.0000001f);
            const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

            const float* cov3D;
            if (cov3D_precomp != nullptr) {
                cov3D = cov3D_precomp + (size_t)idx * 6;
            } else {
                cov3d_from_scale_rot(sc_in, prm.scale_modifier, q_in, g.cov3D + (size_t)idx * 6);
                cov3D = g.cov3D + (size_t)idx * 6;
            }
            const float3 cov = cov2d_fwd(p_orig, prm.focal_x, prm.focal_y, prm.tan_fovx, prm.tan_fovy, cov3D, viewmatrix);

            // forward.cu:235-239
            const float det = (cov.x * cov.z - cov.y * cov.y);
            if (det == 0.0f) break;
            const float det_inv = 1.f / det;
            const float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};

            // forward.cu:242-246 — conic without the low-pass; the 0.3 literals are double (quirk Q4)
            const float det_ori = ((cov.x - 0.3) * (cov.z - 0.3) - cov.y * cov.y);
            if (det_ori == 0.0f) break;
            const float det_inv_ori = 1.f / det_ori;
            const float3 conic_ori = {(float)((cov.z - 0.3) * det_inv_ori), -cov.y * det_inv_ori,
                                      (float)((cov.x - 0.3) * det_inv_ori)};

            // forward.cu:249-259 — per-Gaussian pixel size, stored BEFORE the filters (quirk Q5)
            float pixel_size = level_set_px(occ, conic_ori.x, conic_ori.z);
            pixel_size /= prm.scale_modifier;
            out_px = pixel_size;

            float rel_min_pixel_size = 1.0f;
            if (mn > 0) rel_min_pixel_size = pixel_size / mn;
            float rel_max_pixel_size = 1.0f;
            if (mx > 0) rel_max_pixel_size = pixel_size / mx;

            // what the backward's independent re-check of the two filters (backward.cu:399-412, 649-660, quirk Q2) will
            // find for this Gaussian, kept in the record for the blend backward
            filter_bits = ((rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) ? 1.0f : 0.0f) +
                          ((rel_max_pixel_size > 2.0f && !base) ? 2.0f : 0.0f);

            // forward.cu:270-289 — filter_large is nested inside filter_small (quirk Q1)
            if (prm.filter_small) {
                if (rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) break;
                if (prm.filter_large) {
                    if (rel_max_pixel_size > 2.0f && !base) break;
                }
            }

            // forward.cu:356-371 — screen extent and tile rectangle
            const float mid = 0.5f * (cov.x + cov.z);
            const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
            const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
            const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
            const float2 point_image = {ndc2pix(p_proj.x, prm.W), ndc2pix(p_proj.y, prm.H)};
            uint2 rect_min, rect_max;
            get_rect(point_image, (int)my_radius, rect_min, rect_max, prm.gx, prm.gy);
            if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) break;

            // forward.cu:384-389 — helper data for the next stages, as one 64-byte splat record.
            // blend_px is the per-instance pixel size the reference blend recomputes per tile from the
            // LOW-PASSED conic, not divided by scale_modifier (quirk Q6, forward.cu:499-506); hoisted here.
            const float blend_px = level_set_px(occ, conic.x, conic.z);
            out_radius = (int)my_radius;
            out_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
            rec0 = make_float4(point_image.x, point_image.y, p_view.z, blend_px);
            rec1 = make_float4(conic.x, conic.y, conic.z, occ);
            // binning inputs (common.cuh): depth key and tile rectangle
            depth_key = __float_as_uint(p_view.z);
            rendered = true;
            out_rect = make_ushort4((unsigned short)rect_min.x, (unsigned short)rect_min.y,
                                    (unsigned short)rect_max.x, (unsigned short)rect_max.y);
        } while (0);
    }

    // forward.cu:375-381 — colour of the rendered Gaussians
    v3 rgb{0.f, 0.f, 0.f};
    uint8_t cl = 0;
    if (STAGED) {
        float* const wrows = s_rows + (threadIdx.x >> 5) * 32 * SH_ROW_STRIDE;
        const unsigned rmask = __ballot_sync(0xFFFFFFFFu, rendered);
        if (rmask) {
            stage_sh_rows(wrows, shs + (size_t)(idx - lane) * SH_ROW_FLOATS, rmask, lane);
            stage_wait_all();
        }
        __syncwarp();
        if (rendered) {
            const v3 cam{prm.campos[0], prm.campos[1], prm.campos[2]};
            v3 dir = v3{p_orig.x, p_orig.y, p_orig.z} - cam;
            dir = dir / length(dir);
            const ShDir sd = make_shdir(dir);
            const int ncoef = (prm.D + 1) * (prm.D + 1);
            const float4* row4 = reinterpret_cast<const float4*>(wrows + lane * SH_ROW_STRIDE);
            v3 result{0.f, 0.f, 0.f};
#pragma unroll
            for (int grp = 0; grp < 4; ++grp) {
                if (4 * grp < ncoef) {
                    const float4 q[3] = {row4[3 * grp], row4[3 * grp + 1], row4[3 * grp + 2]};
                    const float* f = reinterpret_cast<const float*>(q);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const int k = 4 * grp + kk;
                        const v3 shk{f[3 * kk], f[3 * kk + 1], f[3 * kk + 2]};
                        if (k == 0) result = kSH0 * shk;                        // dc_delta_interp == 0 (quirk Q9)
                        else if (k < ncoef) result = result + sh_weight(k, sd) * shk;
                    }
                }
            }
            result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
            cl = (uint8_t)((result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0));
            rgb = {fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
        }
    } else if (rendered) {
        if (colors_precomp == nullptr) {
            v3 sh[16];
            const float* f = shs + (size_t)idx * prm.M * 3;
            const int need = (prm.D + 1) * (prm.D + 1);
#pragma unroll
            for (int k = 0; k < 16; ++k)
                sh[k] = (k < need && k < prm.M) ? v3{f[3 * k], f[3 * k + 1], f[3 * k + 2]} : v3{0.f, 0.f, 0.f};
            const v3 cam{prm.campos[0], prm.campos[1], prm.campos[2]};
            rgb = sh_to_rgb(prm.D, sh, v3{p_orig.x, p_orig.y, p_orig.z}, cam, cl);
        } else {
            rgb = {colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]};
        }
    }
    if (rendered) {
        float4* rec = g.rec + REC_F4 * (size_t)idx;
        rec[0] = rec0;
        rec[1] = rec1;
        rec[2] = make_float4(rgb.x, rgb.y, rgb.z, filter_bits);
        rec[3] = splat_footprint(rec1.x, rec1.y, rec1.z, rec1.w);
        float4* ga = reinterpret_cast<float4*>(g.gacc) + 3 * (size_t)idx;      // blend-backward accumulators start at zero
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        ga[0] = z4; ga[1] = z4; ga[2] = z4;
        g.clamped[idx] = cl;
    }
    if (valid) {
        radii[idx] = out_radius;
        pixel_sizes[idx] = out_px;
        g.tiles_touched[idx] = out_tiles;
        g.rect[idx] = out_rect;
        g.dkeys[0][idx] = depth_key;          // the sort's first pass takes the index itself as the value
    }
    // number of rendered Gaussians (they sort to the front of the depth order): one atomic per warp
    const unsigned vis = __ballot_sync(0xFFFFFFFFu, rendered);
    if (lane == 0 && vis) atomicAdd(&g.meta[1], (uint32_t)__popc(vis));
}

// ------------------------------------------------------------------------------------------------------------------
// View group (msgs_forward_group_begin): the S (view, scale) units of ONE view differ only in the image
// size.  Everything that does not depend on the resolution — the per-Gaussian input loads, the frustum test, the
// projection, cov3D, the SH colour (192 B of every survivor's row) and the depth key — is done once; the
// resolution-dependent tail (EWA cov2D with that scale's focal length, conic, pixel size, the MS-GS filters, radius,
// tile rectangle, splat record) runs per scale with the expression trees of preprocess_fwd_kernel, so every unit's
// outputs are bit-identical to a separate msgs_forward of that unit.  One depth key per Gaussian (rendered at ANY scale
// of the group): the depth order of a view is the same at every resolution, so the group sorts once and every unit's
// binning walks that order (Gaussians a unit does not render have a zero-area rectangle there and place nothing).
// ------------------------------------------------------------------------------------------------------------------
template <bool STAGED>
__global__ void __launch_bounds__(PF_THREADS, STAGED ? 6 : 4)
preprocess_fwd_group_kernel(FwdParams prm, const __grid_constant__ GroupParams gp, const float* __restrict__ means3D,
                            const float* __restrict__ scales, const float* __restrict__ rotations,
                            const float* __restrict__ opacities, const float* __restrict__ shs,
                            const float* __restrict__ cov3D_precomp, const float* __restrict__ colors_precomp,
                            const float* __restrict__ max_px, const float* __restrict__ min_px,
                            const uint8_t* __restrict__ base_mask, uint32_t* __restrict__ dkeys,
                            uint32_t* __restrict__ dvals) {
    __shared__ __align__(16) float s_rows[STAGED ? (PF_THREADS / 32) * 32 * SH_ROW_STRIDE : 4];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = idx < prm.P;
    const int S = gp.S;
    uint32_t rendered_at = 0;             // bit s: rendered by unit s
    uint32_t fbits_at = 0;                // bits 2s, 2s+1: the backward's filter re-check outcomes at scale s (quirk Q2)
    bool in_front = false;
    float3 p_orig = {0.f, 0.f, 0.f};
    float3 p_proj = {0.f, 0.f, 0.f};
    float view_z = 0.f;
    float c3[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float occ = 0.f, mn = 0.f, mx = 0.f;
    bool base = false;
    // the view matrix in registers: the per-scale loop below stores through pointers the compiler must assume may alias
    // prm.view, which would make it reload the matrix (a dozen dependent L1 round trips) in every iteration
    float vm[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) vm[k] = prm.view[k];
    if (valid) {
        // every resolution-independent input is requested up front (one round trip instead of a chain of dependent ones)
        p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        occ = opacities[idx];
        mn = min_px[idx];
        mx = max_px[idx];
        base = base_mask[idx] != 0;
        v3 sc{0.f, 0.f, 0.f};
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; ++k) c3[k] = cov3D_precomp[(size_t)idx * 6 + k];
        } else {
            sc = v3{scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
            q = reinterpret_cast<const float4*>(rotations)[idx];
        }
        if (prm.raw) {
            occ = act_sigmoid(occ);
            sc = v3{act_exp(sc.x), act_exp(sc.y), act_exp(sc.z)};
            q = act_normalize(q);
        }
        // auxiliary.h:144-169 in_frustum: near plane only (quirk Q11)
        const float3 p_view = xform4x3(p_orig, vm);
        view_z = p_view.z;
        if (p_view.z <= 0.2f) {
            if (prm.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
        } else {
            in_front = true;
            const float4 p_hom = xform4x4(p_orig, prm.proj);
            const float p_w = 1.0f / (p_hom.w + 0.0000001f);
            p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
            if (cov3D_precomp == nullptr) cov3d_from_scale_rot(sc, prm.scale_modifier, q, c3);
        }
    }

#pragma unroll 1
    for (int s = 0; s < S; ++s) {
        const GroupScale& u = gp.sc[s];
        int out_radius = 0;
        float out_px = 0.0f;
        uint32_t out_tiles = 0;
        ushort4 out_rect = make_ushort4(0, 0, 0, 0);
        if (in_front) {
            do {
                const float3 cov = cov2d_fwd(p_orig, u.focal_x, u.focal_y, prm.tan_fovx, prm.tan_fovy, c3, vm);
                const float det = (cov.x * cov.z - cov.y * cov.y);
                if (det == 0.0f) break;
                const float det_inv = 1.f / det;
                const float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
                const float det_ori = ((cov.x - 0.3) * (cov.z - 0.3) - cov.y * cov.y);
                if (det_ori == 0.0f) break;
                const float det_inv_ori = 1.f / det_ori;
                const float3 conic_ori = {(float)((cov.z - 0.3) * det_inv_ori), -cov.y * det_inv_ori,
                                          (float)((cov.x - 0.3) * det_inv_ori)};
                float pixel_size = level_set_px(occ, conic_ori.x, conic_ori.z);
                pixel_size /= prm.scale_modifier;
                out_px = pixel_size;
                float rel_min_pixel_size = 1.0f;
                if (mn > 0) rel_min_pixel_size = pixel_size / mn;
                float rel_max_pixel_size = 1.0f;
                if (mx > 0) rel_max_pixel_size = pixel_size / mx;
                const uint32_t fb = ((rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) ? 1u : 0u) |
                                    ((rel_max_pixel_size > 2.0f && !base) ? 2u : 0u);
                if (prm.filter_small) {
                    if (rel_min_pixel_size < 0.5f && pixel_size < 2.0f && !base) break;
                    if (prm.filter_large) {
                        if (rel_max_pixel_size > 2.0f && !base) break;
                    }
                }
                const float mid = 0.5f * (cov.x + cov.z);
                const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
                const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
                const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
                const float2 point_image = {ndc2pix(p_proj.x, u.W), ndc2pix(p_proj.y, u.H)};
                uint2 rect_min, rect_max;
                get_rect(point_image, (int)my_radius, rect_min, rect_max, u.gx, u.gy);
                if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) break;

                const float blend_px = level_set_px(occ, conic.x, conic.z);
                out_radius = (int)my_radius;
                out_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
                out_rect = make_ushort4((unsigned short)rect_min.x, (unsigned short)rect_min.y,
                                        (unsigned short)rect_max.x, (unsigned short)rect_max.y);
                rendered_at |= 1u << s;
                fbits_at |= fb << (2 * s);
                float4* rec = u.rec + REC_F4 * (size_t)idx;
                rec[0] = make_float4(point_image.x, point_image.y, view_z, blend_px);
                rec[1] = make_float4(conic.x, conic.y, conic.z, occ);
                rec[3] = splat_footprint(conic.x, conic.y, conic.z, occ);
                float4* ga = reinterpret_cast<float4*>(u.gacc) + 3 * (size_t)idx;
                const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                ga[0] = z4; ga[1] = z4; ga[2] = z4;
                if (cov3D_precomp == nullptr) {
                    float2* c2 = reinterpret_cast<float2*>(u.cov3D + (size_t)idx * 6);
                    c2[0] = make_float2(c3[0], c3[1]); c2[1] = make_float2(c3[2], c3[3]); c2[2] = make_float2(c3[4], c3[5]);
                }
            } while (0);
        }
        if (valid) {
            u.radii[idx] = out_radius;
            u.pixel_sizes[idx] = out_px;
            u.tiles_touched[idx] = out_tiles;
            u.rect[idx] = out_rect;
        }
    }

    // forward.cu:375-381 — colour, once for all the scales that render the Gaussian
    const bool rendered = rendered_at != 0u;
    v3 rgb{0.f, 0.f, 0.f};
    uint8_t cl = 0;
    if (STAGED) {
        float* const wrows = s_rows + (threadIdx.x >> 5) * 32 * SH_ROW_STRIDE;
        const unsigned rmask = __ballot_sync(0xFFFFFFFFu, rendered);
        if (rmask) {
            stage_sh_rows(wrows, shs + (size_t)(idx - lane) * SH_ROW_FLOATS, rmask, lane);
            stage_wait_all();
        }
        __syncwarp();
        if (rendered) {
            const v3 cam{prm.campos[0], prm.campos[1], prm.campos[2]};
            v3 dir = v3{p_orig.x, p_orig.y, p_orig.z} - cam;
            dir = dir / length(dir);
            const ShDir sd = make_shdir(dir);
            const int ncoef = (prm.D + 1) * (prm.D + 1);
            const float4* row4 = reinterpret_cast<const float4*>(wrows + lane * SH_ROW_STRIDE);
            v3 result{0.f, 0.f, 0.f};
#pragma unroll
            for (int grp = 0; grp < 4; ++grp) {
                if (4 * grp < ncoef) {
                    const float4 q4[3] = {row4[3 * grp], row4[3 * grp + 1], row4[3 * grp + 2]};
                    const float* f = reinterpret_cast<const float*>(q4);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const int k = 4 * grp + kk;
                        const v3 shk{f[3 * kk], f[3 * kk + 1], f[3 * kk + 2]};
                        if (k == 0) result = kSH0 * shk;                        // dc_delta_interp == 0 (quirk Q9)
                        else if (k < ncoef) result = result + sh_weight(k, sd) * shk;
                    }
                }
            }
            result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
            cl = (uint8_t)((result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0));
            rgb = {fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
        }
    } else if (rendered) {
        if (colors_precomp == nullptr) {
            v3 sh[16];
            const float* f = shs + (size_t)idx * prm.M * 3;
            const int need = (prm.D + 1) * (prm.D + 1);
#pragma unroll
            for (int k = 0; k < 16; ++k)
                sh[k] = (k < need && k < prm.M) ? v3{f[3 * k], f[3 * k + 1], f[3 * k + 2]} : v3{0.f, 0.f, 0.f};
            const v3 cam{prm.campos[0], prm.campos[1], prm.campos[2]};
            rgb = sh_to_rgb(prm.D, sh, v3{p_orig.x, p_orig.y, p_orig.z}, cam, cl);
        } else {
            rgb = {colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]};
        }
    }
    if (rendered) {
#pragma unroll 1
        for (int s = 0; s < S; ++s) {
            if (!((rendered_at >> s) & 1u)) continue;
            const GroupScale& u = gp.sc[s];
            u.rec[REC_F4 * (size_t)idx + 2] = make_float4(rgb.x, rgb.y, rgb.z, (float)((fbits_at >> (2 * s)) & 3u));
            u.clamped[idx] = cl;
        }
    }
    if (valid) {
        dkeys[idx] = rendered ? __float_as_uint(view_z) : 0xFFFFFFFFu;      // value = index: generated by the sort's first pass
        (void)dvals;
    }
    // Gaussians rendered by any unit of the group (the front of the shared depth order): every unit's meta[1]
    const unsigned vis = __ballot_sync(0xFFFFFFFFu, rendered);
    if (vis && lane < S) atomicAdd(&gp.sc[lane].meta[1], (uint32_t)__popc(vis));
}

void launch_preprocess_fwd_group(const FwdParams& prm, const GroupParams& gp, const float* means3D, const float* scales,
                                 const float* rotations, const float* opacities, const float* shs,
                                 const float* cov3D_precomp, const float* colors_precomp, const float* max_px,
                                 const float* min_px, const uint8_t* base_mask, uint32_t* dkeys, uint32_t* dvals,
                                 cudaStream_t st) {
    const unsigned grid = (unsigned)((prm.P + PF_THREADS - 1) / PF_THREADS);
    if (colors_precomp == nullptr && prm.M == 16)
        preprocess_fwd_group_kernel<true><<<grid, PF_THREADS, 0, st>>>(prm, gp, means3D, scales, rotations, opacities, shs,
                                                                       cov3D_precomp, colors_precomp, max_px, min_px,
                                                                       base_mask, dkeys, dvals);
    else
        preprocess_fwd_group_kernel<false><<<grid, PF_THREADS, 0, st>>>(prm, gp, means3D, scales, rotations, opacities, shs,
                                                                        cov3D_precomp, colors_precomp, max_px, min_px,
                                                                        base_mask, dkeys, dvals);
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    const float3 pv = xform4x3(p, view);
    present[idx] = !(pv.z <= 0.2f);     // auxiliary.h:159 — NaN depth counts as inside, as in the reference
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
    if (P <= 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
}

void launch_preprocess_fwd(const FwdParams& prm, const float* means3D, const float* scales, const float* rotations,
                           const float* opacities, const float* shs, const float* cov3D_precomp,
                           const float* colors_precomp, const float* max_px, const float* min_px,
                           const uint8_t* base_mask, int* radii, float* pixel_sizes, GeomState g, cudaStream_t st) {
    const unsigned grid = (unsigned)((prm.P + PF_THREADS - 1) / PF_THREADS);
    if (colors_precomp == nullptr && prm.M == 16)
        preprocess_fwd_kernel<true><<<grid, PF_THREADS, 0, st>>>(prm, means3D, scales, rotations, opacities, shs, cov3D_precomp,
                                                                 colors_precomp, max_px, min_px, base_mask, radii, pixel_sizes, g);
    else
        preprocess_fwd_kernel<false><<<grid, PF_THREADS, 0, st>>>(prm, means3D, scales, rotations, opacities, shs, cov3D_precomp,
                                                                  colors_precomp, max_px, min_px, base_mask, radii, pixel_sizes, g);
}

}  // namespace msgs
