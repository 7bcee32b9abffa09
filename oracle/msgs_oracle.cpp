// TEST INFRASTRUCTURE — CPU restatement ("oracle") of the MS-GS differentiable Gaussian rasterizer.
//
// PARITY PIN: the reference ships no tests / golden vectors for this path, so this
// restatement is pinned against outputs of the reference CUDA rasterizer itself, produced on a B200 by
// tests/golden/make_golden.py through oracle/_ref (the unmodified reference sources compiled in
// place) and committed under tests/golden/.  tests/test_oracle_golden.py checks the pin.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// this library.  The product path (ms-gs_b200/) never does.
//
// Every function cites the reference lines it restates; paths are relative to
// /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/.
// Arithmetic is plain IEEE fp32 with contraction OFF (the GPU contracts to FMA, so CPU-vs-GPU
// comparisons are toleranced; GPU-vs-GPU comparisons are bit-exact).  Gradient sums can be
// accumulated in fp64 (acc64=1) to serve as an order-independent ground truth.
//
// Build: g++ -O2 -fopenmp -ffp-contract=off -std=c++17 -shared -fPIC (oracle/Makefile).
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <numeric>
#include <string>
#include <limits>

namespace {

constexpr int TILE = 16;                    // config.h:16-17
constexpr float kC0 = 0.28209479177387814f; // auxiliary.h:22-39
constexpr float kC1 = 0.4886025119029199f;
constexpr float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                          -1.0925484305920792f, 0.5462742152960396f};
constexpr float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                          0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                          -0.5900435899266435f};

struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot3(V3 a, V3 b) { V3 t{a.x * b.x, a.y * b.y, a.z * b.z}; return t.x + t.y + t.z; }  // glm dot order

// Column-major 3x3 with glm's product order (glm/detail/type_mat3x3.inl:486-520):
// c[col][row];  (A*B)[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2]
struct M3 { float c[3][3]; };
inline M3 mul(const M3& a, const M3& b) {
    M3 r;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}
inline M3 transpose(const M3& a) {
    M3 r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.c[i][j] = a.c[j][i];
    return r;
}

// CUDA float->int conversion (cvt.rzi.s32.f32) saturates and maps NaN to 0; C++ casts are UB there.
inline int f2i(float v) {
    if (std::isnan(v)) return 0;
    if (v >= 2147483648.0f) return std::numeric_limits<int>::max();
    if (v <= -2147483648.0f) return std::numeric_limits<int>::min();
    return (int)v;
}

// auxiliary.h:41-44 — evaluated in double, narrowed on return.
inline float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

// auxiliary.h:46-56 — tile rectangle from an *integer* radius (Q13).
inline void tile_rect(float px, float py, int radius, int gx, int gy, uint32_t mn[2], uint32_t mx[2]) {
    const float r = (float)radius;
    mn[0] = std::min<uint32_t>((uint32_t)gx, (uint32_t)std::max(0, f2i((px - r) / (float)TILE)));
    mn[1] = std::min<uint32_t>((uint32_t)gy, (uint32_t)std::max(0, f2i((py - r) / (float)TILE)));
    mx[0] = std::min<uint32_t>((uint32_t)gx, (uint32_t)std::max(0, f2i((px + r + (float)(TILE - 1)) / (float)TILE)));
    mx[1] = std::min<uint32_t>((uint32_t)gy, (uint32_t)std::max(0, f2i((py + r + (float)(TILE - 1)) / (float)TILE)));
}

// auxiliary.h:58-77
inline V3 xform43(V3 p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
            m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
inline void xform44(V3 p, const float* m, float o[4]) {
    o[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
    o[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
    o[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
    o[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}

// rasterizer_impl.cu:35-50
inline uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

// forward.cu:122-156 — quaternion NOT normalised (Q3).
inline void cov3d_from_scale_rot(V3 s, float mod, const float* q, float* out6) {
    M3 S{}; S.c[0][0] = mod * s.x; S.c[1][1] = mod * s.y; S.c[2][2] = mod * s.z;
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    M3 R;
    R.c[0][0] = 1.f - 2.f * (y * y + z * z); R.c[0][1] = 2.f * (x * y - r * z); R.c[0][2] = 2.f * (x * z + r * y);
    R.c[1][0] = 2.f * (x * y + r * z); R.c[1][1] = 1.f - 2.f * (x * x + z * z); R.c[1][2] = 2.f * (y * z - r * x);
    R.c[2][0] = 2.f * (x * z - r * y); R.c[2][1] = 2.f * (y * z + r * x); R.c[2][2] = 1.f - 2.f * (x * x + y * y);
    M3 Mx = mul(S, R);
    M3 Sg = mul(transpose(Mx), Mx);
    out6[0] = Sg.c[0][0]; out6[1] = Sg.c[0][1]; out6[2] = Sg.c[0][2];
    out6[3] = Sg.c[1][1]; out6[4] = Sg.c[1][2]; out6[5] = Sg.c[2][2];
}

struct Cov2DTmp { M3 T; M3 Vrk; M3 Wm; V3 t; float txtz, tytz, limx, limy; };

// forward.cu:78-117 (also the recompute at backward.cu:164-199).  Returns low-passed (a, b, c).
inline V3 cov2d(V3 mean, float fx, float fy, float tanx, float tany, const float* cov3D, const float* vm, Cov2DTmp* tmp = nullptr) {
    V3 t = xform43(mean, vm);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = std::min(limx, std::max(-limx, txtz)) * t.z;
    t.y = std::min(limy, std::max(-limy, tytz)) * t.z;
    M3 J{};
    J.c[0][0] = fx / t.z; J.c[0][1] = 0.f; J.c[0][2] = -(fx * t.x) / (t.z * t.z);
    J.c[1][0] = 0.f; J.c[1][1] = fy / t.z; J.c[1][2] = -(fy * t.y) / (t.z * t.z);
    M3 Wm;
    Wm.c[0][0] = vm[0]; Wm.c[0][1] = vm[4]; Wm.c[0][2] = vm[8];
    Wm.c[1][0] = vm[1]; Wm.c[1][1] = vm[5]; Wm.c[1][2] = vm[9];
    Wm.c[2][0] = vm[2]; Wm.c[2][1] = vm[6]; Wm.c[2][2] = vm[10];
    M3 T = mul(Wm, J);
    M3 Vrk;
    Vrk.c[0][0] = cov3D[0]; Vrk.c[0][1] = cov3D[1]; Vrk.c[0][2] = cov3D[2];
    Vrk.c[1][0] = cov3D[1]; Vrk.c[1][1] = cov3D[3]; Vrk.c[1][2] = cov3D[4];
    Vrk.c[2][0] = cov3D[2]; Vrk.c[2][1] = cov3D[4]; Vrk.c[2][2] = cov3D[5];
    M3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
    if (tmp) { tmp->T = T; tmp->Vrk = Vrk; tmp->Wm = Wm; tmp->t = t; tmp->txtz = txtz; tmp->tytz = tytz; tmp->limx = limx; tmp->limy = limy; }
    return {cov.c[0][0] + 0.3f, cov.c[0][1], cov.c[1][1] + 0.3f};
}

// forward.cu:21-75
inline V3 sh_to_rgb(int deg, const float* sh /*M x 3*/, V3 pos, V3 campos, uint8_t* clamped3) {
    V3 dir = pos - campos;
    const float len = std::sqrt(dot3(dir, dir));
    dir = {dir.x / len, dir.y / len, dir.z / len};
    auto S = [&](int i) { return V3{sh[3 * i], sh[3 * i + 1], sh[3 * i + 2]}; };
    V3 res = kC0 * (S(0) + V3{0.f, 0.f, 0.f});   // dc_delta_interp == 0 (Q9)
    if (deg > 0) {
        const float x = dir.x, y = dir.y, z = dir.z;
        res = res - (kC1 * y) * S(1) + (kC1 * z) * S(2) - (kC1 * x) * S(3);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            res = res + (kC2[0] * xy) * S(4) + (kC2[1] * yz) * S(5) + (kC2[2] * (2.0f * zz - xx - yy)) * S(6) +
                  (kC2[3] * xz) * S(7) + (kC2[4] * (xx - yy)) * S(8);
            if (deg > 2) {
                res = res + (kC3[0] * y * (3.0f * xx - yy)) * S(9) + (kC3[1] * xy * z) * S(10) +
                      (kC3[2] * y * (4.0f * zz - xx - yy)) * S(11) +
                      (kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy)) * S(12) +
                      (kC3[4] * x * (4.0f * zz - xx - yy)) * S(13) + (kC3[5] * z * (xx - yy)) * S(14) +
                      (kC3[6] * x * (xx - 3.0f * yy)) * S(15);
            }
        }
    }
    res = {res.x + 0.5f, res.y + 0.5f, res.z + 0.5f};
    clamped3[0] = res.x < 0; clamped3[1] = res.y < 0; clamped3[2] = res.z < 0;
    return {std::max(res.x, 0.0f), std::max(res.y, 0.0f), std::max(res.z, 0.0f)};
}

struct Oracle {
    int P = 0, W = 0, H = 0, gx = 0, gy = 0, R = 0, M = 0, D = 0;
    std::vector<float> depths, means2D, cov3D, conic_opacity, rgb, final_T;
    std::vector<uint8_t> clamped;
    std::vector<uint32_t> tiles_touched, point_offsets, vals_unsorted, vals_sorted, n_contrib, ranges;
    std::vector<uint64_t> keys_unsorted, keys_sorted;
    std::vector<float> dL_dconic, dL_dmean2D_full;  // backward intermediates
    uint64_t sum_contrib = 0;
};

}  // namespace

extern "C" {

void* msgs_oracle_new() { return new Oracle(); }
void msgs_oracle_free(void* h) { delete (Oracle*)h; }

// Restates CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:200-367).  Null pointers signal the
// unused alternative (colors_precomp / cov3D_precomp / shs / scales+rotations) as in the reference.
// occ_multiplier / dc_delta are dead inputs (Q9) and are not taken.  Returns num_rendered.
int msgs_oracle_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* max_px, const float* min_px, const float* opacities,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const uint8_t* base_mask, const float* viewmatrix, const float* projmatrix, const float* campos,
    float tan_fovx, float tan_fovy, int filter_small, int filter_large,
    float* out_color, float* out_acc_px, float* out_depth, int* radii, float* pixel_sizes)
{
    Oracle& o = *(Oracle*)h;
    o.P = P; o.W = W; o.H = H; o.M = M; o.D = D;
    o.gx = (W + TILE - 1) / TILE; o.gy = (H + TILE - 1) / TILE;
    const int gx = o.gx, gy = o.gy, T = gx * gy;
    const float focal_y = H / (2.0f * tan_fovy);   // rasterizer_impl.cu:236-237
    const float focal_x = W / (2.0f * tan_fovx);
    o.depths.assign(P, 0.f); o.means2D.assign(2 * (size_t)P, 0.f); o.cov3D.assign(6 * (size_t)P, 0.f);
    o.conic_opacity.assign(4 * (size_t)P, 0.f); o.rgb.assign(3 * (size_t)P, 0.f); o.clamped.assign(3 * (size_t)P, 0);
    o.tiles_touched.assign(P, 0); o.point_offsets.assign(P, 0);
    const V3 cam{campos[0], campos[1], campos[2]};

    // ---- preprocessCUDA (forward.cu:197-389) ----
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        radii[i] = 0; pixel_sizes[i] = 0; o.tiles_touched[i] = 0;
        const V3 p{means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
        const V3 p_view = xform43(p, viewmatrix);
        if (p_view.z <= 0.2f) continue;                                   // auxiliary.h:157-168 (Q11)
        float ph[4]; xform44(p, projmatrix, ph);
        const float p_w = 1.0f / (ph[3] + 0.0000001f);
        const V3 p_proj{ph[0] * p_w, ph[1] * p_w, ph[2] * p_w};
        const float* c3;
        if (cov3D_precomp) c3 = cov3D_precomp + 6 * (size_t)i;
        else {
            cov3d_from_scale_rot({scales[3 * i], scales[3 * i + 1], scales[3 * i + 2]}, scale_modifier,
                                 rotations + 4 * (size_t)i, &o.cov3D[6 * (size_t)i]);
            c3 = &o.cov3D[6 * (size_t)i];
        }
        const V3 cov = cov2d(p, focal_x, focal_y, tan_fovx, tan_fovy, c3, viewmatrix);
        const float det = cov.x * cov.z - cov.y * cov.y;                  // forward.cu:235-239
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float conic[3] = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
        // forward.cu:242-246 — mixed double/float (Q4)
        const float det_ori = (float)(((double)cov.x - 0.3) * ((double)cov.z - 0.3) - (double)(cov.y * cov.y));
        if (det_ori == 0.0f) continue;
        const float det_inv_ori = 1.f / det_ori;
        const float cori_x = (float)(((double)cov.z - 0.3) * (double)det_inv_ori);
        const float cori_z = (float)(((double)cov.x - 0.3) * (double)det_inv_ori);
        const float occ = opacities[i];
        float level_set = -2.0f * std::log(1.0f / (255.0f * occ));        // forward.cu:251-259
        level_set = std::max(0.0f, level_set);
        const float dx = std::sqrt(level_set / cori_x), dy = std::sqrt(level_set / cori_z);
        float pixel_size = std::fmin(dx, dy);                             // CUDA min(float,float) == fminf
        pixel_size /= scale_modifier;
        pixel_sizes[i] = pixel_size;                                      // written before the filters (Q5)
        float rel_min = 1.0f, rel_max = 1.0f;
        if (min_px[i] > 0) rel_min = pixel_size / min_px[i];
        if (max_px[i] > 0) rel_max = pixel_size / max_px[i];
        if (filter_small) {                                               // forward.cu:270-289 (Q1: nested)
            if (rel_min < 0.5f && pixel_size < 2.0f && !base_mask[i]) continue;
            if (filter_large) { if (rel_max > 2.0f && !base_mask[i]) continue; }
        }
        const float mid = 0.5f * (cov.x + cov.z);                         // forward.cu:356-371
        const float l1 = mid + std::sqrt(std::max(0.1f, mid * mid - det));
        const float l2 = mid - std::sqrt(std::max(0.1f, mid * mid - det));
        const float my_radius = std::ceil(3.f * std::sqrt(std::fmax(l1, l2)));
        const float pix_x = ndc2pix(p_proj.x, W), pix_y = ndc2pix(p_proj.y, H);
        uint32_t mn[2], mx[2];
        tile_rect(pix_x, pix_y, f2i(my_radius), gx, gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        if (!colors_precomp) {
            V3 c = sh_to_rgb(D, shs + 3 * (size_t)M * i, p, cam, &o.clamped[3 * (size_t)i]);
            o.rgb[3 * (size_t)i] = c.x; o.rgb[3 * (size_t)i + 1] = c.y; o.rgb[3 * (size_t)i + 2] = c.z;
        }
        o.depths[i] = p_view.z;
        radii[i] = f2i(my_radius);
        o.means2D[2 * (size_t)i] = pix_x; o.means2D[2 * (size_t)i + 1] = pix_y;
        o.conic_opacity[4 * (size_t)i] = conic[0]; o.conic_opacity[4 * (size_t)i + 1] = conic[1];
        o.conic_opacity[4 * (size_t)i + 2] = conic[2]; o.conic_opacity[4 * (size_t)i + 3] = occ;
        o.tiles_touched[i] = (mx[1] - mn[1]) * (mx[0] - mn[0]);
    }

    // ---- inclusive scan (rasterizer_impl.cu:301-305) ----
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += o.tiles_touched[i]; o.point_offsets[i] = run; }
    const int R = P > 0 ? (int)run : 0;
    o.R = R;
    o.keys_unsorted.assign(R, 0); o.vals_unsorted.assign(R, 0);

    // ---- duplicateWithKeys (rasterizer_impl.cu:70-111) ----
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        uint32_t off = (i == 0) ? 0 : o.point_offsets[i - 1];
        uint32_t mn[2], mx[2];
        tile_rect(o.means2D[2 * (size_t)i], o.means2D[2 * (size_t)i + 1], radii[i], gx, gy, mn, mx);
        uint32_t dbits; std::memcpy(&dbits, &o.depths[i], 4);
        for (uint32_t y = mn[1]; y < mx[1]; ++y)
            for (uint32_t x = mn[0]; x < mx[0]; ++x) {
                uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
                key <<= 32; key |= dbits;
                o.keys_unsorted[off] = key; o.vals_unsorted[off] = (uint32_t)i; ++off;
            }
    }

    // ---- stable sort on the low (32+bit) key bits (rasterizer_impl.cu:324-332) ----
    const uint32_t bit = higher_msb((uint32_t)T);
    const uint64_t mask = (32 + bit >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1ull);
    std::vector<uint32_t> perm(R);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) {
        return (o.keys_unsorted[a] & mask) < (o.keys_unsorted[b] & mask); });
    o.keys_sorted.resize(R); o.vals_sorted.resize(R);
    for (int k = 0; k < R; ++k) { o.keys_sorted[k] = o.keys_unsorted[perm[k]]; o.vals_sorted[k] = o.vals_unsorted[perm[k]]; }

    // ---- identifyTileRanges (rasterizer_impl.cu:116-138, memset :334) ----
    o.ranges.assign(2 * (size_t)T, 0);
    for (int k = 0; k < R; ++k) {
        const uint32_t cur = (uint32_t)(o.keys_sorted[k] >> 32);
        if (k == 0) o.ranges[2 * cur] = 0;
        else {
            const uint32_t prev = (uint32_t)(o.keys_sorted[k - 1] >> 32);
            if (cur != prev) { o.ranges[2 * prev + 1] = k; o.ranges[2 * cur] = k; }
        }
        if (k == R - 1) o.ranges[2 * cur + 1] = R;
    }

    // ---- renderCUDA forward (forward.cu:441-598), restated per pixel ----
    o.final_T.assign((size_t)W * H, 0.f); o.n_contrib.assign((size_t)W * H, 0);
    const float* feat = colors_precomp ? colors_precomp : o.rgb.data();
    uint64_t sum_contrib = 0;
    #pragma omp parallel for schedule(dynamic, 1) reduction(+ : sum_contrib)
    for (int t = 0; t < T; ++t) {
        const int tx = t % gx, ty = t / gx;
        const uint32_t r0 = o.ranges[2 * t], r1 = o.ranges[2 * t + 1];
        // Per-instance blend pixel size from the LOW-PASSED conic, not divided by scale_modifier (Q6; forward.cu:499-506)
        std::vector<float> px_blend(r1 > r0 ? r1 - r0 : 0);
        for (uint32_t k = r0; k < r1; ++k) {
            const uint32_t id = o.vals_sorted[k];
            const float* co = &o.conic_opacity[4 * (size_t)id];
            float ls = -2.0f * std::log(1.0f / (255.0f * co[3])); ls = std::max(0.0f, ls);
            px_blend[k - r0] = std::fmin(std::sqrt(ls / co[0]), std::sqrt(ls / co[2]));
        }
        for (int ly = 0; ly < TILE; ++ly) for (int lx = 0; lx < TILE; ++lx) {
            const int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            const size_t pid = (size_t)W * py + px;
            const float pxf = (float)px, pyf = (float)py;
            float Tr = 1.0f, C[3] = {0, 0, 0}, acc_px = 0, acc_z = 0;
            uint32_t contributor = 0, last = 0;
            for (uint32_t k = r0; k < r1; ++k) {
                ++contributor;
                const uint32_t id = o.vals_sorted[k];
                const float dxp = o.means2D[2 * (size_t)id] - pxf, dyp = o.means2D[2 * (size_t)id + 1] - pyf;
                const float* co = &o.conic_opacity[4 * (size_t)id];
                const float power = -0.5f * (co[0] * dxp * dxp + co[2] * dyp * dyp) - co[1] * dxp * dyp;
                if (power > 0.0f) continue;
                const float alpha = std::fmin(0.99f, co[3] * std::exp(power));
                if (alpha < 1.0f / 255.0f) continue;
                acc_px += px_blend[k - r0] * Tr * alpha;                 // before the termination test (Q7)
                acc_z += o.depths[id] * Tr * alpha;
                const float test_T = Tr * (1 - alpha);
                if (test_T < 0.0001f) break;
                for (int ch = 0; ch < 3; ++ch) C[ch] += feat[3 * (size_t)id + ch] * alpha * Tr;
                Tr = test_T; last = contributor;
            }
            o.final_T[pid] = Tr; o.n_contrib[pid] = last; sum_contrib += last;
            for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pid] = C[ch] + Tr * bg[ch];
            out_acc_px[pid] = acc_px; out_depth[pid] = acc_z;
        }
    }
    o.sum_contrib = sum_contrib;
    return R;
}

// Restates CudaRasterizer::Rasterizer::backward (rasterizer_impl.cu:371-494) on the state kept by
// the preceding msgs_oracle_forward on the same handle.  Output arrays are caller-allocated and are
// zero-filled here (rasterize_points.cu:184-194).  dL_dmean2D is (P,3) with z==0, dL_dconic (P,4)
// with element [2] never written (Q10).  acc64: accumulate the per-pixel sums in double.
void msgs_oracle_backward(void* h, const float* bg,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* pixel_sizes, const float* max_px, const float* min_px,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const uint8_t* base_mask, const float* viewmatrix, const float* projmatrix, const float* campos,
    float tan_fovx, float tan_fovy, const int* radii, int filter_small, int filter_large,
    const float* dL_dpix,
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int acc64)
{
    Oracle& o = *(Oracle*)h;
    const int P = o.P, W = o.W, H = o.H, gx = o.gx, gy = o.gy, T = gx * gy, M = o.M, D = o.D;
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
    std::fill(dL_dmean2D, dL_dmean2D + 3 * (size_t)P, 0.f); std::fill(dL_dconic, dL_dconic + 4 * (size_t)P, 0.f);
    std::fill(dL_dopacity, dL_dopacity + P, 0.f); std::fill(dL_dcolor, dL_dcolor + 3 * (size_t)P, 0.f);
    std::fill(dL_dmean3D, dL_dmean3D + 3 * (size_t)P, 0.f); std::fill(dL_dcov3D, dL_dcov3D + 6 * (size_t)P, 0.f);
    if (M > 0) std::fill(dL_dsh, dL_dsh + 3 * (size_t)M * P, 0.f);
    std::fill(dL_dscale, dL_dscale + 3 * (size_t)P, 0.f); std::fill(dL_drot, dL_drot + 4 * (size_t)P, 0.f);
    const float* feat = colors_precomp ? colors_precomp : o.rgb.data();

    // 9 accumulators per Gaussian: color[3], mean2D[2], conic[3] (x,y,w), opacity
    std::vector<double> acc_d; std::vector<float> acc_f;
    if (acc64) acc_d.assign(9 * (size_t)P, 0.0); else acc_f.assign(9 * (size_t)P, 0.f);
    auto add = [&](size_t slot, float v) {
        if (acc64) {
            #pragma omp atomic
            acc_d[slot] += (double)v;
        } else {
            #pragma omp atomic
            acc_f[slot] += v;
        }
    };
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);   // backward.cu:551-552 (double)

    // ---- renderCUDA backward (backward.cu:499-733), restated per pixel ----
    #pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < T; ++t) {
        const int tx = t % gx, ty = t / gx;
        const uint32_t r0 = o.ranges[2 * t], r1 = o.ranges[2 * t + 1];
        for (int ly = 0; ly < TILE; ++ly) for (int lx = 0; lx < TILE; ++lx) {
            const int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            const size_t pid = (size_t)W * py + px;
            const float pxf = (float)px, pyf = (float)py;
            const float T_final = o.final_T[pid];
            float Tr = T_final;
            const uint32_t last_contributor = o.n_contrib[pid];
            float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0;
            const float dpx[3] = {dL_dpix[pid], dL_dpix[(size_t)H * W + pid], dL_dpix[2 * (size_t)H * W + pid]};
            uint32_t contributor = r1 - r0;
            for (uint32_t k = r1; k-- > r0;) {
                --contributor;
                if (contributor >= last_contributor) continue;
                const uint32_t id = o.vals_sorted[k];
                const float dxp = o.means2D[2 * (size_t)id] - pxf, dyp = o.means2D[2 * (size_t)id + 1] - pyf;
                const float* co = &o.conic_opacity[4 * (size_t)id];
                const float power = -0.5f * (co[0] * dxp * dxp + co[2] * dyp * dyp) - co[1] * dxp * dyp;
                if (power > 0.0f) continue;
                const float G = std::exp(power);
                const float alpha = std::fmin(0.99f, co[3] * G);
                // per-pair filter re-check, independent small/large (Q2; backward.cu:645-660)
                const float ps = pixel_sizes[id];
                float rel_min = 1.0f, rel_max = 1.0f;
                if (min_px[id] > 0) rel_min = ps / min_px[id];
                if (max_px[id] > 0) rel_max = ps / max_px[id];
                if (filter_small && rel_min < 0.5f && ps < 2.0f && !base_mask[id]) continue;
                if (filter_large && rel_max > 2.0f && !base_mask[id]) continue;
                if (alpha < 1.0f / 255.0f) continue;
                Tr = Tr / (1.f - alpha);
                const float dchannel_dcolor = alpha * Tr;
                float dL_dalpha = 0.0f;
                for (int ch = 0; ch < 3; ++ch) {
                    const float c = feat[3 * (size_t)id + ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                    last_color[ch] = c;
                    dL_dalpha += (c - accum_rec[ch]) * dpx[ch];
                    add(9 * (size_t)id + ch, dchannel_dcolor * dpx[ch]);
                }
                dL_dalpha *= Tr;
                last_alpha = alpha;
                float bg_dot = 0;
                for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dpx[ch];
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                const float dL_dG = co[3] * dL_dalpha;
                const float gdx = G * dxp, gdy = G * dyp;
                const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                const float dG_ddely = -gdy * co[2] - gdx * co[1];
                add(9 * (size_t)id + 3, dL_dG * dG_ddelx * ddelx_dx);
                add(9 * (size_t)id + 4, dL_dG * dG_ddely * ddely_dy);
                add(9 * (size_t)id + 5, -0.5f * gdx * dxp * dL_dG);
                add(9 * (size_t)id + 6, -0.5f * gdx * dyp * dL_dG);
                add(9 * (size_t)id + 7, -0.5f * gdy * dyp * dL_dG);
                add(9 * (size_t)id + 8, G * dL_dalpha);
            }
        }
    }
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        auto get = [&](int s) { return acc64 ? (float)acc_d[9 * (size_t)i + s] : acc_f[9 * (size_t)i + s]; };
        for (int ch = 0; ch < 3; ++ch) dL_dcolor[3 * (size_t)i + ch] = get(ch);
        dL_dmean2D[3 * (size_t)i] = get(3); dL_dmean2D[3 * (size_t)i + 1] = get(4);
        dL_dconic[4 * (size_t)i] = get(5); dL_dconic[4 * (size_t)i + 1] = get(6); dL_dconic[4 * (size_t)i + 3] = get(7);
        dL_dopacity[i] = get(8);
    }

    const V3 cam{campos[0], campos[1], campos[2]};
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        if (!(radii[i] > 0)) continue;
        const V3 mean{means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
        const float* c3 = cov3D_precomp ? cov3D_precomp + 6 * (size_t)i : &o.cov3D[6 * (size_t)i];
        // ---- computeCov2DCUDA (backward.cu:155-276) ----
        {
            const float dcx = dL_dconic[4 * (size_t)i], dcy = dL_dconic[4 * (size_t)i + 1], dcz = dL_dconic[4 * (size_t)i + 3];
            Cov2DTmp tm;
            const V3 cv = cov2d(mean, focal_x, focal_y, tan_fovx, tan_fovy, c3, viewmatrix, &tm);
            const float xg = (tm.txtz < -tm.limx || tm.txtz > tm.limx) ? 0.f : 1.f;
            const float yg = (tm.tytz < -tm.limy || tm.tytz > tm.limy) ? 0.f : 1.f;
            const float a = cv.x, b = cv.y, c = cv.z;
            const float denom = a * c - b * b;
            float dL_da = 0, dL_db = 0, dL_dc = 0;
            const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
            const M3& Tm = tm.T; const M3& V = tm.Vrk; const M3& Wm = tm.Wm;
            float* dcov = dL_dcov3D + 6 * (size_t)i;
            if (denom2inv != 0) {
                dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
                dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
                dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
                dcov[0] = (Tm.c[0][0] * Tm.c[0][0] * dL_da + Tm.c[0][0] * Tm.c[1][0] * dL_db + Tm.c[1][0] * Tm.c[1][0] * dL_dc);
                dcov[3] = (Tm.c[0][1] * Tm.c[0][1] * dL_da + Tm.c[0][1] * Tm.c[1][1] * dL_db + Tm.c[1][1] * Tm.c[1][1] * dL_dc);
                dcov[5] = (Tm.c[0][2] * Tm.c[0][2] * dL_da + Tm.c[0][2] * Tm.c[1][2] * dL_db + Tm.c[1][2] * Tm.c[1][2] * dL_dc);
                dcov[1] = 2 * Tm.c[0][0] * Tm.c[0][1] * dL_da + (Tm.c[0][0] * Tm.c[1][1] + Tm.c[0][1] * Tm.c[1][0]) * dL_db + 2 * Tm.c[1][0] * Tm.c[1][1] * dL_dc;
                dcov[2] = 2 * Tm.c[0][0] * Tm.c[0][2] * dL_da + (Tm.c[0][0] * Tm.c[1][2] + Tm.c[0][2] * Tm.c[1][0]) * dL_db + 2 * Tm.c[1][0] * Tm.c[1][2] * dL_dc;
                dcov[4] = 2 * Tm.c[0][2] * Tm.c[0][1] * dL_da + (Tm.c[0][1] * Tm.c[1][2] + Tm.c[0][2] * Tm.c[1][1]) * dL_db + 2 * Tm.c[1][1] * Tm.c[1][2] * dL_dc;
            } else {
                for (int k = 0; k < 6; ++k) dcov[k] = 0;
            }
            auto TV = [&](int r, int col) { return Tm.c[r][0] * V.c[col][0] + Tm.c[r][1] * V.c[col][1] + Tm.c[r][2] * V.c[col][2]; };
            const float dT00 = 2 * TV(0, 0) * dL_da + TV(1, 0) * dL_db;
            const float dT01 = 2 * TV(0, 1) * dL_da + TV(1, 1) * dL_db;
            const float dT02 = 2 * TV(0, 2) * dL_da + TV(1, 2) * dL_db;
            const float dT10 = 2 * TV(1, 0) * dL_dc + TV(0, 0) * dL_db;
            const float dT11 = 2 * TV(1, 1) * dL_dc + TV(0, 1) * dL_db;
            const float dT12 = 2 * TV(1, 2) * dL_dc + TV(0, 2) * dL_db;
            const float dJ00 = Wm.c[0][0] * dT00 + Wm.c[0][1] * dT01 + Wm.c[0][2] * dT02;
            const float dJ02 = Wm.c[2][0] * dT00 + Wm.c[2][1] * dT01 + Wm.c[2][2] * dT02;
            const float dJ11 = Wm.c[1][0] * dT10 + Wm.c[1][1] * dT11 + Wm.c[1][2] * dT12;
            const float dJ12 = Wm.c[2][0] * dT10 + Wm.c[2][1] * dT11 + Wm.c[2][2] * dT12;
            const float tz = 1.f / tm.t.z, tz2 = tz * tz, tz3 = tz2 * tz;
            const float dtx = xg * -focal_x * tz2 * dJ02;
            const float dty = yg * -focal_y * tz2 * dJ12;
            const float dtz = -focal_x * tz2 * dJ00 - focal_y * tz2 * dJ11 + (2 * focal_x * tm.t.x) * tz3 * dJ02 + (2 * focal_y * tm.t.y) * tz3 * dJ12;
            const float* vm = viewmatrix;                                  // transformVec4x3Transpose, auxiliary.h:89-97
            dL_dmean3D[3 * (size_t)i] = vm[0] * dtx + vm[1] * dty + vm[2] * dtz;
            dL_dmean3D[3 * (size_t)i + 1] = vm[4] * dtx + vm[5] * dty + vm[6] * dtz;
            dL_dmean3D[3 * (size_t)i + 2] = vm[8] * dtx + vm[9] * dty + vm[10] * dtz;
        }
        // ---- preprocessCUDA backward (backward.cu:378-468) ----
        const float ps = pixel_sizes[i];
        float rel_min = 1.0f, rel_max = 1.0f;
        if (min_px[i] > 0) rel_min = ps / min_px[i];
        if (max_px[i] > 0) rel_max = ps / max_px[i];
        if (filter_small && rel_min < 0.5f && ps < 2.0f && !base_mask[i]) continue;   // independent tests (Q2)
        if (filter_large && rel_max > 2.0f && !base_mask[i]) continue;
        {
            const float* pj = projmatrix;
            float mh[4]; xform44(mean, pj, mh);
            const float m_w = 1.0f / (mh[3] + 0.0000001f);
            const float mul1 = (pj[0] * mean.x + pj[4] * mean.y + pj[8] * mean.z + pj[12]) * m_w * m_w;
            const float mul2 = (pj[1] * mean.x + pj[5] * mean.y + pj[9] * mean.z + pj[13]) * m_w * m_w;
            const float gx2 = dL_dmean2D[3 * (size_t)i], gy2 = dL_dmean2D[3 * (size_t)i + 1];
            dL_dmean3D[3 * (size_t)i] += (pj[0] * m_w - pj[3] * mul1) * gx2 + (pj[1] * m_w - pj[3] * mul2) * gy2;
            dL_dmean3D[3 * (size_t)i + 1] += (pj[4] * m_w - pj[7] * mul1) * gx2 + (pj[5] * m_w - pj[7] * mul2) * gy2;
            dL_dmean3D[3 * (size_t)i + 2] += (pj[8] * m_w - pj[11] * mul1) * gx2 + (pj[9] * m_w - pj[11] * mul2) * gy2;
        }
        if (shs) {   // SH backward (backward.cu:20-139)
            const V3 dir_orig = mean - cam;
            const float len = std::sqrt(dot3(dir_orig, dir_orig));
            const V3 dir{dir_orig.x / len, dir_orig.y / len, dir_orig.z / len};
            const float* sh = shs + 3 * (size_t)M * i;
            auto S = [&](int k) { return V3{sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]}; };
            V3 g{dL_dcolor[3 * (size_t)i], dL_dcolor[3 * (size_t)i + 1], dL_dcolor[3 * (size_t)i + 2]};
            g.x *= o.clamped[3 * (size_t)i] ? 0 : 1; g.y *= o.clamped[3 * (size_t)i + 1] ? 0 : 1; g.z *= o.clamped[3 * (size_t)i + 2] ? 0 : 1;
            V3 dx{0, 0, 0}, dy{0, 0, 0}, dz{0, 0, 0};
            const float x = dir.x, y = dir.y, z = dir.z;
            float* out = dL_dsh + 3 * (size_t)M * i;
            auto put = [&](int k, float w) { out[3 * k] = w * g.x; out[3 * k + 1] = w * g.y; out[3 * k + 2] = w * g.z; };
            put(0, kC0);
            if (D > 0) {
                put(1, -kC1 * y); put(2, kC1 * z); put(3, -kC1 * x);
                dx = -kC1 * S(3); dy = -kC1 * S(1); dz = kC1 * S(2);
                if (D > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    put(4, kC2[0] * xy); put(5, kC2[1] * yz); put(6, kC2[2] * (2.f * zz - xx - yy));
                    put(7, kC2[3] * xz); put(8, kC2[4] * (xx - yy));
                    dx = dx + ((kC2[0] * y) * S(4) + (kC2[2] * 2.f * -x) * S(6) + (kC2[3] * z) * S(7) + (kC2[4] * 2.f * x) * S(8));
                    dy = dy + ((kC2[0] * x) * S(4) + (kC2[1] * z) * S(5) + (kC2[2] * 2.f * -y) * S(6) + (kC2[4] * 2.f * -y) * S(8));
                    dz = dz + ((kC2[1] * y) * S(5) + (kC2[2] * 2.f * 2.f * z) * S(6) + (kC2[3] * x) * S(7));
                    if (D > 2) {
                        put(9, kC3[0] * y * (3.f * xx - yy)); put(10, kC3[1] * xy * z);
                        put(11, kC3[2] * y * (4.f * zz - xx - yy)); put(12, kC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                        put(13, kC3[4] * x * (4.f * zz - xx - yy)); put(14, kC3[5] * z * (xx - yy));
                        put(15, kC3[6] * x * (xx - 3.f * yy));
                        dx = dx + (kC3[0] * S(9) * 3.f * 2.f * xy + kC3[1] * S(10) * yz + kC3[2] * S(11) * -2.f * xy +
                                   kC3[3] * S(12) * -3.f * 2.f * xz + kC3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                                   kC3[5] * S(14) * 2.f * xz + kC3[6] * S(15) * 3.f * (xx - yy));
                        dy = dy + (kC3[0] * S(9) * 3.f * (xx - yy) + kC3[1] * S(10) * xz + kC3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
                                   kC3[3] * S(12) * -3.f * 2.f * yz + kC3[4] * S(13) * -2.f * xy + kC3[5] * S(14) * -2.f * yz +
                                   kC3[6] * S(15) * -3.f * 2.f * xy);
                        dz = dz + (kC3[1] * S(10) * xy + kC3[2] * S(11) * 4.f * 2.f * yz + kC3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
                                   kC3[4] * S(13) * 4.f * 2.f * xz + kC3[5] * S(14) * (xx - yy));
                    }
                }
            }
            const V3 ddir{dot3(dx, g), dot3(dy, g), dot3(dz, g)};
            // dnormvdv (auxiliary.h:107-117)
            const V3 v = dir_orig;
            const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
            const float inv32 = 1.0f / std::sqrt(sum2 * sum2 * sum2);
            dL_dmean3D[3 * (size_t)i] += ((+sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * inv32;
            dL_dmean3D[3 * (size_t)i + 1] += (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * inv32;
            dL_dmean3D[3 * (size_t)i + 2] += (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * inv32;
        }
        if (scales) {   // cov3D backward (backward.cu:283-343); no quaternion-normalisation backward (Q3)
            const float* q = rotations + 4 * (size_t)i;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            M3 Rm;
            Rm.c[0][0] = 1.f - 2.f * (y * y + z * z); Rm.c[0][1] = 2.f * (x * y - r * z); Rm.c[0][2] = 2.f * (x * z + r * y);
            Rm.c[1][0] = 2.f * (x * y + r * z); Rm.c[1][1] = 1.f - 2.f * (x * x + z * z); Rm.c[1][2] = 2.f * (y * z - r * x);
            Rm.c[2][0] = 2.f * (x * z - r * y); Rm.c[2][1] = 2.f * (y * z + r * x); Rm.c[2][2] = 1.f - 2.f * (x * x + y * y);
            const V3 s{scale_modifier * scales[3 * i], scale_modifier * scales[3 * i + 1], scale_modifier * scales[3 * i + 2]};
            M3 S{}; S.c[0][0] = s.x; S.c[1][1] = s.y; S.c[2][2] = s.z;
            const M3 Mx = mul(S, Rm);
            const float* d = dL_dcov3D + 6 * (size_t)i;
            M3 dSig;
            dSig.c[0][0] = d[0]; dSig.c[0][1] = 0.5f * d[1]; dSig.c[0][2] = 0.5f * d[2];
            dSig.c[1][0] = 0.5f * d[1]; dSig.c[1][1] = d[3]; dSig.c[1][2] = 0.5f * d[4];
            dSig.c[2][0] = 0.5f * d[2]; dSig.c[2][1] = 0.5f * d[4]; dSig.c[2][2] = d[5];
            M3 M2 = Mx; for (auto& col : M2.c) for (auto& e : col) e = e * 2.0f;   // 2.0f * M (glm scalar*mat)
            const M3 dM = mul(M2, dSig);
            const M3 Rt = transpose(Rm);
            M3 dMt = transpose(dM);
            auto col = [](const M3& m, int k) { return V3{m.c[k][0], m.c[k][1], m.c[k][2]}; };
            dL_dscale[3 * (size_t)i] = dot3(col(Rt, 0), col(dMt, 0));
            dL_dscale[3 * (size_t)i + 1] = dot3(col(Rt, 1), col(dMt, 1));
            dL_dscale[3 * (size_t)i + 2] = dot3(col(Rt, 2), col(dMt, 2));
            for (int k = 0; k < 3; ++k) { dMt.c[0][k] *= s.x; dMt.c[1][k] *= s.y; dMt.c[2][k] *= s.z; }
            float* dq = dL_drot + 4 * (size_t)i;
            dq[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
            dq[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
            dq[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
            dq[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
        }
    }
}

// markVisible / checkFrustum (rasterizer_impl.cu:54-66, 141-153)
void msgs_oracle_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present)
{
    (void)projmatrix;
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        const V3 pv = xform43({means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]}, viewmatrix);
        present[i] = pv.z > 0.2f;
    }
}

// Intermediate state accessor: returns pointer and element count; element size implied by name.
const void* msgs_oracle_buffer(void* h, const char* name, uint64_t* count)
{
    Oracle& o = *(Oracle*)h;
    const std::string n(name);
#define RET(v) do { *count = (v).size(); return (v).data(); } while (0)
    if (n == "depths") RET(o.depths);
    if (n == "means2D") RET(o.means2D);
    if (n == "cov3D") RET(o.cov3D);
    if (n == "conic_opacity") RET(o.conic_opacity);
    if (n == "rgb") RET(o.rgb);
    if (n == "clamped") RET(o.clamped);
    if (n == "tiles_touched") RET(o.tiles_touched);
    if (n == "point_offsets") RET(o.point_offsets);
    if (n == "keys_unsorted") RET(o.keys_unsorted);
    if (n == "vals_unsorted") RET(o.vals_unsorted);
    if (n == "keys_sorted") RET(o.keys_sorted);
    if (n == "vals_sorted") RET(o.vals_sorted);
    if (n == "ranges") RET(o.ranges);
    if (n == "final_T") RET(o.final_T);
    if (n == "n_contrib") RET(o.n_contrib);
#undef RET
    *count = 0;
    return nullptr;
}

uint64_t msgs_oracle_sum_contrib(void* h) { return ((Oracle*)h)->sum_contrib; }

}  // extern "C"
