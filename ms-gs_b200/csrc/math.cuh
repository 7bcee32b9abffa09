// Small fp32 vector/matrix helpers whose operator definitions and operand ORDER reproduce the ones the
// reference reaches through its vendored glm 0.9.9.9 (third_party/glm/glm/detail/type_mat3x3.inl:486-520,
// type_vec3.inl, func_geometric.inl) so that nvcc builds the same expression trees — and therefore the same
// FMA contraction — as for the reference kernels.  radii / tiles / depth keys / pixel_sizes depend on it
// bit for bit (SURVEY.md §7.3 hard part 1).  Nothing here is copied from glm: these are the textbook
// definitions written in glm's evaluation order.
#pragma once
#include <cuda_runtime.h>

namespace msgs {

struct v3 { float x, y, z; };
__device__ __forceinline__ v3 operator+(v3 a, v3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ v3 operator-(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ v3 operator*(float s, v3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ v3 operator*(v3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ v3 operator/(v3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ v3& operator+=(v3& a, v3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
__device__ __forceinline__ float dot(v3 a, v3 b) { v3 t{a.x * b.x, a.y * b.y, a.z * b.z}; return t.x + t.y + t.z; }
__device__ __forceinline__ float length(v3 a) { return sqrtf(dot(a, a)); }

// column-major 3x3: c[col][row]
struct m3 { float c[3][3]; };
__device__ __forceinline__ m3 make_m3(float x0, float y0, float z0, float x1, float y1, float z1,
                                      float x2, float y2, float z2) {
    m3 m;
    m.c[0][0] = x0; m.c[0][1] = y0; m.c[0][2] = z0;
    m.c[1][0] = x1; m.c[1][1] = y1; m.c[1][2] = z1;
    m.c[2][0] = x2; m.c[2][1] = y2; m.c[2][2] = z2;
    return m;
}
__device__ __forceinline__ m3 operator*(const m3& a, const m3& b) {
    m3 r;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}
__device__ __forceinline__ m3 operator*(float s, const m3& a) {
    m3 r;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) r.c[i][j] = a.c[i][j] * s;
    return r;
}
__device__ __forceinline__ m3 transpose(const m3& a) {
    m3 r;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) r.c[i][j] = a.c[j][i];
    return r;
}
__device__ __forceinline__ v3 col(const m3& a, int k) { return {a.c[k][0], a.c[k][1], a.c[k][2]}; }

// auxiliary.h:58-77 (row-vector convention: matrix[0],[4],[8],[12] multiply x,y,z,1)
__device__ __forceinline__ float3 xform4x3(const float3& p, const float* m) {
    float3 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
    return t;
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* m) {
    float4 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
                m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
    return t;
}
__device__ __forceinline__ float3 xformVec4x3T(const float3& p, const float* m) {   // auxiliary.h:89-97
    float3 t = {m[0] * p.x + m[1] * p.y + m[2] * p.z,
                m[4] * p.x + m[5] * p.y + m[6] * p.z,
                m[8] * p.x + m[9] * p.y + m[10] * p.z};
    return t;
}

// auxiliary.h:41-44 — evaluated in double because of the 1.0 / 0.5 literals (quirk Q4)
__device__ __forceinline__ float ndc2pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// auxiliary.h:46-56 — tile rectangle from the INT-truncated radius (quirk Q13)
__device__ __forceinline__ void get_rect(float2 p, int max_radius, uint2& rmin, uint2& rmax, unsigned gx, unsigned gy) {
    rmin = {min(gx, (unsigned)max((int)0, (int)((p.x - max_radius) / TILE_X))),
            min(gy, (unsigned)max((int)0, (int)((p.y - max_radius) / TILE_Y)))};
    rmax = {min(gx, (unsigned)max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X))),
            min(gy, (unsigned)max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)))};
}

// SH basis constants (real SH up to degree 3; same numeric literals as auxiliary.h:22-39)
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

// ---- real SH basis (degree <= 3) evaluated per coefficient ------------------------------------------------
// Direction-dependent quantities shared by all coefficients.
struct ShDir { float x, y, z, xx, yy, zz, xy, yz, xz; };
__device__ __forceinline__ ShDir make_shdir(v3 d) {
    return {d.x, d.y, d.z, d.x * d.x, d.y * d.y, d.z * d.z, d.x * d.y, d.y * d.z, d.x * d.z};
}
// weight of coefficient k in the colour sum, with the reference's association order (forward.cu:33-66):
// colour = sum_k sh_weight(k) * sh[k]  (+0.5)
__device__ __forceinline__ float sh_weight(int k, const ShDir& d) {
    switch (k) {
        case 0: return kSH0;
        case 1: return -(kSH1 * d.y);
        case 2: return kSH1 * d.z;
        case 3: return -(kSH1 * d.x);
        case 4: return kSH2[0] * d.xy;
        case 5: return kSH2[1] * d.yz;
        case 6: return kSH2[2] * (2.0f * d.zz - d.xx - d.yy);
        case 7: return kSH2[3] * d.xz;
        case 8: return kSH2[4] * (d.xx - d.yy);
        case 9: return kSH3[0] * d.y * (3.0f * d.xx - d.yy);
        case 10: return kSH3[1] * d.xy * d.z;
        case 11: return kSH3[2] * d.y * (4.0f * d.zz - d.xx - d.yy);
        case 12: return kSH3[3] * d.z * (2.0f * d.zz - 3.0f * d.xx - 3.0f * d.yy);
        case 13: return kSH3[4] * d.x * (4.0f * d.zz - d.xx - d.yy);
        case 14: return kSH3[5] * d.z * (d.xx - d.yy);
        default: return kSH3[6] * d.x * (d.xx - 3.0f * d.yy);
    }
}
// d sh_weight(k) / d (x, y, z)  (backward.cu:66-124)
__device__ __forceinline__ v3 sh_weight_grad(int k, const ShDir& d) {
    switch (k) {
        case 0: return {0.f, 0.f, 0.f};
        case 1: return {0.f, -kSH1, 0.f};
        case 2: return {0.f, 0.f, kSH1};
        case 3: return {-kSH1, 0.f, 0.f};
        case 4: return {kSH2[0] * d.y, kSH2[0] * d.x, 0.f};
        case 5: return {0.f, kSH2[1] * d.z, kSH2[1] * d.y};
        case 6: return {kSH2[2] * 2.f * -d.x, kSH2[2] * 2.f * -d.y, kSH2[2] * 2.f * 2.f * d.z};
        case 7: return {kSH2[3] * d.z, 0.f, kSH2[3] * d.x};
        case 8: return {kSH2[4] * 2.f * d.x, kSH2[4] * 2.f * -d.y, 0.f};
        case 9: return {kSH3[0] * 3.f * 2.f * d.xy, kSH3[0] * 3.f * (d.xx - d.yy), 0.f};
        case 10: return {kSH3[1] * d.yz, kSH3[1] * d.xz, kSH3[1] * d.xy};
        case 11: return {kSH3[2] * -2.f * d.xy, kSH3[2] * (-3.f * d.yy + 4.f * d.zz - d.xx), kSH3[2] * 4.f * 2.f * d.yz};
        case 12: return {kSH3[3] * -3.f * 2.f * d.xz, kSH3[3] * -3.f * 2.f * d.yz, kSH3[3] * 3.f * (2.f * d.zz - d.xx - d.yy)};
        case 13: return {kSH3[4] * (-3.f * d.xx + 4.f * d.zz - d.yy), kSH3[4] * -2.f * d.xy, kSH3[4] * 4.f * 2.f * d.xz};
        case 14: return {kSH3[5] * 2.f * d.xz, kSH3[5] * -2.f * d.yz, kSH3[5] * (d.xx - d.yy)};
        default: return {kSH3[6] * 3.f * (d.xx - d.yy), kSH3[6] * -3.f * 2.f * d.xy, 0.f};
    }
}

// ---- warp-cooperative staging of SH rows (M = 16: 48 floats = 192 B per Gaussian) -----------------------------
// Per-thread reads of a 192-byte row touch 32 different lines per instruction; instead the warp copies the rows of
// its 32 Gaussians with fully coalesced 16-byte chunks (cp.async global -> shared, no registers), and each lane
// then reads its own row from shared memory.  Row stride 52 floats: 16-byte aligned and conflict-free for LDS.128.
constexpr int SH_ROW_FLOATS = 48;
constexpr int SH_ROW_STRIDE = 52;
__device__ __forceinline__ void cp_async16_ca(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
// rows whose bit is set in row_mask are copied; gsrc points at the row of the warp's first Gaussian
__device__ __forceinline__ void stage_sh_rows(float* srows, const float* gsrc, unsigned row_mask, int lane) {
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;            // 16-byte chunk 0..383
        const int row = (c * 2731) >> 15;        // c / 12 for c < 384
        const int col = c - row * 12;
        if ((row_mask >> row) & 1u) cp_async16_ca(srows + row * SH_ROW_STRIDE + col * 4, gsrc + (size_t)c * 4);
    }
    asm volatile("cp.async.commit_group;\n" ::);
}
// the same for 32 rows anywhere in the array: lane L's row is number `row_id` (rows of unset mask bits are skipped)
__device__ __forceinline__ void stage_sh_rows_gather(float* srows, const float* shs, int row_id, unsigned row_mask, int lane) {
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const int c = it * 32 + lane;            // 16-byte chunk 0..383
        const int row = (c * 2731) >> 15;        // c / 12 for c < 384
        const int col = c - row * 12;
        const int rid = __shfl_sync(0xFFFFFFFFu, row_id, row);
        if ((row_mask >> row) & 1u)
            cp_async16_ca(srows + row * SH_ROW_STRIDE + col * 4, shs + (size_t)rid * SH_ROW_FLOATS + col * 4);
    }
    asm volatile("cp.async.commit_group;\n" ::);
}
__device__ __forceinline__ void stage_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

}  // namespace msgs
