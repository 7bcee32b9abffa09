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

}  // namespace msgs
