// gaussmath.cuh — per-Gaussian projection math shared by the forward and backward preprocess kernels.
// Formulas follow HR/cuda_rasterizer/forward.cu:25-215 and auxiliary.h:53-102; the evaluation ORDER of every
// expression is kept (left-to-right sums, glm's column-major products) because radii / rects / tile counts are
// integer outputs behind these FP32 chains and must be bit-identical to the reference build.
#pragma once
#include "common.cuh"

namespace lodgs {

struct v3 {
    float x, y, z;
};
__device__ __forceinline__ v3 operator*(float s, v3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ v3 operator+(v3 a, v3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ v3 operator-(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

__device__ const float kSH0 = 0.28209479177387814f;
__device__ const float kSH1 = 0.4886025119029199f;
__device__ const float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                                  0.5462742152960396f};
__device__ const float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                  -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

__device__ __forceinline__ float3 xform4x3(const float3 p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform4x4(const float3 p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
// ndc -> pixel, evaluated in double like auxiliary.h:53-56 (the literals there are double).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle of a splat with per-axis pixel extent `ext` (auxiliary.h:70-80).
__device__ __forceinline__ void tile_rect(const float2 p, int2 ext, uint2& rmin, uint2& rmax, unsigned gx, unsigned gy) {
    rmin = {min(gx, (unsigned)max((int)0, (int)((p.x - ext.x) / TILE_X))), min(gy, (unsigned)max((int)0, (int)((p.y - ext.y) / TILE_Y)))};
    rmax = {min(gx, (unsigned)max((int)0, (int)((p.x + ext.x + TILE_X - 1) / TILE_X))),
            min(gy, (unsigned)max((int)0, (int)((p.y + ext.y + TILE_Y - 1) / TILE_Y)))};
}

// Rotation matrix of an (unnormalised, as in forward.cu:190) quaternion, glm column layout.
__device__ __forceinline__ mat3 quat_to_mat(float r, float x, float y, float z) {
    return mat3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// Sigma = (S R)^T (S R), upper triangle (forward.cu:181-215).
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q, float* cov6) {
    mat3 S(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S[0][0] = mod * sx;
    S[1][1] = mod * sy;
    S[2][2] = mod * sz;
    mat3 R = quat_to_mat(q.x, q.y, q.z, q.w);
    mat3 M = mul(S, R);
    mat3 Sigma = mul(transpose(M), M);
    cov6[0] = Sigma[0][0];
    cov6[1] = Sigma[0][1];
    cov6[2] = Sigma[0][2];
    cov6[3] = Sigma[1][1];
    cov6[4] = Sigma[1][2];
    cov6[5] = Sigma[2][2];
}

struct Cov2DCtx {
    float3 t;       // view-space mean with the 1.3*fov clamp applied to x,y
    float txtz, tytz;
    mat3 T;         // W * J
};

// EWA projection of the 3D covariance (forward.cu:141-176). Returns (cov[0][0], cov[0][1], cov[1][1]).
__device__ __forceinline__ float3 project_cov(const float3 mean, float fx, float fy, float tan_fovx, float tan_fovy,
                                              const float* cov6, const float* view, Cov2DCtx* ctx) {
    float3 t = xform4x3(mean, view);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;
    mat3 J(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z), 0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z), 0, 0, 0);
    mat3 Wm(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    mat3 T = mul(Wm, J);
    mat3 Vrk(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
    mat3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
    if (ctx) {
        ctx->t = t;
        ctx->txtz = txtz;
        ctx->tytz = tytz;
        ctx->T = T;
    }
    return {cov[0][0], cov[0][1], cov[1][1]};
}

// SH -> RGB for coefficients already in registers (sh[k] = k-th coefficient, RGB triple), direction `d`
// normalised. Same term order as forward.cu:25-77.
__device__ __forceinline__ v3 sh_to_rgb(int deg, const v3* sh, float x, float y, float z) {
    v3 result = kSH0 * sh[0];
    if (deg > 0) {
        result = result - kSH1 * y * sh[1] + kSH1 * z * sh[2] - kSH1 * x * sh[3];
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            result = result + kSH2[0] * xy * sh[4] + kSH2[1] * yz * sh[5] + kSH2[2] * (2.0f * zz - xx - yy) * sh[6] +
                     kSH2[3] * xz * sh[7] + kSH2[4] * (xx - yy) * sh[8];
            if (deg > 2) {
                result = result + kSH3[0] * y * (3.0f * xx - yy) * sh[9] + kSH3[1] * xy * z * sh[10] +
                         kSH3[2] * y * (4.0f * zz - xx - yy) * sh[11] + kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                         kSH3[4] * x * (4.0f * zz - xx - yy) * sh[13] + kSH3[5] * z * (xx - yy) * sh[14] +
                         kSH3[6] * x * (xx - 3.0f * yy) * sh[15];
            }
        }
    }
    return result;
}

}  // namespace lodgs
