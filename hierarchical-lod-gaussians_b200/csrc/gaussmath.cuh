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

// Explicitly rounded FP32 ops: the per-Gaussian chain below decides integer outputs (radii, rects, tile counts), so the
// placement of every fused multiply-add is pinned with intrinsics instead of being left to the compiler's
// contraction heuristics. The placement is the one nvcc/ptxas 12.9 choose for the reference's glm expressions
// (HR/cuda_rasterizer/forward.cu:141-215): nvcc's fma.rn in the PTX (tools/ptx_expr.py) plus the mul/add pairs ptxas
// fuses afterwards (cuobjdump -sass of oracle/_ref), which is what makes these outputs bit-identical to oracle/_ref.
#define FM(a, b) __fmul_rn((a), (b))
#define FA(a, b) __fadd_rn((a), (b))
#define FS(a, b) __fsub_rn((a), (b))
#define FF(a, b, c) __fmaf_rn((a), (b), (c))

// Sigma = (S R)^T (S R), upper triangle (forward.cu:181-215). q = (r,x,y,z), not normalised (forward.cu:190).
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* cov6) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float sx = FM(mod, s0), sy = FM(mod, s1), sz = FM(mod, s2);
    const float yy = FM(y, y), zz = FM(z, z);
    // ptxas fuses the second product of each off-diagonal pair (SASS of the reference build): xy -+ rz, ry +- xz, yz -+ rx
    const float rz = FM(r, z), xz = FM(x, z), rx = FM(r, x);
    const float a00 = FA(yy, zz), a11 = FF(x, x, zz), a22 = FF(x, x, yy);
    const float b01 = FF(x, y, -rz), b10 = FF(x, y, rz), b02 = FF(r, y, xz), b20 = FF(y, -r, xz), b12 = FF(y, z, -rx), b21 = FF(y, z, rx);
    const float R00 = FS(1.0f, FA(a00, a00)), R01 = FA(b01, b01), R02 = FA(b02, b02);
    const float R10 = FA(b10, b10), R11 = FS(1.0f, FA(a11, a11)), R12 = FA(b12, b12);
    const float R20 = FA(b20, b20), R21 = FA(b21, b21), R22 = FS(1.0f, FA(a22, a22));
    // M = S * R with S = diag(sx,sy,sz); the products with S's literal zeros stay (they are not foldable under IEEE rules)
    const float z0 = FM(R00, 0.0f), z1 = FM(R11, 0.0f), z2 = FM(R21, 0.0f);
    const float M00 = FF(R02, 0.0f, FF(R01, 0.0f, FM(sx, R00)));
    const float M01 = FF(R02, 0.0f, FF(sy, R01, z0));
    const float M02 = FF(R02, sz, FF(R01, 0.0f, z0));
    const float M10 = FF(R12, 0.0f, FF(sx, R10, z1));
    const float M11 = FF(R12, 0.0f, FF(R10, 0.0f, FM(sy, R11)));
    const float M12 = FF(R12, sz, FF(R10, 0.0f, z1));
    const float M20 = FF(R22, 0.0f, FF(sx, R20, z2));
    const float M21 = FF(R22, 0.0f, FF(R20, 0.0f, FM(sy, R21)));
    const float M22 = FF(R22, sz, FF(R20, 0.0f, z2));
    cov6[0] = FF(M02, M02, FF(M00, M00, FM(M01, M01)));
    cov6[1] = FF(M12, M02, FF(M10, M00, FM(M11, M01)));
    cov6[2] = FF(M22, M02, FF(M20, M00, FM(M21, M01)));
    cov6[3] = FF(M12, M12, FF(M10, M10, FM(M11, M11)));
    cov6[4] = FF(M22, M12, FF(M20, M10, FM(M21, M11)));
    cov6[5] = FF(M22, M22, FF(M20, M20, FM(M21, M21)));
}

struct Cov2DCtx {
    float3 t;       // view-space mean with the 1.3*fov clamp applied to x,y
    float txtz, tytz;
    mat3 T;         // W * J (third column is zero and never read)
};

// EWA projection of the 3D covariance (forward.cu:141-176). Returns (cov[0][0], cov[0][1], cov[1][1]).
__device__ __forceinline__ float3 project_cov(const float3 mean, float fx, float fy, float tan_fovx, float tan_fovy,
                                              const float* cov6, const float* view, Cov2DCtx* ctx) {
    const float v0 = view[0], v1 = view[1], v2 = view[2], v4 = view[4], v5 = view[5], v6 = view[6], v8 = view[8], v9 = view[9],
                v10 = view[10];
    const float tx = FA(view[12], FF(mean.z, v8, FF(mean.x, v0, FM(mean.y, v4))));
    const float ty = FA(view[13], FF(mean.z, v9, FF(mean.x, v1, FM(mean.y, v5))));
    const float tz = FA(view[14], FF(mean.z, v10, FF(mean.x, v2, FM(mean.y, v6))));
    const float limx = FM(tan_fovx, 1.3f), limy = FM(tan_fovy, 1.3f);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    const float cx = fminf(limx, fmaxf(-limx, txtz)), cy = fminf(limy, fmaxf(-limy, tytz));
    const float ntz = -tz, tz2 = FM(tz, tz);
    const float J00 = __fdiv_rn(fx, tz), J02 = __fdiv_rn(FM(fx, FM(cx, ntz)), tz2);
    const float J11 = __fdiv_rn(fy, tz), J12 = __fdiv_rn(FM(fy, FM(cy, ntz)), tz2);
    // T = W * J, W = view rotation (glm columns (v0,v4,v8) (v1,v5,v9) (v2,v6,v10)), J = ((J00,0,J02),(0,J11,J12),0)
    const float T00 = FF(v2, J02, FF(v0, J00, FM(v1, 0.0f)));
    const float T01 = FF(v6, J02, FF(v4, J00, FM(v5, 0.0f)));
    const float T02 = FF(J02, v10, FF(v8, J00, FM(v9, 0.0f)));
    const float T10 = FF(v2, J12, FF(v0, 0.0f, FM(J11, v1)));
    const float T11 = FF(v6, J12, FF(v4, 0.0f, FM(J11, v5)));
    const float T12 = FF(J12, v10, FF(v8, 0.0f, FM(J11, v9)));
    const float c0 = cov6[0], c1 = cov6[1], c2 = cov6[2], c3 = cov6[3], c4 = cov6[4], c5 = cov6[5];
    // A = T^T * Vrk^T, cov = A * T
    const float A00 = FF(T02, c2, FF(T00, c0, FM(T01, c1))), A01 = FF(T12, c2, FF(T10, c0, FM(T11, c1)));
    const float A10 = FF(T02, c4, FF(T00, c1, FM(T01, c3))), A11 = FF(T12, c4, FF(T10, c1, FM(T11, c3)));
    const float A20 = FF(T02, c5, FF(T00, c2, FM(T01, c4))), A21 = FF(T12, c5, FF(T10, c2, FM(T11, c4)));
    const float cov00 = FF(T02, A20, FF(T00, A00, FM(T01, A10)));
    const float cov01 = FF(T02, A21, FF(T00, A01, FM(T01, A11)));
    const float cov11 = FF(T12, A21, FF(T10, A01, FM(T11, A11)));
    if (ctx) {
        ctx->t = {FM(cx, tz), FM(cy, tz), tz};
        ctx->txtz = txtz;
        ctx->tytz = tytz;
        ctx->T = mat3(T00, T01, T02, T10, T11, T12, 0.f, 0.f, 0.f);
    }
    return {cov00, cov01, cov11};
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
