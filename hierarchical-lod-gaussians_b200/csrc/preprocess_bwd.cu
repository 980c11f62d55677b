// preprocess_bwd.cu — per-Gaussian backward: conic/opacity -> 2D covariance -> 3D covariance -> scale/rotation,
// screen-space mean -> 3D mean, colour -> SH, and the LOD redirect of the mean gradient to the parent.
//
// Replaces computeCov2DCUDA (HR/cuda_rasterizer/backward.cu:147-326) + preprocessCUDA bwd (:398-495) +
// computeCov3D bwd (:330-393) + computeColorFromSH bwd (:23-142). The reference runs two kernels that both
// re-read the parameters and communicate through dL_dcov3D / dL_dmeans in global memory; here it is ONE pass:
// every intermediate stays in registers, each input is read once and each output written once. The mean
// gradient is accumulated with atomics because LOD children add (1-t) * their gradient to the parent's row
// (backward.cu:485-487) — the caller zero-fills the outputs, so for non-LOD entries this equals a plain store.
#include "gaussmath.cuh"
#include "state.cuh"

namespace lodgs {

struct PreBwdParams {
    int P, D, M;
    const int32_t* indices;
    const int32_t* parent_indices;
    const float* ts;
    const float* means3D;
    const int32_t* radii;
    const float* shs;
    const uint8_t* clamped;
    const float* opacities;
    const float* scales;
    const float* rotations;
    float scale_modifier;
    const float* cov3Ds;  // [P,6] by list index (geometry state) or cov3D_precomp
    int cov3D_by_global;  // 1 when cov3Ds is cov3D_precomp (indexed by global id)
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    const float* dL_dmean2D;
    const float* dL_dconic;
    const float* dL_dinvdepth;
    const float* dL_dcolor;
    float* dL_dopacity;
    float* dL_dmean3D;
    float* dL_dcov3D;
    float* dL_dsh;   // may be null when dL_drgb is given: the SH gradient is then not materialised (lodgs_sh_adam_from_rgb rebuilds it)
    float* dL_drgb;  // optional [fullP,3]: colour gradient after the clamp mask, the factor dL_dsh[k] = basis_k(dir) * dL_drgb is built from
    float* dL_dscale;
    float* dL_drot;
    int sh_vec4;
    unsigned flags;  // LODGS_RASTER_*
    int write_all;  // no index list and no parents: every row [0,P) of dL_dmean3D / dL_dcov3D / dL_dscale / dL_drot is stored (no RMW)
};

__device__ __forceinline__ float sq(float x) { return x * x; }

template <bool STAGED>
__device__ __forceinline__ void preprocess_bwd_one(const PreBwdParams& p, int t_idx, const float* view, const float* proj, float4* sh_row,
                                                   uint64_t* bar);

// STAGED (no index list, SH rows a multiple of 16 bytes): the CTA's 256 SH rows are one contiguous block of the parameter
// array, so they arrive in shared memory by ONE bulk copy (TMA) issued before the covariance math, and the dL_dsh rows
// leave the same buffer by ONE bulk store — instead of 12+12 strided 16-byte accesses per thread (192-byte stride).
template <bool STAGED>
__global__ void __launch_bounds__(256, 3) preprocess_bwd_kernel(const PreBwdParams p) {
    extern __shared__ __align__(128) float4 sh_stage[];  // [256][3M/4] when STAGED
    __shared__ float s_view[16], s_proj[16];
    __shared__ uint64_t s_bar;
    const int row4 = (p.M * 3) / 4;  // float4 per SH row
    const int tile_base = blockIdx.x * blockDim.x;
    const int tile_cnt = min((int)blockDim.x, p.P - tile_base);
    if (threadIdx.x < 16) s_view[threadIdx.x] = __ldg(p.viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = __ldg(p.projmatrix + threadIdx.x - 16);
    if (STAGED && threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        const unsigned bytes = (unsigned)tile_cnt * (unsigned)row4 * 16u;
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(sh_stage, p.shs + (size_t)tile_base * p.M * 3, bytes, &s_bar);
    }
    __syncthreads();
    const int t_idx = tile_base + threadIdx.x;
    const bool active = t_idx < p.P && (__ldg(p.radii + t_idx) > 0);
    if (p.write_all && !active && t_idx < p.P) {
        // culled Gaussians get their zero gradients here, so the caller need not zero-fill these four tensors
        float* m = p.dL_dmean3D + 3 * (size_t)t_idx;
        m[0] = m[1] = m[2] = 0.f;
        float2* c = reinterpret_cast<float2*>(p.dL_dcov3D + 6 * (size_t)t_idx);
        c[0] = c[1] = c[2] = make_float2(0.f, 0.f);
        if (p.scales) {
            float* sc = p.dL_dscale + 3 * (size_t)t_idx;
            sc[0] = sc[1] = sc[2] = 0.f;
            reinterpret_cast<float4*>(p.dL_drot)[t_idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (p.dL_drgb) {
            float* c3 = p.dL_drgb + 3 * (size_t)t_idx;
            c3[0] = c3[1] = c3[2] = 0.f;
        }
    }
    if (STAGED) {
        if (!active) {
            // culled rows still travel with the block store: they carry the zeros the caller's zero-fill put there
            mbar_wait(&s_bar, 0);
            if (t_idx < p.P && p.dL_dsh)
                for (int k = 0; k < row4; k++) sh_stage[threadIdx.x * row4 + k] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            preprocess_bwd_one<STAGED>(p, t_idx, s_view, s_proj, sh_stage + threadIdx.x * row4, &s_bar);
        }
        fence_async_smem();
        __syncthreads();
        if (threadIdx.x == 0 && p.dL_dsh) {
            bulk_s2g(p.dL_dsh + (size_t)tile_base * p.M * 3, sh_stage, (unsigned)tile_cnt * (unsigned)row4 * 16u);
            bulk_commit_wait_all();
        }
    } else {
        if (active) preprocess_bwd_one<STAGED>(p, t_idx, s_view, s_proj, nullptr, nullptr);
    }
}

template <bool STAGED>
__device__ __forceinline__ void preprocess_bwd_one(const PreBwdParams& p, int t_idx, const float* view, const float* proj, float4* sh_row,
                                                   uint64_t* bar) {
    const int idx = p.indices ? __ldg(p.indices + t_idx) : t_idx;

    // ---------------- 2D covariance / conic / AA-opacity backward (backward.cu:165-325) ----------------
    float cov6[6];
    {
        const float* c = p.cov3Ds + 6 * (size_t)(p.cov3D_by_global ? idx : t_idx);
#pragma unroll
        for (int k = 0; k < 6; k++) cov6[k] = __ldg(c + k);
    }
    const float3 mean = {__ldg(p.means3D + 3 * (size_t)idx), __ldg(p.means3D + 3 * (size_t)idx + 1), __ldg(p.means3D + 3 * (size_t)idx + 2)};
    const float4 dconic4 = __ldg(reinterpret_cast<const float4*>(p.dL_dconic) + idx);
    const float3 dL_dconic = {dconic4.x, dconic4.y, dconic4.w};
    Cov2DCtx ctx;
    const float3 cov2 = project_cov(mean, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov6, view, &ctx);
    const float3 t = ctx.t;
    const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
    const float x_grad_mul = ctx.txtz < -limx || ctx.txtz > limx ? 0 : 1;
    const float y_grad_mul = ctx.tytz < -limy || ctx.tytz > limy ? 0 : 1;
    const mat3& T = ctx.T;
    const mat3 Vrk(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
    const mat3 Wm(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);

    float c_xx = cov2.x, c_xy = cov2.y, c_yy = cov2.z;
    constexpr float h_var = 0.3f;
    const float det_cov = c_xx * c_yy - c_xy * c_xy;
    c_xx += h_var;
    c_yy += h_var;
    const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
    // without antialiasing (AR/cuda_rasterizer/backward.cu:210-229) opacity is not compensated: scaling 1, no gradient through it
    const bool aa = !(p.flags & LODGS_RASTER_NO_ANTIALIAS);
    const float h_convolution_scaling = aa ? sqrtf(max(0.000025f, det_cov / det_cov_plus_h_cov)) : 1.0f;
    const float dL_dopacity_v = p.dL_dopacity[idx];
    const float d_h_convolution_scaling = dL_dopacity_v * __ldg(p.opacities + idx);
    float dL_dopacity_out = dL_dopacity_v * h_convolution_scaling;
    const float d_inside_root =
        (!aa || (det_cov / det_cov_plus_h_cov) <= 0.000025f) ? 0.f : d_h_convolution_scaling / (2 * h_convolution_scaling);

    float dL_dc_xx, dL_dc_xy, dL_dc_yy;
    {
        const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
        const float denom_f = d_inside_root / sq(w * w + w * (x + y) + x * y - z * z);
        dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
        dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
        dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
    }
    const float denom = c_xx * c_yy - c_xy * c_xy;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6];
    if (denom2inv != 0) {
        dL_dc_xx += denom2inv * (-c_yy * c_yy * dL_dconic.x + 2 * c_xy * c_yy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.z);
        dL_dc_yy += denom2inv * (-c_xx * c_xx * dL_dconic.z + 2 * c_xx * c_xy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.x);
        dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dL_dconic.x - (denom + 2 * c_xy * c_xy) * dL_dconic.y + c_xx * c_xy * dL_dconic.z);
        dcov[0] = (T[0][0] * T[0][0] * dL_dc_xx + T[0][0] * T[1][0] * dL_dc_xy + T[1][0] * T[1][0] * dL_dc_yy);
        dcov[3] = (T[0][1] * T[0][1] * dL_dc_xx + T[0][1] * T[1][1] * dL_dc_xy + T[1][1] * T[1][1] * dL_dc_yy);
        dcov[5] = (T[0][2] * T[0][2] * dL_dc_xx + T[0][2] * T[1][2] * dL_dc_xy + T[1][2] * T[1][2] * dL_dc_yy);
        dcov[1] = 2 * T[0][0] * T[0][1] * dL_dc_xx + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_dc_xy + 2 * T[1][0] * T[1][1] * dL_dc_yy;
        dcov[2] = 2 * T[0][0] * T[0][2] * dL_dc_xx + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_dc_xy + 2 * T[1][0] * T[1][2] * dL_dc_yy;
        dcov[4] = 2 * T[0][2] * T[0][1] * dL_dc_xx + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_dc_xy + 2 * T[1][1] * T[1][2] * dL_dc_yy;
    } else {
#pragma unroll
        for (int i = 0; i < 6; i++) dcov[i] = 0;
    }
#pragma unroll
    for (int i = 0; i < 6; i++) p.dL_dcov3D[6 * (size_t)idx + i] = dcov[i];

    const float dL_dT00 = 2 * (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_dc_xx +
                          (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc_xy;
    const float dL_dT01 = 2 * (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_dc_xx +
                          (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc_xy;
    const float dL_dT02 = 2 * (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_dc_xx +
                          (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc_xy;
    const float dL_dT10 = 2 * (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc_yy +
                          (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_dc_xy;
    const float dL_dT11 = 2 * (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc_yy +
                          (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_dc_xy;
    const float dL_dT12 = 2 * (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc_yy +
                          (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_dc_xy;
    const float dL_dJ00 = Wm[0][0] * dL_dT00 + Wm[0][1] * dL_dT01 + Wm[0][2] * dL_dT02;
    const float dL_dJ02 = Wm[2][0] * dL_dT00 + Wm[2][1] * dL_dT01 + Wm[2][2] * dL_dT02;
    const float dL_dJ11 = Wm[1][0] * dL_dT10 + Wm[1][1] * dL_dT11 + Wm[1][2] * dL_dT12;
    const float dL_dJ12 = Wm[2][0] * dL_dT10 + Wm[2][1] * dL_dT11 + Wm[2][2] * dL_dT12;
    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;
    const float h_x = p.focal_x, h_y = p.focal_y;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    if (p.dL_dinvdepth) dL_dtz -= __ldg(p.dL_dinvdepth + idx) / (t.z * t.z);
    // transformVec4x3Transpose
    float3 dmean = {view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz, view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz,
                    view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz};

    // ---------------- screen-space mean -> 3D mean (backward.cu:429-447) ----------------
    {
        const float4 m_hom = xform4x4(mean, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        const float gx = __ldg(p.dL_dmean2D + 3 * (size_t)idx), gy = __ldg(p.dL_dmean2D + 3 * (size_t)idx + 1);
        float3 a;
        a.x = (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
        a.y = (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
        a.z = (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
        dmean.x += a.x; dmean.y += a.y; dmean.z += a.z;
    }

    int parent_id = -1;
    if (p.parent_indices) parent_id = __ldg(p.parent_indices + t_idx);
    const bool is_child = parent_id != -1;

    // ---------------- colour -> SH (backward.cu:23-142) ----------------
    if (p.shs) {
        const float cx = __ldg(p.campos), cy = __ldg(p.campos + 1), cz = __ldg(p.campos + 2);
        const float3 dir_orig = {mean.x - cx, mean.y - cy, mean.z - cz};
        const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
        const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
        const uint8_t cl = __ldg(p.clamped + t_idx);
        v3 dRGB = {__ldg(p.dL_dcolor + 3 * (size_t)idx), __ldg(p.dL_dcolor + 3 * (size_t)idx + 1), __ldg(p.dL_dcolor + 3 * (size_t)idx + 2)};
        dRGB.x *= (cl & 1) ? 0 : 1;
        dRGB.y *= (cl & 2) ? 0 : 1;
        dRGB.z *= (cl & 4) ? 0 : 1;
        if (p.dL_drgb && !is_child) {
            float* c3 = p.dL_drgb + 3 * (size_t)idx;
            c3[0] = dRGB.x; c3[1] = dRGB.y; c3[2] = dRGB.z;
        }
        const int deg = p.D;
        const int ncoef = (deg + 1) * (deg + 1);
        v3 sh[16];
        {
            float* f = reinterpret_cast<float*>(sh);
            const float* base = p.shs + (size_t)idx * p.M * 3;
            if (STAGED) {
                mbar_wait(bar, 0);
#pragma unroll
                for (int k = 0; k < 12; k++)
                    if (k * 4 < ncoef * 3) {
                        float4 a = sh_row[k];
                        f[4 * k] = a.x; f[4 * k + 1] = a.y; f[4 * k + 2] = a.z; f[4 * k + 3] = a.w;
                    }
            } else if (p.sh_vec4) {
#pragma unroll
                for (int k = 0; k < 12; k++)
                    if (k * 4 < ncoef * 3) {
                        float4 a = __ldg(reinterpret_cast<const float4*>(base) + k);
                        f[4 * k] = a.x; f[4 * k + 1] = a.y; f[4 * k + 2] = a.z; f[4 * k + 3] = a.w;
                    }
            } else {
#pragma unroll
                for (int k = 0; k < 48; k++)
                    if (k < ncoef * 3) f[k] = __ldg(base + k);
            }
        }
        float w[16];  // dRGB/dsh_k
        v3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
        w[0] = kSH0;
        if (deg > 0) {
            w[1] = -kSH1 * y; w[2] = kSH1 * z; w[3] = -kSH1 * x;
            dRGBdx = -kSH1 * sh[3];
            dRGBdy = -kSH1 * sh[1];
            dRGBdz = kSH1 * sh[2];
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                w[4] = kSH2[0] * xy; w[5] = kSH2[1] * yz; w[6] = kSH2[2] * (2.f * zz - xx - yy); w[7] = kSH2[3] * xz; w[8] = kSH2[4] * (xx - yy);
                dRGBdx = dRGBdx + (kSH2[0] * y * sh[4] + kSH2[2] * 2.f * -x * sh[6] + kSH2[3] * z * sh[7] + kSH2[4] * 2.f * x * sh[8]);
                dRGBdy = dRGBdy + (kSH2[0] * x * sh[4] + kSH2[1] * z * sh[5] + kSH2[2] * 2.f * -y * sh[6] + kSH2[4] * 2.f * -y * sh[8]);
                dRGBdz = dRGBdz + (kSH2[1] * y * sh[5] + kSH2[2] * 2.f * 2.f * z * sh[6] + kSH2[3] * x * sh[7]);
                if (deg > 2) {
                    w[9] = kSH3[0] * y * (3.f * xx - yy); w[10] = kSH3[1] * xy * z; w[11] = kSH3[2] * y * (4.f * zz - xx - yy);
                    w[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = kSH3[4] * x * (4.f * zz - xx - yy);
                    w[14] = kSH3[5] * z * (xx - yy); w[15] = kSH3[6] * x * (xx - 3.f * yy);
                    dRGBdx = dRGBdx + (kSH3[0] * 3.f * 2.f * xy * sh[9] + kSH3[1] * yz * sh[10] + kSH3[2] * -2.f * xy * sh[11] +
                                       kSH3[3] * -3.f * 2.f * xz * sh[12] + kSH3[4] * (-3.f * xx + 4.f * zz - yy) * sh[13] +
                                       kSH3[5] * 2.f * xz * sh[14] + kSH3[6] * 3.f * (xx - yy) * sh[15]);
                    dRGBdy = dRGBdy + (kSH3[0] * 3.f * (xx - yy) * sh[9] + kSH3[1] * xz * sh[10] + kSH3[2] * (-3.f * yy + 4.f * zz - xx) * sh[11] +
                                       kSH3[3] * -3.f * 2.f * yz * sh[12] + kSH3[4] * -2.f * xy * sh[13] + kSH3[5] * -2.f * yz * sh[14] +
                                       kSH3[6] * -3.f * 2.f * xy * sh[15]);
                    dRGBdz = dRGBdz + (kSH3[1] * xy * sh[10] + kSH3[2] * 4.f * 2.f * yz * sh[11] + kSH3[3] * 3.f * (2.f * zz - xx - yy) * sh[12] +
                                       kSH3[4] * 4.f * 2.f * xz * sh[13] + kSH3[5] * (xx - yy) * sh[14]);
                }
            }
        }
        // dL_dsh: LOD children's own SH gradient is discarded by the reference (backward.cu:489-493)
        if (!p.dL_dsh) {
            // not materialised: the consumer rebuilds w[k] * dRGB from dL_drgb (csrc/sh_adam.cu)
        } else if (STAGED) {
            // the row is rebuilt in place (this thread is its only reader and writer); coefficients above the active degree are 0
            float o[48];
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const float wk = k < ncoef ? w[k] : 0.f;
                o[3 * k + 0] = wk * dRGB.x;
                o[3 * k + 1] = wk * dRGB.y;
                o[3 * k + 2] = wk * dRGB.z;
            }
            const int row4 = (p.M * 3) / 4;
#pragma unroll
            for (int k = 0; k < 12; k++)
                if (k < row4) sh_row[k] = make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
            for (int k = 12; k < row4; k++) sh_row[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (!is_child) {
            float* out = p.dL_dsh + (size_t)idx * p.M * 3;
#pragma unroll
            for (int k = 0; k < 16; k++)
                if (k < ncoef) {
                    out[3 * k + 0] = w[k] * dRGB.x;
                    out[3 * k + 1] = w[k] * dRGB.y;
                    out[3 * k + 2] = w[k] * dRGB.z;
                }
        }
        const float3 dL_ddir = {dRGBdx.x * dRGB.x + dRGBdx.y * dRGB.y + dRGBdx.z * dRGB.z, dRGBdy.x * dRGB.x + dRGBdy.y * dRGB.y + dRGBdy.z * dRGB.z,
                                dRGBdz.x * dRGB.x + dRGBdz.y * dRGB.y + dRGBdz.z * dRGB.z};
        // dnormvdv (auxiliary.h:133-143)
        const float3 vv = dir_orig;
        const float sum2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmean.x += ((+sum2 - vv.x * vv.x) * dL_ddir.x - vv.y * vv.x * dL_ddir.y - vv.z * vv.x * dL_ddir.z) * invsum32;
        dmean.y += (-vv.x * vv.y * dL_ddir.x + (sum2 - vv.y * vv.y) * dL_ddir.y - vv.z * vv.y * dL_ddir.z) * invsum32;
        dmean.z += (-vv.x * vv.z * dL_ddir.x - vv.y * vv.z * dL_ddir.y + (sum2 - vv.z * vv.z) * dL_ddir.z) * invsum32;
    }

    // ---------------- 3D covariance -> scale / rotation (backward.cu:330-393) ----------------
    if (p.scales && !is_child) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const mat3 R = quat_to_mat(r, x, y, z);
        const float sx = p.scale_modifier * __ldg(p.scales + 3 * (size_t)idx), sy = p.scale_modifier * __ldg(p.scales + 3 * (size_t)idx + 1),
                    sz = p.scale_modifier * __ldg(p.scales + 3 * (size_t)idx + 2);
        mat3 S(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
        S[0][0] = sx; S[1][1] = sy; S[2][2] = sz;
        const mat3 M = mul(S, R);
        const mat3 dL_dSigma(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4], 0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
        mat3 twoM;
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int k = 0; k < 3; k++) twoM[c][k] = 2.0f * M[c][k];
        const mat3 dL_dM = mul(twoM, dL_dSigma);
        const mat3 Rt = transpose(R);
        mat3 dL_dMt = transpose(dL_dM);
        float* ds = p.dL_dscale + 3 * (size_t)idx;
        ds[0] = Rt[0][0] * dL_dMt[0][0] + Rt[0][1] * dL_dMt[0][1] + Rt[0][2] * dL_dMt[0][2];
        ds[1] = Rt[1][0] * dL_dMt[1][0] + Rt[1][1] * dL_dMt[1][1] + Rt[1][2] * dL_dMt[1][2];
        ds[2] = Rt[2][0] * dL_dMt[2][0] + Rt[2][1] * dL_dMt[2][1] + Rt[2][2] * dL_dMt[2][2];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            dL_dMt[0][k] *= sx;
            dL_dMt[1][k] *= sy;
            dL_dMt[2][k] *= sz;
        }
        float4 dq;
        dq.x = 2 * z * (dL_dMt[0][1] - dL_dMt[1][0]) + 2 * y * (dL_dMt[2][0] - dL_dMt[0][2]) + 2 * x * (dL_dMt[1][2] - dL_dMt[2][1]);
        dq.y = 2 * y * (dL_dMt[1][0] + dL_dMt[0][1]) + 2 * z * (dL_dMt[2][0] + dL_dMt[0][2]) + 2 * r * (dL_dMt[1][2] - dL_dMt[2][1]) - 4 * x * (dL_dMt[2][2] + dL_dMt[1][1]);
        dq.z = 2 * x * (dL_dMt[1][0] + dL_dMt[0][1]) + 2 * r * (dL_dMt[2][0] - dL_dMt[0][2]) + 2 * z * (dL_dMt[1][2] + dL_dMt[2][1]) - 4 * y * (dL_dMt[2][2] + dL_dMt[0][0]);
        dq.w = 2 * r * (dL_dMt[0][1] - dL_dMt[1][0]) + 2 * x * (dL_dMt[2][0] + dL_dMt[0][2]) + 2 * y * (dL_dMt[1][2] + dL_dMt[2][1]) - 4 * z * (dL_dMt[1][1] + dL_dMt[0][0]);
        reinterpret_cast<float4*>(p.dL_drot)[idx] = dq;
    }

    // ---------------- outputs, with the LOD redirect (backward.cu:458-494) ----------------
    if (is_child) {
        const float tt = __ldg(p.ts + t_idx);
        p.dL_dopacity[idx] = 0.f;
        atomicAdd(p.dL_dmean3D + 3 * (size_t)parent_id + 0, (1.0f - tt) * dmean.x);
        atomicAdd(p.dL_dmean3D + 3 * (size_t)parent_id + 1, (1.0f - tt) * dmean.y);
        atomicAdd(p.dL_dmean3D + 3 * (size_t)parent_id + 2, (1.0f - tt) * dmean.z);
    } else {
        p.dL_dopacity[idx] = dL_dopacity_out;
        if (p.write_all) {
            p.dL_dmean3D[3 * (size_t)idx + 0] = dmean.x;
            p.dL_dmean3D[3 * (size_t)idx + 1] = dmean.y;
            p.dL_dmean3D[3 * (size_t)idx + 2] = dmean.z;
        } else {
            atomicAdd(p.dL_dmean3D + 3 * (size_t)idx + 0, dmean.x);
            atomicAdd(p.dL_dmean3D + 3 * (size_t)idx + 1, dmean.y);
            atomicAdd(p.dL_dmean3D + 3 * (size_t)idx + 2, dmean.z);
        }
    }
}

int launch_preprocess_bwd(const PreBwdParams& p, cudaStream_t s) {
    const size_t smem = 256 * (size_t)p.M * 12;
    // staged SH rows (one bulk load per CTA, and one bulk store of dL_dsh when that gradient is materialised)
    const bool staged = p.dL_dsh ? lodgs_backward_overwrites_dsh(p.M, p.indices != nullptr, p.parent_indices != nullptr, p.shs, p.dL_dsh) != 0
                                 : lodgs_backward_overwrites_dsh(p.M, p.indices != nullptr, p.parent_indices != nullptr, p.shs, p.shs) != 0;
    if (staged) {
        static std::atomic<size_t> attr_smem[MAX_DEVICES];
        const int dslot = device_slot();
        if (smem > attr_smem[dslot].load(std::memory_order_acquire)) {
            LODGS_CUDA(cudaFuncSetAttribute(preprocess_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_smem[dslot].store(smem, std::memory_order_release);
        }
        preprocess_bwd_kernel<true><<<(p.P + 255) / 256, 256, smem, s>>>(p);
    } else {
        preprocess_bwd_kernel<false><<<(p.P + 255) / 256, 256, 0, s>>>(p);
    }
    LODGS_LAUNCHED("preprocess_bwd_kernel");
    return LODGS_OK;
}

}  // namespace lodgs

namespace lodgs {
int run_preprocess_bwd(int P, int D, int M, const int32_t* indices, const int32_t* parent_indices, const float* ts, const float* means3D,
                       const int32_t* radii, const float* shs, const uint8_t* clamped, const float* opacities, const float* scales,
                       const float* rotations, float scale_modifier, const float* cov3Ds, int cov3D_by_global, const float* viewmatrix,
                       const float* projmatrix, const float* campos, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                       const float* dL_dmean2D, const float* dL_dconic, const float* dL_dinvdepth, const float* dL_dcolor,
                       float* dL_dopacity, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                       float* dL_drgb, unsigned flags, cudaStream_t s) {
    PreBwdParams p;
    p.dL_drgb = dL_drgb;
    p.flags = flags;
    p.P = P; p.D = D; p.M = M; p.indices = indices; p.parent_indices = parent_indices; p.ts = ts; p.means3D = means3D; p.radii = radii;
    p.shs = shs; p.clamped = clamped; p.opacities = opacities; p.scales = scales; p.rotations = rotations; p.scale_modifier = scale_modifier;
    p.cov3Ds = cov3Ds; p.cov3D_by_global = cov3D_by_global; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix; p.campos = campos;
    p.focal_x = focal_x; p.focal_y = focal_y; p.tan_fovx = tan_fovx; p.tan_fovy = tan_fovy; p.dL_dmean2D = dL_dmean2D; p.dL_dconic = dL_dconic;
    p.dL_dinvdepth = dL_dinvdepth; p.dL_dcolor = dL_dcolor; p.dL_dopacity = dL_dopacity; p.dL_dmean3D = dL_dmean3D; p.dL_dcov3D = dL_dcov3D;
    p.dL_dsh = dL_dsh; p.dL_dscale = dL_dscale; p.dL_drot = dL_drot;
    p.write_all = lodgs_backward_overwrites_geometry(indices != nullptr, parent_indices != nullptr);
    p.sh_vec4 = (shs && ((uintptr_t)shs & 15) == 0 && (M * 3) % 4 == 0) ? 1 : 0;
    if (parent_indices && !ts) { set_error("lodgs_backward: parent_indices without ts"); return LODGS_E_ARG; }
    return launch_preprocess_bwd(p, s);
}
}  // namespace lodgs

// 1 when lodgs_backward takes the staged path of preprocess_bwd_kernel, which writes EVERY row of dL_dsh[0..P) (zeros for
// culled Gaussians): the caller may then skip the zero-fill of that tensor (192 B per Gaussian at M = 16).
// 1 when lodgs_backward stores every row [0,P) of dL_dmean3D and dL_dcov3D, and of dL_dscale / dL_drot when scales are given
// (zeros for culled Gaussians): without an index list no other row exists and without parents nothing is scattered.
extern "C" int lodgs_backward_overwrites_geometry(int has_indices, int has_parents) { return (!has_indices && !has_parents) ? 1 : 0; }

extern "C" int lodgs_backward_overwrites_dsh(int M, int has_indices, int has_parents, const float* shs, const float* dL_dsh) {
    return (shs && dL_dsh && !has_indices && !has_parents && M > 0 && M <= 16 && (M * 12) % 16 == 0 && ((uintptr_t)shs & 15) == 0 &&
            ((uintptr_t)dL_dsh & 15) == 0) ? 1 : 0;
}
