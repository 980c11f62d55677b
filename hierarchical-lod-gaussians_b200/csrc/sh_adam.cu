// sh_adam.cu — sparse Adam on the SH coefficients driven by the COLOUR gradient (north_star item 5: the preprocess backward and the
// sparse Adam step fused).
//
// The backward of computeColorFromSH (HR/cuda_rasterizer/backward.cu:23-142) forms dL_dsh[k] = basis_k(dir) * dL_drgb, an outer product
// of the 3-float colour gradient with basis values that depend only on the Gaussian's mean and the camera centre. Materialising it
// costs 12*M bytes written by the backward and read again by the optimizer (scene/OurAdam.py:251-354 after backward.cu:398-495, or
// AR's adamUpdateCUDA, AR/cuda_rasterizer/adam.cu:9-38), and in multi-view data-parallel training an all-reduce over 12*M bytes per
// Gaussian. Here the optimizer rebuilds the row from dL_drgb (lodgs_backward_rgb): per selected row one thread evaluates the basis for
// every view and accumulates the 3*M gradient values in registers (views in index order, so every replica adds in the same order and
// the replicas stay bit-identical), the row goes through shared memory, and all threads of the CTA then apply the element-wise update
// with coalesced accesses to parameter and moments. Traffic: 24 B per updated scalar + 12 B per (row, view).
// The update arithmetic is that of adam_rows_kernel (adam.cu), for both conventions (OurAdam with bias correction, AR without).
#include "gaussmath.cuh"
#include <cmath>

namespace lodgs {

constexpr int SHA_ROWS = 128;      // rows per CTA = threads per CTA
constexpr int SHA_STRIDE = 52;     // floats per staged row: 16-byte aligned rows, 13 float4 apart = conflict-free 128-bit row-wise writes and element-wise reads

struct ShAdamParams {
    long long n_rows;              // selected rows (length of `relevant`, or all rows)
    int D, n_views, ncoef;         // active degree; views; (D+1)^2
    const float* dL_drgb;          // [n_views, P, 3]
    long long view_stride;         // floats between views (3 * P)
    const float* campos;           // [n_views, 3]
    const float* means3D;          // [P, 3]
    const int64_t* relevant;
    const uint8_t* visible;
    float *dc, *dc_m, *dc_v;       // rows of 3 floats, dc_stride floats apart
    float *rest, *rest_m, *rest_v; // rows of rest_floats floats, rest_stride floats apart
    int dc_stride, rest_stride, rest_floats;
    float step_dc, step_rest;      // -lr / bc1 (bias correction) or lr (AR)
    float b1, b2, omb1, omb2, eps, bc2_sqrt;
};

__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float (&w)[16]) {
    w[0] = kSH0;
    if (deg > 0) {
        w[1] = -kSH1 * y; w[2] = kSH1 * z; w[3] = -kSH1 * x;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            w[4] = kSH2[0] * xy; w[5] = kSH2[1] * yz; w[6] = kSH2[2] * (2.f * zz - xx - yy); w[7] = kSH2[3] * xz; w[8] = kSH2[4] * (xx - yy);
            if (deg > 2) {
                w[9] = kSH3[0] * y * (3.f * xx - yy); w[10] = kSH3[1] * xy * z; w[11] = kSH3[2] * y * (4.f * zz - xx - yy);
                w[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = kSH3[4] * x * (4.f * zz - xx - yy);
                w[14] = kSH3[5] * z * (xx - yy); w[15] = kSH3[6] * x * (xx - 3.f * yy);
            }
        }
    }
}

// One element of the update (adam_rows_kernel's arithmetic, adam.cu): returns the new parameter, m and v are updated in place.
template <bool BIAS>
__device__ __forceinline__ float sh_adam_one(const ShAdamParams& p, float par, float g, float step, float& m1, float& v1) {
    if (BIAS) {
        m1 = __fmaf_rn(g, p.omb1, __fmul_rn(m1, p.b1));
        v1 = __fmaf_rn(__fmul_rn(g, g), p.omb2, __fmul_rn(v1, p.b2));
        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v1), p.bc2_sqrt), p.eps);
        return __fmaf_rn(step, __fdiv_rn(m1, denom), par);
    }
    m1 = p.b1 * m1 + p.omb1 * g;
    v1 = p.b2 * v1 + p.omb2 * g * g;
    return par + -step * m1 / (sqrtf(v1) + p.eps);
}

// VEC: dc and rest are the two parts of ONE [P,16,3] block (row = 48 contiguous floats, 16-byte aligned), the layout both trainers use:
// phase B then moves parameter and moments as float4 (3 x 16 B per element group and thread in flight instead of 3 x 4 B).
template <bool BIAS, bool VEC>
__global__ void __launch_bounds__(SHA_ROWS, VEC ? 6 : 8) sh_adam_rgb_kernel(const ShAdamParams p) {
    __shared__ __align__(16) float s_g[SHA_ROWS * SHA_STRIDE];
    __shared__ long long s_row[SHA_ROWS];
    const long long r0 = (long long)blockIdx.x * SHA_ROWS;
    const int tid = threadIdx.x;
    // ---- phase A: one thread per row rebuilds the gradient row sum_v basis(dir_v) (x) dL_drgb_v, accumulating in shared memory (a
    // register-resident row costs 48 registers and with them the occupancy the memory-bound phase B lives on)
    {
        const long long r = r0 + tid;
        long long row = -1;
        if (r < p.n_rows) {
            row = p.relevant ? __ldg(p.relevant + r) : r;
            if (p.visible && __ldg(p.visible + row) == 0) row = -1;
        }
        s_row[tid] = row;
        if (row >= 0) {
            float* g = s_g + tid * SHA_STRIDE;
            const float mx = __ldg(p.means3D + 3 * row), my = __ldg(p.means3D + 3 * row + 1), mz = __ldg(p.means3D + 3 * row + 2);
            bool first = true;
            for (int v = 0; v < p.n_views; v++) {
                const float* c3 = p.dL_drgb + (size_t)v * p.view_stride + 3 * row;
                const float cr = __ldg(c3), cg = __ldg(c3 + 1), cb = __ldg(c3 + 2);
                if (cr == 0.f && cg == 0.f && cb == 0.f) continue;   // not seen (or fully clamped) in this view: adds exact zeros
                const float dx = mx - __ldg(p.campos + 3 * v), dy = my - __ldg(p.campos + 3 * v + 1), dz = mz - __ldg(p.campos + 3 * v + 2);
                const float len = sqrtf(dx * dx + dy * dy + dz * dz);   // backward.cu:28-31 / preprocess_bwd.cu
                float w[16];
                sh_basis(p.D, dx / len, dy / len, dz / len, w);
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if (k >= p.ncoef) w[k] = 0.f;
                auto val = [&](int j) { const int k = j / 3, ch = j - 3 * k; return w[k] * (ch == 0 ? cr : (ch == 1 ? cg : cb)); };   // j is a constant after unrolling
                float4* g4 = reinterpret_cast<float4*>(g);
#pragma unroll
                for (int q = 0; q < 12; q++) {
                    float4 a = make_float4(val(4 * q), val(4 * q + 1), val(4 * q + 2), val(4 * q + 3));
                    if (!first) {
                        const float4 o = g4[q];
                        a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
                    }
                    g4[q] = a;
                }
                first = false;
            }
            if (first) {
                float4* g4 = reinterpret_cast<float4*>(g);
#pragma unroll
                for (int q = 0; q < 12; q++) g4[q] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    __syncthreads();
    // ---- phase B: element-wise update, consecutive threads on consecutive floats of a row; U independent elements per trip: indices
    // first, then all 3U loads, then arithmetic and stores (the phase is bound by memory latency)
    if (VEC) {
        constexpr int UV = 3;   // 12 float4 per row and 128 rows per CTA: every thread takes 12 groups, 3 at a time
        for (int e0 = tid; e0 < SHA_ROWS * 12; e0 += UV * SHA_ROWS) {
            long long off[UV];
            float4 gg[UV], pp[UV], mm[UV], vv[UV];
            int qq[UV];
#pragma unroll
            for (int u = 0; u < UV; u++) {
                const int e = e0 + u * SHA_ROWS, lr = e / 12, q = e - lr * 12;
                const long long row = s_row[lr];
                off[u] = row >= 0 ? row * 48 + 4 * q : -1;
                qq[u] = q;
                gg[u] = *reinterpret_cast<const float4*>(s_g + lr * SHA_STRIDE + 4 * q);
            }
#pragma unroll
            for (int u = 0; u < UV; u++)
                if (off[u] >= 0) {
                    mm[u] = *reinterpret_cast<const float4*>(p.dc_m + off[u]);
                    vv[u] = *reinterpret_cast<const float4*>(p.dc_v + off[u]);
                    pp[u] = *reinterpret_cast<const float4*>(p.dc + off[u]);
                }
#pragma unroll
            for (int u = 0; u < UV; u++) {
                if (off[u] < 0) continue;
                const float s012 = qq[u] == 0 ? p.step_dc : p.step_rest;   // floats 0..2 of a row are the dc coefficients
                float4 o;
                o.x = sh_adam_one<BIAS>(p, pp[u].x, gg[u].x, s012, mm[u].x, vv[u].x);
                o.y = sh_adam_one<BIAS>(p, pp[u].y, gg[u].y, s012, mm[u].y, vv[u].y);
                o.z = sh_adam_one<BIAS>(p, pp[u].z, gg[u].z, s012, mm[u].z, vv[u].z);
                o.w = sh_adam_one<BIAS>(p, pp[u].w, gg[u].w, p.step_rest, mm[u].w, vv[u].w);
                *reinterpret_cast<float4*>(p.dc + off[u]) = o;
                *reinterpret_cast<float4*>(p.dc_m + off[u]) = mm[u];
                *reinterpret_cast<float4*>(p.dc_v + off[u]) = vv[u];
            }
        }
        return;
    }
    const int row_floats = 3 + p.rest_floats;
    const int total = SHA_ROWS * row_floats;
    constexpr int U = 4;
    for (int e0 = tid; e0 < total; e0 += U * SHA_ROWS) {
        long long idx[U];     // element offset; bit 62 marks the dc tensor, -1 = skip
        float pp[U], mm[U], vv[U], gg[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int e = e0 + u * SHA_ROWS;
            idx[u] = -1;
            if (e < total) {
                const int lr = row_floats == 48 ? e / 48 : e / row_floats, j = e - lr * row_floats;
                const long long row = s_row[lr];
                if (row >= 0) {
                    idx[u] = j < 3 ? ((row * p.dc_stride + j) | (1ll << 62)) : (row * p.rest_stride + (j - 3));
                    gg[u] = s_g[lr * SHA_STRIDE + j];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++)
            if (idx[u] >= 0) {
                const bool is_dc = (idx[u] >> 62) & 1;
                const long long i = idx[u] & ~(1ll << 62);
                mm[u] = (is_dc ? p.dc_m : p.rest_m)[i];
                vv[u] = (is_dc ? p.dc_v : p.rest_v)[i];
                pp[u] = (is_dc ? p.dc : p.rest)[i];
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (idx[u] < 0) continue;
            const bool is_dc = (idx[u] >> 62) & 1;
            const long long i = idx[u] & ~(1ll << 62);
            const float step = is_dc ? p.step_dc : p.step_rest;
            float m1 = mm[u], v1 = vv[u];
            float* P_ = is_dc ? p.dc : p.rest;
            P_[i] = sh_adam_one<BIAS>(p, pp[u], gg[u], step, m1, v1);
            (is_dc ? p.dc_m : p.rest_m)[i] = m1;
            (is_dc ? p.dc_v : p.rest_v)[i] = v1;
        }
    }
}

}  // namespace lodgs

using namespace lodgs;

extern "C" int lodgs_sh_adam_from_rgb(int64_t rows, int D, int n_views, const float* dL_drgb, int64_t view_stride, const float* campos,
                                      const float* means3D, const int64_t* relevant, int64_t n_relevant, const uint8_t* visible, float* dc,
                                      float* dc_m, float* dc_v, int dc_stride, float* rest, float* rest_m, float* rest_v, int rest_stride,
                                      int rest_floats, double lr_dc, double lr_rest, double beta1, double beta2, double eps, int64_t step,
                                      int bias_correction, void* stream) {
    if (rows < 0 || D < 0 || D > 3 || n_views < 1 || !dL_drgb || !campos || !means3D || !dc || !dc_m || !dc_v || dc_stride < 3 || n_relevant < 0 ||
        rest_floats < 0 || rest_floats > 45 || (rest_floats && (!rest || !rest_m || !rest_v || rest_stride < rest_floats))) {
        set_error("lodgs_sh_adam_from_rgb: bad argument");
        return LODGS_E_ARG;
    }
    const int ncoef = (D + 1) * (D + 1);
    if (3 * ncoef > 3 + rest_floats) { set_error("lodgs_sh_adam_from_rgb: degree %d needs %d floats per row, the tensors hold %d", D, 3 * ncoef, 3 + rest_floats); return LODGS_E_ARG; }
    ShAdamParams p;
    p.n_rows = relevant ? n_relevant : rows;
    if (p.n_rows == 0) return LODGS_OK;
    p.D = D; p.n_views = n_views; p.ncoef = ncoef; p.dL_drgb = dL_drgb; p.view_stride = view_stride; p.campos = campos; p.means3D = means3D;
    p.relevant = relevant; p.visible = visible; p.dc = dc; p.dc_m = dc_m; p.dc_v = dc_v; p.rest = rest; p.rest_m = rest_m; p.rest_v = rest_v;
    p.dc_stride = dc_stride; p.rest_stride = rest_stride; p.rest_floats = rest_floats;
    p.b1 = (float)beta1; p.b2 = (float)beta2; p.eps = (float)eps;
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const unsigned blocks = (unsigned)((p.n_rows + SHA_ROWS - 1) / SHA_ROWS);
    // one fused [P,16,3] block per tensor (dc = floats 0..2 of a 48-float row, rest = the other 45), rows 16-byte aligned: vector path
    static const bool no_vec = getenv("LODGS_SH_ADAM_SCALAR") != nullptr;   // A/B
    const bool vec = !no_vec && rest_floats == 45 && dc_stride == 48 && rest_stride == 48 && rest == dc + 3 && rest_m == dc_m + 3 && rest_v == dc_v + 3 &&
                     ((uintptr_t)dc % 16) == 0 && ((uintptr_t)dc_m % 16) == 0 && ((uintptr_t)dc_v % 16) == 0;
    if (bias_correction) {
        if (step < 1) { set_error("lodgs_sh_adam_from_rgb: step must be >= 1 with bias correction"); return LODGS_E_ARG; }
        const double bc1 = 1.0 - std::pow(beta1, (double)step), bc2 = 1.0 - std::pow(beta2, (double)step);   // OurAdam.py:331-337
        p.step_dc = (float)(-(lr_dc / bc1)); p.step_rest = (float)(-(lr_rest / bc1));
        p.omb1 = (float)(1.0 - beta1); p.omb2 = (float)(1.0 - beta2); p.bc2_sqrt = (float)std::sqrt(bc2);
        if (vec) sh_adam_rgb_kernel<true, true><<<blocks, SHA_ROWS, 0, s>>>(p);
        else sh_adam_rgb_kernel<true, false><<<blocks, SHA_ROWS, 0, s>>>(p);
    } else {
        p.step_dc = (float)lr_dc; p.step_rest = (float)lr_rest;
        p.omb1 = 1.0f - (float)beta1; p.omb2 = 1.0f - (float)beta2; p.bc2_sqrt = 1.0f;   // AR adam.cu:26-33 evaluates 1 - beta in float
        if (vec) sh_adam_rgb_kernel<false, true><<<blocks, SHA_ROWS, 0, s>>>(p);
        else sh_adam_rgb_kernel<false, false><<<blocks, SHA_ROWS, 0, s>>>(p);
    }
    LODGS_LAUNCHED("sh_adam_rgb_kernel");
    return LODGS_OK;
}
