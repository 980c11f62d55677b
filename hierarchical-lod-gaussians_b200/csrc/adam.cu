// adam.cu — sparse Adam step in one pass over the selected rows.
//
// Replaces scene/OurAdam.py Adam.step(relevant) -> _single_tensor_adam (:251-354): gather rows, ~10 eager
// elementwise kernels, scatter back, per parameter group. Here: one kernel per parameter tensor, 28 B per
// updated scalar (read p,g,m,v; write p,m,v), rows selected by an index list (OurAdam), a visibility mask
// (AR adamUpdateCUDA, AR/cuda_rasterizer/adam.cu:9-38) or all rows (_single_tensor_adam2, :357-452).
// Update order and rounding follow the torch ops: m = m*b1 + g*(1-b1); v = v*b2 + (g*g)*(1-b2);
// denom = sqrt(v)/sqrt(bc2) + eps; p = p + (-step_size) * (m/denom).
#include "common.cuh"
#include <cmath>

namespace lodgs {

template <bool BIAS>
__global__ void __launch_bounds__(256) adam_rows_kernel(float* __restrict__ param, const float* __restrict__ grad,
                                                       float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq,
                                                       const int64_t* __restrict__ relevant, const uint8_t* __restrict__ visible,
                                                       long long n_rows, int width, unsigned long long magic, float lr_or_neg_step,
                                                       float b1, float b2, float omb1, float omb2, float eps, float bc2_sqrt) {
    // row of a flat element index by one 64-bit multiply-high with magic = floor(2^64 / width) + 1 (exact while
    // index * width < 2^64) instead of a 64-bit division per element
    const long long total = n_rows * width;
    // Four independent elements per thread and trip: all index lookups first, then all loads, then arithmetic and stores, so
    // that 16 loads per thread are in flight (the kernel is bound by memory latency, not by bandwidth or issue).
    constexpr int U = 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long e0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; e0 < total; e0 += U * stride) {
        long long idx[U];
        float g[U], m[U], v[U], p[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const long long e = e0 + u * stride;
            idx[u] = -1;
            if (e < total) {
                const long long r = width == 1 ? e : (long long)__umul64hi((unsigned long long)e, magic);
                const int c = (int)(e - r * width);
                long long row = r;
                bool on = true;
                if (relevant) row = __ldg(relevant + r);
                else if (visible) on = __ldg(visible + r) != 0;
                if (on) idx[u] = row * width + c;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (idx[u] >= 0) {
                g[u] = __ldg(grad + idx[u]);
                m[u] = exp_avg[idx[u]];
                v[u] = exp_avg_sq[idx[u]];
                p[u] = param[idx[u]];
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (idx[u] < 0) continue;
            const long long i = idx[u];
            float mm = m[u], vv = v[u];
            if (BIAS) {
                // torch: exp_avg.mul_(beta1).add_(grad, alpha=1-beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
                mm = __fmaf_rn(g[u], omb1, __fmul_rn(mm, b1));
                vv = __fmaf_rn(__fmul_rn(g[u], g[u]), omb2, __fmul_rn(vv, b2));
                const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vv), bc2_sqrt), eps);
                param[i] = __fmaf_rn(lr_or_neg_step, __fdiv_rn(mm, denom), p[u]);
            } else {
                // AR adam.cu:26-33: no bias correction
                mm = b1 * mm + omb1 * g[u];
                vv = b2 * vv + omb2 * g[u] * g[u];
                param[i] = p[u] + -lr_or_neg_step * mm / (sqrtf(vv) + eps);
            }
            exp_avg[i] = mm;
            exp_avg_sq[i] = vv;
        }
    }
}

}  // namespace lodgs

using namespace lodgs;

extern "C" int lodgs_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const int64_t* relevant,
                               int64_t n_relevant, const uint8_t* visible, int64_t rows, int width, double lr, double beta1, double beta2,
                               double eps, int64_t step, int bias_correction, void* stream) {
    if (!param || !grad || !exp_avg || !exp_avg_sq || rows < 0 || width <= 0 || n_relevant < 0) {
        set_error("lodgs_adam_step: bad argument");
        return LODGS_E_ARG;
    }
    const long long n_rows = relevant ? n_relevant : rows;
    if (n_rows == 0) return LODGS_OK;
    const long long total = n_rows * width;
    int blocks = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    const unsigned long long magic = width == 1 ? 0ull : ~0ull / (unsigned long long)width + 1ull;
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    if (bias_correction) {
        if (step < 1) { set_error("lodgs_adam_step: step must be >= 1 with bias correction"); return LODGS_E_ARG; }
        // bias terms in double on the host, exactly like the Python floats in OurAdam.py:331-337
        const double bc1 = 1.0 - std::pow(beta1, (double)step);
        const double bc2 = 1.0 - std::pow(beta2, (double)step);
        const double step_size = lr / bc1;
        adam_rows_kernel<true><<<blocks, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, relevant, visible, n_rows, width, magic,
                                                      (float)(-step_size), (float)beta1, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2),
                                                      (float)eps, (float)std::sqrt(bc2));
    } else {
        adam_rows_kernel<false><<<blocks, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, relevant, visible, n_rows, width, magic, (float)lr,
                                                       (float)beta1, (float)beta2, 1.0f - (float)beta1, 1.0f - (float)beta2, (float)eps, 1.0f);
    }
    LODGS_LAUNCHED("adam_rows_kernel");
    return LODGS_OK;
}

// All parameter tensors of one optimizer step in one call: the per-tensor launches are issued back to back from C, so the step
// costs one foreign call instead of one per tensor (the five small tensors of a Gaussian model take ~10 us of GPU time each and
// were bound by the Python-side cost of their calls). Every tensor shares the row list / visibility mask; rows, width, lr, betas,
// eps and step are per tensor.
extern "C" int lodgs_adam_step_multi(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avgs,
                                     float* const* exp_avg_sqs, const int64_t* relevant, int64_t n_relevant, const uint8_t* visible,
                                     const int64_t* rows, const int32_t* widths, const double* lrs, const double* betas1, const double* betas2,
                                     const double* epss, const int64_t* steps, int bias_correction, void* stream) {
    if (n_tensors < 0 || (n_tensors > 0 && (!params || !grads || !exp_avgs || !exp_avg_sqs || !rows || !widths || !lrs || !betas1 || !betas2 || !epss ||
                                            !steps))) {
        set_error("lodgs_adam_step_multi: bad argument");
        return LODGS_E_ARG;
    }
    for (int t = 0; t < n_tensors; t++) {
        const int rc = lodgs_adam_step(params[t], grads[t], exp_avgs[t], exp_avg_sqs[t], relevant, n_relevant, visible, rows[t], widths[t], lrs[t],
                                       betas1[t], betas2[t], epss[t], steps[t], bias_correction, stream);
        if (rc) return rc;
    }
    return LODGS_OK;
}
