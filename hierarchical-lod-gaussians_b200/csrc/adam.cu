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
                                                       long long n_rows, int width, float lr_or_neg_step, float b1, float b2,
                                                       float omb1, float omb2, float eps, float bc2_sqrt) {
    const long long total = n_rows * width;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / width;
        const int c = (int)(e - r * width);
        long long row = r;
        if (relevant) row = __ldg(relevant + r);
        else if (visible && !__ldg(visible + r)) continue;
        const long long i = row * width + c;
        const float g = __ldg(grad + i);
        float m = exp_avg[i], v = exp_avg_sq[i];
        if (BIAS) {
            // torch: exp_avg.mul_(beta1).add_(grad, alpha=1-beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
            m = __fmaf_rn(g, omb1, __fmul_rn(m, b1));
            v = __fmaf_rn(__fmul_rn(g, g), omb2, __fmul_rn(v, b2));
            const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
            param[i] = __fmaf_rn(lr_or_neg_step, __fdiv_rn(m, denom), param[i]);
        } else {
            // AR adam.cu:26-33: no bias correction
            m = b1 * m + omb1 * g;
            v = b2 * v + omb2 * g * g;
            param[i] += -lr_or_neg_step * m / (sqrtf(v) + eps);
        }
        exp_avg[i] = m;
        exp_avg_sq[i] = v;
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
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    if (bias_correction) {
        if (step < 1) { set_error("lodgs_adam_step: step must be >= 1 with bias correction"); return LODGS_E_ARG; }
        // bias terms in double on the host, exactly like the Python floats in OurAdam.py:331-337
        const double bc1 = 1.0 - std::pow(beta1, (double)step);
        const double bc2 = 1.0 - std::pow(beta2, (double)step);
        const double step_size = lr / bc1;
        adam_rows_kernel<true><<<blocks, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, relevant, visible, n_rows, width,
                                                      (float)(-step_size), (float)beta1, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2),
                                                      (float)eps, (float)std::sqrt(bc2));
    } else {
        adam_rows_kernel<false><<<blocks, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, relevant, visible, n_rows, width, (float)lr,
                                                       (float)beta1, (float)beta2, 1.0f - (float)beta1, 1.0f - (float)beta2, (float)eps, 1.0f);
    }
    LODGS_LAUNCHED("adam_rows_kernel");
    return LODGS_OK;
}
