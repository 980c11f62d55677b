// relocation.cu — opacity / scale of a Gaussian that is split into N copies (MCMC relocation).
//
// Replaces UTILS::ComputeRelocation (HR/cuda_rasterizer/utils.cu:6-51, bound as _C.compute_relocation,
// HR/rasterize_points.cu ComputeRelocationCUDA): equation (9) of "3D Gaussian Splatting as Markov Chain Monte Carlo",
//   o_new = 1 - (1 - o)^(1/N),   s_new = s * o / sum_{i=1..N} sum_{k=0..i-1} binom(i-1,k) (-1)^k o_new^(k+1) / sqrt(k+1).
// One thread per Gaussian like the reference; the double sum is O(N^2) with N <= n_max (51 in the training scripts), so
// the kernel is latency-bound on the binomial table, which is staged in shared memory here. The mixed float/double
// arithmetic of the reference expression is kept term by term (float powi, double sign/sqrt quotient, float term and
// running sum) so that the results compare bit for bit.
#include "state.cuh"

namespace lodgs {

constexpr int RELOC_THREADS = 256;
constexpr int RELOC_MAX_TABLE = 64 * 64;  // binomial table entries staged in shared memory (n_max <= 64)

__global__ void __launch_bounds__(RELOC_THREADS) relocation_kernel(int P, const float* __restrict__ opacity_old,
                                                                   const float* __restrict__ scale_old, const int32_t* __restrict__ N,
                                                                   const float* __restrict__ binoms, int n_max, int staged,
                                                                   float* __restrict__ opacity_new, float* __restrict__ scale_new) {
    __shared__ float s_binoms[RELOC_MAX_TABLE];
    if (staged) {
        for (int i = threadIdx.x; i < n_max * n_max; i += RELOC_THREADS) s_binoms[i] = __ldg(binoms + i);
        __syncthreads();
    }
    const float* table = staged ? s_binoms : binoms;
    const int idx = blockIdx.x * RELOC_THREADS + threadIdx.x;
    if (idx >= P) return;
    const int n = __ldg(N + idx);
    const float o = __ldg(opacity_old + idx);
    const float o_new = 1.0f - powf(1.0f - o, 1.0f / n);
    opacity_new[idx] = o_new;
    float denom_sum = 0.0f;
    for (int i = 1; i <= n; ++i) {
        const float* row = table + (size_t)(i - 1) * n_max;
        for (int k = 0; k < i; ++k) {
            const double sign = (k & 1) ? -1.0 : 1.0;               // pow(-1, k) of the reference, exact
            const double weight = sign / sqrt((double)(k + 1));      // double quotient (the int arguments promote)
            const float term = (float)(weight * (double)pow(o_new, k + 1));  // float powi, widened for the product, stored as float
            denom_sum = __fmaf_rn(row[k], term, denom_sum);          // the reference's mul+add pair contracts to one FMA
        }
    }
    const float coeff = o / denom_sum;
#pragma unroll
    for (int c = 0; c < 3; ++c) scale_new[3 * (size_t)idx + c] = coeff * __ldg(scale_old + 3 * (size_t)idx + c);
}

}  // namespace lodgs

using namespace lodgs;

extern "C" int lodgs_compute_relocation(int P, const float* opacity_old, const float* scale_old, const int32_t* N, const float* binoms,
                                        int n_max, float* opacity_new, float* scale_new, void* stream) {
    if (P < 0 || n_max <= 0) { set_error("lodgs_compute_relocation: bad size argument"); return LODGS_E_ARG; }
    if (P == 0) return LODGS_OK;
    if (!opacity_old || !scale_old || !N || !binoms || !opacity_new || !scale_new) {
        set_error("lodgs_compute_relocation: null pointer argument");
        return LODGS_E_ARG;
    }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const int staged = n_max * n_max <= RELOC_MAX_TABLE ? 1 : 0;
    relocation_kernel<<<(P + RELOC_THREADS - 1) / RELOC_THREADS, RELOC_THREADS, 0, s>>>(P, opacity_old, scale_old, N, binoms, n_max, staged,
                                                                                      opacity_new, scale_new);
    LODGS_LAUNCHED("relocation_kernel");
    return LODGS_OK;
}
