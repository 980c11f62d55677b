// morton.cu — 63-bit Morton codes of the Gaussian positions (used to order Gaussians spatially before the hierarchy is built).
//
// Replaces MORTON::getMortonCode (GH/morton.cu:9-60), bound as gaussian_hierarchy._C.get_morton_indices
// (GH/torch/torch_interface.cpp GetMortonCode, called by scene/gaussian_model.py:574): every coordinate is normalised to the
// box [min, max], scaled by 2^21, truncated to a 64-bit integer, and bits 0..20 of x, y, z are interleaved (x lowest).
// The reference interleaves with a 63-step shift/or loop; here the low 21 bits are spread with the five-step magic-number
// dilation, which yields the same code for every input (bits above 20 are ignored by the loop too, so a coordinate equal to
// max wraps to 0 in both). The arithmetic before the conversion is the reference's: subtract, IEEE divide, multiply by 2^21.
// HBM-bound: 12 B read + 8 B written per point; a warp's 32 points are one contiguous 384-byte span.
#include "state.cuh"

namespace lodgs {

constexpr int MORTON_THREADS = 256;

__device__ __forceinline__ unsigned long long dilate21(long long v) {
    unsigned long long x = (unsigned long long)v & 0x1fffffull;
    x = (x | x << 32) & 0x001f00000000ffffull;
    x = (x | x << 16) & 0x001f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(MORTON_THREADS) morton_kernel(const float* __restrict__ xyz, const float* __restrict__ mn3,
                                                               const float* __restrict__ mx3, int64_t* __restrict__ codes, int P) {
    const int idx = blockIdx.x * MORTON_THREADS + threadIdx.x;
    if (idx >= P) return;
    const float mnx = __ldg(mn3), mny = __ldg(mn3 + 1), mnz = __ldg(mn3 + 2);
    const float bx = __fsub_rn(__ldg(mx3), mnx), by = __fsub_rn(__ldg(mx3 + 1), mny), bz = __fsub_rn(__ldg(mx3 + 2), mnz);
    const float* p = xyz + 3 * (size_t)idx;
    const float x = __fmul_rn(__fdiv_rn(__fsub_rn(__ldg(p), mnx), bx), 2097152.0f);
    const float y = __fmul_rn(__fdiv_rn(__fsub_rn(__ldg(p + 1), mny), by), 2097152.0f);
    const float z = __fmul_rn(__fdiv_rn(__fsub_rn(__ldg(p + 2), mnz), bz), 2097152.0f);
    const unsigned long long code = dilate21((long long)x) | dilate21((long long)y) << 1 | dilate21((long long)z) << 2;
    codes[idx] = (int64_t)code;
}

}  // namespace lodgs

using namespace lodgs;

extern "C" int lodgs_get_morton_codes(const float* xyz, const float* min3_dev, const float* max3_dev, int64_t* codes, int P, void* stream) {
    if (P < 0) { set_error("lodgs_get_morton_codes: bad size argument"); return LODGS_E_ARG; }
    if (P == 0) return LODGS_OK;
    if (!xyz || !min3_dev || !max3_dev || !codes) { set_error("lodgs_get_morton_codes: null pointer argument"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    morton_kernel<<<(P + MORTON_THREADS - 1) / MORTON_THREADS, MORTON_THREADS, 0, s>>>(xyz, min3_dev, max3_dev, codes, P);
    LODGS_LAUNCHED("morton_kernel");
    return LODGS_OK;
}
