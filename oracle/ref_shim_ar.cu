// ref_shim_ar.cu — TEST INFRASTRUCTURE ONLY. C-ABI shim around the UNMODIFIED sources of the reference's second rasterizer
// (submodules/alt-rasterizer/cuda_rasterizer/{forward,backward,rasterizer_impl,adam,utils}.cu), compiled in place from
// /root/reference by oracle/Makefile into oracle/_ref/libref_ar.so (a separate library: it defines the same
// CudaRasterizer symbols as the hierarchy rasterizer). Used by tests/test_gpu_alt_rasterizer.py only.
//   CudaRasterizer::Rasterizer::forward/backward  (AR/cuda_rasterizer/rasterizer.h:33-96)
//   ADAM::adamUpdate (AR/cuda_rasterizer/adam.h), UTILS::ComputeRelocation (AR/cuda_rasterizer/utils.h)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <tuple>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/adam.h"
#include "cuda_rasterizer/utils.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {
struct Buf {
    char* p = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (p) cudaFree(p);
            cap = n + n / 4 + 256;
            if (cudaMalloc(&p, cap) != cudaSuccess) { p = nullptr; cap = 0; fprintf(stderr, "ref_shim_ar: cudaMalloc(%zu) failed\n", n); }
        }
        return p;
    }
};
Buf g_geom, g_bin, g_img, g_sample;
std::function<char*(size_t)> fn(Buf& b) { return [&b](size_t n) { return b.get(n); }; }
}  // namespace

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API int refar_forward(int P, int D, int M, const float* background, int W, int H, const float* means3D, const float* dc,
                          const float* shs, const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color, float* out_invdepth,
                          int antialiasing, int* radii, int debug, int* num_buckets) {
    auto t = CudaRasterizer::Rasterizer::forward(fn(g_geom), fn(g_bin), fn(g_img), fn(g_sample), P, D, M, background, W, H, means3D, dc, shs,
                                                 colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
                                                 projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered != 0, out_color, out_invdepth,
                                                 antialiasing != 0, radii, debug != 0);
    if (num_buckets) *num_buckets = std::get<1>(t);
    return std::get<0>(t);
}

REF_API void refar_backward(int P, int D, int M, int R, int B, const float* background, int W, int H, const float* means3D, const float* dc,
                            const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                            float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                            const float* dL_dpix, const float* dL_invdepths, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                            float* dL_dcolor, float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_ddc, float* dL_dsh,
                            float* dL_dscale, float* dL_drot, int antialiasing, int debug) {
    CudaRasterizer::Rasterizer::backward(P, D, M, R, B, background, W, H, means3D, dc, shs, colors_precomp, opacities, scales, scale_modifier,
                                         rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, g_geom.p,
                                         g_bin.p, g_img.p, g_sample.p, dL_dpix, dL_invdepths, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                                         dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_ddc, dL_dsh, dL_dscale, dL_drot, antialiasing != 0,
                                         debug != 0);
}

REF_API void refar_adam(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const bool* visible, float lr, float b1,
                        float b2, float eps, unsigned N, unsigned M) {
    ADAM::adamUpdate(param, grad, exp_avg, exp_avg_sq, visible, lr, b1, b2, eps, N, M);
}

REF_API void refar_relocation(int P, float* opacity_old, float* scale_old, int* N, float* binoms, int n_max, float* opacity_new,
                              float* scale_new) {
    UTILS::ComputeRelocation(P, opacity_old, scale_old, N, binoms, n_max, opacity_new, scale_new);
}

// Device pointers into the reference's own scratch buffers of the LAST forward (layouts from its fromChunk functions,
// AR/cuda_rasterizer/rasterizer_impl.cu:249-305), for tests/gpu_diag_ar.py: out[0..] = ranges, point_list, n_contrib, max_contrib,
// bucket_offsets, bucket_to_tile, sampled_T, means2D, conic_opacity, rgb, accum_alpha (final T), sampled_ar, pixel_colors, depths.
REF_API void refar_internals(int P, int W, int H, int R, int B, void** out) {
    char* c = g_geom.p;
    auto geom = CudaRasterizer::GeometryState::fromChunk(c, P);
    c = g_bin.p;
    auto bin = CudaRasterizer::BinningState::fromChunk(c, R);
    c = g_img.p;
    auto img = CudaRasterizer::ImageState::fromChunk(c, (size_t)W * H);
    c = g_sample.p;
    auto smp = CudaRasterizer::SampleState::fromChunk(c, B);
    out[0] = img.ranges; out[1] = bin.point_list; out[2] = img.n_contrib; out[3] = img.max_contrib; out[4] = img.bucket_offsets;
    out[5] = smp.bucket_to_tile; out[6] = smp.T; out[7] = geom.means2D; out[8] = geom.conic_opacity; out[9] = geom.rgb;
    out[10] = img.accum_alpha; out[11] = smp.ar; out[12] = img.pixel_colors; out[13] = geom.depths;
}

REF_API int refar_copy(void* dst, const void* src, size_t nbytes) { return (int)cudaMemcpy(dst, src, nbytes, cudaMemcpyDeviceToDevice); }

REF_API void refar_free(void) {
    for (Buf* b : {&g_geom, &g_bin, &g_img, &g_sample}) {
        if (b->p) cudaFree(b->p);
        b->p = nullptr;
        b->cap = 0;
    }
}
