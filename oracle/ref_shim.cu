// ref_shim.cu — TEST INFRASTRUCTURE ONLY. C-ABI shim around the UNMODIFIED reference CUDA sources, which are
// compiled in place from /root/reference by oracle/Makefile into oracle/_ref/libref_lodgs.so (never copied
// into this repo). It exists so that the GPU parity tests and bench.py --impl reference can run the
// reference's own kernels on the same B200:
//   CudaRasterizer::Rasterizer::forward/backward  (submodules/hierarchy-rasterizer/cuda_rasterizer/rasterizer.h:20-109)
//   Switching::expandToSize / getTsIndexed        (submodules/gaussianhierarchy/runtime_switching.h:17-66)
// Scratch buffers are owned here (cudaMalloc, grown on demand) and their arrays are exposed through the
// reference's own fromChunk layout (rasterizer_impl.h) for integer-exact comparisons.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"
#include "runtime_switching.h"
#include "morton.h"

namespace {
struct Buf {
    char* p = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (p) cudaFree(p);
            cap = n + n / 4 + 256;
            if (cudaMalloc(&p, cap) != cudaSuccess) { p = nullptr; cap = 0; fprintf(stderr, "ref_shim: cudaMalloc(%zu) failed\n", n); }
        }
        return p;
    }
};
Buf g_geom, g_bin, g_img;
std::function<char*(size_t)> fn(Buf& b) { return [&b](size_t n) { return b.get(n); }; }
}  // namespace

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API int ref_forward(int P, int D, int M, const float* background, int W, int H, const int* indices, const int* parent_indices,
                        const float* ts, const int* kids, const float* means3D, const float* shs, const float* colors_precomp,
                        const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                        const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                        float tan_fovx, float tan_fovy, int prefiltered, float* out_color, float* out_invdepth, int* radii, int* rects,
                        int debug, int* seen) {
    return CudaRasterizer::Rasterizer::forward(fn(g_geom), fn(g_bin), fn(g_img), P, D, M, background, W, H, indices, parent_indices, ts,
                                               kids, means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations,
                                               cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered != 0,
                                               out_color, out_invdepth, radii, rects, nullptr, nullptr, debug != 0, seen);
}

REF_API void ref_backward(int P, int D, int M, int R, const float* background, int W, int H, const int* indices,
                          const int* parent_indices, const float* ts, const int* kids, const float* means3D, const float* shs,
                          const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tan_fovx, float tan_fovy, const int* radii, const float* dL_dpix,
                          const float* dL_invdepths, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                          float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                          int debug) {
    CudaRasterizer::Rasterizer::backward(P, D, M, R, background, W, H, indices, parent_indices, ts, kids, means3D, shs, colors_precomp,
                                         opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
                                         tan_fovx, tan_fovy, radii, g_geom.p, g_bin.p, g_img.p, dL_dpix, dL_invdepths, dL_dmean2D,
                                         dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
                                         dL_drot, debug != 0);
}

// Device pointers of the reference's state arrays (valid after ref_forward with the same P / R / W*H).
// geom: depths, clamped, p_clamped, internal_radii, means2D, cov3D, conic_opacity, rgb, tiles_touched, point_offsets
REF_API void ref_geom_ptrs(int P, void** out10) {
    char* c = g_geom.p;
    CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c, P);
    out10[0] = g.depths; out10[1] = g.clamped; out10[2] = g.p_clamped; out10[3] = g.internal_radii; out10[4] = g.means2D;
    out10[5] = g.cov3D; out10[6] = g.conic_opacity; out10[7] = g.rgb; out10[8] = g.tiles_touched; out10[9] = g.point_offsets;
}
// binning: keys_unsorted, values_unsorted, keys_sorted, point_list
REF_API void ref_binning_ptrs(int R, void** out4) {
    char* c = g_bin.p;
    CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c, R);
    out4[0] = b.point_list_keys_unsorted; out4[1] = b.point_list_unsorted; out4[2] = b.point_list_keys; out4[3] = b.point_list;
}
// image: accum_alpha (final T), n_contrib, ranges
REF_API void ref_image_ptrs(int WH, void** out3) {
    char* c = g_img.p;
    CudaRasterizer::ImageState s = CudaRasterizer::ImageState::fromChunk(c, WH);
    out3[0] = s.accum_alpha; out3[1] = s.n_contrib; out3[2] = s.ranges;
}

REF_API int ref_expand_to_size(int N, float target_size, int* nodes, float* boxes, float* viewpoint_dev, float zx, float zy, float zz,
                               int* render_indices, int* parent_indices, int* nodes_for_render_indices) {
    return Switching::expandToSize(N, target_size, nodes, boxes, viewpoint_dev, zx, zy, zz, render_indices, nullptr, parent_indices,
                                   nodes_for_render_indices);
}

REF_API void ref_get_ts_indexed(int K, int* indices, float target_size, int* nodes, float* boxes, float vx, float vy, float vz, float zx,
                                float zy, float zz, float* ts, int* kids, void* stream) {
    Switching::getTsIndexed(K, indices, target_size, nodes, boxes, vx, vy, vz, zx, zy, zz, ts, kids, stream);
}

REF_API int ref_expand_to_size_dynamic(int N, float target_size, int* nodes, float* positions, float* scales, float* viewpoint_dev, float zx,
                                       float zy, float zz, int* render_indices, int* parent_indices, int* nodes_for_render_indices) {
    return Switching::expandToSizeDynamic(N, target_size, nodes, positions, scales, viewpoint_dev, zx, zy, zz, render_indices, nullptr,
                                          parent_indices, nodes_for_render_indices);
}

REF_API void ref_get_ts_indexed_dynamic(int K, int* indices, float target_size, int* nodes, float* positions, float* scales, float vx,
                                        float vy, float vz, float zx, float zy, float zz, float* ts, int* kids, void* stream) {
    Switching::getTsIndexedDynamic(K, indices, target_size, nodes, positions, scales, vx, vy, vz, zx, zy, zz, ts, kids, stream);
}

// Switching::getSPTCut returns two device allocations; _run keeps them until _fetch copies them into the caller's buffers.
static int* g_spt_cut = nullptr;
static int* g_spt_counts = nullptr;
REF_API int ref_get_spt_cut_run(int S, int* gaussian_indices, int* SPT_starts, float* SPT_max, float* SPT_min, int* SPT_indices,
                                float* SPT_distances) {
    if (g_spt_cut) cudaFree(g_spt_cut);
    if (g_spt_counts) cudaFree(g_spt_counts);
    g_spt_cut = g_spt_counts = nullptr;
    return Switching::getSPTCut(S, gaussian_indices, SPT_starts, SPT_max, SPT_min, SPT_indices, SPT_distances, &g_spt_cut, &g_spt_counts);
}
REF_API int ref_get_spt_cut_fetch(int K, int S, int* cut_out, int* counts_out) {
    cudaError_t e = cudaSuccess;
    if (K > 0) e = cudaMemcpy(cut_out, g_spt_cut, (size_t)K * sizeof(int), cudaMemcpyDeviceToDevice);
    if (e == cudaSuccess && S > 0) e = cudaMemcpy(counts_out, g_spt_counts, (size_t)S * sizeof(int), cudaMemcpyDeviceToDevice);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (g_spt_cut) cudaFree(g_spt_cut);
    if (g_spt_counts) cudaFree(g_spt_counts);
    g_spt_cut = g_spt_counts = nullptr;
    return (int)e;
}

REF_API void ref_get_morton_code(float* xyz, float* mn, float* mx, int64_t* codes, int P) { MORTON::getMortonCode(xyz, mn, mx, codes, P); }

REF_API int ref_copy_d2d(void* dst, const void* src, size_t nbytes) {
    cudaError_t e = cudaMemcpy(dst, src, nbytes, cudaMemcpyDeviceToDevice);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    return (int)e;
}

REF_API void ref_free(void) {
    for (Buf* b : {&g_geom, &g_bin, &g_img}) {
        if (b->p) cudaFree(b->p);
        b->p = nullptr;
        b->cap = 0;
    }
}
