// raster_api.cu — C ABI orchestration of the forward render stage and the layout accessors.
// Mirrors the tail of CudaRasterizer::Rasterizer::forward (HR/cuda_rasterizer/rasterizer_impl.cu:334-400).
#include "state.cuh"

namespace lodgs {
int launch_emit_balanced(int P, int R, const GeomState& g, unsigned gx, uint32_t* tiles, uint32_t* vals, cudaStream_t s);
int sort_pairs32_device(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t* keys_tmp,
                        uint32_t* vals_tmp, int n, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                        const char* hist_name);
int launch_tile_ranges(int R, const BinningState& b, uint2* ranges, size_t ntiles, cudaStream_t s);
int launch_tile_order(const uint2* ranges, uint32_t* order, size_t ntiles, cudaStream_t s);
int launch_lod_patch(int P, const GeomState& g, const float* ts, const int32_t* kids, cudaStream_t s);
int launch_sorted_keys(int R, const BinningState& b, const GeomState& g, uint64_t* keys, cudaStream_t s);
int sort_pairs_device(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, uint64_t* keys_tmp,
                      uint32_t* vals_tmp, int n, int end_bit, void* ws, cudaStream_t s);
int launch_blend_fwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                     const float* bg, float* out_color, float* out_invdepth, int32_t* seen, cudaStream_t s);

// Same as getHigherMsb (rasterizer_impl.cu:35-50): number of bits needed for the tile id.
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step;
        else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}
}  // namespace lodgs

using namespace lodgs;

// The binning buffer additionally holds the sort's ping-pong pair; it is carved after the public arrays.
struct BinningFull {
    BinningState b;
    uint32_t* keys_tmp;
    uint32_t* vals_tmp;
    size_t bytes;
};
static BinningFull binning_full(void* chunk, size_t R) {
    BinningFull f;
    f.b = binning_from_chunk(chunk, R);
    char* base = (char*)chunk;
    size_t off = f.b.bytes;
    f.keys_tmp = carve_arr<uint32_t>(base, off, R);
    f.vals_tmp = carve_arr<uint32_t>(base, off, R);
    f.bytes = align_up(off, 128);
    return f;
}

extern "C" size_t lodgs_image_bytes(int W, int H) { return image_from_chunk(nullptr, W, H).bytes + 128; }
extern "C" size_t lodgs_binning_bytes(int R) { return binning_full(nullptr, R < 0 ? 0 : R).bytes + 128; }

extern "C" int lodgs_binning_layout(int R, size_t* offsets5) {
    BinningState b = binning_from_chunk(nullptr, R);
    for (int i = 0; i < 5; i++) offsets5[i] = b.offsets[i];
    return LODGS_OK;
}
extern "C" int lodgs_image_layout(int W, int H, size_t* offsets3) {
    ImageState s = image_from_chunk(nullptr, W, H);
    for (int i = 0; i < 3; i++) offsets3[i] = s.offsets[i];
    return LODGS_OK;
}

extern "C" int lodgs_forward_render(int P, int W, int H, int R, const float* background, const float* ts, const int32_t* kids,
                                    const float* colors_precomp, const int32_t* radii, void* geom_buffer, void* binning_buffer,
                                    size_t binning_bytes, void* image_buffer, size_t image_bytes, float* out_color,
                                    float* out_invdepth, int32_t* seen, void* stream) {
    if (P < 0 || W <= 0 || H <= 0 || R < 0) { set_error("lodgs_forward_render: bad size argument"); return LODGS_E_ARG; }
    if (!background || !out_color || !image_buffer) { set_error("lodgs_forward_render: null pointer argument"); return LODGS_E_ARG; }
    if (image_bytes < lodgs_image_bytes(W, H) - 128) { set_error("lodgs_forward_render: image buffer too small"); return LODGS_E_WORKSPACE; }
    if (((uintptr_t)image_buffer & 127) != 0) { set_error("lodgs_forward_render: image buffer must be 128-byte aligned"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    ImageState img = image_from_chunk(image_buffer, W, H);
    const unsigned gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    (void)colors_precomp;  // already folded into the blend records by stage 1
    BinningFull bf;
    bf.b = binning_from_chunk(nullptr, 0);
    const SplatRec* recs = nullptr;
    if (R > 0) {
        if (!geom_buffer || !binning_buffer || !radii || P == 0) { set_error("lodgs_forward_render: null pointer argument"); return LODGS_E_ARG; }
        if (binning_bytes < lodgs_binning_bytes(R) - 128) { set_error("lodgs_forward_render: binning buffer too small"); return LODGS_E_WORKSPACE; }
        if (((uintptr_t)binning_buffer & 127) != 0) { set_error("lodgs_forward_render: binning buffer must be 128-byte aligned"); return LODGS_E_ARG; }
        GeomState g = geom_from_chunk(geom_buffer, P);
        bf = binning_full(binning_buffer, R);
        // stage 1 left the Gaussians ordered by (depth bits, index) in g.order and their tile counts scanned in that order
        int rc = launch_emit_balanced(P, R, g, gx, bf.b.tiles_unsorted, bf.b.vals_unsorted, s);
        if (rc) return rc;
        const int bit = (int)higher_msb(gx * gy);
        rc = sort_pairs32_device(bf.b.tiles_unsorted, bf.b.vals_unsorted, bf.b.tiles_sorted, bf.b.point_list, bf.keys_tmp, bf.vals_tmp, R, 0,
                                 bit, bf.b.sort_ws, s, "sort_tile_pass_kernel", "sort_tile_hist_kernel");
        if (rc) return rc;
        rc = launch_tile_ranges(R, bf.b, img.ranges, (size_t)gx * gy, s);
        if (rc) return rc;
        rc = launch_tile_order(img.ranges, img.tile_order, (size_t)gx * gy, s);
        if (rc) return rc;
        if (ts != nullptr && kids != nullptr) {
            rc = launch_lod_patch(P, g, ts, kids, s);
            if (rc) return rc;
        }
        recs = g.recs;
    } else {
        // The reference returns before rendering when nothing is visible (rasterizer_impl.cu:332-333), leaving the
        // zero-filled image; the blend below runs with empty ranges and writes the background instead. Keep the
        // reference behaviour: leave out_color untouched.
        return LODGS_OK;
    }
    const bool lod = ts != nullptr;  // kids == NULL with ts: stage 1 (lodgs_forward_preprocess_lod) already stored t and 1/kids
    return launch_blend_fwd(lod, out_invdepth != nullptr, img, bf.b, recs, W, H, background, out_color, out_invdepth, seen, s);
}

// ------------------------------------------------------------------------------------------------ backward
namespace lodgs {
struct PreBwdParams;
int launch_blend_bwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                     const float* bg, float bg_mult, float null_thr, const int32_t* indices, const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, cudaStream_t s);
int run_preprocess_bwd(int P, int D, int M, const int32_t* indices, const int32_t* parent_indices, const float* ts, const float* means3D,
                       const int32_t* radii, const float* shs, const uint8_t* clamped, const float* opacities, const float* scales,
                       const float* rotations, float scale_modifier, const float* cov3Ds, int cov3D_by_global, const float* viewmatrix,
                       const float* projmatrix, const float* campos, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                       const float* dL_dmean2D, const float* dL_dconic, const float* dL_dinvdepth, const float* dL_dcolor,
                       float* dL_dopacity, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                       float* dL_drgb, unsigned flags, cudaStream_t s);
}  // namespace lodgs

extern "C" int lodgs_backward(int P, int fullP, int D, int M, int R, int W, int H, const float* background, const int32_t* indices,
                              const int32_t* parent_indices, const float* ts, const int32_t* kids, const float* means3D,
                              const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                              float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                              const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int32_t* radii,
                              void* geom_buffer, void* binning_buffer, void* image_buffer, const float* dL_dpix,
                              const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                              float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                              void* stream) {
    return lodgs_backward_ex(P, fullP, D, M, R, W, H, background, indices, parent_indices, ts, kids, means3D, shs, colors_precomp, opacities,
                             scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
                             geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_dinvdepth_pix, dL_dmean2D, dL_dconic, dL_dopacity,
                             dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, 0u, stream);
}

extern "C" int lodgs_backward_ex(int P, int fullP, int D, int M, int R, int W, int H, const float* background, const int32_t* indices,
                                 const int32_t* parent_indices, const float* ts, const int32_t* kids, const float* means3D,
                                 const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                                 float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                 const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int32_t* radii,
                                 void* geom_buffer, void* binning_buffer, void* image_buffer, const float* dL_dpix,
                                 const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                 float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                 unsigned flags, void* stream) {
    return lodgs_backward_rgb(P, fullP, D, M, R, W, H, background, indices, parent_indices, ts, kids, means3D, shs, colors_precomp, opacities,
                              scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
                              geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_dinvdepth_pix, dL_dmean2D, dL_dconic, dL_dopacity,
                              dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, flags, stream);
}

extern "C" int lodgs_backward_rgb(int P, int fullP, int D, int M, int R, int W, int H, const float* background, const int32_t* indices,
                                  const int32_t* parent_indices, const float* ts, const int32_t* kids, const float* means3D,
                                  const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                                  float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                  const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int32_t* radii,
                                  void* geom_buffer, void* binning_buffer, void* image_buffer, const float* dL_dpix,
                                  const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                  float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                  float* dL_drgb, unsigned flags, void* stream) {
    (void)colors_precomp;
    if (P < 0 || fullP < 0 || W <= 0 || H <= 0 || R < 0 || D < 0 || D > 3) { set_error("lodgs_backward: bad size argument"); return LODGS_E_ARG; }
    if (P == 0) return LODGS_OK;
    if (!background || !means3D || !opacities || !viewmatrix || !projmatrix || !campos || !radii || !geom_buffer || !image_buffer || !dL_dpix ||
        !dL_dmean2D || !dL_dconic || !dL_dopacity || !dL_dcolor || !dL_dmean3D || !dL_dcov3D) {
        set_error("lodgs_backward: null pointer argument");
        return LODGS_E_ARG;
    }
    if ((dL_dinvdepth_pix == nullptr) != (dL_dinvdepth == nullptr)) { set_error("lodgs_backward: dL_dinvdepth_pix and dL_dinvdepth go together"); return LODGS_E_ARG; }
    if (shs && !dL_dsh && !dL_drgb) { set_error("lodgs_backward: dL_dsh missing"); return LODGS_E_ARG; }
    if (scales && (!rotations || !dL_dscale || !dL_drot)) { set_error("lodgs_backward: scale/rotation gradient buffers missing"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    GeomState g = geom_from_chunk(geom_buffer, P);
    ImageState img = image_from_chunk(image_buffer, W, H);
    if (R > 0) {
        if (!binning_buffer) { set_error("lodgs_backward: null binning buffer"); return LODGS_E_ARG; }
        BinningState bin = binning_from_chunk(binning_buffer, R);
        const bool lod = ts != nullptr && kids != nullptr;
        int rc = launch_blend_bwd(lod, dL_dinvdepth_pix != nullptr, img, bin, g.recs, W, H, background,
                                  (flags & LODGS_RASTER_BACKGROUND_TWICE) ? 2.0f : 1.0f,
                                  (flags & LODGS_RASTER_CLAMPED_ALPHA_GRAD) ? __builtin_huge_valf() : 0.99f, indices, dL_dpix, dL_dinvdepth_pix,
                                  dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth, s);
        if (rc) return rc;
    }
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
    const float* cov3Ds = cov3D_precomp ? cov3D_precomp : g.cov3D;
    return run_preprocess_bwd(P, D, M, indices, parent_indices, ts, means3D, radii, shs, g.clamped, opacities, scales, rotations,
                              scale_modifier, cov3Ds, cov3D_precomp ? 1 : 0, viewmatrix, projmatrix, campos, focal_x, focal_y, tan_fovx,
                              tan_fovy, dL_dmean2D, dL_dconic, dL_dinvdepth, dL_dcolor, dL_dopacity, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
                              dL_drot, dL_drgb, flags, s);
}

namespace lodgs {
int launch_blend_stats(bool lod, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                       unsigned long long* counters, cudaStream_t s);
}
extern "C" int lodgs_blend_stats(int P, int W, int H, int R, int lod, void* geom_buffer, void* binning_buffer, void* image_buffer,
                                 uint64_t* counters3_dev, void* stream) {
    if (P <= 0 || W <= 0 || H <= 0 || R <= 0 || !geom_buffer || !binning_buffer || !image_buffer || !counters3_dev) {
        set_error("lodgs_blend_stats: bad argument");
        return LODGS_E_ARG;
    }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    return launch_blend_stats(lod != 0, image_from_chunk(image_buffer, W, H), binning_from_chunk(binning_buffer, R),
                              geom_from_chunk(geom_buffer, P).recs, W, H, (unsigned long long*)counters3_dev, s);
}

extern "C" int lodgs_sorted_keys(int P, int R, void* geom_buffer, void* binning_buffer, uint64_t* keys_out, void* stream) {
    if (P <= 0 || R <= 0 || !geom_buffer || !binning_buffer || !keys_out) { set_error("lodgs_sorted_keys: bad argument"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    return launch_sorted_keys(R, binning_from_chunk(binning_buffer, R), geom_from_chunk(geom_buffer, P), keys_out, s);
}
