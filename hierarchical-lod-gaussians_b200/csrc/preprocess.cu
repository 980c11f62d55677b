// preprocess.cu — forward per-Gaussian stage: LOD parent interpolation, projection, EWA covariance -> conic,
// SH -> RGB, tile rectangle, fused with the prefix sum of tile counts.
//
// Replaces preprocessCUDA<3> (HR/cuda_rasterizer/forward.cu:218-445) and the total of the following
// cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:322): num_rendered is accumulated by one atomic per CTA, the
// per-Gaussian offsets are produced later in DEPTH order (binning.cu), which is the order instances are emitted in. SH coefficients (192 B per Gaussian at
// degree 3) and quaternions move as 16-byte vector loads.
#include "gaussmath.cuh"
#include <cstdlib>
#include "state.cuh"
#include <cfloat>

namespace lodgs {

int sort_pairs32_device(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t* keys_tmp,
                        uint32_t* vals_tmp, int n, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                        const char* hist_name);

int launch_scan_sorted(int P, const int* n_dev, const GeomState& g, cudaStream_t s);
int launch_depth_compact(int P, const GeomState& g, cudaStream_t s);
int sort_pairs32_device_n(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t* keys_tmp,
                          uint32_t* vals_tmp, int n, const int* n_dev, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                          const char* hist_name);

struct PreParams {
    int P, D, M, W, H;
    const int32_t* indices;
    const int32_t* parent_indices;
    const float* ts;
    const int32_t* kids;   // sibling counts: when given, the record's last two slots are stored as (t, 1/kids) right here
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int32_t* radii;
    unsigned grid_x, grid_y;
    int sh_vec4;  // 1 when shs is 16-byte aligned and 3*M is a multiple of 4
    int staged;   // 1 when the SH rows of a tile are staged through shared memory by one bulk copy
    unsigned flags;  // LODGS_RASTER_* (include/lodgs.h): alt-rasterizer conventions
};

// Load the (D+1)^2 RGB coefficient triples of Gaussian `idx`, lerped with parent `pidx` by t when has_parent
// (interp(), forward.cu:79-85: t*a + (1-t)*b per component).
__device__ __forceinline__ void load_sh(const PreParams& p, int idx, int pidx, bool has_parent, float t, v3* sh, const float4* sh_row,
                                        uint64_t* bar) {
    const int ncoef = (p.D + 1) * (p.D + 1);
    const int nfloat = ncoef * 3;
    float* f = reinterpret_cast<float*>(sh);
    if (sh_row) {  // staged by a bulk copy (see preprocess_fwd_kernel); never combined with LOD interpolation
        mbar_wait(bar, 0);
#pragma unroll
        for (int k = 0; k < 12; k++)
            if (k * 4 < nfloat) {
                const float4 a = sh_row[k];
                f[k * 4 + 0] = a.x; f[k * 4 + 1] = a.y; f[k * 4 + 2] = a.z; f[k * 4 + 3] = a.w;
            }
        return;
    }
    const float* base = p.shs + (size_t)idx * p.M * 3;
    const float* pbase = p.shs + (size_t)pidx * p.M * 3;
    const float omt = 1.0f - t;
    if (p.sh_vec4) {
#pragma unroll
        for (int k = 0; k < 12; k++) {
            if (k * 4 < nfloat) {
                float4 a = __ldg(reinterpret_cast<const float4*>(base) + k);
                if (has_parent) {
                    float4 b = __ldg(reinterpret_cast<const float4*>(pbase) + k);
                    a.x = t * a.x + omt * b.x;
                    a.y = t * a.y + omt * b.y;
                    a.z = t * a.z + omt * b.z;
                    a.w = t * a.w + omt * b.w;
                }
                f[k * 4 + 0] = a.x; f[k * 4 + 1] = a.y; f[k * 4 + 2] = a.z; f[k * 4 + 3] = a.w;
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < 48; k++) {
            if (k < nfloat) {
                float a = __ldg(base + k);
                if (has_parent) a = t * a + omt * __ldg(pbase + k);
                f[k] = a;
            }
        }
    }
}

// Everything preprocessCUDA does for one Gaussian; returns tiles_touched (0 = culled).
__device__ __forceinline__ uint32_t preprocess_one(const PreParams& p, const GeomState& g, int t_idx, const float* view,
                                                   const float* proj, const float4* sh_row, uint64_t* bar, uint2& box,
                                                   unsigned& depth_bits) {
    const int r_idx = p.indices ? __ldg(p.indices + t_idx) : t_idx;
    p.radii[t_idx] = 0;

    int p_idx = 0;
    float t = 0;
    bool has_parent = false;
    float3 p_orig = {__ldg(p.means3D + 3 * (size_t)r_idx), __ldg(p.means3D + 3 * (size_t)r_idx + 1), __ldg(p.means3D + 3 * (size_t)r_idx + 2)};
    const float3 mean_own = p_orig;
    if (p.parent_indices) {
        p_idx = __ldg(p.parent_indices + t_idx);
        if (p_idx != -1) {
            has_parent = true;
            t = __ldg(p.ts + t_idx);
        }
    }
    if (has_parent) {
        float3 pa = {__ldg(p.means3D + 3 * (size_t)p_idx), __ldg(p.means3D + 3 * (size_t)p_idx + 1), __ldg(p.means3D + 3 * (size_t)p_idx + 2)};
        p_orig = {t * p_orig.x + (1.0f - t) * pa.x, t * p_orig.y + (1.0f - t) * pa.y, t * p_orig.z + (1.0f - t) * pa.z};
    }

    float4 p_hom = xform4x4(p_orig, proj);
    float p_w = 1.0f / (p_hom.w + 0.0000001f);
    float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
    float3 p_view = xform4x3(p_orig, view);
    if (p_view.z <= 0.2f) return 0;
    // box cull with the default +-FLT_MAX box (rasterizer_impl.cu:262-268): only non-finite means are dropped
    if (p_orig.x < -FLT_MAX || p_orig.y < -FLT_MAX || p_orig.z < -FLT_MAX || p_orig.x > FLT_MAX || p_orig.y > FLT_MAX ||
        p_orig.z > FLT_MAX)
        return 0;

    float cov6[6];
    if (p.cov3D_precomp == nullptr) {
        float sx = __ldg(p.scales + 3 * (size_t)r_idx), sy = __ldg(p.scales + 3 * (size_t)r_idx + 1), sz = __ldg(p.scales + 3 * (size_t)r_idx + 2);
        float4 rot = __ldg(reinterpret_cast<const float4*>(p.rotations) + r_idx);
        if (has_parent) {
            float px = __ldg(p.scales + 3 * (size_t)p_idx), py = __ldg(p.scales + 3 * (size_t)p_idx + 1), pz = __ldg(p.scales + 3 * (size_t)p_idx + 2);
            sx = t * sx + (1.0f - t) * px;
            sy = t * sy + (1.0f - t) * py;
            sz = t * sz + (1.0f - t) * pz;
            float4 o = __ldg(reinterpret_cast<const float4*>(p.rotations) + p_idx);
            float dot = (rot.x * o.x + rot.y * o.y) + (rot.z * o.z + rot.w * o.w);
            if (dot < 0.0) { o.x = -o.x; o.y = -o.y; o.z = -o.z; o.w = -o.w; }
            rot = {t * rot.x + (1.0f - t) * o.x, t * rot.y + (1.0f - t) * o.y, t * rot.z + (1.0f - t) * o.z, t * rot.w + (1.0f - t) * o.w};
        }
        cov3d_from_scale_rot(sx, sy, sz, p.scale_modifier, rot, cov6);
    } else {
#pragma unroll
        for (int k = 0; k < 6; k++) cov6[k] = __ldg(p.cov3D_precomp + 6 * (size_t)r_idx + k);
    }
#pragma unroll
    for (int k = 0; k < 6; k++) g.cov3D[6 * (size_t)t_idx + k] = cov6[k];

    float3 cov = project_cov(p_orig, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov6, view, nullptr);

    // (forward.cu:362-392) same pinned rounding as gaussmath.cuh: the determinants are single FMAs in the reference SASS
    constexpr float h_var = 0.3f;
    const float cov01sq = FM(cov.y, cov.y);
    const float det_cov = FF(cov.x, cov.z, -cov01sq);
    cov.x = FA(cov.x, h_var);
    cov.z = FA(cov.z, h_var);
    const float det_cov_plus_h_cov = FF(cov.x, cov.z, -cov01sq);
    // AR/cuda_rasterizer/forward.cu:226-229: the alt-rasterizer applies the opacity compensation only with antialiasing=True
    const float h_convolution_scaling =
        (p.flags & LODGS_RASTER_NO_ANTIALIAS) ? 1.0f : __fsqrt_rn(fmaxf(__fdiv_rn(det_cov, det_cov_plus_h_cov), 0.000025f));
    const float det = det_cov_plus_h_cov;
    if (det == 0.0f) return 0;
    const float det_inv = __frcp_rn(det);
    const float3 conic = {FM(cov.z, det_inv), FM(det_inv, -cov.y), FM(cov.x, det_inv)};

    const float mid = FM(FA(cov.x, cov.z), 0.5f);
    const float disc = __fsqrt_rn(fmaxf(FF(mid, mid, -det), 0.1f));
    const float lambda1 = FA(mid, disc);
    const float lambda2 = FS(mid, disc);
    const float my_radius = ceilf(FM(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.f));
    float2 point_image = {ndc_to_pix(p_proj.x, p.W), ndc_to_pix(p_proj.y, p.H)};
    uint2 rmin, rmax;
    const int2 my_rect = {(int)ceilf(FM(__fsqrt_rn(cov.x), 3.f)), (int)ceilf(FM(__fsqrt_rn(cov.z), 3.f))};
    g.rects[t_idx] = my_rect;
    // AR/cuda_rasterizer/forward.cu:249 bounds the splat by the square of its largest 3-sigma radius instead of the per-axis rectangle
    const int2 ext = (p.flags & LODGS_RASTER_SQUARE_RECT) ? make_int2((int)my_radius, (int)my_radius) : my_rect;
    tile_rect(point_image, ext, rmin, rmax, p.grid_x, p.grid_y);
    if ((rmax.x - rmin.x) * (rmax.y - rmin.y) == 0) return 0;
    box = make_uint2(rmin.x | (rmin.y << 16), rmax.x | (rmax.y << 16));

    float3 rgb;
    if (p.colors_precomp == nullptr) {
        // view direction from the UN-interpolated mean of r_idx (forward.cu:92: means[idx])
        float cx = __ldg(p.campos), cy = __ldg(p.campos + 1), cz = __ldg(p.campos + 2);
        float dx = mean_own.x - cx, dy = mean_own.y - cy, dz = mean_own.z - cz;
        float len = sqrtf(dx * dx + dy * dy + dz * dz);
        dx = dx / len; dy = dy / len; dz = dz / len;
        v3 sh[16];
        load_sh(p, r_idx, p_idx, has_parent, t, sh, sh_row, bar);
        v3 result = sh_to_rgb(p.D, sh, dx, dy, dz);
        result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
        uint8_t cl = (result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0);
        g.clamped[t_idx] = cl;
        rgb = {max(result.x, 0.0f), max(result.y, 0.0f), max(result.z, 0.0f)};
    } else {
        rgb = {__ldg(p.colors_precomp + 3 * (size_t)t_idx), __ldg(p.colors_precomp + 3 * (size_t)t_idx + 1),
               __ldg(p.colors_precomp + 3 * (size_t)t_idx + 2)};
    }

    depth_bits = __float_as_uint(p_view.z);
    p.radii[t_idx] = (int)my_radius;
    float opacity = __ldg(p.opacities + r_idx);
    if (has_parent) opacity = t * opacity + (1.0f - t) * __ldg(p.opacities + p_idx);
    const float opac = FM(h_convolution_scaling, opacity);
    // The blend record (48 B). Without LOD the t slot carries a conservative cut-off for the quadratic form: power < cut
    // => opacity * exp(power) < 1/255, so the blend kernels skip the exponential for such pairs (margin >> the 2-ulp error
    // of expf); with LOD, lod_patch_kernel overwrites the last two slots with t and 1/kids.
    const float lim = -logf(255.0f * opac);
    float4* out = reinterpret_cast<float4*>(g.recs + t_idx);
    out[0] = {point_image.x, point_image.y, conic.x, conic.y};
    out[1] = {conic.z, opac, rgb.x, rgb.y};
    if (p.kids) out[2] = {rgb.z, __frcp_rn(p_view.z), __ldg(p.ts + t_idx), 1.0f / (float)__ldg(p.kids + t_idx)};
    else out[2] = {rgb.z, __frcp_rn(p_view.z), lim - 1e-4f * (1.0f + fabsf(lim)), 1.0f};  // rcp.rn == the reference's 1/depth per pair
    return (rmax.y - rmin.y) * (rmax.x - rmin.x);
}

__global__ void __launch_bounds__(PRE_BLOCK) preprocess_fwd_kernel(const PreParams p, const GeomState g) {
    __shared__ float s_view[16], s_proj[16];
    __shared__ uint64_t s_bar;
    extern __shared__ __align__(128) float4 sh_stage[];  // [256][3M/4] when p.staged
    const int row4 = (p.M * 3) / 4;
    const int tile = blockIdx.x;
    if (threadIdx.x == 0) {
        if (p.staged) {
            // the tile's SH rows are one contiguous block: a single bulk copy (TMA) that lands while the projection math runs
            mbar_init(&s_bar, 1);
            mbar_fence_init();
            const int cnt = min(PRE_BLOCK, p.P - tile * PRE_BLOCK);
            const unsigned bytes = (unsigned)cnt * (unsigned)row4 * 16u;
            mbar_expect_tx(&s_bar, bytes);
            bulk_g2s(sh_stage, p.shs + (size_t)tile * PRE_BLOCK * p.M * 3, bytes, &s_bar);
        }
    }
    if (threadIdx.x < 16) s_view[threadIdx.x] = __ldg(p.viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = __ldg(p.projmatrix + threadIdx.x - 16);
    __syncthreads();
    const int t_idx = tile * PRE_BLOCK + threadIdx.x;
    unsigned tiles = 0;
    if (t_idx < p.P) {
        uint2 box = make_uint2(0u, 0u);
        unsigned depth_bits = 0;
        tiles = preprocess_one(p, g, t_idx, s_view, s_proj, p.staged ? sh_stage + threadIdx.x * row4 : nullptr, &s_bar, box, depth_bits);
        g.tiles_touched[t_idx] = tiles;
        g.tile_box[t_idx] = tiles ? box : make_uint2(0u, 0u);
        // sort key of the depth-ordered emission: culled Gaussians go to the end of the order and emit nothing
        g.depth_keys[t_idx] = tiles ? depth_bits : 0xffffffffu;
    }
    if (p.staged && threadIdx.x == 0) mbar_wait(&s_bar, 0);  // the copy must have landed before the CTA may retire
    // num_rendered = sum of the tile counts: one atomic per CTA. The per-Gaussian prefix sum of the reference
    // (InclusiveSum, rasterizer_impl.cu:322) is not needed here — instances are emitted in depth order from a scan in
    // THAT order (scan_sorted_kernel) — and a fused look-back scan chained all 3907 CTAs of this kernel for nothing.
    const unsigned lane = threadIdx.x & 31;
    unsigned sum = tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0 && sum) atomicAdd(&g.hdr->num_rendered, sum);
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means, const float* __restrict__ view, uint8_t* present) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float3 m = {__ldg(means + 3 * (size_t)i), __ldg(means + 3 * (size_t)i + 1), __ldg(means + 3 * (size_t)i + 2)};
    float z = view[2] * m.x + view[6] * m.y + view[10] * m.z + view[14];
    present[i] = !(z <= 0.2f);
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_geom_bytes(int P) { return geom_from_chunk(nullptr, P < 0 ? 0 : P).bytes + 128; }

extern "C" int lodgs_geom_layout(int P, size_t* offsets9) {
    GeomState g = geom_from_chunk(nullptr, P);
    for (int i = 0; i < 9; i++) offsets9[i] = g.offsets[i];
    return LODGS_OK;
}

extern "C" int lodgs_forward_preprocess(int P, int D, int M, int W, int H, const int32_t* indices, const int32_t* parent_indices,
                                        const float* ts, const float* means3D, const float* shs, const float* colors_precomp,
                                        const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                                        const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                        const float* campos, float tan_fovx, float tan_fovy, int32_t* radii, void* geom_buffer,
                                        size_t geom_bytes, int32_t* num_rendered_host, void* stream) {
    return lodgs_forward_preprocess_ex(P, D, M, W, H, indices, parent_indices, ts, means3D, shs, colors_precomp, opacities, scales,
                                       scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
                                       geom_buffer, geom_bytes, num_rendered_host, 0u, stream);
}

extern "C" int lodgs_forward_preprocess_ex(int P, int D, int M, int W, int H, const int32_t* indices, const int32_t* parent_indices,
                                           const float* ts, const float* means3D, const float* shs, const float* colors_precomp,
                                           const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                                           const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                           const float* campos, float tan_fovx, float tan_fovy, int32_t* radii, void* geom_buffer,
                                           size_t geom_bytes, int32_t* num_rendered_host, unsigned flags, void* stream) {
    return lodgs_forward_preprocess_lod(P, D, M, W, H, indices, parent_indices, ts, nullptr, means3D, shs, colors_precomp, opacities, scales,
                                        scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
                                        geom_buffer, geom_bytes, num_rendered_host, flags, stream);
}

extern "C" int lodgs_forward_preprocess_lod(int P, int D, int M, int W, int H, const int32_t* indices, const int32_t* parent_indices,
                                            const float* ts, const int32_t* kids, const float* means3D, const float* shs,
                                            const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                                            const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                            const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, int32_t* radii,
                                            void* geom_buffer, size_t geom_bytes, int32_t* num_rendered_host, unsigned flags, void* stream) {
    if (kids && !ts) { set_error("lodgs_forward_preprocess_lod: kids without ts"); return LODGS_E_ARG; }
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) { set_error("lodgs_forward_preprocess: bad size argument"); return LODGS_E_ARG; }
    if (P == 0) { if (num_rendered_host) *num_rendered_host = 0; return LODGS_OK; }
    if (!means3D || !opacities || !viewmatrix || !projmatrix || !campos || !radii || !geom_buffer) {
        set_error("lodgs_forward_preprocess: null pointer argument"); return LODGS_E_ARG;
    }
    if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("lodgs_forward_preprocess: provide exactly one of shs / colors_precomp"); return LODGS_E_ARG; }
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) { set_error("lodgs_forward_preprocess: provide exactly one of scales+rotations / cov3D_precomp"); return LODGS_E_ARG; }
    if (shs && M < (D + 1) * (D + 1)) { set_error("lodgs_forward_preprocess: M=%d too small for degree %d", M, D); return LODGS_E_ARG; }
    if (parent_indices && !ts) { set_error("lodgs_forward_preprocess: parent_indices without ts"); return LODGS_E_ARG; }
    if (geom_bytes < lodgs_geom_bytes(P) - 128) { set_error("lodgs_forward_preprocess: geometry buffer too small"); return LODGS_E_WORKSPACE; }
    if (((uintptr_t)geom_buffer & 127) != 0) { set_error("lodgs_forward_preprocess: geometry buffer must be 128-byte aligned"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    GeomState g = geom_from_chunk(geom_buffer, P);
    PreParams p;
    p.P = P; p.D = D; p.M = M; p.W = W; p.H = H;
    p.indices = indices; p.parent_indices = parent_indices; p.ts = ts; p.kids = kids; p.means3D = means3D; p.shs = shs;
    p.colors_precomp = colors_precomp; p.opacities = opacities; p.scales = scales; p.scale_modifier = scale_modifier;
    p.rotations = rotations; p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix;
    p.campos = campos; p.tan_fovx = tan_fovx; p.tan_fovy = tan_fovy;
    p.focal_y = H / (2.0f * tan_fovy);
    p.focal_x = W / (2.0f * tan_fovx);
    p.radii = radii;
    p.flags = flags;
    p.grid_x = (W + TILE_X - 1) / TILE_X; p.grid_y = (H + TILE_Y - 1) / TILE_Y;
    p.sh_vec4 = (shs && ((uintptr_t)shs & 15) == 0 && (M * 3) % 4 == 0) ? 1 : 0;
    if (rotations && ((uintptr_t)rotations & 15) != 0) { set_error("lodgs_forward_preprocess: rotations must be 16-byte aligned"); return LODGS_E_ARG; }
    LODGS_CUDA(cudaMemsetAsync(geom_buffer, 0, g.zero_bytes, s));
    int tiles = (P + PRE_BLOCK - 1) / PRE_BLOCK;
    const size_t stage_bytes = (size_t)PRE_BLOCK * M * 12;
    p.staged = (shs && !indices && !parent_indices && p.sh_vec4 && (M * 12) % 16 == 0 && M <= 16) ? 1 : 0;
    if (p.staged) {
        static std::atomic<size_t> attr_smem[MAX_DEVICES];
        const int dslot = device_slot();
        if (stage_bytes > attr_smem[dslot].load(std::memory_order_acquire)) {
            LODGS_CUDA(cudaFuncSetAttribute(preprocess_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stage_bytes));
            attr_smem[dslot].store(stage_bytes, std::memory_order_release);
        }
    }
    preprocess_fwd_kernel<<<tiles, PRE_BLOCK, p.staged ? stage_bytes : 0, s>>>(p, g);
    LODGS_LAUNCHED("preprocess_fwd_kernel");
    // The count goes back to the host (the caller sizes the binning buffer from it, rasterizer_impl.cu:330-333) while the
    // GPU already orders the Gaussians by depth for the emission stage: the host waits on an event recorded right after
    // the count is published, not on the stream. The count is published by a one-thread kernel that stores it into a
    // MAPPED pinned word instead of a 4-byte cudaMemcpyAsync (publish_words, common.cuh): a device-to-host copy is queued on the
    // D2H copy engine behind whatever bulk transfer the application has in flight (measured in bench.py's end-to-end
    // pipeline: up to 5 ms behind the previous step's results), a store over PCIe is not.
    if (num_rendered_host) {
        const int prc = publish_words(reinterpret_cast<const int32_t*>(&g.hdr->num_rendered), 1, s);
        if (prc) return prc;
    }
    // Render-list mode: the list is a hierarchy cut, not view-culled, so most entries carry no depth key; the sort then runs on the
    // compacted visible entries (their number stays on the device). LODGS_DEPTH_COMPACT=0 / 2 switches it off / on for every call (A/B).
    static const int compact_mode = getenv("LODGS_DEPTH_COMPACT") ? atoi(getenv("LODGS_DEPTH_COMPACT")) : 1;
    const bool compact = compact_mode == 2 || (compact_mode == 1 && indices != nullptr);
    int rc;
    if (compact) {
        rc = launch_depth_compact(P, g, s);
        if (rc) return rc;
        const int* n_vis = &g.hdr->n_visible;
        // input = output arrays on purpose (state.cuh): with 32 key bits there are four passes, tmp <- in, out <- tmp, tmp <- out, out <- tmp
        rc = sort_pairs32_device_n(g.compact_keys, g.compact_vals, g.depth_keys_sorted, g.order, g.sort_tmp_k, g.sort_tmp_v, P, n_vis, 0, 32, g.sort_ws, s,
                                   "sort_depth_pass_kernel", "sort_depth_hist_kernel");
        if (rc) return rc;
        rc = launch_scan_sorted(P, n_vis, g, s);
    } else {
        rc = sort_pairs32_device(g.depth_keys, nullptr, g.depth_keys_sorted, g.order, g.sort_tmp_k, g.sort_tmp_v, P, 0, 32, g.sort_ws, s, "sort_depth_pass_kernel",
                                 "sort_depth_hist_kernel");
        if (rc) return rc;
        rc = launch_scan_sorted(P, nullptr, g, s);
    }
    if (rc) return rc;
    if (num_rendered_host) {
        const int32_t* w = wait_words();
        if (!w) { set_error("lodgs_forward_preprocess: waiting for the instance count failed"); return LODGS_E_CUDA; }
        *num_rendered_host = w[0];
    }
    return LODGS_OK;
}

extern "C" int lodgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                                  void* stream) {
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("lodgs_mark_visible: bad argument"); return LODGS_E_ARG; }
    if (P == 0) return LODGS_OK;
    prof_stage_begin((cudaStream_t)stream);
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, means3D, viewmatrix, present);
    LODGS_LAUNCHED("mark_visible_kernel");
    return LODGS_OK;
}
