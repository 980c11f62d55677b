// ar_buckets.cu — the alt-rasterizer's `num_buckets` (second value of its 9-tuple), exactly.
//
// The reference's second rasterizer drops, per tile, the (Gaussian, tile) instances whose largest alpha inside the tile stays below
// 1/255 (AR/cuda_rasterizer/rasterizer_impl.cu:56-100 max_contrib_power_rect_gaussian_float, :143-187 duplicateWithKeys) and then
// counts ceil(list length / 32) "buckets" per tile (:223-232 perTileBucketCount, summed by the scan at :464-469). The blend kernels
// here skip those pairs per pixel and need no per-bucket state, so the count only has to be REPORTED: one thread per sorted instance
// evaluates the reference's tile test (same expression tree, so nvcc contracts it the same way) and bumps its tile's counter; a
// second kernel turns the per-tile counts into buckets.
#include "state.cuh"

namespace lodgs {

__device__ __forceinline__ float ar_opacity_factor(const float dx, const float dy, const float4 co) {
    return 0.5f * (co.x * dx * dx + co.z * dy * dy) + co.y * dx * dy;
}

// max_contrib_power_rect_gaussian_float<15, 15>: smallest power (= largest alpha) of the Gaussian over the tile's pixel rectangle
__device__ __forceinline__ float ar_min_power_in_tile(const float4 co, const float2 mean, const float2 rect_min, const float2 rect_max) {
    const float x_min_diff = rect_min.x - mean.x;
    const float x_left = x_min_diff > 0.0f;
    const float not_in_x_range = x_left + (mean.x > rect_max.x);
    const float y_min_diff = rect_min.y - mean.y;
    const float y_above = y_min_diff > 0.0f;
    const float not_in_y_range = y_above + (mean.y > rect_max.y);
    float max_contrib_power = 0.0f;
    if ((not_in_y_range + not_in_x_range) > 0.0f) {
        const float px = x_left * rect_min.x + (1.0f - x_left) * rect_max.x;
        const float py = y_above * rect_min.y + (1.0f - y_above) * rect_max.y;
        const float dx = copysign(float(TILE_X - 1), x_min_diff);
        const float dy = copysign(float(TILE_Y - 1), y_min_diff);
        const float diffx = mean.x - px;
        const float diffy = mean.y - py;
        const float rcp_dxdxcox = __frcp_rn((TILE_X - 1) * (TILE_X - 1) * co.x);
        const float rcp_dydycoz = __frcp_rn((TILE_Y - 1) * (TILE_Y - 1) * co.z);
        const float tx = not_in_y_range * __saturatef((dx * co.x * diffx + dx * co.y * diffy) * rcp_dxdxcox);
        const float ty = not_in_x_range * __saturatef((dy * co.y * diffx + dy * co.z * diffy) * rcp_dydycoz);
        const float2 max_pos = {px + tx * dx, py + ty * dy};
        const float2 max_pos_diff = {mean.x - max_pos.x, mean.y - max_pos.y};
        max_contrib_power = ar_opacity_factor(max_pos_diff.x, max_pos_diff.y, co);
    }
    return max_contrib_power;
}

__global__ void ar_tile_count_kernel(int R, unsigned gx, const uint32_t* __restrict__ tiles_sorted, const uint32_t* __restrict__ point_list,
                                     const SplatRec* __restrict__ recs, unsigned* __restrict__ tile_count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const unsigned tile = __ldg(tiles_sorted + i);
    const SplatRec& r = recs[__ldg(point_list + i)];
    const float4 co = {r.conx, r.cony, r.conz, r.opac};
    const float2 xy = {r.x, r.y};
    const unsigned x = tile % gx, y = tile / gx;
    const float2 tile_min = {(float)(x * TILE_X), (float)(y * TILE_Y)};
    const float2 tile_max = {(float)((x + 1) * TILE_X - 1), (float)((y + 1) * TILE_Y - 1)};
    const float opacity_threshold = 1.0f / 255.0f;
    const float opacity_factor_threshold = logf(co.w / opacity_threshold);
    if (ar_min_power_in_tile(co, xy, tile_min, tile_max) <= opacity_factor_threshold) atomicAdd(tile_count + tile, 1u);
}

__global__ void ar_bucket_sum_kernel(int tiles, const unsigned* __restrict__ tile_count, int* __restrict__ out2) {
    __shared__ unsigned long long s_b[256], s_k[256];
    unsigned long long b = 0, k = 0;
    for (int t = threadIdx.x; t < tiles; t += 256) { const unsigned c = tile_count[t]; b += (c + 31u) / 32u; k += c; }
    s_b[threadIdx.x] = b; s_k[threadIdx.x] = k;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_b[threadIdx.x] += s_b[threadIdx.x + o]; s_k[threadIdx.x] += s_k[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out2[0] = (int)s_b[0]; out2[1] = (int)s_k[0]; }
}

}  // namespace lodgs

using namespace lodgs;

extern "C" int lodgs_ar_bucket_count(int P, int W, int H, int R, void* geom_buffer, void* binning_buffer, void* ws, size_t ws_bytes,
                                     int32_t* out2_host, void* stream) {
    if (P <= 0 || W <= 0 || H <= 0 || R < 0 || !geom_buffer || !ws || !out2_host) { set_error("lodgs_ar_bucket_count: bad argument"); return LODGS_E_ARG; }
    const unsigned gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const size_t need = align_up((size_t)gx * gy * 4, 128) + 128;
    if (ws_bytes < need) { set_error("lodgs_ar_bucket_count: workspace too small (%zu < %zu)", ws_bytes, need); return LODGS_E_WORKSPACE; }
    if (R == 0) { out2_host[0] = out2_host[1] = 0; return LODGS_OK; }
    if (!binning_buffer) { set_error("lodgs_ar_bucket_count: null binning buffer"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    unsigned* counts = (unsigned*)ws;
    int* out2 = (int*)((char*)ws + align_up((size_t)gx * gy * 4, 128));
    LODGS_CUDA(cudaMemsetAsync(counts, 0, (size_t)gx * gy * 4, s));
    GeomState g = geom_from_chunk(geom_buffer, P);
    BinningState b = binning_from_chunk(binning_buffer, R);
    ar_tile_count_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, gx, b.tiles_sorted, b.point_list, g.recs, counts);
    LODGS_LAUNCHED("ar_tile_count_kernel");
    ar_bucket_sum_kernel<<<1, 256, 0, s>>>((int)(gx * gy), counts, out2);
    LODGS_LAUNCHED("ar_bucket_sum_kernel");
    int rc = publish_words(out2, 2, s);
    if (rc) return rc;
    const int32_t* h = wait_words();
    if (!h) { set_error("lodgs_ar_bucket_count: reading the result back failed"); return LODGS_E_CUDA; }
    out2_host[0] = h[0]; out2_host[1] = h[1];
    return LODGS_OK;
}
