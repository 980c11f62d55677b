// binning.cu — (Gaussian,tile) instance generation, tile ranges, sorted splat records.
//
// Replaces duplicateWithKeys (HR/cuda_rasterizer/rasterizer_impl.cu:70-115) and identifyTileRanges (:120-142),
// and adds the record gather that lets the blend kernels stage batches with 1-D bulk (TMA) copies.
//
// emit_keys: the reference walks each Gaussian's tile rectangle serially in one thread, so one large splat
// stalls its warp and every store is an isolated 8-byte scatter. Here a warp flattens the rectangles of its 32
// Gaussians into one run of output slots (their offsets are consecutive in point_offsets), finds each slot's
// owner by a shuffle binary search and writes 32 consecutive keys/values per instruction. Slot -> (x,y) order
// inside a rectangle is row-major exactly as the reference, so the unsorted arrays are bit-identical.
#include "gaussmath.cuh"
#include "state.cuh"

namespace lodgs {

__global__ void __launch_bounds__(256) emit_keys_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                                                       const uint32_t* __restrict__ offsets, const int32_t* __restrict__ radii,
                                                       const int2* __restrict__ rects, unsigned grid_x, unsigned grid_y,
                                                       uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const unsigned lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int base = warp_global * 32;
    if (base >= P) return;
    const int idx = base + lane;
    unsigned count = 0, rminx = 0, rminy = 0, rw = 1, depth_bits = 0;
    if (idx < P && __ldg(radii + idx) > 0) {
        uint2 rmin, rmax;
        tile_rect(__ldg(means2D + idx), __ldg(rects + idx), rmin, rmax, grid_x, grid_y);
        rminx = rmin.x; rminy = rmin.y;
        rw = rmax.x - rmin.x;
        count = rw * (rmax.y - rmin.y);
        depth_bits = __float_as_uint(__ldg(depths + idx));
        if (rw == 0) rw = 1;
    }
    // first output slot of this warp = offsets[base-1]; slots of the 32 Gaussians are consecutive
    const unsigned warp_off = (base == 0) ? 0u : __ldg(offsets + base - 1);
    unsigned incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += v;
    }
    const unsigned excl = incl - count;
    const unsigned total = __shfl_sync(0xffffffffu, incl, 31);
    for (unsigned s0 = 0; s0 < total; s0 += 32) {
        const unsigned s = s0 + lane;
        int lo = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            int probe = lo + step;
            unsigned v = __shfl_sync(0xffffffffu, excl, probe & 31);
            if (probe < 32 && v <= s) lo = probe;
        }
        const unsigned o_excl = __shfl_sync(0xffffffffu, excl, lo);
        const unsigned o_minx = __shfl_sync(0xffffffffu, rminx, lo);
        const unsigned o_miny = __shfl_sync(0xffffffffu, rminy, lo);
        const unsigned o_w = __shfl_sync(0xffffffffu, rw, lo);
        const unsigned o_depth = __shfl_sync(0xffffffffu, depth_bits, lo);
        if (s < total) {
            const unsigned k = s - o_excl;
            const unsigned y = o_miny + k / o_w, x = o_minx + k % o_w;
            uint64_t key = (uint64_t)(y * grid_x + x);
            key <<= 32;
            key |= o_depth;
            keys[warp_off + s] = key;
            vals[warp_off + s] = (uint32_t)(base + lo);
        }
    }
}

__global__ void tile_ranges_kernel(int L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    uint32_t cur = (uint32_t)(__ldg(keys + idx) >> 32);
    if (idx == 0) {
        ranges[cur].x = 0;
    } else {
        uint32_t prev = (uint32_t)(__ldg(keys + idx - 1) >> 32);
        if (cur != prev) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
        // inside the else like the reference (:139-140): a list of exactly one instance keeps range.y == 0
        if (idx == L - 1) ranges[cur].y = L;
    }
}

// Gather per-Gaussian blend inputs into sorted order: one 48-byte record per instance.
__global__ void __launch_bounds__(256) build_records_kernel(int R, int P, const uint32_t* __restrict__ point_list,
                                                           const float2* __restrict__ means2D,
                                                           const float4* __restrict__ conic_opacity, const float* __restrict__ colors,
                                                           const float* __restrict__ depths, const float* __restrict__ ts,
                                                           const int32_t* __restrict__ kids, SplatRec* __restrict__ recs) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t id = __ldg(point_list + i);
    const float2 xy = __ldg(means2D + id);
    const float4 co = __ldg(conic_opacity + id);
    float4 a = {xy.x, xy.y, co.x, co.y};
    float4 b = {co.z, co.w, __ldg(colors + 3 * (size_t)id), __ldg(colors + 3 * (size_t)id + 1)};
    float4 c = {__ldg(colors + 3 * (size_t)id + 2), __frcp_rn(__ldg(depths + id)), 1.0f, 1.0f};  // rcp.rn == the reference's 1/depth per pair
    if (ts != nullptr && kids != nullptr) {
        c.z = __ldg(ts + id);
        c.w = 1.0f / (float)__ldg(kids + id);
    }
    float4* out = reinterpret_cast<float4*>(recs + i);
    out[0] = a;
    out[1] = b;
    out[2] = c;
}

int launch_emit_keys(int P, const GeomState& g, const int32_t* radii, unsigned gx, unsigned gy, uint64_t* keys, uint32_t* vals,
                     cudaStream_t s) {
    emit_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.means2D, g.depths, g.point_offsets, radii, g.rects, gx, gy, keys, vals);
    LODGS_LAUNCHED("emit_keys_kernel");
    return LODGS_OK;
}
int launch_tile_ranges(int R, const uint64_t* keys, uint2* ranges, size_t ntiles, cudaStream_t s) {
    LODGS_CUDA(cudaMemsetAsync(ranges, 0, ntiles * sizeof(uint2), s));
    if (R > 0) {
        tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, keys, ranges);
        LODGS_LAUNCHED("tile_ranges_kernel");
    }
    return LODGS_OK;
}
int launch_build_records(int R, int P, const BinningState& b, const GeomState& g, const float* colors, const float* ts,
                         const int32_t* kids, cudaStream_t s) {
    build_records_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, P, b.point_list, g.means2D, g.conic_opacity, colors, g.depths, ts, kids,
                                                        b.recs);
    LODGS_LAUNCHED("build_records_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
