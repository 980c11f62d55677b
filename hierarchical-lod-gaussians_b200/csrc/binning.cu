// binning.cu — (Gaussian,tile) instance generation, tile ranges, sorted splat records.
//
// Replaces duplicateWithKeys (HR/cuda_rasterizer/rasterizer_impl.cu:70-115) and identifyTileRanges (:120-142),
// and adds the record gather that lets the blend kernels stage batches with 1-D bulk (TMA) copies.
//
// The reference walks each Gaussian's tile rectangle serially in one thread (one large splat stalls its warp, every
// store is an isolated 8-byte scatter) and then radix-sorts 64-bit keys in 6 passes. See emit_tiles_kernel for the
// depth-ordered emission used here.
#include "gaussmath.cuh"
#include "state.cuh"

namespace lodgs {

// Inclusive scan of tiles_touched in depth order: sorted_offsets[pos] = sum_{q<=pos} tiles_touched[order[q]].
// Single pass: 256-entry tiles in launch order (ticket), block scan, decoupled look-back.
__global__ void __launch_bounds__(PRE_BLOCK) scan_sorted_kernel(int P, const uint32_t* __restrict__ order,
                                                              const uint32_t* __restrict__ tiles_touched,
                                                              uint32_t* __restrict__ sorted_offsets, unsigned long long* state,
                                                              int* ticket) {
    __shared__ int s_tile;
    __shared__ unsigned s_warp_sum[PRE_BLOCK / 32];
    __shared__ unsigned long long s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const int pos = tile * PRE_BLOCK + threadIdx.x;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned v = 0;
    if (pos < P) v = __ldg(tiles_touched + __ldg(order + pos));
    unsigned incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += t;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned w = (lane < PRE_BLOCK / 32) ? s_warp_sum[lane] : 0;
        unsigned wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (unsigned)o) wi += t;
        }
        if (lane < PRE_BLOCK / 32) s_warp_sum[lane] = wi - w;
        const unsigned block_total = __shfl_sync(0xffffffffu, wi, PRE_BLOCK / 32 - 1);
        unsigned long long excl = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(state, pack_state(2, block_total));
        } else {
            if (lane == 0) st_relaxed(state + tile, pack_state(1, block_total));
            excl = lookback_warp(state, tile);
            if (lane == 0) st_relaxed(state + tile, pack_state(2, excl + block_total));
        }
        if (lane == 0) s_prefix = excl;
    }
    __syncthreads();
    if (pos < P) sorted_offsets[pos] = (unsigned)s_prefix + s_warp_sum[warp] + incl;
}

// Instance emission in depth order. The reference's duplicateWithKeys (rasterizer_impl.cu:70-115) emits 64-bit tile|depth
// keys in index order and lets a 6-pass radix sort order them; here the P Gaussians are sorted by depth bits first (cheap:
// P << R), instances are emitted in that order carrying only their tile id, and a stable sort on the tile id alone
// (2 passes at 1080p) finishes the job: (tile, depth, index) order, i.e. exactly the reference's sorted list.
// A warp flattens the rectangles of 32 consecutive Gaussians of the order into one run of output slots (their offsets are
// consecutive in sorted_offsets), finds each slot's owner by a shuffle binary search and writes 32 consecutive entries
// per instruction; slot -> (x,y) inside a rectangle is row-major like the reference.
__global__ void __launch_bounds__(256) emit_tiles_kernel(int P, const uint32_t* __restrict__ order, const float2* __restrict__ means2D,
                                                        const uint32_t* __restrict__ sorted_offsets, const uint32_t* __restrict__ tiles_touched,
                                                        const int2* __restrict__ rects, unsigned grid_x, unsigned grid_y,
                                                        uint32_t* __restrict__ tiles_out, uint32_t* __restrict__ vals) {
    const unsigned lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int base = warp_global * 32;
    if (base >= P) return;
    const int pos = base + lane;
    unsigned count = 0, rminx = 0, rminy = 0, rw = 1, gid = 0;
    if (pos < P) {
        gid = __ldg(order + pos);
        if (__ldg(tiles_touched + gid) > 0) {
            uint2 rmin, rmax;
            tile_rect(__ldg(means2D + gid), __ldg(rects + gid), rmin, rmax, grid_x, grid_y);
            rminx = rmin.x; rminy = rmin.y;
            rw = rmax.x - rmin.x;
            count = rw * (rmax.y - rmin.y);
            if (rw == 0) rw = 1;
        }
    }
    // first output slot of this warp = sorted_offsets[base-1]; slots of the 32 Gaussians are consecutive
    const unsigned warp_off = (base == 0) ? 0u : __ldg(sorted_offsets + base - 1);
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
        const unsigned o_gid = __shfl_sync(0xffffffffu, gid, lo);
        if (s < total) {
            const unsigned k = s - o_excl;
            const unsigned y = o_miny + k / o_w, x = o_minx + k % o_w;
            tiles_out[warp_off + s] = y * grid_x + x;
            vals[warp_off + s] = o_gid;
        }
    }
}

__global__ void tile_ranges_kernel(int L, const uint32_t* __restrict__ tiles, uint2* __restrict__ ranges) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t cur = __ldg(tiles + idx);
    if (idx == 0) {
        ranges[cur].x = 0;
    } else {
        const uint32_t prev = __ldg(tiles + idx - 1);
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
                                                           const uint32_t* __restrict__ tiles_sorted, uint64_t* __restrict__ keys,
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
    const float depth = __ldg(depths + id);
    keys[i] = ((uint64_t)__ldg(tiles_sorted + i) << 32) | (uint64_t)__float_as_uint(depth);  // the reference's sorted tile|depth key
    float4 c = {__ldg(colors + 3 * (size_t)id + 2), __frcp_rn(depth), 1.0f, 1.0f};  // rcp.rn == the reference's 1/depth per pair
    if (ts != nullptr && kids != nullptr) {
        c.z = __ldg(ts + id);
        c.w = 1.0f / (float)__ldg(kids + id);
    }
    float4* out = reinterpret_cast<float4*>(recs + i);
    out[0] = a;
    out[1] = b;
    out[2] = c;
}

int launch_scan_sorted(int P, const GeomState& g, cudaStream_t s) {
    scan_sorted_kernel<<<(P + PRE_BLOCK - 1) / PRE_BLOCK, PRE_BLOCK, 0, s>>>(P, g.order, g.tiles_touched, g.sorted_offsets, g.scan2_state,
                                                                           &g.hdr->ticket2);
    LODGS_LAUNCHED("scan_sorted_kernel");
    return LODGS_OK;
}
int launch_emit_tiles(int P, const GeomState& g, unsigned gx, unsigned gy, uint32_t* tiles, uint32_t* vals, cudaStream_t s) {
    emit_tiles_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.order, g.means2D, g.sorted_offsets, g.tiles_touched, g.rects, gx, gy, tiles, vals);
    LODGS_LAUNCHED("emit_tiles_kernel");
    return LODGS_OK;
}
int launch_tile_ranges(int R, const uint32_t* keys, uint2* ranges, size_t ntiles, cudaStream_t s) {
    LODGS_CUDA(cudaMemsetAsync(ranges, 0, ntiles * sizeof(uint2), s));
    if (R > 0) {
        tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, keys, ranges);
        LODGS_LAUNCHED("tile_ranges_kernel");
    }
    return LODGS_OK;
}
int launch_build_records(int R, int P, const BinningState& b, const GeomState& g, const float* colors, const float* ts,
                         const int32_t* kids, cudaStream_t s) {
    build_records_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, P, b.point_list, b.tiles_sorted, b.keys, g.means2D, g.conic_opacity, colors,
                                                        g.depths, ts, kids, b.recs);
    LODGS_LAUNCHED("build_records_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
