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
#include <cstdlib>

namespace lodgs {

// Instance emission in depth order. The reference's duplicateWithKeys (rasterizer_impl.cu:70-115) emits 64-bit tile|depth
// keys in index order and lets a 6-pass radix sort order them; here the P Gaussians are sorted by depth bits first (cheap:
// P << R), instances are emitted in that order carrying only their tile id, and a stable sort on the tile id alone
// (2 passes at 1080p) finishes the job: (tile, depth, index) order, i.e. exactly the reference's sorted list.
//
//
// Two kernels. scan_sorted_kernel: a CTA takes SCAN_TILE consecutive positions of the depth order, gathers their 8-byte
// tile boxes and scans the tile counts (warp scans + decoupled look-back across CTAs, 32 predecessors per step) into
// sorted_offsets. emit_balanced_kernel: work is split by OUTPUT slots, not by Gaussians — in depth order the nearest
// Gaussians come first and cover hundreds of tiles each, so a per-Gaussian split leaves a few CTAs with most of the work.
// A CTA owns EMIT_SLOTS consecutive output slots, finds the range of owning positions with two warp-wide 32-ary searches
// in sorted_offsets, stages that range's offsets in shared memory, and every thread resolves its slots by a binary search
// there. Slot -> (x,y) inside a rectangle is row-major like the reference.
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_TILE = SCAN_ITEMS * SCAN_THREADS;

__global__ void __launch_bounds__(SCAN_THREADS) scan_sorted_kernel(int P, const int* __restrict__ n_dev, const uint32_t* __restrict__ order,
                                                                  const uint2* __restrict__ tile_box, uint32_t* __restrict__ sorted_offsets,
                                                                  unsigned long long* state, int* ticket) {
    // n_dev != NULL: only the first *n_dev positions of `order` are valid (compacted depth sort); the others count zero tiles
    const int n_valid = n_dev ? min(P, __ldg(n_dev)) : P;
    __shared__ int s_tile;
    __shared__ unsigned s_warp_total[SCAN_THREADS / 32];
    __shared__ unsigned long long s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int wbase = tile * SCAN_TILE + (int)warp * (32 * SCAN_ITEMS);  // this warp owns 256 consecutive positions
    unsigned gid[SCAN_ITEMS], cnt[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int pos = wbase + k * 32 + (int)lane;
        gid[k] = pos < n_valid ? __ldg(order + pos) : 0u;
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int pos = wbase + k * 32 + (int)lane;
        const uint2 b = pos < n_valid ? __ldg(tile_box + gid[k]) : make_uint2(0u, 0u);
        cnt[k] = ((b.y & 0xffffu) - (b.x & 0xffffu)) * ((b.y >> 16) - (b.x >> 16));
    }
    unsigned incl[SCAN_ITEMS];
    unsigned run = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        unsigned v = cnt[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= (unsigned)o) v += t;
        }
        incl[k] = run + v;
        run += __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) s_warp_total[warp] = run;
    __syncthreads();
    if (warp == 0) {
        unsigned block_total = 0;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; w++) block_total += s_warp_total[w];
        unsigned long long ex = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(state, pack_state(2, block_total));
        } else {
            if (lane == 0) st_relaxed(state + tile, pack_state(1, block_total));
            ex = lookback_warp(state, tile);
            if (lane == 0) st_relaxed(state + tile, pack_state(2, ex + block_total));
        }
        if (lane == 0) s_prefix = ex;
    }
    __syncthreads();
    unsigned warp_off = (unsigned)s_prefix;
    for (unsigned w = 0; w < warp; w++) warp_off += s_warp_total[w];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int pos = wbase + k * 32 + (int)lane;
        if (pos < P) sorted_offsets[pos] = warp_off + incl[k];
    }
}

// Render-list mode (the list comes from a hierarchy cut and is not view-culled: most entries fail the frustum test): the depth sort runs on
// the visible entries only. One pass keeps {depth key, list position} of every entry with a key, in list order (block scan + decoupled
// look-back, like the scan above), and leaves their number in the header; the sort and the scan above then work on that many elements.
__global__ void __launch_bounds__(SCAN_THREADS) depth_compact_kernel(int P, const uint32_t* __restrict__ depth_keys, uint32_t* __restrict__ keys_out,
                                                                    uint32_t* __restrict__ vals_out, unsigned long long* state, int* ticket,
                                                                    int* __restrict__ n_out) {
    __shared__ int s_tile;
    __shared__ unsigned s_warp_total[SCAN_THREADS / 32];
    __shared__ unsigned long long s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int wbase = tile * SCAN_TILE + (int)warp * (32 * SCAN_ITEMS);
    uint32_t key[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int pos = wbase + k * 32 + (int)lane;
        key[k] = pos < P ? __ldg(depth_keys + pos) : 0xffffffffu;
    }
    unsigned before[SCAN_ITEMS];   // kept entries of this warp before item k of this lane
    unsigned run = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, key[k] != 0xffffffffu);
        before[k] = run + __popc(m & ((1u << lane) - 1u));
        run += __popc(m);
    }
    if (lane == 0) s_warp_total[warp] = run;
    __syncthreads();
    if (warp == 0) {
        unsigned block_total = 0;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; w++) block_total += s_warp_total[w];
        unsigned long long ex = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(state, pack_state(2, block_total));
        } else {
            if (lane == 0) st_relaxed(state + tile, pack_state(1, block_total));
            ex = lookback_warp(state, tile);
            if (lane == 0) st_relaxed(state + tile, pack_state(2, ex + block_total));
        }
        if (lane == 0) {
            s_prefix = ex;
            if (tile == (P + SCAN_TILE - 1) / SCAN_TILE - 1) *n_out = (int)(ex + block_total);
        }
    }
    __syncthreads();
    unsigned warp_off = (unsigned)s_prefix;
    for (unsigned w = 0; w < warp; w++) warp_off += s_warp_total[w];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (key[k] != 0xffffffffu) {
            const unsigned dst = warp_off + before[k];
            keys_out[dst] = key[k];
            vals_out[dst] = (uint32_t)(wbase + k * 32 + (int)lane);
        }
    }
}

constexpr int EMIT_THREADS = 256;
constexpr int EMIT_PER_THREAD = 8;
constexpr int EMIT_SLOTS = EMIT_THREADS * EMIT_PER_THREAD;  // output slots per CTA; at most EMIT_SLOTS+1... owners, see below

// Smallest position p in [0,P) with incl[p] > slot (incl is non-decreasing, incl[P-1] > slot guaranteed): 32-ary search by
// one warp, 4 dependent loads for P = 1 M instead of 20.
__device__ __forceinline__ unsigned warp_upper_bound(const uint32_t* __restrict__ incl, unsigned P, unsigned slot, unsigned lane) {
    unsigned lo = 0, n = P;  // answer in [lo, lo+n)
    while (n > 1) {
        const unsigned step = (n + 31) / 32;
        const unsigned probe = lo + min(n, (lane + 1) * step) - 1;  // last element of this lane's chunk
        const bool gt = __ldg(incl + probe) > slot;
        const unsigned m = __ballot_sync(0xffffffffu, gt);
        const unsigned c = (unsigned)(__ffs(m) - 1);  // first chunk whose last element exceeds slot
        const unsigned nlo = lo + min(n, c * step);
        n = min(n, (c + 1) * step) - min(n, c * step);
        lo = nlo;
    }
    return lo;
}

__global__ void __launch_bounds__(EMIT_THREADS) emit_balanced_kernel(int P, unsigned R, const uint32_t* __restrict__ order,
                                                                    const uint2* __restrict__ tile_box,
                                                                    const uint32_t* __restrict__ sorted_offsets, unsigned grid_x,
                                                                    uint32_t* __restrict__ tiles_out, uint32_t* __restrict__ vals) {
    __shared__ unsigned s_off[EMIT_SLOTS + 2];  // exclusive offset of every owner position in this CTA's range (+ sentinel)
    __shared__ unsigned s_lo, s_hi;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned slot0 = blockIdx.x * EMIT_SLOTS;
    const unsigned slot1 = min(R, slot0 + EMIT_SLOTS);  // exclusive
    if (warp == 0) {
        const unsigned v = warp_upper_bound(sorted_offsets, (unsigned)P, slot0, lane);
        if (lane == 0) s_lo = v;
    } else if (warp == 1) {
        const unsigned v = warp_upper_bound(sorted_offsets, (unsigned)P, slot1 - 1, lane);
        if (lane == 0) s_hi = v;
    }
    __syncthreads();
    const unsigned lo = s_lo, hi = s_hi;  // owners lo..hi; every visible Gaussian owns >= 1 slot, so hi - lo + 1 <= EMIT_SLOTS
    const unsigned nown = min(hi - lo + 1, (unsigned)EMIT_SLOTS + 1u);  // (the clamp only guards against NaN-poisoned input)
    for (unsigned i = tid; i <= nown; i += EMIT_THREADS) {
        const unsigned p = lo + i;  // s_off[i] = exclusive offset of position lo+i = inclusive offset of lo+i-1
        s_off[i] = (p == 0) ? 0u : __ldg(sorted_offsets + min(p, (unsigned)P) - 1);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < EMIT_PER_THREAD; k++) {
        const unsigned s = slot0 + k * EMIT_THREADS + tid;
        if (s >= slot1) break;
        // largest i with s_off[i] <= s
        unsigned a = 0, len = nown;
        while (len > 1) {
            const unsigned half = len >> 1;
            if (s_off[a + half] <= s) { a += half; len -= half; } else { len = half; }
        }
        const unsigned gid = __ldg(order + lo + a);
        const uint2 b = __ldg(tile_box + gid);
        const unsigned w = max((b.y & 0xffffu) - (b.x & 0xffffu), 1u);
        const unsigned q = s - s_off[a];
        const unsigned qy = q / w, qx = q - qy * w;
        tiles_out[s] = ((b.x >> 16) + qy) * grid_x + (b.x & 0xffffu) + qx;
        vals[s] = gid;
    }
}

// identifyTileRanges (rasterizer_impl.cu:120-142) over the sorted tile ids; ranges were cleared before the launch.
constexpr int RANGE_ITEMS = 4;
__global__ void __launch_bounds__(256) tile_ranges_kernel(int R, const uint32_t* __restrict__ tiles_sorted, uint2* __restrict__ ranges) {
    const int i0 = (blockIdx.x * blockDim.x + threadIdx.x) * RANGE_ITEMS;
    if (i0 >= R) return;
    uint32_t t[RANGE_ITEMS + 1];
    t[0] = i0 ? __ldg(tiles_sorted + i0 - 1) : 0u;
    if (i0 + RANGE_ITEMS <= R && (reinterpret_cast<uintptr_t>(tiles_sorted) & 15) == 0) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(tiles_sorted + i0));
        t[1] = q.x; t[2] = q.y; t[3] = q.z; t[4] = q.w;
    } else {
#pragma unroll
        for (int k = 0; k < RANGE_ITEMS; k++) t[k + 1] = i0 + k < R ? __ldg(tiles_sorted + i0 + k) : 0u;
    }
#pragma unroll
    for (int k = 0; k < RANGE_ITEMS; k++) {
        const int i = i0 + k;
        if (i >= R) break;
        const uint32_t cur = t[k + 1];
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = t[k];
            if (cur != prev) {
                ranges[prev].y = i;
                ranges[cur].x = i;
            }
            // inside the else like the reference (:139-140): a list of exactly one instance keeps range.y == 0
            if (i == R - 1) ranges[cur].y = R;
        }
    }
}

// Blend CTAs take the tiles with the longest lists first (the grid's hardware order is CTA index order): a view that puts most
// instances into a band of tiles (a slab seen from its edge) otherwise starts its longest lists in the middle of the launch and ends
// on them with most SMs idle. Four classes by list length relative to the longest (>= 1/2, >= 1/4, >= 1/8, rest), tile-id order kept
// inside a class (a stable partition: neighbouring tiles share splat records in L2; a full sort by length was measured slower on the
// eval path for that reason). One CTA, every thread owns a contiguous run of tiles. A tile's result does not depend on when it runs.
constexpr int ORDER_CLASSES = 4;
constexpr int LONG_TILES_LIMIT = 2048;   // = LONG_TILES_MAX of the blend kernels
__device__ __forceinline__ int order_class(unsigned len, unsigned mx) {
    return len >= (mx >> 1) ? 0 : len >= (mx >> 2) ? 1 : len >= (mx >> 3) ? 2 : 3;
}
__global__ void __launch_bounds__(1024) tile_order_kernel(const uint2* __restrict__ ranges, uint32_t* __restrict__ order, int ntiles, int split_long) {
    __shared__ unsigned s_max, s_total, s_warp[ORDER_CLASSES][32], s_base[ORDER_CLASSES];
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_max = 0; s_total = 0; }
    __syncthreads();
    const int per = (ntiles + 1023) / 1024, lo = min(ntiles, (int)tid * per), hi = min(ntiles, lo + per);
    unsigned m = 0, sum = 0;
    for (int t = lo; t < hi; t++) {
        const uint2 r = ranges[t];
        m = max(m, r.y - r.x);
        sum += r.y - r.x;
    }
    m = __reduce_max_sync(0xffffffffu, m);
    sum = __reduce_add_sync(0xffffffffu, sum);
    if (lane == 0) { atomicMax(&s_max, m); atomicAdd(&s_total, sum); }
    __syncthreads();
    const unsigned mx = max(s_max, 8u);
    unsigned cnt[ORDER_CLASSES] = {0, 0, 0, 0}, incl[ORDER_CLASSES];
    for (int t = lo; t < hi; t++) {
        const uint2 r = ranges[t];
        const int c = order_class(r.y - r.x, mx);
#pragma unroll
        for (int k = 0; k < ORDER_CLASSES; k++) cnt[k] += (c == k);
    }
#pragma unroll
    for (int k = 0; k < ORDER_CLASSES; k++) {
        unsigned v = cnt[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= (unsigned)o) v += t;
        }
        incl[k] = v;
        if (lane == 31) s_warp[k][warp] = v;
    }
    __syncthreads();
    if (warp < ORDER_CLASSES) {
        const unsigned w = s_warp[warp][lane];
        unsigned wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (unsigned)o) wi += t;
        }
        s_warp[warp][lane] = wi - w;
        if (lane == 31) s_base[warp] = wi;   // class total
    }
    __syncthreads();
    unsigned pos[ORDER_CLASSES], base = 0;
#pragma unroll
    for (int k = 0; k < ORDER_CLASSES; k++) {
        pos[k] = base + s_warp[k][warp] + incl[k] - cnt[k];
        base += s_base[k];
    }
    for (int t = lo; t < hi; t++) {
        const uint2 r = ranges[t];
        const int c = order_class(r.y - r.x, mx);
#pragma unroll
        for (int k = 0; k < ORDER_CLASSES; k++)
            if (c == k) order[pos[k]++] = (uint32_t)t;
    }
    // LONG tiles (blend_fwd.cu / blend_bwd.cu take each of them with two CTAs of two pixels per thread instead of one CTA of four): the
    // launch is latency-bound on its longest lists when the longest alone is more than 1/1024 of all instances (a B200 keeps ~1500-2000
    // blend CTAs resident, so every tile starts at once and the kernel lasts as long as its slowest tile). Then every tile within a factor
    // of eight of the longest (classes 0-2, the first entries of `order`) is a candidate for the critical path — the longest lists often
    // end early by saturation, the medium ones are walked to the end — and gets the shorter per-splat chain; at most LONG_TILES_LIMIT.
    if (tid == 0) {
        const bool skewed = split_long && s_max >= 256u && (unsigned long long)s_max * 1024ull > (unsigned long long)s_total;
        unsigned n = 0;
        if (skewed) {
            const unsigned c01 = s_base[0] + s_base[1], c012 = c01 + s_base[2];
            n = c012 <= (unsigned)LONG_TILES_LIMIT ? c012 : c01 <= (unsigned)LONG_TILES_LIMIT ? c01 : min(s_base[0], (unsigned)LONG_TILES_LIMIT);
        }
        order[ntiles] = n;
    }
}

// LOD rendering: the last two slots of every blend record become the interpolation weight and 1/kids.
__global__ void __launch_bounds__(256) lod_patch_kernel(int P, const float* __restrict__ ts, const int32_t* __restrict__ kids,
                                                       SplatRec* __restrict__ recs) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    *reinterpret_cast<float2*>(&recs[i].t) = make_float2(__ldg(ts + i), 1.0f / (float)__ldg(kids + i));
}

// The reference's sorted 64-bit key list (tile << 32 | depth bits), rebuilt from the products of the tile sort. Parity tests only.
__global__ void __launch_bounds__(256) sorted_keys_kernel(int R, const uint32_t* __restrict__ tiles_sorted,
                                                         const uint32_t* __restrict__ point_list, const uint32_t* __restrict__ depth_keys,
                                                         uint64_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    keys[i] = ((uint64_t)__ldg(tiles_sorted + i) << 32) | (uint64_t)__ldg(depth_keys + __ldg(point_list + i));
}

int launch_scan_sorted(int P, const int* n_dev, const GeomState& g, cudaStream_t s) {
    scan_sorted_kernel<<<(P + SCAN_TILE - 1) / SCAN_TILE, SCAN_THREADS, 0, s>>>(P, n_dev, g.order, g.tile_box, g.sorted_offsets, g.scan2_state,
                                                                               &g.hdr->ticket2);
    LODGS_LAUNCHED("scan_sorted_kernel");
    return LODGS_OK;
}
int launch_depth_compact(int P, const GeomState& g, cudaStream_t s) {
    // look-back states and ticket of the (retired) fused scan of the preprocess kernel: cleared with the header at the start of the forward
    depth_compact_kernel<<<(P + SCAN_TILE - 1) / SCAN_TILE, SCAN_THREADS, 0, s>>>(P, g.depth_keys, g.compact_keys, g.compact_vals, g.scan_state,
                                                                                 &g.hdr->ticket, &g.hdr->n_visible);
    LODGS_LAUNCHED("depth_compact_kernel");
    return LODGS_OK;
}
int launch_emit_balanced(int P, int R, const GeomState& g, unsigned gx, uint32_t* tiles, uint32_t* vals, cudaStream_t s) {
    emit_balanced_kernel<<<(R + EMIT_SLOTS - 1) / EMIT_SLOTS, EMIT_THREADS, 0, s>>>(P, (unsigned)R, g.order, g.tile_box, g.sorted_offsets, gx,
                                                                                   tiles, vals);
    LODGS_LAUNCHED("emit_balanced_kernel");
    return LODGS_OK;
}
int launch_tile_ranges(int R, const BinningState& b, uint2* ranges, size_t ntiles, cudaStream_t s) {
    LODGS_CUDA(cudaMemsetAsync(ranges, 0, ntiles * sizeof(uint2), s));
    const int threads = (R + RANGE_ITEMS - 1) / RANGE_ITEMS;
    tile_ranges_kernel<<<(threads + 255) / 256, 256, 0, s>>>(R, b.tiles_sorted, ranges);
    LODGS_LAUNCHED("tile_ranges_kernel");
    return LODGS_OK;
}
int launch_tile_order(const uint2* ranges, uint32_t* order, size_t ntiles, cudaStream_t s) {
    static const int split_long = getenv("LODGS_LONG_TILES") ? atoi(getenv("LODGS_LONG_TILES")) : 1;   // A/B: 0 = every tile as one CTA
    tile_order_kernel<<<1, 1024, 0, s>>>(ranges, order, (int)ntiles, split_long);
    LODGS_LAUNCHED("tile_order_kernel");
    return LODGS_OK;
}
int launch_lod_patch(int P, const GeomState& g, const float* ts, const int32_t* kids, cudaStream_t s) {
    lod_patch_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, ts, kids, g.recs);
    LODGS_LAUNCHED("lod_patch_kernel");
    return LODGS_OK;
}
int launch_sorted_keys(int R, const BinningState& b, const GeomState& g, uint64_t* keys, cudaStream_t s) {
    sorted_keys_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.tiles_sorted, b.point_list, g.depth_keys, keys);
    LODGS_LAUNCHED("sorted_keys_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
