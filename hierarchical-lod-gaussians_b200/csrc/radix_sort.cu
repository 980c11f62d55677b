// radix_sort.cu — hand-written onesweep LSD radix sort of (u64 key, u32 value) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs (HR/cuda_rasterizer/rasterizer_impl.cu:358-363). Stable, 8-bit
// digits over bits [0,end_bit): one histogram pass over the keys builds all digit histograms at once, then one
// kernel per digit does rank -> decoupled look-back -> scatter, reading and writing every pair exactly once
// per pass. Stability (and therefore bit-identical output to CUB) comes from ranking keys in index order:
// warp-striped tiles, per-warp ballot ranking, warps and tiles chained in order.
//
// The kernels are templated on the key type: u64 keys serve the exported lodgs_sort_pairs (the reference's tile|depth
// sort, kept for the CUB cross-check), u32 keys serve the two sorts of the forward pass (Gaussians by depth bits, then
// instances by tile id — see raster_api.cu). vals_in == nullptr means "value = element index" (saves the iota pass).
#include "state.cuh"

namespace lodgs {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = SORT_TILE / SORT_THREADS;  // 16
constexpr int SORT_WARPS = SORT_THREADS / 32;

// ---- global histograms for all passes in one read of the keys
template <typename KeyT>
__global__ void __launch_bounds__(256) sort_histogram_kernel(const KeyT* __restrict__ keys, int n, const int* __restrict__ n_dev, int passes,
                                                            int begin_bit, unsigned* hist) {
    __shared__ unsigned s_hist[SORT_MAX_PASSES * SORT_RADIX];
    if (n_dev) n = min(n, __ldg(n_dev));   // element count known only on the device (n = the capacity the grid was sized for)
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31;
    // warp-uniform trip count so the shuffles below are always executed by full warps; HIST_UNROLL independent loads are in
    // flight per thread (one load per iteration leaves the kernel latency-bound)
    constexpr int HIST_UNROLL = 8;
    const int stride = gridDim.x * blockDim.x * HIST_UNROLL;
    for (int base = (blockIdx.x * blockDim.x + (threadIdx.x & ~31)) * HIST_UNROLL; base < n; base += stride) {
        KeyT keyv[HIST_UNROLL];
#pragma unroll
        for (int u = 0; u < HIST_UNROLL; u++) {
            const int i = base + u * 32 + lane;
            keyv[u] = i < n ? __ldg(keys + i) : (KeyT)0;
        }
#pragma unroll
        for (int u = 0; u < HIST_UNROLL; u++) {
            const int i = base + u * 32 + lane;
            const bool valid = i < n;
            const KeyT key = keyv[u];
            const unsigned vmask = __ballot_sync(0xffffffffu, valid);
            if (vmask == 0u) break;
            for (int p = 0; p < passes; p++) {
                const unsigned d = (unsigned)(key >> (begin_bit + 8 * p)) & 255u;
                const unsigned d0 = __shfl_sync(0xffffffffu, d, __ffs(vmask) - 1);
                // degenerate digits (tile high byte, float exponent) are warp-uniform almost always: one add
                if (__all_sync(0xffffffffu, !valid || d == d0)) {
                    if (lane == 0) atomicAdd(&s_hist[p * SORT_RADIX + d0], __popc(vmask));
                } else if (valid) {
                    atomicAdd(&s_hist[p * SORT_RADIX + d], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
        unsigned v = s_hist[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// ---- exclusive scan of each pass' 256-bin histogram (one block per pass)
__global__ void __launch_bounds__(SORT_RADIX) sort_scan_hist_kernel(unsigned* hist) {
    __shared__ unsigned s_w[8];
    unsigned* h = hist + blockIdx.x * SORT_RADIX;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned v = h[threadIdx.x];
    unsigned incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    unsigned woff = 0;
    for (unsigned w = 0; w < warp; w++) woff += s_w[w];
    h[threadIdx.x] = woff + incl - v;
}

template <typename KeyT>
struct SortSmem {
    KeyT keys[SORT_TILE];
    uint32_t vals[SORT_TILE];
    unsigned warp_hist[SORT_WARPS][SORT_RADIX];  // per-warp digit counts -> exclusive offsets across warps
    unsigned digit_start[SORT_RADIX];            // exclusive scan of the tile's digit counts
    unsigned out_off[SORT_RADIX];                // global base of digit d minus digit_start[d]
    unsigned scan_w[SORT_WARPS];
    unsigned scan_h[SORT_WARPS];
    int tile;
};

template <typename KeyT>
__global__ void __launch_bounds__(SORT_THREADS, sizeof(KeyT) == 4 ? 4 : 2) sort_onesweep_kernel(const KeyT* __restrict__ keys_in,
                                                                    const uint32_t* __restrict__ vals_in,
                                                                    KeyT* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                    int n, const int* __restrict__ n_dev, int shift,
                                                                    const unsigned* __restrict__ hist_base,
                                                                    int* ticket, unsigned* lookback /* [tiles][256] */) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<KeyT>& sm = *reinterpret_cast<SortSmem<KeyT>*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (n_dev) n = min(n, __ldg(n_dev));   // device-side count: the tiles beyond it leave at once (nobody looks back at a LATER tile)
    if (tid == 0) sm.tile = atomicAdd(ticket, 1);
    for (int i = tid; i < SORT_WARPS * SORT_RADIX; i += SORT_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int tile = sm.tile;
    const long long tile_base = (long long)tile * SORT_TILE;
    if (tile_base >= (long long)n) return;
    const int tile_n = (int)min((long long)SORT_TILE, (long long)n - tile_base);

    // ---- load, warp-striped: item k of lane l in warp w is element w*512 + k*32 + l of the tile
    KeyT key[SORT_ITEMS];
    unsigned rank[SORT_ITEMS];
    const int wbase = warp * (32 * SORT_ITEMS);
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const int e = wbase + k * 32 + lane;
        if (e < tile_n) {
            key[k] = __ldg(keys_in + tile_base + e);
        } else {
            key[k] = ~(KeyT)0;
        }
    }
    // ---- rank inside the warp, item by item (index order => stable). Per group of RANK_GROUP items: the peer masks
    // are computed first; then the leader lane of every digit group bumps the warp's
    // running per-digit count with ONE shared-memory atomic per item whose return value is not consumed until all
    // atomics of the group are in flight (same-warp shared atomics execute in issue order; __syncwarp orders them
    // formally) — a chain of SORT_ITEMS / RANK_GROUP round trips instead of SORT_ITEMS LDS -> STS -> SHFL ones.
    unsigned* wh = sm.warp_hist[warp];
    constexpr int RANK_GROUP = 8;
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int g = 0; g < SORT_ITEMS; g += RANK_GROUP) {
        unsigned old[RANK_GROUP];
#pragma unroll
        for (int k = g; k < g + RANK_GROUP; k++) {
            const int e = wbase + k * 32 + lane;
            const bool valid = e < tile_n;
            const unsigned d = (unsigned)(key[k] >> shift) & 255u;
            // peers = lanes holding the same digit, from 8 ballots (one per digit bit). MATCH.ANY does the same in one
            // instruction but its cost grows with the number of distinct values in the warp (measured: ~half of the
            // kernel's stall samples sat behind it on high-entropy digits); 8 VOTE + 8 LOP3 issue at full rate.
            unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
            for (int b = 0; b < 8; b++) {
                const bool bit = (d >> b) & 1u;
                const unsigned bal = __ballot_sync(0xffffffffu, bit);
                peers &= bit ? bal : ~bal;
            }
            rank[k] = valid ? peers : (1u << lane);  // invalid lanes form private groups and never touch the counters
        }
#pragma unroll
        for (int k = g; k < g + RANK_GROUP; k++) {
            const int e = wbase + k * 32 + lane;
            const bool valid = e < tile_n;
            const unsigned d = (unsigned)(key[k] >> shift) & 255u;
            const unsigned mask = rank[k];
            old[k - g] = 0;
            if ((mask & lt_mask) == 0u && valid) old[k - g] = atomicAdd(&wh[d], (unsigned)__popc(mask));  // lowest lane of the group
            __syncwarp();
        }
#pragma unroll
        for (int k = g; k < g + RANK_GROUP; k++) {
            const unsigned mask = rank[k];
            rank[k] = __shfl_sync(0xffffffffu, old[k - g], __ffs(mask) - 1) + __popc(mask & lt_mask);
        }
    }
    __syncthreads();

    // ---- per digit (thread d): exclusive offsets across warps, tile count, look-back
    {
        const unsigned d = tid;
        unsigned running = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            unsigned c = sm.warp_hist[w][d];
            sm.warp_hist[w][d] = running;
            running += c;
        }
        const unsigned count = running;
        // block exclusive scan over digits for the local reorder
        unsigned incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (unsigned)o) incl += t;
        }
        if (lane == 31) sm.scan_w[warp] = incl;
        // decoupled look-back along the tile chain for this digit; state = {flag:2 | value:30}
        unsigned* my = lookback + (size_t)tile * SORT_RADIX + d;
        unsigned excl = 0;
        if (tile == 0) {
            atomicExch(my, (2u << 30) | count);
        } else {
            atomicExch(my, (1u << 30) | count);
            // LB independent loads per step: the chain of dependent L2 round trips, not HBM, bounds this kernel when every
            // tile of a wave publishes its aggregate at the same moment (measured: ~sqrt(2*tiles) steps per wave)
            constexpr int LB = 8;
            bool done = false;
            for (int t = tile - 1; t >= 0 && !done; t -= LB) {
                unsigned sv[LB];
#pragma unroll
                for (int j = 0; j < LB; j++) {
                    const int idx = t - j;
                    sv[j] = idx >= 0 ? *((const volatile unsigned*)(lookback + (size_t)idx * SORT_RADIX + d)) : (2u << 30);
                }
#pragma unroll
                for (int j = 0; j < LB; j++) {
                    if (done) break;
                    const int idx = t - j;
                    while ((sv[j] >> 30) == 0) sv[j] = *((const volatile unsigned*)(lookback + (size_t)idx * SORT_RADIX + d));
                    excl += sv[j] & 0x3fffffffu;
                    done = (sv[j] >> 30) == 2u;
                }
            }
            atomicExch(my, (2u << 30) | (excl + count));
        }
        // global base of digit d = exclusive scan of the pass' raw histogram, done here by every CTA (256 values) instead
        // of a separate launch
        const unsigned hraw = __ldg(hist_base + d);
        unsigned hincl = hraw;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, hincl, o);
            if (lane >= (unsigned)o) hincl += t;
        }
        if (lane == 31) sm.scan_h[warp] = hincl;
        __syncthreads();
        unsigned woff = 0, hoff = 0;
        for (unsigned w = 0; w < warp; w++) {
            woff += sm.scan_w[w];
            hoff += sm.scan_h[w];
        }
        const unsigned dstart = woff + incl - count;
        sm.digit_start[d] = dstart;
        sm.out_off[d] = (hoff + hincl - hraw) + excl - dstart;  // modular arithmetic: dst = out_off[d] + local index
    }
    __syncthreads();

    // ---- reorder inside the tile through shared memory; values are only read now (16 fewer live registers while ranking)
    uint32_t val[SORT_ITEMS];
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const int e = wbase + k * 32 + lane;
        val[k] = (e < tile_n && vals_in) ? __ldg(vals_in + tile_base + e) : (uint32_t)(tile_base + e);
    }
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const int e = wbase + k * 32 + lane;
        if (e < tile_n) {
            const unsigned d = (unsigned)(key[k] >> shift) & 255u;
            const unsigned pos = sm.digit_start[d] + sm.warp_hist[warp][d] + rank[k];
            sm.keys[pos] = key[k];
            sm.vals[pos] = val[k];
        }
    }
    __syncthreads();
    // ---- scatter: runs of equal digits go to consecutive global addresses
    for (int i = tid; i < tile_n; i += SORT_THREADS) {
        const KeyT kk = sm.keys[i];
        const unsigned d = (unsigned)(kk >> shift) & 255u;
        const unsigned dst = sm.out_off[d] + (unsigned)i;
        keys_out[dst] = kk;
        vals_out[dst] = sm.vals[i];
    }
}

template <typename KeyT>
int sort_pairs_device_t(const KeyT* keys_in, const uint32_t* vals_in, KeyT* keys_out, uint32_t* vals_out, KeyT* keys_tmp,
                        uint32_t* vals_tmp, int n, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                        const char* hist_name, const int* n_dev = nullptr) {
    if (n <= 0) return LODGS_OK;
    int passes = (end_bit - begin_bit + 7) / 8;
    if (passes < 1) passes = 1;
    if (passes > SORT_MAX_PASSES) passes = SORT_MAX_PASSES;
    SortWs w = sort_from_chunk(ws, n);
    const size_t tiles = sort_tiles(n);
    LODGS_CUDA(cudaMemsetAsync(ws, 0, (size_t)((char*)(w.lookback + (size_t)passes * tiles * SORT_RADIX) - (char*)ws), s));
    static std::atomic<bool> attr_set[MAX_DEVICES];
    const int dslot = device_slot();
    if (!attr_set[dslot].load(std::memory_order_acquire)) {
        LODGS_CUDA(cudaFuncSetAttribute(sort_onesweep_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem<KeyT>)));
        attr_set[dslot].store(true, std::memory_order_release);
    }
    int hist_blocks = (int)min((size_t)148 * 4, ((size_t)n + 2047) / 2048);
    sort_histogram_kernel<KeyT><<<hist_blocks, 256, 0, s>>>(keys_in, n, n_dev, passes, begin_bit, w.hist);
    LODGS_LAUNCHED(hist_name);
    const KeyT* kin = keys_in;
    const uint32_t* vin = vals_in;
    for (int p = 0; p < passes; p++) {
        const bool to_out = ((passes - 1 - p) % 2) == 0;
        KeyT* kout = to_out ? keys_out : keys_tmp;
        uint32_t* vout = to_out ? vals_out : vals_tmp;
        sort_onesweep_kernel<KeyT><<<(unsigned)tiles, SORT_THREADS, sizeof(SortSmem<KeyT>), s>>>(
            kin, vin, kout, vout, n, n_dev, begin_bit + 8 * p, w.hist + p * SORT_RADIX, w.tickets + p, w.lookback + (size_t)p * tiles * SORT_RADIX);
        LODGS_LAUNCHED(pass_name);
        kin = kout;
        vin = vout;
    }
    return LODGS_OK;
}

int sort_pairs_device(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, uint64_t* keys_tmp,
                      uint32_t* vals_tmp, int n, int end_bit, void* ws, cudaStream_t s) {
    return sort_pairs_device_t<uint64_t>(keys_in, vals_in, keys_out, vals_out, keys_tmp, vals_tmp, n, 0, end_bit, ws, s,
                                         "sort_onesweep_kernel", "sort_histogram_kernel");
}
int sort_pairs32_device(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t* keys_tmp,
                        uint32_t* vals_tmp, int n, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                        const char* hist_name) {
    return sort_pairs_device_t<uint32_t>(keys_in, vals_in, keys_out, vals_out, keys_tmp, vals_tmp, n, begin_bit, end_bit, ws, s, pass_name,
                                         hist_name);
}
// the same with the element count on the device (*n_dev <= n): the launches are sized for n, the work follows *n_dev
int sort_pairs32_device_n(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t* keys_tmp,
                          uint32_t* vals_tmp, int n, const int* n_dev, int begin_bit, int end_bit, void* ws, cudaStream_t s, const char* pass_name,
                          const char* hist_name) {
    return sort_pairs_device_t<uint32_t>(keys_in, vals_in, keys_out, vals_out, keys_tmp, vals_tmp, n, begin_bit, end_bit, ws, s, pass_name,
                                         hist_name, n_dev);
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_sort_workspace_bytes(int n) { return sort_from_chunk(nullptr, n < 0 ? 0 : n).bytes + 128; }

extern "C" int lodgs_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out, uint32_t* values_out,
                                uint64_t* keys_tmp, uint32_t* values_tmp, int n, int end_bit, void* ws, size_t ws_bytes, void* stream) {
    if (n < 0 || end_bit < 1 || end_bit > 64) { set_error("lodgs_sort_pairs: bad size argument"); return LODGS_E_ARG; }
    if (n == 0) return LODGS_OK;
    if (!keys_in || !values_in || !keys_out || !values_out || !keys_tmp || !values_tmp || !ws) { set_error("lodgs_sort_pairs: null pointer"); return LODGS_E_ARG; }
    if (n >= (1 << 30)) { set_error("lodgs_sort_pairs: n must be < 2^30"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_sort_workspace_bytes(n) - 128) { set_error("lodgs_sort_pairs: workspace too small"); return LODGS_E_WORKSPACE; }
    if (((uintptr_t)ws & 127) != 0) { set_error("lodgs_sort_pairs: workspace must be 128-byte aligned"); return LODGS_E_ARG; }
    prof_stage_begin((cudaStream_t)stream);
    return sort_pairs_device(keys_in, values_in, keys_out, values_out, keys_tmp, values_tmp, n, end_bit, ws, (cudaStream_t)stream);
}
