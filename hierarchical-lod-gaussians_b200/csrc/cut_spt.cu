// cut_spt.cu — cut through the selected SPTs (the distance-sorted subtree tables hanging below the upper hierarchy).
//
// Replaces Switching::getSPTCut (GH/runtime_switching.cu:881-994; kernels binary_search_SPTs :784-810 and populate_SPT :812-852),
// bound as gaussian_hierarchy._C.get_spt_cut_cuda (GH/torch/torch_interface.cpp:287-316) and called once per cut update by
// train_post.py:398, hierarchy_viewer.py:338 and scene/gaussian_model.py:152.
//
// An SPT is a run [SPT_starts[t], SPT_starts[t+1]) of entries (gaussian id, min distance, max distance) sorted by max distance,
// descending. For every selected slot i (table SPT_indices[i] seen from distance SPT_distances[i]) the reference
//   1. bisects the run for the first entry whose max distance is <= the slot's distance (the entry at the run start is never
//      probed) -> interval size n_i,
//   2. lays the S intervals end to end, and for every position of that concatenation looks its slot up again by bisecting the
//      EXCLUSIVE prefix sum with a ">=" test. That lookup returns the last slot whose prefix is strictly below the position, so
//      position p of slot i's interval is evaluated as entry (start_i + p - prefix_i) only for p - prefix_i in 1..n_i; the
//      interval's first entry is skipped and the entry one past its end is tested instead (position 0 is the exception: slot 0,
//      entry 0). The last non-empty slot loses its final position because the concatenation ends there,
//   3. keeps an entry when min distance < the slot's distance, counts kept entries per slot, drops kept entries whose gaussian id
//      is 0 (DeviceSelect::If(!= 0)), and returns the surviving ids in position order plus the exclusive prefix sum of the counts.
// The output is reproduced entry for entry, including 2.'s shifted window, because train_post.py indexes render lists with the
// returned counts. The reference spends ~10 cudaMalloc/cudaFree and ~10 device synchronisations per call on it; here it is four
// launches and one 12-byte read-back: a thread per slot bisects, a warp per slot counts, one block scans the per-slot totals, a
// warp per slot writes its ids at the scanned offset. An entry index at or past n_entries (possible only for the one-past-the-end probe of a
// table that ends the arrays, where the reference reads out of bounds) counts as not kept.
#include "state.cuh"

namespace lodgs {

constexpr int SPT_THREADS = 256;
constexpr int SPT_WARPS = SPT_THREADS / 32;
constexpr int SPT_SCAN_THREADS = 1024;

struct SptHeader {
    int last_nonempty;  // largest slot with a non-empty interval, -1 if none (atomicMax)
    int total;          // surviving ids
    int positions;      // length of the concatenation (sum of the interval sizes)
};

// binary_search_SPTs (:784-810): the probe sequence is kept as is, so unsorted input gives the reference's answer too.
__device__ __forceinline__ int spt_interval(const int32_t* __restrict__ starts, const float* __restrict__ spt_max, int table, float distance) {
    const int first = __ldg(starts + table);
    int low = first, high = __ldg(starts + table + 1);
    int pivot = (low + high) / 2;
    while (high - low > 1) {
        if (__ldg(spt_max + pivot) > distance) low = pivot; else high = pivot;
        pivot = (low + high) / 2;
    }
    return high - first;
}

// Window of entry offsets slot i is evaluated on (see 2. above): [lo, hi], empty when hi < lo.
__device__ __forceinline__ void spt_window(int i, int size, int last_nonempty, int& lo, int& hi) {
    lo = i == 0 ? 0 : 1;
    hi = size - (i == last_nonempty ? 1 : 0);
    if (last_nonempty < 0) hi = lo - 1;                  // nothing anywhere: the concatenation is empty
    else if (size <= 0) hi = i == 0 ? 0 : lo - 1;        // position 0 still lands on slot 0, entry 0
}

__global__ void __launch_bounds__(SPT_THREADS) spt_search_kernel(int S, const int32_t* __restrict__ starts, const float* __restrict__ spt_max,
                                                                const int32_t* __restrict__ spt_indices, const float* __restrict__ spt_distances,
                                                                int32_t* __restrict__ sizes, SptHeader* __restrict__ hdr) {
    const int i = blockIdx.x * SPT_THREADS + threadIdx.x;
    if (i >= S) return;
    const int size = spt_interval(starts, spt_max, __ldg(spt_indices + i), __ldg(spt_distances + i));
    sizes[i] = size;
    if (size > 0) atomicMax(&hdr->last_nonempty, i);
}

// A warp per slot. WRITE = false: kept entries -> hits[i], kept entries with a non-zero id -> emitted[i].
// WRITE = true: the non-zero ids go to cut[offset[i] ...] in entry order.
template <bool WRITE>
__global__ void __launch_bounds__(SPT_THREADS) spt_slot_kernel(int S, const int32_t* __restrict__ gaussian_indices, int n_entries,
                                                              const int32_t* __restrict__ starts, const float* __restrict__ spt_min,
                                                              const int32_t* __restrict__ spt_indices, const float* __restrict__ spt_distances,
                                                              const int32_t* __restrict__ sizes, const SptHeader* __restrict__ hdr,
                                                              int32_t* __restrict__ hits, int32_t* __restrict__ emitted,
                                                              const int32_t* __restrict__ offsets, int32_t* __restrict__ cut, int capacity) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * SPT_WARPS + (threadIdx.x >> 5);
    if (i >= S) return;
    int lo, hi;
    spt_window(i, sizes[i], hdr->last_nonempty, lo, hi);
    const int first = __ldg(starts + __ldg(spt_indices + i));
    const float distance = __ldg(spt_distances + i);
    int n_hits = 0, n_out = 0;
    int base = WRITE ? offsets[i] : 0;
    for (int o = lo + lane; o - lane <= hi; o += 32) {
        const long long e = (long long)first + o;
        bool kept = false;
        int id = 0;
        if (o <= hi && e >= 0 && e < n_entries && __ldg(spt_min + e) < distance) {
            kept = true;
            id = __ldg(gaussian_indices + e);
        }
        const unsigned m_kept = __ballot_sync(0xffffffffu, kept), m_out = __ballot_sync(0xffffffffu, kept && id != 0);
        if (WRITE) {
            const int at = base + __popc(m_out & ((1u << lane) - 1u));
            if (kept && id != 0 && at < capacity) cut[at] = id;
            base += __popc(m_out);
        } else {
            n_hits += __popc(m_kept);
            n_out += __popc(m_out);
        }
    }
    if (!WRITE && lane == 0) {
        hits[i] = n_hits;
        emitted[i] = n_out;
    }
}

// One block: exclusive prefix sums of hits (in place; the second return value of the reference) and of emitted (-> offsets).
__global__ void __launch_bounds__(SPT_SCAN_THREADS) spt_scan_kernel(int S, int32_t* __restrict__ hits, int32_t* __restrict__ emitted,
                                                                   const int32_t* __restrict__ sizes, SptHeader* __restrict__ hdr) {
    __shared__ int s_h[SPT_SCAN_THREADS], s_e[SPT_SCAN_THREADS], s_p[SPT_SCAN_THREADS];
    const int t = threadIdx.x;
    const int per = (S + SPT_SCAN_THREADS - 1) / SPT_SCAN_THREADS;
    const int b = min(S, t * per), e = min(S, b + per);
    int h = 0, m = 0, p = 0;
    for (int i = b; i < e; i++) { h += hits[i]; m += emitted[i]; p += sizes[i]; }
    s_h[t] = h; s_e[t] = m; s_p[t] = p;
    __syncthreads();
    for (int d = 1; d < SPT_SCAN_THREADS; d <<= 1) {
        const int ah = t >= d ? s_h[t - d] : 0, ae = t >= d ? s_e[t - d] : 0, ap = t >= d ? s_p[t - d] : 0;
        __syncthreads();
        s_h[t] += ah; s_e[t] += ae; s_p[t] += ap;
        __syncthreads();
    }
    int run_h = s_h[t] - h, run_e = s_e[t] - m;
    for (int i = b; i < e; i++) {
        const int ch = hits[i], ce = emitted[i];
        hits[i] = run_h;
        emitted[i] = run_e;
        run_h += ch;
        run_e += ce;
    }
    if (t == SPT_SCAN_THREADS - 1) { hdr->total = s_e[t]; hdr->positions = s_p[t]; }
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_spt_cut_workspace_bytes(int S) {
    return align_up(sizeof(SptHeader), 128) + 2 * align_up((size_t)(S > 0 ? S : 1) * sizeof(int32_t), 128);
}

static inline int32_t* spt_sizes(void* ws) { return (int32_t*)((char*)ws + align_up(sizeof(SptHeader), 128)); }
static inline int32_t* spt_offsets(void* ws, int S) { return (int32_t*)((char*)spt_sizes(ws) + align_up((size_t)(S > 0 ? S : 1) * sizeof(int32_t), 128)); }

extern "C" int lodgs_spt_cut_count(int S, const int32_t* gaussian_indices, int n_entries, const int32_t* SPT_starts, const float* SPT_max,
                                   const float* SPT_min, const int32_t* SPT_indices, const float* SPT_distances, int32_t* counts_out,
                                   void* ws, size_t ws_bytes, int32_t* count_host, void* stream) {
    if (S < 0 || n_entries < 0) { set_error("lodgs_spt_cut_count: bad size argument"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    if (S == 0) {
        if (count_host) *count_host = 0;
        return LODGS_OK;
    }
    if (!gaussian_indices || !SPT_starts || !SPT_max || !SPT_min || !SPT_indices || !SPT_distances || !counts_out || !ws) {
        set_error("lodgs_spt_cut_count: null pointer argument");
        return LODGS_E_ARG;
    }
    if (ws_bytes < lodgs_spt_cut_workspace_bytes(S)) { set_error("lodgs_spt_cut_count: workspace too small"); return LODGS_E_WORKSPACE; }
    prof_stage_begin(s);
    SptHeader* hdr = (SptHeader*)ws;
    int32_t* sizes = spt_sizes(ws);
    int32_t* offsets = spt_offsets(ws, S);
    LODGS_CUDA(cudaMemsetAsync(&hdr->last_nonempty, 0xff, sizeof(int), s));  // -1
    LODGS_CUDA(cudaMemsetAsync(&hdr->total, 0, 2 * sizeof(int), s));
    spt_search_kernel<<<(S + SPT_THREADS - 1) / SPT_THREADS, SPT_THREADS, 0, s>>>(S, SPT_starts, SPT_max, SPT_indices, SPT_distances, sizes, hdr);
    LODGS_LAUNCHED("spt_search_kernel");
    spt_slot_kernel<false><<<(S + SPT_WARPS - 1) / SPT_WARPS, SPT_THREADS, 0, s>>>(S, gaussian_indices, n_entries, SPT_starts, SPT_min, SPT_indices,
                                                                                 SPT_distances, sizes, hdr, counts_out, offsets, nullptr, nullptr, 0);
    LODGS_LAUNCHED("spt_slot_kernel<count>");
    spt_scan_kernel<<<1, SPT_SCAN_THREADS, 0, s>>>(S, counts_out, offsets, sizes, hdr);
    LODGS_LAUNCHED("spt_scan_kernel");
    if (count_host) {
        const int prc = publish_words(reinterpret_cast<const int32_t*>(hdr), 3, s);  // mapped pinned words, not a D2H copy
        if (prc) return prc;
        const int32_t* h = wait_words();
        if (!h) { set_error("lodgs_spt_cut_count: waiting for the count failed"); return LODGS_E_CUDA; }
        *count_host = h[1];  // SptHeader::total
    }
    return LODGS_OK;
}

extern "C" int lodgs_spt_cut_write(int S, const int32_t* gaussian_indices, int n_entries, const int32_t* SPT_starts, const float* SPT_min,
                                   const int32_t* SPT_indices, const float* SPT_distances, const void* ws, int32_t* cut_out, int capacity,
                                   void* stream) {
    if (S < 0 || n_entries < 0 || capacity < 0) { set_error("lodgs_spt_cut_write: bad size argument"); return LODGS_E_ARG; }
    if (S == 0 || capacity == 0) return LODGS_OK;
    if (!gaussian_indices || !SPT_starts || !SPT_min || !SPT_indices || !SPT_distances || !ws || !cut_out) {
        set_error("lodgs_spt_cut_write: null pointer argument");
        return LODGS_E_ARG;
    }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const SptHeader* hdr = (const SptHeader*)ws;
    spt_slot_kernel<true><<<(S + SPT_WARPS - 1) / SPT_WARPS, SPT_THREADS, 0, s>>>(S, gaussian_indices, n_entries, SPT_starts, SPT_min, SPT_indices,
                                                                                SPT_distances, spt_sizes((void*)ws), hdr, nullptr, nullptr,
                                                                                spt_offsets((void*)ws, S), cut_out, capacity);
    LODGS_LAUNCHED("spt_slot_kernel<write>");
    return LODGS_OK;
}
