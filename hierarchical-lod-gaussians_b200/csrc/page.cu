// page.cu — CPU <-> GPU paging of Gaussian parameters and Adam moments between training iterations.
//
// Replaces train_post.py:440-479: per tensor (6 parameters + 12 moment tensors) the reference runs `storage[write_back_indices] =
// gpu[write_back_mask].to(cpu)`, `gpu = gpu[keep_mask]`, `gpu = cat(gpu, storage[load_indices].cuda())` — three boolean-mask gathers (each a
// scan + copy + host sync for the size), an advanced-index gather on pageable host memory and two pageable transfers.
// Here the mask is scanned ONCE per iteration (lodgs_page_plan: destination slot of every resident row, kept or evicted, order preserved
// as boolean-mask indexing preserves it), every tensor then needs one pass (lodgs_page_apply: kept rows and the freshly loaded rows go to
// the new resident tensor, evicted rows to a contiguous staging buffer) and ONE transfer each way between pinned staging buffers; the
// host side scatters / gathers whole rows of the big storage tensors with plain memcpy over the host cores (lodgs_host_*_rows).
#include "common.cuh"
#include <cstring>

namespace lodgs {

constexpr int PG_BLOCK = 1024;

__global__ void pg_scan_block_kernel(const uint8_t* __restrict__ keep, long long n, int* __restrict__ rank, int* __restrict__ block_sums) {
    __shared__ int sh[PG_BLOCK];
    const long long i = (long long)blockIdx.x * PG_BLOCK + threadIdx.x;
    const int f = (i < n && keep[i] != 0) ? 1 : 0;
    sh[threadIdx.x] = f;
    __syncthreads();
    for (int off = 1; off < PG_BLOCK; off <<= 1) {
        const int v = threadIdx.x >= off ? sh[threadIdx.x - off] : 0;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    if (i < n) rank[i] = f ? (sh[threadIdx.x] - 1) : -1 - (int)(threadIdx.x + 1 - sh[threadIdx.x] - 1);   // kept: rank among kept; evicted: -1 - rank among evicted (block-local)
    if (threadIdx.x == PG_BLOCK - 1) block_sums[blockIdx.x] = sh[threadIdx.x];
}

// exclusive scan of the per-block kept counts (one CTA; nb <= a few ten thousand), totals to counts[0..1]
__global__ void pg_scan_sums_kernel(int nb, long long n, int* block_sums, int* counts) {
    __shared__ int sh[PG_BLOCK];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nb; b0 += PG_BLOCK) {
        const int b = b0 + threadIdx.x;
        const int v = b < nb ? block_sums[b] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < PG_BLOCK; off <<= 1) {
            const int t = threadIdx.x >= off ? sh[threadIdx.x - off] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        if (b < nb) block_sums[b] = carry + sh[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == PG_BLOCK - 1) carry += sh[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) { counts[0] = carry; counts[1] = (int)(n - carry); }
}

__global__ void pg_scan_add_kernel(long long n, int* __restrict__ rank, const int* __restrict__ block_sums) {
    const long long i = (long long)blockIdx.x * PG_BLOCK + threadIdx.x;
    if (i >= n) return;
    const int kept_before = block_sums[blockIdx.x];
    const int evicted_before = (int)((long long)blockIdx.x * PG_BLOCK - kept_before);
    const int r = rank[i];
    rank[i] = r >= 0 ? r + kept_before : r - evicted_before;
}

// one thread per float of the resident tensor: kept rows -> out[rank], evicted rows -> evict[-1 - rank]; then the loaded rows behind the kept
__global__ void pg_apply_kernel(const float* __restrict__ src, long long n, int width, const int* __restrict__ dst, float* __restrict__ out,
                                float* __restrict__ evict, const float* __restrict__ loaded, long long n_load, long long n_keep) {
    const long long total = (n + n_load) * width;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long row = e / width;
        const int c = (int)(e - row * width);
        if (row < n) {
            const int d = __ldg(dst + row);
            const float v = src[e];
            if (d >= 0) out[(long long)d * width + c] = v;
            else if (evict) evict[(long long)(-1 - d) * width + c] = v;
        } else {
            out[(n_keep + (row - n)) * width + c] = loaded[(row - n) * width + c];
        }
    }
}

// ---- scratch of the allocation-free variant (lodgs_page_plan_static): the per-block kept counts of up to PG_STATIC_MAX_ROWS rows and the
// two totals, held in the module's own device memory (every GPU that loads the library keeps these 64.4 KiB until exit).
constexpr long long PG_STATIC_MAX_ROWS = 16ll << 20;                                   // 16 777 216 resident rows
constexpr int PG_STATIC_WORDS = 32 + (int)(PG_STATIC_MAX_ROWS / PG_BLOCK) + 2;          // [0..1] totals, [32..] block sums
static __device__ __align__(256) int g_pg_static_scratch[PG_STATIC_WORDS];

__global__ void pg_copy_counts_kernel(const int* __restrict__ counts, int32_t* __restrict__ out) {
    if (threadIdx.x < 2) out[threadIdx.x] = counts[threadIdx.x];
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_page_workspace_bytes(int64_t n) { return align_up(((size_t)(n < 0 ? 0 : n) / PG_BLOCK + 2) * 4, 128) + 256; }

extern "C" int lodgs_page_plan(const uint8_t* keep_mask, int64_t n, int32_t* dst, void* ws, size_t ws_bytes, int32_t* counts_host, void* stream) {
    if (n < 0 || n >= (1ll << 31) || (n > 0 && (!keep_mask || !dst)) || !ws) { set_error("lodgs_page_plan: bad argument"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_page_workspace_bytes(n)) { set_error("lodgs_page_plan: workspace too small"); return LODGS_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    int* counts = (int*)ws;                       // [0] kept, [1] evicted
    int* sums = (int*)((char*)ws + 128);
    const int nb = (int)((n + PG_BLOCK - 1) / PG_BLOCK);
    if (n > 0) {
        pg_scan_block_kernel<<<nb, PG_BLOCK, 0, s>>>(keep_mask, n, dst, sums);
        LODGS_LAUNCHED("pg_scan_block_kernel");
    }
    pg_scan_sums_kernel<<<1, PG_BLOCK, 0, s>>>(nb, n, sums, counts);
    LODGS_LAUNCHED("pg_scan_sums_kernel");
    if (n > 0) {
        pg_scan_add_kernel<<<nb, PG_BLOCK, 0, s>>>(n, dst, sums);
        LODGS_LAUNCHED("pg_scan_add_kernel");
    }
    if (counts_host) {
        int rc = publish_words(counts, 2, s);
        if (rc) return rc;
        const int32_t* h = wait_words();
        if (!h) { set_error("lodgs_page_plan: reading the counts back failed"); return LODGS_E_CUDA; }
        counts_host[0] = h[0]; counts_host[1] = h[1];
    }
    return LODGS_OK;
}

extern "C" int64_t lodgs_page_plan_static_max_rows(void) { return (int64_t)PG_STATIC_MAX_ROWS; }

// lodgs_page_plan without a caller workspace and without any allocation: the scan's scratch is g_pg_static_scratch. In host code that name
// is only the host-side stand-in of the symbol; its device address is looked up once per device (with that device current) and cached.
// The totals go to a caller-owned DEVICE buffer (no mapped-pinned read-back: that path sets up an event and pinned words on first use).
extern "C" int lodgs_page_plan_static(const uint8_t* keep_mask, int64_t n, int32_t* dst, int32_t* counts_dev, void* stream) {
    if (n < 0 || (n > 0 && (!keep_mask || !dst)) || !counts_dev) { set_error("lodgs_page_plan_static: bad argument"); return LODGS_E_ARG; }
    if (n > PG_STATIC_MAX_ROWS) {
        set_error("lodgs_page_plan_static: %lld rows exceed the static scratch's maximum of %lld (use lodgs_page_plan with a caller workspace)",
                  (long long)n, PG_STATIC_MAX_ROWS);
        return LODGS_E_WORKSPACE;
    }
    static std::atomic<int*> scratch_of[MAX_DEVICES];
    const int dslot = device_slot();
    int* scratch = scratch_of[dslot].load(std::memory_order_acquire);
    if (!scratch) {
        LODGS_CUDA(cudaGetSymbolAddress((void**)&scratch, g_pg_static_scratch));
        scratch_of[dslot].store(scratch, std::memory_order_release);
    }
    cudaStream_t s = (cudaStream_t)stream;
    int* counts = scratch;        // [0] kept, [1] evicted — written by pg_scan_sums_kernel on every call
    int* sums = scratch + 32;     // one word per block of this call, written before it is read: nothing has to start at zero
    const int nb = (int)((n + PG_BLOCK - 1) / PG_BLOCK);
    if (n > 0) pg_scan_block_kernel<<<nb, PG_BLOCK, 0, s>>>(keep_mask, n, dst, sums);
    pg_scan_sums_kernel<<<1, PG_BLOCK, 0, s>>>(nb, n, sums, counts);
    if (n > 0) pg_scan_add_kernel<<<nb, PG_BLOCK, 0, s>>>(n, dst, sums);
    pg_copy_counts_kernel<<<1, 32, 0, s>>>(counts, counts_dev);
    for (int i = 0; i < (n > 0 ? 4 : 2); i++) count_launch();
    LODGS_CUDA(cudaGetLastError());
    return LODGS_OK;
}

extern "C" int lodgs_page_apply(const float* src, int64_t n, int width, const int32_t* dst, float* out, float* evict, const float* loaded,
                                int64_t n_load, int64_t n_keep, void* stream) {
    if (n < 0 || width <= 0 || n_load < 0 || n_keep < 0 || (n > 0 && (!src || !dst)) || ((n_keep > 0 || n_load > 0) && !out) || (n_load > 0 && !loaded)) {
        set_error("lodgs_page_apply: bad argument");
        return LODGS_E_ARG;
    }
    const long long total = (n + n_load) * width;
    if (total == 0) return LODGS_OK;
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const int blocks = (int)std::min<long long>((total + 255) / 256, 148LL * 32);
    pg_apply_kernel<<<blocks, 256, 0, s>>>(src, n, width, dst, out, evict, loaded, n_load, n_keep);
    LODGS_LAUNCHED("pg_apply_kernel");
    return LODGS_OK;
}

// storage[indices[i], :] = staging[i, :]   (train_post.py:440-445 `gaussians._xyz[write_back_indices, :] = ...`); rows of `row_bytes`
extern "C" int lodgs_host_scatter_rows(void* storage, int64_t storage_rows, int64_t row_bytes, const int64_t* indices, int64_t n, const void* staging) {
    if (n < 0 || row_bytes <= 0 || (n > 0 && (!storage || !indices || !staging))) { set_error("lodgs_host_scatter_rows: bad argument"); return LODGS_E_ARG; }
    for (int64_t i = 0; i < n; i++)
        if (indices[i] < 0 || indices[i] >= storage_rows) { set_error("lodgs_host_scatter_rows: index %lld out of range", (long long)indices[i]); return LODGS_E_ARG; }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) memcpy((char*)storage + indices[i] * row_bytes, (const char*)staging + i * row_bytes, (size_t)row_bytes);
    return LODGS_OK;
}

// staging[i, :] = storage[indices[i], :]   (train_post.py:456-461 `gaussians._xyz[load_from_disk_indices]`)
extern "C" int lodgs_host_gather_rows(const void* storage, int64_t storage_rows, int64_t row_bytes, const int64_t* indices, int64_t n, void* staging) {
    if (n < 0 || row_bytes <= 0 || (n > 0 && (!storage || !indices || !staging))) { set_error("lodgs_host_gather_rows: bad argument"); return LODGS_E_ARG; }
    for (int64_t i = 0; i < n; i++)
        if (indices[i] < 0 || indices[i] >= storage_rows) { set_error("lodgs_host_gather_rows: index %lld out of range", (long long)indices[i]); return LODGS_E_ARG; }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) memcpy((char*)staging + i * row_bytes, (const char*)storage + indices[i] * row_bytes, (size_t)row_bytes);
    return LODGS_OK;
}
