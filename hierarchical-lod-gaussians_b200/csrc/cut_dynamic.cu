// cut_dynamic.cu — LOD cut and interpolation weights on the DYNAMIC hierarchy format (sm_100a).
//
// Replaces Switching::expandToSizeDynamic (GH/runtime_switching.cu:748-782: markNodesForSizeDynamic :533-582 + cub
// InclusiveSum + putRenderIndicesDynamic :95-108) and Switching::getTsIndexedDynamic (computeTsIndexedDynamic :637-684).
// Nodes are `HierarchyNode` records of 6 x i32 {depth, parent, child_count, first_child, next_sibling, max_side_length}
// (GH/types.h:60-79) — one Gaussian per node, so a cut entry is the node id itself — with per-node positions [N,3] and
// scales [N,3]. The size of a node is max(scale) / |viewpoint - position| (computeSizeGPUDynamic :222-232) and nodes
// outside the view cone (cos(angle to -zdir) >= -0.5, is_in_frustum :165-186) are dropped.
//
// Same single-pass design as cut_flat_kernel (cut.cu): one kernel reads 48 B per node (24 B record + 12 B position +
// 12 B scale; parents' 24 B come from L2), counts are 0/1 so the warp scan is a ballot + popc, tiles of 4096 nodes chain
// through a decoupled look-back, entries are scattered in ascending node-id order. The reference uses 3 kernels, a CUB
// scan and two N-sized device_vectors allocated per call, and reads the count back with a blocking memcpy.
#include "common.cuh"
#include <cfloat>

namespace lodgs {

// layout shared with cut.cu (first words of the cut workspace)
struct CutHeaderDyn {
    int count;
    int ticket;
    int overflow;
    int pad[5];
};

constexpr int DYN_THREADS = 512;
constexpr int DYN_ROWS = 8;
constexpr int DYN_TILE = DYN_THREADS * DYN_ROWS;
constexpr int DYN_WARPS = DYN_THREADS / 32;
constexpr int DYN_PARTS = DYN_ROWS * DYN_WARPS;  // (row, warp) partial sums per tile

struct DynGeom {
    float dx, dy, dz, norm;
};

// |viewpoint - position|, IEEE sqrt, with the association the reference SASS uses — which differs between its two kernels:
// markNodesForSizeDynamic (viewpoint read through a pointer) compiles to d2 = fma(dz,dz, fma(dy,dy, dx*dx)),
// computeTsIndexedDynamic (viewpoint passed by value) to d2 = fma(dz,dz, fma(dx,dx, dy*dy)).
template <bool Y_FIRST>
__device__ __forceinline__ DynGeom dyn_geom(const float* __restrict__ positions, int i, float vx, float vy, float vz) {
    DynGeom g;
    g.dx = __fsub_rn(vx, __ldg(positions + 3 * (size_t)i));
    g.dy = __fsub_rn(vy, __ldg(positions + 3 * (size_t)i + 1));
    g.dz = __fsub_rn(vz, __ldg(positions + 3 * (size_t)i + 2));
    const float d2 = Y_FIRST ? __fmaf_rn(g.dz, g.dz, __fmaf_rn(g.dx, g.dx, __fmul_rn(g.dy, g.dy)))
                             : __fmaf_rn(g.dz, g.dz, __fmaf_rn(g.dy, g.dy, __fmul_rn(g.dx, g.dx)));
    g.norm = __fsqrt_rn(d2);
    return g;
}

// computeSizeGPUDynamic (:222-232): max(scale) / distance; a negative distance cannot occur, 0 gives inf/NaN as there.
__device__ __forceinline__ float dyn_size(const float* __restrict__ scales, int i, float norm) {
    if (norm < 0.0f) return 0.0f;
    const float s = fmaxf(__ldg(scales + 3 * (size_t)i), fmaxf(__ldg(scales + 3 * (size_t)i + 1), __ldg(scales + 3 * (size_t)i + 2)));
    return __fdiv_rn(s, norm);
}

// is_in_frustum (:165-186): cosine between the normalised (viewpoint - position) and zdir below -0.5 (IEEE divides,
// dot product as x*zx -> fma(y,zy,.) -> fma(z,zz,.) like the reference SASS).
__device__ __forceinline__ bool dyn_in_cone(const DynGeom& g, float zx, float zy, float zz) {
    const float nx = __fdiv_rn(g.dx, g.norm), ny = __fdiv_rn(g.dy, g.norm), nz = __fdiv_rn(g.dz, g.norm);
    const float c = __fmaf_rn(nz, zz, __fmaf_rn(ny, zy, __fmul_rn(nx, zx)));
    return c < -0.5f;
}

__global__ void __launch_bounds__(DYN_THREADS, 2) cut_dynamic_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ positions,
                                                                    const float* __restrict__ scales, int N, float tau,
                                                                    const float* __restrict__ viewpoint, float zx, float zy, float zz,
                                                                    int32_t* __restrict__ render_indices,
                                                                    int32_t* __restrict__ parent_indices, int32_t* __restrict__ node_ids,
                                                                    int capacity, CutHeaderDyn* hdr, unsigned long long* tile_state) {
    __shared__ int s_tile;
    __shared__ int s_part[DYN_PARTS];
    __shared__ unsigned long long s_prefix;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&hdr->ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const int base = tile * DYN_TILE;
    const int cnt = min(DYN_TILE, N - base);
    const float vx = __ldg(viewpoint), vy = __ldg(viewpoint + 1), vz = __ldg(viewpoint + 2);

    unsigned sel_bits = 0;       // bit r: this thread's node of row r is on the cut
    int parent[DYN_ROWS];
    unsigned below[DYN_ROWS];    // selected nodes of lower lanes in the same (row, warp)
#pragma unroll
    for (int r = 0; r < DYN_ROWS; r++) {
        const int j = r * DYN_THREADS + (int)tid;
        bool sel = false;
        parent[r] = -1;
        if (j < cnt) {
            const int i = base + j;
            const int2 dp = __ldg(reinterpret_cast<const int2*>(nodes + (size_t)i * 6));  // depth, parent (24-byte records: 8-byte aligned)
            const int child_count = __ldg(nodes + (size_t)i * 6 + 2);
            parent[r] = dp.y;
            const DynGeom g = dyn_geom<false>(positions, i, vx, vy, vz);
            if (dyn_in_cone(g, zx, zy, zz) && dp.x >= 0) {  // skybox nodes (depth < 0) are added elsewhere (:556-560)
                const float size = dyn_size(scales, i, g.norm);
                if (size >= tau && child_count == 0) {
                    sel = true;  // leaf that is still too coarse
                } else if (dp.y >= 0) {
                    const DynGeom pg = dyn_geom<false>(positions, dp.y, vx, vy, vz);
                    const float psize = dyn_size(scales, dp.y, pg.norm);
                    sel = psize >= tau && size < tau;  // first node below the threshold on its path
                }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, sel);
        below[r] = __popc(m & ((1u << lane) - 1u));
        if (sel) sel_bits |= 1u << r;
        if (lane == 0) s_part[r * DYN_WARPS + warp] = __popc(m);
    }
    __syncthreads();
    if (warp == 0) {
        constexpr int PPL = DYN_PARTS / 32;
        int w[PPL], wsum = 0;
#pragma unroll
        for (int k = 0; k < PPL; k++) {
            w[k] = s_part[lane * PPL + k];
            wsum += w[k];
        }
        int wi = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (unsigned)o) wi += t;
        }
        int run = wi - wsum;
#pragma unroll
        for (int k = 0; k < PPL; k++) {
            s_part[lane * PPL + k] = run;
            run += w[k];
        }
        const int block_total = __shfl_sync(0xffffffffu, wi, 31);
        unsigned long long excl = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(tile_state, pack_state(2, (unsigned long long)block_total));
        } else {
            if (lane == 0) st_relaxed(tile_state + tile, pack_state(1, (unsigned long long)block_total));
            excl = lookback_warp(tile_state, tile);
            if (lane == 0) st_relaxed(tile_state + tile, pack_state(2, excl + (unsigned long long)block_total));
        }
        if (lane == 0) {
            s_prefix = excl;
            if ((long long)base + DYN_TILE >= N) {
                const long long total = (long long)excl + block_total;
                hdr->count = (int)total;
                if (total > capacity) hdr->overflow = 1;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < DYN_ROWS; r++) {
        if (sel_bits & (1u << r)) {
            const long long o = (long long)s_prefix + s_part[r * DYN_WARPS + warp] + below[r];
            if (o < capacity) {
                const int id = base + r * DYN_THREADS + (int)tid;
                render_indices[o] = id;
                if (node_ids) node_ids[o] = id;
                // putRenderIndicesDynamic (:104-107) leaves the caller's value in place for root entries
                if (parent_indices && parent[r] != -1) parent_indices[o] = parent[r];
            }
        }
    }
}

// computeTsIndexedDynamic (:637-684)
__global__ void __launch_bounds__(256) cut_weights_dynamic_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ positions,
                                                                 const float* __restrict__ scales, int K,
                                                                 const int32_t* __restrict__ indices, float vx, float vy, float vz, float tau,
                                                                 float* __restrict__ ts, int32_t* __restrict__ num_siblings) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= K) return;
    const int node_id = __ldg(indices + idx);
    const int parent = __ldg(nodes + (size_t)node_id * 6 + 1);
    float t = 1.0f;
    int sib = 1;
    if (parent >= 0) {
        const float parentsize = dyn_size(scales, parent, dyn_geom<true>(positions, parent, vx, vy, vz).norm);
        if (!(parentsize > __fmul_rn(2.0f, tau))) {
            const float size = dyn_size(scales, node_id, dyn_geom<true>(positions, node_id, vx, vy, vz).norm);
            const float start = fmaxf(__fmul_rn(0.5f, parentsize), size);
            const float diff = __fsub_rn(parentsize, start);
            if (!(diff <= 0)) {
                const float tdiff = fmaxf(0.0f, __fsub_rn(tau, start));
                t = fmaxf(__fsub_rn(1.0f, __fdiv_rn(tdiff, diff)), 0.0f);
            }
        }
        sib = __ldg(nodes + (size_t)parent * 6 + 2);
    }
    ts[idx] = t;
    num_siblings[idx] = sib;
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_cut_workspace_bytes(int N);

extern "C" int lodgs_expand_to_size_dynamic(const int32_t* nodes, const float* positions, const float* scales, int N, float target_size,
                                            const float* viewpoint_dev, float zx, float zy, float zz, int32_t* render_indices,
                                            int32_t* parent_indices, int32_t* nodes_for_render_indices, int capacity, void* ws,
                                            size_t ws_bytes, int32_t* count_host, void* stream) {
    if (!nodes || !positions || !scales || !viewpoint_dev || !render_indices || !ws || N < 0 || capacity < 0) {
        set_error("lodgs_expand_to_size_dynamic: bad argument");
        return LODGS_E_ARG;
    }
    if (ws_bytes < lodgs_cut_workspace_bytes(N)) { set_error("lodgs_expand_to_size_dynamic: workspace too small"); return LODGS_E_WORKSPACE; }
    if (((uintptr_t)nodes & 7) != 0) { set_error("lodgs_expand_to_size_dynamic: nodes must be 8-byte aligned"); return LODGS_E_ARG; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    CutHeaderDyn* hdr = (CutHeaderDyn*)ws;
    unsigned long long* states = (unsigned long long*)((char*)ws + align_up(sizeof(CutHeaderDyn), 128));
    const int tiles = (N + DYN_TILE - 1) / DYN_TILE;
    LODGS_CUDA(cudaMemsetAsync(ws, 0, align_up(sizeof(CutHeaderDyn), 128) + (size_t)(tiles + 1) * 8, s));
    if (N == 0) {
        if (count_host) { *count_host = 0; LODGS_CUDA(cudaStreamSynchronize(s)); }
        return LODGS_OK;
    }
    cut_dynamic_kernel<<<tiles, DYN_THREADS, 0, s>>>(nodes, positions, scales, N, target_size, viewpoint_dev, zx, zy, zz, render_indices,
                                                    parent_indices, nodes_for_render_indices, capacity, hdr, states);
    LODGS_LAUNCHED("cut_dynamic_kernel");
    if (count_host) {
        const int prc = publish_words(reinterpret_cast<const int32_t*>(hdr), 3, s);  // mapped pinned words, not a D2H copy
        if (prc) return prc;
        const int32_t* h = wait_words();
        if (!h) { set_error("lodgs_expand_to_size_dynamic: waiting for the count failed"); return LODGS_E_CUDA; }
        *count_host = h[0];
        if (h[2]) { set_error("lodgs_expand_to_size_dynamic: cut has %d entries, output capacity is %d", h[0], capacity); return LODGS_E_WORKSPACE; }
    }
    return LODGS_OK;
}

extern "C" int lodgs_get_interpolation_weights_dynamic(const int32_t* node_indices, int K, float target_size, const int32_t* nodes,
                                                       const float* positions, const float* scales, float vx, float vy, float vz,
                                                       float* ts, int32_t* num_siblings, void* stream) {
    if (K < 0 || (K > 0 && (!node_indices || !nodes || !positions || !scales || !ts || !num_siblings))) {
        set_error("lodgs_get_interpolation_weights_dynamic: bad argument");
        return LODGS_E_ARG;
    }
    if (K == 0) return LODGS_OK;
    prof_stage_begin((cudaStream_t)stream);
    cut_weights_dynamic_kernel<<<(K + 255) / 256, 256, 0, (cudaStream_t)stream>>>(nodes, positions, scales, K, node_indices, vx, vy, vz,
                                                                                 target_size, ts, num_siblings);
    LODGS_LAUNCHED("cut_weights_dynamic_kernel");
    return LODGS_OK;
}
