// cut.cu — LOD cut selection and interpolation weights (sm_100a).
//
// Replaces GH/runtime_switching.cu: markNodesForSize (:495-529) + cub InclusiveSum + putRenderIndices (:68-93)
// [expandToSize :997-1030] and computeTsIndexed (:588-634) [getTsIndexed :687-714].
//
// Design (B200): the reference runs 3 kernels + a CUB scan over all N nodes with two N-sized temporaries
// allocated per call. Here the flat rule is ONE kernel: each 512-node tile evaluates the size test, does a
// block scan of the render counts, resolves its global offset with a decoupled look-back (single-pass chained
// scan) and scatters its entries — 60 B read per node, 12 B written per cut entry, nothing else touches HBM.
// The traversal mode visits only nodes whose parent passes the size test (level-synchronous frontier, one warp
// per frontier chunk, siblings in adjacent lanes), marks selected nodes in an N-bit map and compacts the map in
// ascending node-id order so that the output is bit-identical to the flat rule.
#include "common.cuh"
#include <cfloat>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace lodgs {

struct NodeRec {
    int depth, parent, start, count_leafs, count_merged, start_children, count_children;
};

// Projected size of a node's AABB seen from `v` (computeSizeGPU, runtime_switching.cu:210-219 with
// inboxCUDA :136-144 and pointboxdistCUDA :189-204). The squared distance keeps the reference's compiled
// association t = dy*dy; t = fma(dx,dx,t); t = fma(dz,dz,t) (SURVEY top, SASS of markNodesForSize), with IEEE
// sqrt and divide, so `size >= tau` decisions are bit-identical.
__device__ __forceinline__ float node_size(const float4 bmin, const float4 bmax, float vx, float vy, float vz) {
    bool inside = (vx >= bmin.x && vx <= bmax.x) && (vy >= bmin.y && vy <= bmax.y) && (vz >= bmin.z && vz <= bmax.z);
    if (inside) return FLT_MAX;
    float cx = fmaxf(bmin.x, fminf(bmax.x, vx));
    float cy = fmaxf(bmin.y, fminf(bmax.y, vy));
    float cz = fmaxf(bmin.z, fminf(bmax.z, vz));
    float dx = __fsub_rn(vx, cx), dy = __fsub_rn(vy, cy), dz = __fsub_rn(vz, cz);
    float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    return __fdiv_rn(bmin.w, __fsqrt_rn(d2));
}

__device__ __forceinline__ NodeRec load_node(const int32_t* nodes, int i) {
    const int32_t* p = nodes + (size_t)i * 7;
    NodeRec n;
    n.depth = __ldg(p + 0); n.parent = __ldg(p + 1); n.start = __ldg(p + 2); n.count_leafs = __ldg(p + 3);
    n.count_merged = __ldg(p + 4); n.start_children = __ldg(p + 5); n.count_children = __ldg(p + 6);
    return n;
}

// Render count of node i under the flat rule (markNodesForSize :507-522).
__device__ __forceinline__ int render_count(const NodeRec& n, float size, const float* boxes, float tau, float vx, float vy,
                                            float vz) {
    if (size >= tau) return n.count_leafs;
    if (n.parent != -1) {
        const float4* pb = reinterpret_cast<const float4*>(boxes) + (size_t)n.parent * 2;
        float psize = node_size(__ldg(pb), __ldg(pb + 1), vx, vy, vz);
        if (psize >= tau) return n.count_leafs + (n.depth != 0 ? n.count_merged : 0);
    }
    return 0;
}

struct CutHeader {
    int count;      // total number of cut entries (valid after the kernel)
    int ticket;     // dynamic tile id dispenser
    int overflow;   // set when count > capacity
    int monotone;   // lodgs_cut_prepare result
    int frontier_n[3];
    int pad[1];
};

constexpr int CUT_BLOCK = 512;

// Flat rule, one pass. A tile is FLAT_TILE = 4096 consecutive nodes handled by 512 threads in 8 rows: thread t owns
// nodes t, t+512, ... (box loads are fully coalesced 16-byte loads, the 28-byte records of a warp are 896 contiguous
// bytes served from L1 after the first touch; output rows are coalesced). Counts are scanned in node order, the tile's
// global offset comes from a decoupled look-back, entries are scattered in ascending node-id order.
// Tile size matters more than anything else here: every tile publishes/polls look-back states that live on a few L2
// lines, and with 512-node tiles that polling (not HBM) bounded the kernel at ~60 tiles/us whatever the staging scheme
// (TMA-staged records, persistent prefetching CTAs and wider look-back windows were all measured, profiles/).
constexpr int FLAT_THREADS = 512;
constexpr int FLAT_ROWS = 8;
constexpr int FLAT_TILE = FLAT_THREADS * FLAT_ROWS;
constexpr int FLAT_PARTS = FLAT_ROWS * (FLAT_THREADS / 32);  // 64 (row, warp) partial sums per tile

__global__ void __launch_bounds__(FLAT_THREADS, 2) cut_flat_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ boxes,
                                                                  int N, float tau, const float* __restrict__ viewpoint,
                                                                  int32_t* __restrict__ render_indices,
                                                                  int32_t* __restrict__ parent_indices,
                                                                  int32_t* __restrict__ node_ids, int capacity, CutHeader* hdr,
                                                                  unsigned long long* tile_state) {
    __shared__ int s_tile;
    __shared__ int s_part[FLAT_PARTS];  // per (row, warp) totals -> exclusive offsets
    __shared__ unsigned long long s_prefix;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&hdr->ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const int base = tile * FLAT_TILE;
    const int cnt = min(FLAT_TILE, N - base);
    const float vx = __ldg(viewpoint), vy = __ldg(viewpoint + 1), vz = __ldg(viewpoint + 2);

    int count[FLAT_ROWS], start[FLAT_ROWS], pgauss[FLAT_ROWS], incl[FLAT_ROWS];
#pragma unroll
    for (int r = 0; r < FLAT_ROWS; r++) {
        const int j = r * FLAT_THREADS + (int)tid;
        count[r] = 0; start[r] = 0; pgauss[r] = -1;
        if (j < cnt) {
            const int i = base + j;
            const float4* b = reinterpret_cast<const float4*>(boxes) + (size_t)i * 2;
            const float4 bmin = __ldg(b), bmax = __ldg(b + 1);
            const NodeRec n = load_node(nodes, i);
            // the parent's Gaussian id is needed by every emitted entry: fetch it now, off the critical path
            if (n.parent != -1) pgauss[r] = __ldg(nodes + (size_t)n.parent * 7 + 2);
            const float size = node_size(bmin, bmax, vx, vy, vz);
            count[r] = render_count(n, size, boxes, tau, vx, vy, vz);
            start[r] = n.start;
        }
        int v = count[r];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= (unsigned)o) v += t;
        }
        incl[r] = v;
        if (lane == 31) s_part[r * (FLAT_THREADS / 32) + warp] = v;
    }
    __syncthreads();
    if (warp == 0) {
        // FLAT_PARTS partial sums in node order: lane l scans entries [l*PPL, (l+1)*PPL)
        constexpr int PPL = FLAT_PARTS / 32;
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
            if ((long long)base + FLAT_TILE >= N) {  // last tile: total count
                const long long total = (long long)excl + block_total;
                hdr->count = (int)total;
                if (total > capacity) hdr->overflow = 1;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < FLAT_ROWS; r++) {
        if (count[r] > 0) {
            const long long off = (long long)s_prefix + s_part[r * (FLAT_THREADS / 32) + warp] + (incl[r] - count[r]);
            const int id = base + r * FLAT_THREADS + (int)tid;
            for (int k = 0; k < count[r]; k++) {
                const long long o = off + k;
                if (o < capacity) {
                    render_indices[o] = start[r] + k;
                    if (parent_indices) parent_indices[o] = pgauss[r];
                    if (node_ids) node_ids[o] = id;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Traversal mode. Frontier entries are node ids whose PARENT passed (size >= tau), plus the root(s).
// A node with size >= tau is "expanded": it contributes count_leafs and pushes its children; otherwise it
// contributes count_leafs + (depth != 0) * count_merged (its parent passed by construction). Under monotone
// sizes (lodgs_cut_prepare) nodes never reached have size < tau and parent size < tau => flat count 0.
// Selected nodes get their count stored in a sparse way: bit in `mark` (N bits); counts are recomputed from
// the node record during compaction (count is a pure function of the record and of which branch was taken,
// recorded in a second bitmap `expanded`).
__global__ void __launch_bounds__(256) cut_traverse_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ boxes,
                                                          float tau, const float* __restrict__ viewpoint, int* frontier_a,
                                                          int* frontier_b, CutHeader* hdr, unsigned* mark,
                                                          unsigned* expanded) {
    cg::grid_group grid = cg::this_grid();
    const float vx = __ldg(viewpoint), vy = __ldg(viewpoint + 1), vz = __ldg(viewpoint + 2);
    const unsigned lane = threadIdx.x & 31;
    volatile int* fn = hdr->frontier_n;  // 3 rotating counters: level L reads [L%3], appends to [(L+1)%3], clears [(L+2)%3]
    int* fin = frontier_a;
    int* fout = frontier_b;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        frontier_a[0] = 0;  // single root = node 0 (validated by lodgs_cut_prepare)
        fn[0] = 1;
    }
    grid.sync();
    for (int level = 0;; level++) {
        const int n_in = fn[level % 3];
        if (n_in == 0) break;
        int* next_count = const_cast<int*>(&fn[(level + 1) % 3]);
        if (blockIdx.x == 0 && threadIdx.x == 0) fn[(level + 2) % 3] = 0;
        for (int base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31; base < n_in; base += gridDim.x * blockDim.x) {
            int f = base + lane;
            int nchild = 0, cstart = 0;
            if (f < n_in) {
                int i = fin[f];
                NodeRec n = load_node(nodes, i);
                const float4* b = reinterpret_cast<const float4*>(boxes) + (size_t)i * 2;
                float size = node_size(__ldg(b), __ldg(b + 1), vx, vy, vz);
                bool pass = size >= tau;
                int count = pass ? n.count_leafs : (n.parent != -1 ? n.count_leafs + (n.depth != 0 ? n.count_merged : 0) : 0);
                if (count != 0) atomicOr(mark + (i >> 5), 1u << (i & 31));
                if (pass) {
                    atomicOr(expanded + (i >> 5), 1u << (i & 31));
                    nchild = n.count_children;
                    cstart = n.start_children;
                    if (cstart < 0) nchild = 0;
                }
            }
            // warp-cooperative append of all children: one atomic per warp, slots filled by all lanes
            int incl = nchild;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= (unsigned)o) incl += t;
            }
            int total = __shfl_sync(0xffffffffu, incl, 31);
            if (total == 0) continue;
            int wbase = 0;
            if (lane == 0) wbase = atomicAdd(next_count, total);
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            int excl = incl - nchild;
            for (int s0 = 0; s0 < total; s0 += 32) {
                int s = s0 + lane;
                // owner lane = last lane whose exclusive offset <= s (binary search over the warp's offsets)
                int lo = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    int probe = lo + step;
                    int v = __shfl_sync(0xffffffffu, excl, probe & 31);
                    if (probe < 32 && v <= s) lo = probe;
                }
                // skip back over lanes with zero children that share the same exclusive offset: the last lane with
                // excl <= s is the owner only if it has children covering s; lanes with nchild==0 at the tail have
                // excl == total > s, lanes in the middle with nchild==0 have excl equal to the next owner's excl, and
                // "last lane with excl <= s" picks the highest such lane, which is the one that owns s.
                int owner_excl = __shfl_sync(0xffffffffu, excl, lo);
                int owner_start = __shfl_sync(0xffffffffu, cstart, lo);
                if (s < total) fout[wbase + s] = owner_start + (s - owner_excl);
            }
        }
        grid.sync();
        int* t = fin; fin = fout; fout = t;
    }
}

// Compaction of the bitmap in ascending node id order: each thread owns one 32-bit word of `mark`.
__global__ void __launch_bounds__(CUT_BLOCK) cut_compact_kernel(const int32_t* __restrict__ nodes, int N,
                                                               const unsigned* __restrict__ mark,
                                                               const unsigned* __restrict__ expanded,
                                                               int32_t* __restrict__ render_indices,
                                                               int32_t* __restrict__ parent_indices,
                                                               int32_t* __restrict__ node_ids, int capacity, CutHeader* hdr,
                                                               unsigned long long* tile_state) {
    __shared__ int s_tile;
    __shared__ int s_warp_sum[CUT_BLOCK / 32];
    __shared__ unsigned long long s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(&hdr->ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const int nwords = (N + 31) >> 5;
    const int w = tile * CUT_BLOCK + threadIdx.x;
    unsigned bits = (w < nwords) ? __ldg(mark + w) : 0u;
    unsigned ebits = (bits != 0u) ? __ldg(expanded + w) : 0u;
    int count = 0;
    for (unsigned b = bits; b; b &= b - 1) {
        int i = (w << 5) + (__ffs(b) - 1);
        const int32_t* p = nodes + (size_t)i * 7;
        int c = __ldg(p + 3);
        if (!((ebits >> (i & 31)) & 1u)) c += (__ldg(p + 0) != 0 ? __ldg(p + 4) : 0);
        count += c;
    }
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += t;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int ws = (lane < CUT_BLOCK / 32) ? s_warp_sum[lane] : 0;
        int wi = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (unsigned)o) wi += t;
        }
        if (lane < CUT_BLOCK / 32) s_warp_sum[lane] = wi - ws;
        int block_total = __shfl_sync(0xffffffffu, wi, CUT_BLOCK / 32 - 1);
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
            if ((long long)tile * CUT_BLOCK + CUT_BLOCK >= nwords) {
                long long total = (long long)excl + block_total;
                hdr->count = (int)total;
                if (total > capacity) hdr->overflow = 1;
            }
        }
    }
    __syncthreads();
    if (count > 0) {
        long long off = (long long)s_prefix + s_warp_sum[warp] + (incl - count);
        for (unsigned b = bits; b; b &= b - 1) {
            int i = (w << 5) + (__ffs(b) - 1);
            NodeRec n = load_node(nodes, i);
            int c = n.count_leafs;
            if (!((ebits >> (i & 31)) & 1u)) c += (n.depth != 0 ? n.count_merged : 0);
            int parentgaussian = -1;
            if (n.parent != -1) parentgaussian = __ldg(nodes + (size_t)n.parent * 7 + 2);
            for (int k = 0; k < c; k++, off++) {
                if (off < capacity) {
                    render_indices[off] = n.start + k;
                    if (parent_indices) parent_indices[off] = parentgaussian;
                    if (node_ids) node_ids[off] = i;
                }
            }
        }
    }
}

// Monotonicity check for the traversal (SURVEY §8a "Cut semantics"): box(parent) ⊇ box(child) and
// w(parent) >= w(child) for every edge, parent id < ... (no ordering assumption needed).
__global__ void cut_check_monotone_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ boxes, int N,
                                          CutHeader* hdr) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int parent = __ldg(nodes + (size_t)i * 7 + 1);
    // child blocks must not overlap: every member of this node's block [start_children, +count_children) names THIS node as its
    // parent (a node listed by two parents would enter the frontier twice and could overflow its N entries)
    {
        int cs = __ldg(nodes + (size_t)i * 7 + 5), cc = __ldg(nodes + (size_t)i * 7 + 6);
        if (cc > 0) {
            if (cs < 0 || (long long)cs + cc > N) hdr->monotone = 0;
            else
                for (int c = cs; c < cs + cc; c++)
                    if (__ldg(nodes + (size_t)c * 7 + 1) != i) { hdr->monotone = 0; break; }
        }
    }
    if (parent < 0) {
        if (i != 0) hdr->monotone = 0;  // the traversal starts from node 0 as the single root
        return;
    }
    if (i == 0) { hdr->monotone = 0; return; }
    const float4* b = reinterpret_cast<const float4*>(boxes) + (size_t)i * 2;
    const float4* pb = reinterpret_cast<const float4*>(boxes) + (size_t)parent * 2;
    float4 cmin = __ldg(b), cmax = __ldg(b + 1), pmin = __ldg(pb), pmax = __ldg(pb + 1);
    bool ok = pmin.x <= cmin.x && pmin.y <= cmin.y && pmin.z <= cmin.z && pmax.x >= cmax.x && pmax.y >= cmax.y &&
              pmax.z >= cmax.z && pmin.w >= cmin.w && cmin.w >= 0.0f;
    // the child must also be listed inside its parent's child block, otherwise the traversal cannot reach it
    int ps = __ldg(nodes + (size_t)parent * 7 + 5), pc = __ldg(nodes + (size_t)parent * 7 + 6);
    ok = ok && (i >= ps && i < ps + pc);
    if (!ok) hdr->monotone = 0;
}

__global__ void cut_weights_kernel(const int32_t* __restrict__ nodes, const float* __restrict__ boxes, int K,
                                   const int32_t* __restrict__ indices, float vx, float vy, float vz, float tau,
                                   float* __restrict__ ts, int32_t* __restrict__ num_siblings) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= K) return;
    int node_id = __ldg(indices + idx);
    int parent = __ldg(nodes + (size_t)node_id * 7 + 1);
    float t;
    int sib = 1;
    if (parent == -1) {
        t = 1.0f;
    } else {
        const float4* pb = reinterpret_cast<const float4*>(boxes) + (size_t)parent * 2;
        float parentsize = node_size(__ldg(pb), __ldg(pb + 1), vx, vy, vz);
        if (parentsize > __fmul_rn(2.0f, tau)) {
            t = 1.0f;
        } else {
            const float4* b = reinterpret_cast<const float4*>(boxes) + (size_t)node_id * 2;
            float size = node_size(__ldg(b), __ldg(b + 1), vx, vy, vz);
            float start = fmaxf(__fmul_rn(0.5f, parentsize), size);
            float diff = __fsub_rn(parentsize, start);
            if (diff <= 0) {
                t = 1.0f;
            } else {
                float tdiff = fmaxf(0.0f, __fsub_rn(tau, start));
                t = fmaxf(__fsub_rn(1.0f, __fdiv_rn(tdiff, diff)), 0.0f);
            }
        }
        sib = __ldg(nodes + (size_t)parent * 7 + 6);
    }
    ts[idx] = t;
    num_siblings[idx] = sib;
}

}  // namespace lodgs

using namespace lodgs;

static size_t cut_states(int N) { return (size_t)(N + CUT_BLOCK - 1) / CUT_BLOCK + 1; }  // CUT_BLOCK <= FLAT_TILE: enough for both scans

extern "C" size_t lodgs_cut_workspace_bytes(int N) {
    if (N < 0) N = 0;
    size_t words = ((size_t)N + 31) / 32;
    // header | tile states (flat) | mark bitmap | expanded bitmap | 2 frontiers of N ints
    return align_up(sizeof(CutHeader), 128) + align_up(cut_states(N) * 8, 128) + 2 * align_up(words * 4, 128) +
           2 * align_up((size_t)N * 4, 128) + 128;
}

namespace {
struct CutWs {
    CutHeader* hdr;
    unsigned long long* states;
    unsigned* mark;
    unsigned* expanded;
    int* frontier[2];
    size_t head_bytes;  // header + states + bitmaps (what has to be zeroed per call)
};
CutWs carve(void* ws, int N) {
    char* p = (char*)ws;
    CutWs c;
    size_t words = ((size_t)N + 31) / 32;
    c.hdr = (CutHeader*)p; p += align_up(sizeof(CutHeader), 128);
    c.states = (unsigned long long*)p; p += align_up(cut_states(N) * 8, 128);
    c.mark = (unsigned*)p; p += align_up(words * 4, 128);
    c.expanded = (unsigned*)p; p += align_up(words * 4, 128);
    c.head_bytes = (size_t)(p - (char*)ws);
    c.frontier[0] = (int*)p; p += align_up((size_t)N * 4, 128);
    c.frontier[1] = (int*)p;
    return c;
}
}  // namespace

extern "C" int lodgs_cut_prepare(const int32_t* nodes, const float* boxes, int N, void* ws, size_t ws_bytes,
                                 int32_t* monotone_host, void* stream) {
    if (!nodes || !boxes || !ws || !monotone_host || N < 0) { set_error("lodgs_cut_prepare: bad argument"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_cut_workspace_bytes(N)) { set_error("lodgs_cut_prepare: workspace too small"); return LODGS_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    CutWs c = carve(ws, N);
    int one = 1;
    LODGS_CUDA(cudaMemcpyAsync(&c.hdr->monotone, &one, sizeof(int), cudaMemcpyHostToDevice, s));
    if (N > 0) {
        cut_check_monotone_kernel<<<(N + 255) / 256, 256, 0, s>>>(nodes, boxes, N, c.hdr);
        LODGS_LAUNCHED("cut_check_monotone_kernel");
    }
    LODGS_CUDA(cudaMemcpyAsync(monotone_host, &c.hdr->monotone, sizeof(int), cudaMemcpyDeviceToHost, s));
    LODGS_CUDA(cudaStreamSynchronize(s));
    return LODGS_OK;
}

extern "C" int lodgs_expand_to_size(const int32_t* nodes, const float* boxes, int N, float target_size,
                                    const float* viewpoint_dev, int32_t* render_indices, int32_t* parent_indices,
                                    int32_t* nodes_for_render_indices, int capacity, void* ws, size_t ws_bytes, int mode,
                                    int32_t* count_host, void* stream) {
    if (!nodes || !boxes || !viewpoint_dev || !render_indices || !ws || N < 0 || capacity < 0) {
        set_error("lodgs_expand_to_size: bad argument");
        return LODGS_E_ARG;
    }
    if (ws_bytes < lodgs_cut_workspace_bytes(N)) { set_error("lodgs_expand_to_size: workspace too small (%zu < %zu)", ws_bytes, lodgs_cut_workspace_bytes(N)); return LODGS_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    CutWs c = carve(ws, N);
    if (N == 0) {
        LODGS_CUDA(cudaMemsetAsync(c.hdr, 0, sizeof(CutHeader), s));
        if (count_host) { *count_host = 0; LODGS_CUDA(cudaStreamSynchronize(s)); }
        return LODGS_OK;
    }
    if (mode == 0) {
        LODGS_CUDA(cudaMemsetAsync(ws, 0, align_up(sizeof(CutHeader), 128) + cut_states(N) * 8, s));
        int tiles = (N + FLAT_TILE - 1) / FLAT_TILE;
        cut_flat_kernel<<<tiles, FLAT_THREADS, 0, s>>>(nodes, boxes, N, target_size, viewpoint_dev, render_indices, parent_indices,
                                                   nodes_for_render_indices, capacity, c.hdr, c.states);
        LODGS_LAUNCHED("cut_flat_kernel");
    } else {
        LODGS_CUDA(cudaMemsetAsync(ws, 0, c.head_bytes, s));
        // One cooperative launch sweeps all levels (grid-wide sync between levels), sized to be co-resident.
        static std::atomic<int> grid_of[MAX_DEVICES];   // co-resident CTAs of the cooperative launch, per device
        const int dslot = device_slot();
        int coop_grid = grid_of[dslot].load(std::memory_order_acquire);
        if (!coop_grid) {
            int dev = 0, num_sms = 0, blocks_per_sm = 0;
            LODGS_CUDA(cudaGetDevice(&dev));
            LODGS_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
            LODGS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, cut_traverse_kernel, 256, 0));
            if (blocks_per_sm > 4) blocks_per_sm = 4;
            coop_grid = num_sms * blocks_per_sm;
            grid_of[dslot].store(coop_grid, std::memory_order_release);
        }
        const int32_t* a_nodes = nodes; const float* a_boxes = boxes; float a_tau = target_size; const float* a_vp = viewpoint_dev;
        int* fa = c.frontier[0]; int* fb = c.frontier[1]; CutHeader* hdr = c.hdr; unsigned* mk = c.mark; unsigned* ex = c.expanded;
        void* args[] = {&a_nodes, &a_boxes, &a_tau, &a_vp, &fa, &fb, &hdr, &mk, &ex};
        LODGS_CUDA(cudaLaunchCooperativeKernel((void*)cut_traverse_kernel, dim3(coop_grid), dim3(256), args, 0, s));
        LODGS_LAUNCHED("cut_traverse_kernel");
        int nwords = (N + 31) / 32;
        int tiles = (nwords + CUT_BLOCK - 1) / CUT_BLOCK;
        cut_compact_kernel<<<tiles, CUT_BLOCK, 0, s>>>(nodes, N, c.mark, c.expanded, render_indices, parent_indices,
                                                      nodes_for_render_indices, capacity, c.hdr, c.states);
        LODGS_LAUNCHED("cut_compact_kernel");
    }
    if (count_host) {
        const int prc = publish_words(reinterpret_cast<const int32_t*>(c.hdr), 3, s);  // mapped pinned words, not a D2H copy
        if (prc) return prc;
        const int32_t* h = wait_words();
        if (!h) { set_error("lodgs_expand_to_size: waiting for the count failed"); return LODGS_E_CUDA; }
        *count_host = h[0];
        if (h[2]) { set_error("lodgs_expand_to_size: cut has %d entries, output capacity is %d", h[0], capacity); return LODGS_E_WORKSPACE; }
    }
    return LODGS_OK;
}

extern "C" int lodgs_get_interpolation_weights(const int32_t* node_indices, int K, float target_size, const int32_t* nodes,
                                               const float* boxes, float vx, float vy, float vz, float* ts,
                                               int32_t* num_siblings, void* stream) {
    if (K < 0 || (K > 0 && (!node_indices || !nodes || !boxes || !ts || !num_siblings))) {
        set_error("lodgs_get_interpolation_weights: bad argument");
        return LODGS_E_ARG;
    }
    if (K == 0) return LODGS_OK;
    prof_stage_begin((cudaStream_t)stream);
    cut_weights_kernel<<<(K + 255) / 256, 256, 0, (cudaStream_t)stream>>>(nodes, boxes, K, node_indices, vx, vy, vz, target_size,
                                                                         ts, num_siblings);
    LODGS_LAUNCHED("cut_weights_kernel");
    return LODGS_OK;
}
