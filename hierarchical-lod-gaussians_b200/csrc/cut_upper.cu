// cut_upper.cu — the cut of the "upper tree" that precedes the SPT cut, as ONE level-synchronous kernel.
//
// Replaces, per frame / training iteration (train_post.py:323-342,370-394; hierarchy_viewer.py:262-345; scene/gaussian_model.py:105-156):
//   GaussianModel.cut_hierarchy_on_condition   scene/gaussian_model.py:354-394   level-synchronous BFS written with ~8 torch ops per level
//                                                                               (index, mask, cat), each with its own launches and the
//                                                                               `while len(stack)` host synchronisation per level
//   GaussianModel.frustum_cull_spheres         scene/gaussian_model.py:80-103    per-plane signed distance of bounding spheres (the cull)
//   the LOD condition                          train_post.py:338                 min_distance_squared[i] > |campos - xyz[i]|^2 * multiplier
//   the split of the cut that feeds get_spt_cut_cuda   train_post.py:349-356 / gaussian_model.py:139-150
//
// One cooperative launch walks all levels (grid-wide sync between them). The reference's OUTPUT ORDER is reproduced, because the render
// lists built from it are order-dependent: per level the stack is filtered by the cull, its leaves are appended to the cut (stack
// order), then the nodes failing the condition (stack order); the next stack is [first children of the remaining nodes, in order] followed
// by [their next siblings, in order]. Ordered compaction is a two-pass scan per level: every CTA classifies a contiguous chunk of the
// stack and ranks its elements (ballots), the per-CTA totals are combined after a grid sync, then every element is written to its slot.
// The derived lists (SPT slots with their distances, directly rendered leaves) are ranked in the same passes.
#include "common.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace lodgs {

constexpr int UC_THREADS = 256;
constexpr int UC_NCAT = 5;   // leaf-cut, condition-cut, expand, SPT slot, direct leaf

struct UcHeader {
    int n_cut, n_spt, n_direct, overflow;
};

struct UcParams {
    const int32_t* nodes; int N; int root;
    const float* xyz; const float* bounds; const float* planes; int n_planes; const float* min_dist_sq; const float* campos; float mult;
    int32_t* cut; int capacity; int32_t* spt_indices; int32_t* spt_nodes; float* spt_dist; int32_t* direct;
    int* frontier[2]; int* rank_a; int* rank_s; int* rank_d; int* cta_counts; UcHeader* hdr;
};

// torch.sum over a 3-element row ON CUDA — where the reference evaluates these expressions — adds (a0 + a2) + a1 (two strided partial
// accumulators; measured against torch 2.11 on a B200, profiles/r2_torch_sum3_probe.txt); torch's CPU reduction adds (a0 + a1) + a2.
__device__ __forceinline__ float sum3(float a0, float a1, float a2) { return __fadd_rn(__fadd_rn(a0, a2), a1); }

// class of a stack element: 0 culled, 1 leaf (to the cut), 2 condition false (to the cut), 3 expand; flags: bit 4 SPT slot, bit 5 direct
__device__ __forceinline__ int uc_classify(const UcParams& p, int i, float cx, float cy, float cz) {
    const float x = __ldg(p.xyz + 3 * (size_t)i), y = __ldg(p.xyz + 3 * (size_t)i + 1), z = __ldg(p.xyz + 3 * (size_t)i + 2);
    if (p.bounds) {   // frustum_cull_spheres: culled when the sphere lies entirely behind one plane
        const float r = __ldg(p.bounds + i);
        for (int k = 0; k < p.n_planes; k++) {
            const float4 pl = __ldg(reinterpret_cast<const float4*>(p.planes) + k);
            const float sd = __fadd_rn(sum3(__fmul_rn(x, pl.x), __fmul_rn(y, pl.y), __fmul_rn(z, pl.z)), pl.w);
            if (__fadd_rn(sd, r) < 0.f) return 0;
        }
    }
    const int child_count = __ldg(p.nodes + (size_t)i * 6 + 2);
    if (child_count == 0) {
        const int fc = __ldg(p.nodes + (size_t)i * 6 + 3);
        return 1 | (fc >= 0 ? 16 : 0) | (fc <= 0 ? 32 : 0);   // the reference tests `>= 0` and `<= 0`: a 0 lands in both lists
    }
    const float dx = __fsub_rn(cx, x), dy = __fsub_rn(cy, y), dz = __fsub_rn(cz, z);
    const float dist2 = sum3(__fmul_rn(dx, dx), __fmul_rn(dy, dy), __fmul_rn(dz, dz));
    return (__ldg(p.min_dist_sq + i) > __fmul_rn(dist2, p.mult)) ? 3 : 2;
}

__global__ void __launch_bounds__(UC_THREADS) upper_cut_kernel(const UcParams p) {
    cg::grid_group grid = cg::this_grid();
    __shared__ int s_warp[UC_THREADS / 32][UC_NCAT];
    __shared__ int s_run[UC_NCAT];
    __shared__ int s_base[UC_NCAT];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float cx = __ldg(p.campos), cy = __ldg(p.campos + 1), cz = __ldg(p.campos + 2);
    if (blockIdx.x == 0 && threadIdx.x == 0) p.frontier[0][0] = p.root;
    grid.sync();
    // every CTA keeps the same running totals: they follow from the per-CTA counts all CTAs read
    int n_in = (p.N > 0 && p.root >= 0 && p.root < p.N) ? 1 : 0, n_cut = 0, n_spt = 0, n_dir = 0;
    bool overflow = false;
    for (int level = 0; n_in > 0; level++) {
        const int* fin = p.frontier[level & 1];
        int* fout = p.frontier[(level + 1) & 1];
        const int chunk = (n_in + gridDim.x - 1) / gridDim.x;
        const int c0 = min(n_in, (int)blockIdx.x * chunk), c1 = min(n_in, c0 + chunk);
        // ---- pass 1: classify this CTA's chunk in stack order, rank every element inside the CTA
        if (threadIdx.x < UC_NCAT) s_run[threadIdx.x] = 0;
        __syncthreads();
        for (int t0 = c0; t0 < c1; t0 += UC_THREADS) {
            const int f = t0 + threadIdx.x;
            const int cls = f < c1 ? uc_classify(p, fin[f], cx, cy, cz) : 0;
            const int kind = cls & 15;
            const bool flag[UC_NCAT] = {kind == 1, kind == 2, kind == 3, (cls & 16) != 0, (cls & 32) != 0};
            int pre[UC_NCAT];
#pragma unroll
            for (int c = 0; c < UC_NCAT; c++) {
                const unsigned b = __ballot_sync(0xffffffffu, flag[c]);
                pre[c] = __popc(b & ((1u << lane) - 1u));
                if (lane == 0) s_warp[warp][c] = __popc(b);
            }
            __syncthreads();
            int tile_tot[UC_NCAT];
#pragma unroll
            for (int c = 0; c < UC_NCAT; c++) {
                int before = 0, tot = 0;
#pragma unroll
                for (int w = 0; w < UC_THREADS / 32; w++) {
                    const int v = s_warp[w][c];
                    before += w < (int)warp ? v : 0;
                    tot += v;
                }
                pre[c] += before + s_run[c];
                tile_tot[c] = tot;
            }
            if (f < c1) {
                p.rank_a[f] = (kind << 28) | (kind == 1 ? pre[0] : kind == 2 ? pre[1] : kind == 3 ? pre[2] : 0);
                p.rank_s[f] = (cls & 16) ? pre[3] : -1;
                p.rank_d[f] = (cls & 32) ? pre[4] : -1;
            }
            __syncthreads();
            if (threadIdx.x < UC_NCAT) s_run[threadIdx.x] += tile_tot[threadIdx.x];
            __syncthreads();
        }
        if (threadIdx.x < UC_NCAT) p.cta_counts[blockIdx.x * UC_NCAT + threadIdx.x] = s_run[threadIdx.x];
        grid.sync();
        // ---- combine: exclusive prefix of the per-CTA totals for this CTA, and the level totals
        if (threadIdx.x < UC_NCAT) {
            int before = 0, tot = 0;
            for (int b = 0; b < (int)gridDim.x; b++) {
                const int v = p.cta_counts[b * UC_NCAT + threadIdx.x];
                before += b < (int)blockIdx.x ? v : 0;
                tot += v;
            }
            s_base[threadIdx.x] = before;
            s_run[threadIdx.x] = tot;
        }
        __syncthreads();
        const int totL = s_run[0], totC = s_run[1], totE = s_run[2], totS = s_run[3], totD = s_run[4];
        // ---- pass 2: every element goes to its slot
        for (int f = c0 + threadIdx.x; f < c1; f += UC_THREADS) {
            const int ra = p.rank_a[f];
            const int kind = (ra >> 28) & 15, r = ra & 0x0fffffff;
            const int node = fin[f];
            if (kind == 1 || kind == 2) {
                const int slot = n_cut + (kind == 1 ? s_base[0] + r : totL + s_base[1] + r);
                if (slot < p.capacity) p.cut[slot] = node;
                if (kind == 1) {
                    const int rs = p.rank_s[f], rd = p.rank_d[f];
                    if (rs >= 0) {
                        const int s = n_spt + s_base[3] + rs;
                        if (s < p.capacity) {
                            if (p.spt_indices) p.spt_indices[s] = __ldg(p.nodes + (size_t)node * 6 + 3);
                            if (p.spt_nodes) p.spt_nodes[s] = node;
                            if (p.spt_dist) {   // (upper_tree_xyz[i] - camera_position).pow(2).sum(1).sqrt() * distance_multiplier
                                const float x = __ldg(p.xyz + 3 * (size_t)node), y = __ldg(p.xyz + 3 * (size_t)node + 1), z = __ldg(p.xyz + 3 * (size_t)node + 2);
                                const float dx = __fsub_rn(x, cx), dy = __fsub_rn(y, cy), dz = __fsub_rn(z, cz);
                                p.spt_dist[s] = __fmul_rn(__fsqrt_rn(sum3(__fmul_rn(dx, dx), __fmul_rn(dy, dy), __fmul_rn(dz, dz))), p.mult);
                            }
                        }
                    }
                    if (rd >= 0 && p.direct) {
                        const int s = n_dir + s_base[4] + rd;
                        if (s < p.capacity) p.direct[s] = __ldg(p.nodes + (size_t)node * 6 + 5);
                    }
                }
            } else if (kind == 3) {   // first_children = hierarchy[stack, 3]; second_children = hierarchy[first_children, 4]
                const int first = __ldg(p.nodes + (size_t)node * 6 + 3);
                const int second = (first >= 0 && first < p.N) ? __ldg(p.nodes + (size_t)first * 6 + 4) : -1;
                fout[s_base[2] + r] = first;
                fout[totE + s_base[2] + r] = second;
            }
        }
        n_cut += totL + totC;
        n_spt += totS;
        n_dir += totD;
        overflow = overflow || n_cut > p.capacity;
        n_in = 2 * totE;
        if (n_in > p.N) { overflow = true; n_in = 0; }   // a malformed table (cycle / shared children) cannot outgrow the frontier buffers
        grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        p.hdr->n_cut = n_cut; p.hdr->n_spt = n_spt; p.hdr->n_direct = n_dir; p.hdr->overflow = overflow ? 1 : 0;
    }
}

struct UcWs { UcHeader* hdr; int* frontier[2]; int* rank_a; int* rank_s; int* rank_d; int* cta_counts; size_t bytes; };
constexpr int UC_MAX_GRID = 1024;
static UcWs uc_carve(void* ws, int N) {
    char* q = (char*)ws;
    auto take = [&](size_t bytes) { char* r = q; q += align_up(bytes, 128); return r; };
    UcWs w;
    w.hdr = (UcHeader*)take(sizeof(UcHeader));
    const size_t n = (size_t)(N > 0 ? N : 1) + 2;
    w.frontier[0] = (int*)take(n * 4); w.frontier[1] = (int*)take(n * 4);
    w.rank_a = (int*)take(n * 4); w.rank_s = (int*)take(n * 4); w.rank_d = (int*)take(n * 4);
    w.cta_counts = (int*)take((size_t)UC_MAX_GRID * UC_NCAT * 4);
    w.bytes = (size_t)(q - (char*)ws);
    return w;
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_upper_cut_workspace_bytes(int N) { return uc_carve(nullptr, N < 0 ? 0 : N).bytes + 128; }

extern "C" int lodgs_upper_tree_cut(const int32_t* nodes, int N, int root, const float* xyz, const float* bounds, const float* planes, int n_planes,
                                    const float* min_dist_sq, const float* campos, float distance_multiplier, int32_t* cut, int capacity,
                                    int32_t* spt_indices, int32_t* spt_nodes, float* spt_distances, int32_t* direct, void* ws, size_t ws_bytes,
                                    int32_t* counts_host, void* stream) {
    if (N < 0 || !nodes || !xyz || !min_dist_sq || !campos || !cut || capacity < 0 || !ws || n_planes < 0 || n_planes > 6 || (bounds && n_planes && !planes)) {
        set_error("lodgs_upper_tree_cut: bad argument");
        return LODGS_E_ARG;
    }
    if (((uintptr_t)ws & 127) != 0 || (planes && ((uintptr_t)planes & 15) != 0)) { set_error("lodgs_upper_tree_cut: workspace needs 128-byte, planes 16-byte alignment"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_upper_cut_workspace_bytes(N) - 128) { set_error("lodgs_upper_tree_cut: workspace too small"); return LODGS_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    UcWs w = uc_carve(ws, N);
    UcParams p;
    p.nodes = nodes; p.N = N; p.root = root; p.xyz = xyz; p.bounds = bounds; p.planes = planes; p.n_planes = bounds ? n_planes : 0;
    p.min_dist_sq = min_dist_sq; p.campos = campos; p.mult = distance_multiplier; p.cut = cut; p.capacity = capacity; p.spt_indices = spt_indices;
    p.spt_nodes = spt_nodes; p.spt_dist = spt_distances; p.direct = direct; p.frontier[0] = w.frontier[0]; p.frontier[1] = w.frontier[1];
    p.rank_a = w.rank_a; p.rank_s = w.rank_s; p.rank_d = w.rank_d; p.cta_counts = w.cta_counts; p.hdr = w.hdr;
    static std::atomic<int> grid_of[MAX_DEVICES];
    const int dslot = device_slot();
    int coop = grid_of[dslot].load(std::memory_order_acquire);
    if (!coop) {
        int dev = 0, sms = 0, per_sm = 0;
        LODGS_CUDA(cudaGetDevice(&dev));
        LODGS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        LODGS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, upper_cut_kernel, UC_THREADS, 0));
        if (per_sm > 2) per_sm = 2;
        coop = sms * per_sm;
        if (coop > UC_MAX_GRID) coop = UC_MAX_GRID;
        grid_of[dslot].store(coop, std::memory_order_release);
    }
    // small trees: fewer CTAs make the grid-wide syncs cheaper (a level of a few hundred nodes is one CTA's work)
    int grid = coop;
    const int want = (N + 4 * UC_THREADS - 1) / (4 * UC_THREADS);
    if (want < grid) grid = want < 1 ? 1 : want;
    void* args[] = {(void*)&p};
    LODGS_CUDA(cudaLaunchCooperativeKernel((void*)upper_cut_kernel, dim3(grid), dim3(UC_THREADS), args, 0, s));
    LODGS_LAUNCHED("upper_cut_kernel");
    if (counts_host) {
        int rc = publish_words(&w.hdr->n_cut, 4, s);
        if (rc) return rc;
        const int32_t* h = wait_words();
        if (!h) { set_error("lodgs_upper_tree_cut: reading the counts back failed"); return LODGS_E_CUDA; }
        counts_host[0] = h[0]; counts_host[1] = h[1]; counts_host[2] = h[2];
        if (h[3]) { set_error("lodgs_upper_tree_cut: capacity %d too small for a cut of %d nodes (or malformed node table)", capacity, h[0]); return LODGS_E_WORKSPACE; }
    }
    return LODGS_OK;
}
