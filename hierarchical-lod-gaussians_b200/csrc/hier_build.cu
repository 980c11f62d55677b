// hier_build.cu — GPU construction of the Gaussian hierarchy.
//
// Replaces the reference's single-threaded, recursive CPU creator (GH/mainHierarchyCreator.cpp:149-183):
//   PointbasedKdTreeGenerator::generate  GH/PointbasedKdTreeGenerator.cpp:16-75  (bounds = union of pos +- 3*max(scale) cubes, split the
//                                        longest axis at std::nth_element(pivot = num/2 - 1): left num/2 points, right the rest)
//   ClusterMerger::merge                 GH/ClusterMerger.cpp:23-173             (weights opacity * (s0 s1 + s0 s2 + s1 s2), moment-matched
//                                        mean / SH / covariance, symmetric 3x3 eigen-decomposition -> scale, rotation; w = longest box side)
//   RotationAligner::align               GH/rotation_aligner.cpp:24-114          (child axes permuted / flipped to best match the parent)
//   Writer::makeDynamicHierarchy         GH/writer.cpp:98-183                    (pre-order node ids, one Gaussian per node)
//   Writer::makeHierarchy                GH/writer.cpp:21-96 with the upstream child emission: a node's children take a
//                                        contiguous id block allocated when the node is visited in pre-order
//
// Design. The tree SHAPE depends only on the leaf count (sizes floor/ceil), so every node's range, pre-order rank, internal-node rank and
// both id schemes follow in closed form from its parent's; only the ORDER of the leaves depends on the data. The build is level-synchronous:
// per level one pass reduces the node boxes (warp-aggregated ordered-int atomics), one forms 64-bit keys {node start | orderable
// coordinate on the node's longest axis} and the onesweep radix sort (radix_sort.cu) orders every node's points at once — the left child
// is the first num/2 of them, which is the SET std::nth_element selects (ties between equal coordinates fall as the stable sort leaves
// them; nth_element leaves them unspecified). Merging runs bottom-up and alignment top-down, one thread per node and level, the 48-float
// SH rows by one thread per coefficient. Output Gaussians are in pre-order (gaussian i belongs to dynamic node i; static nodes point
// at them through `start`).
#include "common.cuh"
#include <cfloat>
#include <map>
#include <vector>

namespace lodgs {

int sort_pairs_device(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, uint64_t* keys_tmp,
                      uint32_t* vals_tmp, int n, int end_bit, void* ws, cudaStream_t s);

struct HbNode {
    int start, num;      // leaf positions [start, start + num) in the final leaf order
    int prerank;         // pre-order rank = dynamic node id = index of the node's Gaussian
    int irank;           // rank among INTERNAL nodes in pre-order (-1 for a leaf)
    int sid;             // static-format node id (upstream populateRec)
    int parent;          // slot of the parent (-1 for the root)
    int level;           // distance from the root
    int pad;
};

__device__ __forceinline__ int f2ord(float f) {
    int o = __float_as_int(f);
    return o >= 0 ? o : o ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7fffffff); }

__global__ void hb_init_kernel(int L, uint32_t* perm, int* node_of, HbNode* nodes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    perm[i] = (uint32_t)i;
    node_of[i] = 0;
    if (i == 0) nodes[0] = HbNode{0, L, 0, L > 1 ? 0 : -1, 0, -1, 0, 0};
}

__global__ void hb_box_reset_kernel(int* boxi, int slot0, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int* b = boxi + (size_t)(slot0 + i) * 6;
    const int lo = 0x7f7fffff, hi = (int)0x80800000;   // ordered images of +FLT_MAX and -FLT_MAX (PointbasedKdTreeGenerator.cpp:19-20)
    b[0] = b[1] = b[2] = lo;
    b[3] = b[4] = b[5] = hi;
}

// bounds of every node of the current level: union over its points of [pos - r, pos + r], r = 3 * max(scale)
__global__ void hb_bbox_kernel(int L, int slot0, int slot1, const uint32_t* __restrict__ perm, const int* __restrict__ node_of,
                               const float* __restrict__ pos, const float* __restrict__ scale, int* __restrict__ boxi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int node = i < L ? node_of[i] : -1;
    const bool act = node >= slot0 && node < slot1;
    const unsigned m = __match_any_sync(0xffffffffu, act ? node : -1);
    if (!act) return;
    const uint32_t g = perm[i];
    const float r = 3.0f * fmaxf(fmaxf(scale[3 * (size_t)g], scale[3 * (size_t)g + 1]), scale[3 * (size_t)g + 2]);
    int v[6];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const float p = pos[3 * (size_t)g + c];
        v[c] = f2ord(p - r);
        v[3 + c] = f2ord(p + r);
    }
    const bool leader = (threadIdx.x & 31) == (unsigned)(__ffs(m) - 1);
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const int lo = __reduce_min_sync(m, v[c]), hi = __reduce_max_sync(m, v[3 + c]);
        if (leader) {
            atomicMin(boxi + (size_t)node * 6 + c, lo);
            atomicMax(boxi + (size_t)node * 6 + 3 + c, hi);
        }
    }
}

__global__ void hb_keys_kernel(int L, int slot0, int slot1, const uint32_t* __restrict__ perm, const int* __restrict__ node_of,
                               const HbNode* __restrict__ nodes, const int* __restrict__ boxi, const float* __restrict__ pos,
                               uint64_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const int node = node_of[i];
    uint64_t key = (uint64_t)(uint32_t)i << 32;   // a finished leaf stays where it is
    if (node >= slot0 && node < slot1 && nodes[node].num > 1) {
        const int* b = boxi + (size_t)node * 6;
        int axis = 0;
        float greatest = 0.f;
#pragma unroll
        for (int c = 0; c < 3; c++) {   // PointbasedKdTreeGenerator.cpp:44-53: first axis with the strictly greatest extent
            const float d = ord2f(b[3 + c]) - ord2f(b[c]);
            if (d > greatest) { greatest = d; axis = c; }
        }
        uint32_t bits = __float_as_uint(pos[3 * (size_t)perm[i] + axis]);
        bits = (bits & 0x80000000u) ? ~bits : (bits ^ 0x80000000u);
        key = ((uint64_t)(uint32_t)nodes[node].start << 32) | bits;
    }
    keys[i] = key;
}

// children of the level's internal nodes; child_rank = exclusive count of internal nodes before slot s in this level (null: every node
// of the level is internal, the rank is s itself)
__global__ void hb_children_kernel(int slot0, int n_level, int next_slot0, const int* __restrict__ child_rank, HbNode* __restrict__ nodes,
                                   int* __restrict__ child_base) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_level) return;
    const int node = slot0 + s;
    const HbNode nd = nodes[node];
    if (nd.num <= 1) { child_base[node] = -1; return; }
    const int base = next_slot0 + 2 * (child_rank ? child_rank[s] : s);
    child_base[node] = base;
    const int nl = nd.num / 2, nr = nd.num - nl;   // pivot = num/2 - 1: the left child takes pivot + 1 points
    nodes[base] = HbNode{nd.start, nl, nd.prerank + 1, nl > 1 ? nd.irank + 1 : -1, 1 + 2 * nd.irank, node, nd.level + 1, 0};
    nodes[base + 1] = HbNode{nd.start + nl, nr, nd.prerank + 2 * nl, nr > 1 ? nd.irank + nl : -1, 2 + 2 * nd.irank, node, nd.level + 1, 0};
}

__global__ void hb_assign_kernel(int L, int slot0, int slot1, const HbNode* __restrict__ nodes, const int* __restrict__ child_base,
                                 int* __restrict__ node_of) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const int node = node_of[i];
    if (node < slot0 || node >= slot1) return;
    const int base = child_base[node];
    if (base >= 0) node_of[i] = base + ((i - nodes[node].start) >= nodes[node].num / 2 ? 1 : 0);
}

// ---- exclusive scan of the "is internal" flags of one level (needed only for the one level that mixes sizes 1 and 2)
constexpr int HB_SCAN_BLOCK = 1024;
__global__ void hb_scan_block_kernel(int slot0, int n, const HbNode* __restrict__ nodes, int* __restrict__ out, int* __restrict__ block_sums) {
    __shared__ int sh[HB_SCAN_BLOCK];
    const int i = blockIdx.x * HB_SCAN_BLOCK + threadIdx.x;
    const int f = (i < n && nodes[slot0 + i].num > 1) ? 1 : 0;
    sh[threadIdx.x] = f;
    __syncthreads();
    for (int off = 1; off < HB_SCAN_BLOCK; off <<= 1) {
        const int v = threadIdx.x >= off ? sh[threadIdx.x - off] : 0;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    if (i < n) out[i] = sh[threadIdx.x] - f;
    if (threadIdx.x == HB_SCAN_BLOCK - 1) block_sums[blockIdx.x] = sh[threadIdx.x];
}
__global__ void hb_scan_sums_kernel(int nb, int* block_sums) {   // one thread: nb <= 32768
    int run = 0;
    for (int b = 0; b < nb; b++) { const int v = block_sums[b]; block_sums[b] = run; run += v; }
}
__global__ void hb_scan_add_kernel(int n, int* out, const int* __restrict__ block_sums) {
    const int i = blockIdx.x * HB_SCAN_BLOCK + threadIdx.x;
    if (i < n) out[i] += block_sums[blockIdx.x];
}

// ---- per-Gaussian math of the creator
__device__ __forceinline__ void mat_from_quat(const float* q, float (&R)[3][3]) {   // GH/common.h:102-117, q = (w, x, y, z)
    const float s = q[0], x = q[1], y = q[2], z = q[3];
    R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - s * z); R[0][2] = 2.f * (x * z + s * y);
    R[1][0] = 2.f * (x * y + s * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - s * x);
    R[2][0] = 2.f * (x * z - s * y); R[2][1] = 2.f * (y * z + s * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
}

__device__ __forceinline__ void covariance_of(const float* scale, const float* rot, float* cov) {   // GH/common.h:119-147
    float R[3][3], T[3][3];
    mat_from_quat(rot, R);
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) T[i][j] = R[i][j] * scale[j];
    auto dot = [&](int i, int k) { return T[i][0] * T[k][0] + T[i][1] * T[k][1] + T[i][2] * T[k][2]; };
    cov[0] = dot(0, 0); cov[1] = dot(0, 1); cov[2] = dot(0, 2); cov[3] = dot(1, 1); cov[4] = dot(1, 2); cov[5] = dot(2, 2);
}

// rotation matrix -> quaternion (w, x, y, z): the branch structure Eigen's Quaternion(Matrix3) uses (trace, else largest diagonal element)
__device__ __forceinline__ void quat_from_mat(const float (&m)[3][3], float* q) {
    float c[4];   // x, y, z, w
    float t = m[0][0] + m[1][1] + m[2][2];
    if (t > 0.f) {
        t = sqrtf(t + 1.0f);
        c[3] = 0.5f * t;
        t = 0.5f / t;
        c[0] = (m[2][1] - m[1][2]) * t;
        c[1] = (m[0][2] - m[2][0]) * t;
        c[2] = (m[1][0] - m[0][1]) * t;
    } else {
        int i = 0;
        if (m[1][1] > m[0][0]) i = 1;
        if (m[2][2] > m[i][i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = sqrtf(m[i][i] - m[j][j] - m[k][k] + 1.0f);
        c[i] = 0.5f * t;
        t = 0.5f / t;
        c[3] = (m[k][j] - m[j][k]) * t;
        c[j] = (m[j][i] + m[i][j]) * t;
        c[k] = (m[k][i] + m[i][k]) * t;
    }
    q[0] = c[3]; q[1] = c[0]; q[2] = c[1]; q[3] = c[2];
}

// symmetric 3x3 eigen-decomposition, eigenvalues ascending with matching eigenvector COLUMNS (cyclic Jacobi, float; the same routine as
// oracle/eigen_stub/Eigen/Eigenvalues, which stands in for Eigen's SelfAdjointEigenSolver when the reference creator is compiled)
__device__ void eig_sym3(const float (&min)[3][3], float* vals, float (&vecs)[3][3]) {
    float a[3][3], v[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) { a[i][j] = i <= j ? min[i][j] : min[j][i]; v[i][j] = i == j ? 1.f : 0.f; }
    for (int sweep = 0; sweep < 32; sweep++) {
        const float off = fabsf(a[0][1]) + fabsf(a[0][2]) + fabsf(a[1][2]);
        const float diag = fabsf(a[0][0]) + fabsf(a[1][1]) + fabsf(a[2][2]);
        if (!(off > 1e-9f * diag)) break;
#pragma unroll
        for (int p = 0; p < 2; p++)
#pragma unroll
            for (int q = p + 1; q < 3; q++) {
                if (a[p][q] == 0.f) continue;
                const float theta = (a[q][q] - a[p][p]) / (2.f * a[p][q]);
                const float t = (theta >= 0.f ? 1.f : -1.f) / (fabsf(theta) + sqrtf(theta * theta + 1.f));
                const float c = 1.f / sqrtf(t * t + 1.f), s = t * c;
#pragma unroll
                for (int k = 0; k < 3; k++) { const float akp = a[k][p], akq = a[k][q]; a[k][p] = c * akp - s * akq; a[k][q] = s * akp + c * akq; }
#pragma unroll
                for (int k = 0; k < 3; k++) { const float apk = a[p][k], aqk = a[q][k]; a[p][k] = c * apk - s * aqk; a[q][k] = s * apk + c * aqk; }
#pragma unroll
                for (int k = 0; k < 3; k++) { const float vkp = v[k][p], vkq = v[k][q]; v[k][p] = c * vkp - s * vkq; v[k][q] = s * vkp + c * vkq; }
            }
    }
    int o[3] = {0, 1, 2};
    if (a[o[0]][o[0]] > a[o[1]][o[1]]) { const int t = o[0]; o[0] = o[1]; o[1] = t; }
    if (a[o[1]][o[1]] > a[o[2]][o[2]]) { const int t = o[1]; o[1] = o[2]; o[2] = t; }
    if (a[o[0]][o[0]] > a[o[1]][o[1]]) { const int t = o[0]; o[0] = o[1]; o[1] = t; }
#pragma unroll
    for (int j = 0; j < 3; j++) {
        vals[j] = a[o[j]][o[j]];
#pragma unroll
        for (int i = 0; i < 3; i++) vecs[i][j] = v[i][o[j]];
    }
}

struct HbOut {
    float *pos, *rot, *scale, *opacity, *shs;   // [N,3] [N,4] [N,3] [N] [N,sh]
};

__global__ void hb_leaf_kernel(int L, const uint32_t* __restrict__ perm, const int* __restrict__ node_of, const HbNode* __restrict__ nodes,
                               const float* __restrict__ pos, const float* __restrict__ scale, const float* __restrict__ rot,
                               const float* __restrict__ opacity, HbOut o, float* __restrict__ cov, int* __restrict__ leaf_src) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const size_t r = (size_t)nodes[node_of[i]].prerank, g = perm[i];
    float s[3], q[4];
#pragma unroll
    for (int c = 0; c < 3; c++) { o.pos[3 * r + c] = pos[3 * g + c]; s[c] = scale[3 * g + c]; o.scale[3 * r + c] = s[c]; }
#pragma unroll
    for (int c = 0; c < 4; c++) { q[c] = rot[4 * g + c]; o.rot[4 * r + c] = q[c]; }
    o.opacity[r] = opacity[g];
    covariance_of(s, q, cov + 6 * r);
    leaf_src[r] = (int)g;
}

__global__ void hb_leaf_shs_kernel(int L, int sh, const uint32_t* __restrict__ perm, const int* __restrict__ node_of,
                                   const HbNode* __restrict__ nodes, const float* __restrict__ shs, float* __restrict__ out_shs) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (size_t)L * sh) return;
    const int i = (int)(t / sh), c = (int)(t % sh);
    out_shs[(size_t)nodes[node_of[i]].prerank * sh + c] = shs[(size_t)perm[i] * sh + c];
}

// ClusterMerger::mergeRec for the internal nodes of one level (their children, one level down, are final)
__global__ void hb_merge_kernel(int slot0, int n_level, const HbNode* __restrict__ nodes, const int* __restrict__ child_base,
                                const int* __restrict__ boxi, HbOut o, float* __restrict__ cov, float* __restrict__ wts,
                                float* __restrict__ boxw, int* __restrict__ leaf_src) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_level) return;
    const int node = slot0 + s;
    const HbNode nd = nodes[node];
    if (nd.num <= 1) { boxw[node] = 0.f; return; }
    const size_t r = (size_t)nd.prerank;
    const int cb = child_base[node];
    const size_t rc[2] = {(size_t)nodes[cb].prerank, (size_t)nodes[cb + 1].prerank};
    float w[2], p[2][3];
    float wsum = 0.f;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float* sc = o.scale + 3 * rc[i];
        w[i] = o.opacity[rc[i]] * (sc[0] * sc[1] + sc[0] * sc[2] + sc[1] * sc[2]);   // ClusterMerger.cpp:16-21,52-58
        wsum += w[i];
#pragma unroll
        for (int c = 0; c < 3; c++) p[i][c] = o.pos[3 * rc[i] + c];
    }
    const float a[2] = {w[0] / wsum, w[1] / wsum};
    float mean[3];
#pragma unroll
    for (int c = 0; c < 3; c++) mean[c] = a[0] * p[0][c] + a[1] * p[1][c];
    float cv[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 2; i++) {   // ClusterMerger.cpp:70-83
        const float dx = p[i][0] - mean[0], dy = p[i][1] - mean[1], dz = p[i][2] - mean[2];
        const float* ci = cov + 6 * rc[i];
        cv[0] += a[i] * (ci[0] + dx * dx);
        cv[1] += a[i] * (ci[1] + dy * dx);
        cv[2] += a[i] * (ci[2] + dz * dx);
        cv[3] += a[i] * (ci[3] + dy * dy);
        cv[4] += a[i] * (ci[4] + dz * dy);
        cv[5] += a[i] * (ci[5] + dz * dz);
    }
    float M[3][3] = {{cv[0], cv[1], cv[2]}, {cv[1], cv[3], cv[4]}, {cv[2], cv[4], cv[5]}};
    float ev[3], E[3][3];
    eig_sym3(M, ev, E);
    for (int loops = 0; (ev[0] == 0.f || ev[1] == 0.f || ev[2] == 0.f) && loops < 64; loops++) {   // ClusterMerger.cpp:104-117
#pragma unroll
        for (int d = 0; d < 3; d++) M[d][d] += fmaxf(M[d][d] * 0.0001f, FLT_EPSILON);
        eig_sym3(M, ev, E);
    }
    {   // right-handed frame (ClusterMerger.cpp:119-125)
        const float tx = E[1][0] * E[2][1] - E[2][0] * E[1][1], ty = E[2][0] * E[0][1] - E[0][0] * E[2][1], tz = E[0][0] * E[1][1] - E[1][0] * E[0][1];
        if (tx * E[0][2] + ty * E[1][2] + tz * E[2][2] < 0.f) { E[0][2] = -E[0][2]; E[1][2] = -E[1][2]; E[2][2] = -E[2][2]; }
    }
    const float sc[3] = {sqrtf(fabsf(ev[0])), sqrtf(fabsf(ev[1])), sqrtf(fabsf(ev[2]))};
    float q[4];
    quat_from_mat(E, q);
#pragma unroll
    for (int c = 0; c < 3; c++) { o.pos[3 * r + c] = mean[c]; o.scale[3 * r + c] = sc[c]; }
#pragma unroll
    for (int c = 0; c < 4; c++) o.rot[4 * r + c] = q[c];
    o.opacity[r] = wsum / (sc[0] * sc[1] + sc[0] * sc[2] + sc[1] * sc[2]);
#pragma unroll
    for (int c = 0; c < 6; c++) cov[6 * r + c] = cv[c];
    wts[2 * r] = a[0];
    wts[2 * r + 1] = a[1];
    leaf_src[r] = -1;
    const int* b = boxi + (size_t)node * 6;   // ClusterMerger.cpp:163-168
    const float ex = ord2f(b[3]) - ord2f(b[0]), ey = ord2f(b[4]) - ord2f(b[1]), ez = ord2f(b[5]) - ord2f(b[2]);
    boxw[node] = fmaxf(fmaxf(ex, ey), ez);
}

__global__ void hb_merge_shs_kernel(int slot0, int n_level, int sh, const HbNode* __restrict__ nodes, const int* __restrict__ child_base,
                                    const float* __restrict__ wts, float* __restrict__ out_shs) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (size_t)n_level * sh) return;
    const int node = slot0 + (int)(t / sh), c = (int)(t % sh);
    if (nodes[node].num <= 1) return;
    const size_t r = (size_t)nodes[node].prerank;
    const int cb = child_base[node];
    out_shs[r * sh + c] = wts[2 * r] * out_shs[(size_t)nodes[cb].prerank * sh + c] + wts[2 * r + 1] * out_shs[(size_t)nodes[cb + 1].prerank * sh + c];
}

// RotationAligner::matchExhaustive (rotation_aligner.cpp:24-90) of every node of one level against its (already aligned) parent
__global__ void hb_align_kernel(int slot0, int n_level, const HbNode* __restrict__ nodes, HbOut o) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_level) return;
    const HbNode nd = nodes[slot0 + s];
    if (nd.parent < 0) return;
    const size_t r = (size_t)nd.prerank, rp = (size_t)nodes[nd.parent].prerank;
    float Rref[3][3], Rm[3][3];
    mat_from_quat(o.rot + 4 * rp, Rref);
    mat_from_quat(o.rot + 4 * r, Rm);
    float best_score = 0.f, best[3][3];
    int bi = 0, bj = 1, bk = 2;
    bool found = false;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            if (j == i) continue;
            const int k = 3 - i - j;
            const int col[3] = {i, j, k};
            for (int state = 0; state < 8; state++) {
                float T[3][3];
#pragma unroll
                for (int w = 0; w < 3; w++) {
                    const float sg = (state & (1 << w)) ? -1.f : 1.f;
#pragma unroll
                    for (int row = 0; row < 3; row++) T[row][w] = sg * Rm[row][col[w]];
                }
                const float cx = T[1][0] * T[2][1] - T[2][0] * T[1][1], cy = T[2][0] * T[0][1] - T[0][0] * T[2][1], cz = T[0][0] * T[1][1] - T[1][0] * T[0][1];
                if (cx * T[0][2] + cy * T[1][2] + cz * T[2][2] < 0.f) continue;
                float score = 0.f;
#pragma unroll
                for (int a = 0; a < 3; a++)
#pragma unroll
                    for (int b = 0; b < 3; b++) score += T[a][b] * Rref[a][b];
                if (score > best_score) {
                    best_score = score; bi = i; bj = j; bk = k; found = true;
#pragma unroll
                    for (int a = 0; a < 3; a++)
#pragma unroll
                        for (int b = 0; b < 3; b++) best[a][b] = T[a][b];
                }
            }
        }
    if (!found) return;   // the reference reads an uninitialised matrix here; leaving the Gaussian as it is keeps its covariance
    float q[4];
    quat_from_mat(best, q);
    const float s0 = o.scale[3 * r + bi], s1 = o.scale[3 * r + bj], s2 = o.scale[3 * r + bk];
    o.scale[3 * r] = s0; o.scale[3 * r + 1] = s1; o.scale[3 * r + 2] = s2;
#pragma unroll
    for (int c = 0; c < 4; c++) o.rot[4 * r + c] = q[c];
}

__global__ void hb_flatten_kernel(int N, const HbNode* __restrict__ nodes, const int* __restrict__ child_base, const int* __restrict__ boxi,
                                  const float* __restrict__ boxw, const int* __restrict__ leaf_src, int32_t* __restrict__ dyn,
                                  int32_t* __restrict__ stat, float* __restrict__ boxes) {
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= N) return;
    const HbNode nd = nodes[node];
    const bool inner = nd.num > 1;
    if (dyn) {   // GH/types.h:60-67 {depth, parent, child_count, first_child, next_sibling, max_side_length}; writer.cpp:98-165
        int next = 0;
        if (nd.parent >= 0) {
            const int cb = child_base[nd.parent];
            if (node == cb) next = nodes[cb + 1].prerank;
        }
        int32_t* d = dyn + (size_t)nd.prerank * 6;
        d[0] = nd.level;
        d[1] = nd.parent >= 0 ? nodes[nd.parent].prerank : -1;
        d[2] = inner ? 2 : 0;
        d[3] = nd.prerank + 1;
        d[4] = next;
        d[5] = inner ? -1 : leaf_src[nd.prerank];
    }
    if (stat) {   // GH/types.h:83-92 {depth, parent, start, count_leafs, count_merged, start_children, count_children}
        int32_t* t = stat + (size_t)nd.sid * 7;
        t[0] = nd.num > 1 ? 32 - __clz(nd.num - 1) : 0;   // height above the leaves: max(children) + 1 (PointbasedKdTreeGenerator.cpp:63)
        t[1] = nd.parent >= 0 ? nodes[nd.parent].sid : -1;
        t[2] = nd.prerank;
        t[3] = inner ? 0 : 1;
        t[4] = inner ? 1 : 0;
        t[5] = inner ? 1 + 2 * nd.irank : 1 + 2 * (nd.prerank - nd.start);   // number of node slots allocated when this node is visited
        t[6] = inner ? 2 : 0;
    }
    if (boxes) {
        const int* b = boxi + (size_t)node * 6;
        float* o = boxes + (size_t)nd.sid * 8;
        const float w = boxw[node];
        o[0] = ord2f(b[0]); o[1] = ord2f(b[1]); o[2] = ord2f(b[2]); o[3] = w;
        o[4] = ord2f(b[3]); o[5] = ord2f(b[4]); o[6] = ord2f(b[5]); o[7] = w;
    }
}

struct HbWs {
    HbNode* nodes; int* boxi; int* child_base; int* node_of; uint32_t* perm[3]; uint64_t* keys[3]; void* sort_ws; float* cov; float* wts;
    float* boxw; int* leaf_src; int* scan; int* scan_sums; size_t bytes;
};

static HbWs hb_carve(void* ws, int L) {
    const size_t N = L > 0 ? 2 * (size_t)L - 1 : 0;
    char* p = (char*)ws;
    auto take = [&](size_t bytes) { char* r = p; p += align_up(bytes, 128); return r; };
    HbWs w;
    w.nodes = (HbNode*)take(N * sizeof(HbNode));
    w.boxi = (int*)take(N * 6 * 4);
    w.child_base = (int*)take(N * 4);
    w.node_of = (int*)take((size_t)L * 4);
    for (int i = 0; i < 3; i++) w.perm[i] = (uint32_t*)take((size_t)L * 4);
    for (int i = 0; i < 3; i++) w.keys[i] = (uint64_t*)take((size_t)L * 8);
    w.sort_ws = take(lodgs_sort_workspace_bytes(L));
    w.cov = (float*)take(N * 6 * 4);
    w.wts = (float*)take(N * 2 * 4);
    w.boxw = (float*)take(N * 4);
    w.leaf_src = (int*)take(N * 4);
    w.scan = (int*)take((size_t)L * 4);
    w.scan_sums = (int*)take(((size_t)L / HB_SCAN_BLOCK + 2) * 4);
    w.bytes = (size_t)(p - (char*)ws);
    return w;
}

}  // namespace lodgs

using namespace lodgs;

extern "C" size_t lodgs_hier_build_workspace_bytes(int L) { return L <= 0 ? 128 : hb_carve(nullptr, L).bytes + 128; }

extern "C" int lodgs_hier_build(int L, int sh_floats, const float* pos, const float* scale, const float* rot, const float* opacity,
                                const float* shs, float* out_pos, float* out_rot, float* out_scale, float* out_opacity, float* out_shs,
                                int32_t* dyn_nodes, int32_t* nodes, float* boxes, void* ws, size_t ws_bytes, void* stream) {
    if (L < 1 || sh_floats < 0 || !pos || !scale || !rot || !opacity || (sh_floats && (!shs || !out_shs)) || !out_pos || !out_rot || !out_scale ||
        !out_opacity || !ws) {
        set_error("lodgs_hier_build: bad argument");
        return LODGS_E_ARG;
    }
    if (L >= (1 << 29)) { set_error("lodgs_hier_build: at most 2^29 - 1 leaves"); return LODGS_E_ARG; }
    if (((uintptr_t)ws & 127) != 0) { set_error("lodgs_hier_build: workspace must be 128-byte aligned"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_hier_build_workspace_bytes(L) - 128) { set_error("lodgs_hier_build: workspace too small"); return LODGS_E_WORKSPACE; }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const int N = 2 * L - 1;
    HbWs w = hb_carve(ws, L);
    HbOut o{out_pos, out_rot, out_scale, out_opacity, out_shs};
    const int T = 256;
    auto blocks = [&](size_t n) { return (unsigned)((n + T - 1) / T); };

    // level table from the leaf count alone: sizes at a level are s or s + 1
    struct Lev { int slot0, n; bool all_internal; };
    std::vector<Lev> levels;
    {
        std::map<int, long long> sizes;   // size -> number of nodes
        sizes[L] = 1;
        int slot0 = 0;
        while (!sizes.empty()) {
            long long n = 0, internal = 0;
            for (auto& kv : sizes) { n += kv.second; if (kv.first > 1) internal += kv.second; }
            levels.push_back(Lev{slot0, (int)n, internal == n});
            slot0 += (int)n;
            std::map<int, long long> next;
            for (auto& kv : sizes)
                if (kv.first > 1) { next[kv.first / 2] += kv.second; next[kv.first - kv.first / 2] += kv.second; }
            sizes.swap(next);
        }
        if (slot0 != N) { set_error("lodgs_hier_build: level table does not add up (%d != %d)", slot0, N); return LODGS_E_ARG; }
    }
    int bits_L = 1;
    while ((1ll << bits_L) < L) bits_L++;

    hb_init_kernel<<<blocks(L), T, 0, s>>>(L, w.perm[0], w.node_of, w.nodes);
    LODGS_LAUNCHED("hb_init_kernel");
    uint32_t* perm = w.perm[0];
    uint32_t* perm_alt = w.perm[1];
    for (size_t li = 0; li < levels.size(); li++) {
        const Lev lv = levels[li];
        const int slot1 = lv.slot0 + lv.n;
        hb_box_reset_kernel<<<blocks(lv.n), T, 0, s>>>(w.boxi, lv.slot0, lv.n);
        LODGS_LAUNCHED("hb_box_reset_kernel");
        hb_bbox_kernel<<<blocks(L), T, 0, s>>>(L, lv.slot0, slot1, perm, w.node_of, pos, scale, w.boxi);
        LODGS_LAUNCHED("hb_bbox_kernel");
        const bool last = li + 1 == levels.size();
        if (last) {   // only leaves at the deepest level
            hb_children_kernel<<<blocks(lv.n), T, 0, s>>>(lv.slot0, lv.n, slot1, nullptr, w.nodes, w.child_base);
            LODGS_LAUNCHED("hb_children_kernel");
            break;
        }
        hb_keys_kernel<<<blocks(L), T, 0, s>>>(L, lv.slot0, slot1, perm, w.node_of, w.nodes, w.boxi, pos, w.keys[0]);
        LODGS_LAUNCHED("hb_keys_kernel");
        int rc = sort_pairs_device(w.keys[0], perm, w.keys[1], perm_alt, w.keys[2], w.perm[2], L, 32 + bits_L, w.sort_ws, s);
        if (rc) return rc;
        { uint32_t* t = perm; perm = perm_alt; perm_alt = t; }
        const int* rank = nullptr;
        if (!lv.all_internal) {
            const int nb = (lv.n + HB_SCAN_BLOCK - 1) / HB_SCAN_BLOCK;
            hb_scan_block_kernel<<<nb, HB_SCAN_BLOCK, 0, s>>>(lv.slot0, lv.n, w.nodes, w.scan, w.scan_sums);
            LODGS_LAUNCHED("hb_scan_block_kernel");
            hb_scan_sums_kernel<<<1, 1, 0, s>>>(nb, w.scan_sums);
            LODGS_LAUNCHED("hb_scan_sums_kernel");
            hb_scan_add_kernel<<<nb, HB_SCAN_BLOCK, 0, s>>>(lv.n, w.scan, w.scan_sums);
            LODGS_LAUNCHED("hb_scan_add_kernel");
            rank = w.scan;
        }
        hb_children_kernel<<<blocks(lv.n), T, 0, s>>>(lv.slot0, lv.n, slot1, rank, w.nodes, w.child_base);
        LODGS_LAUNCHED("hb_children_kernel");
        hb_assign_kernel<<<blocks(L), T, 0, s>>>(L, lv.slot0, slot1, w.nodes, w.child_base, w.node_of);
        LODGS_LAUNCHED("hb_assign_kernel");
    }
    hb_leaf_kernel<<<blocks(L), T, 0, s>>>(L, perm, w.node_of, w.nodes, pos, scale, rot, opacity, o, w.cov, w.leaf_src);
    LODGS_LAUNCHED("hb_leaf_kernel");
    if (sh_floats) {
        hb_leaf_shs_kernel<<<blocks((size_t)L * sh_floats), T, 0, s>>>(L, sh_floats, perm, w.node_of, w.nodes, shs, out_shs);
        LODGS_LAUNCHED("hb_leaf_shs_kernel");
    }
    for (int li = (int)levels.size() - 1; li >= 0; li--) {   // bottom-up: a level's children are one level down
        const Lev lv = levels[li];
        hb_merge_kernel<<<blocks(lv.n), T, 0, s>>>(lv.slot0, lv.n, w.nodes, w.child_base, w.boxi, o, w.cov, w.wts, w.boxw, w.leaf_src);
        LODGS_LAUNCHED("hb_merge_kernel");
        if (sh_floats && li + 1 < (int)levels.size()) {
            hb_merge_shs_kernel<<<blocks((size_t)lv.n * sh_floats), T, 0, s>>>(lv.slot0, lv.n, sh_floats, w.nodes, w.child_base, w.wts, out_shs);
            LODGS_LAUNCHED("hb_merge_shs_kernel");
        }
    }
    for (size_t li = 1; li < levels.size(); li++) {          // top-down: the parent is aligned before its children
        hb_align_kernel<<<blocks(levels[li].n), T, 0, s>>>(levels[li].slot0, levels[li].n, w.nodes, o);
        LODGS_LAUNCHED("hb_align_kernel");
    }
    hb_flatten_kernel<<<blocks(N), T, 0, s>>>(N, w.nodes, w.child_base, w.boxi, w.boxw, w.leaf_src, dyn_nodes, nodes, boxes);
    LODGS_LAUNCHED("hb_flatten_kernel");
    return LODGS_OK;
}
