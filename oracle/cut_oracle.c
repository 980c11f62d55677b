/* cut_oracle.c — TEST INFRASTRUCTURE ONLY. Sequential CPU restatement of the reference's size-based LOD cut and
 * interpolation weights (the reference has them only as CUDA kernels):
 *   computeSizeGPU / inboxCUDA / pointboxdistCUDA   GH/runtime_switching.cu:136-144,189-219
 *   markNodesForSize + InclusiveSum + putRenderIndices  :495-529, :1021-1025, :68-93  (a running counter over ascending
 *   node ids reproduces scan + scatter order exactly)
 *   computeTsIndexed                                 :588-634
 * The squared distance uses the association the reference compiles to (t = dy*dy; fma(dx,dx,t); fma(dz,dz,t), SURVEY top)
 * so `size >= tau` decisions are bit-identical to the GPU. Compile with -ffp-contract=off. */
#include "oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <stdint.h>

static float box_size(const float* b, const float* v) {
    const float* mn = b;
    const float* mx = b + 4;
    int inside = 1;
    for (int i = 0; i < 3; i++) inside &= (v[i] >= mn[i] && v[i] <= mx[i]);
    if (inside) return FLT_MAX;
    float d[3];
    for (int i = 0; i < 3; i++) {
        float c = fmaxf(mn[i], fminf(mx[i], v[i]));
        d[i] = v[i] - c;
    }
    float d2 = fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1]));
    return mn[3] / sqrtf(d2);
}

int orc_expand_to_size(const int32_t* nodes, const float* boxes, int N, float tau, const float* viewpoint, int32_t* render_indices,
                       int32_t* parent_indices, int32_t* node_ids) {
    int off = 0;
    for (int i = 0; i < N; i++) {
        const int32_t* n = nodes + (long)i * 7; /* depth parent start count_leafs count_merged start_children count_children */
        float size = box_size(boxes + (long)i * 8, viewpoint);
        int count = 0;
        if (size >= tau) {
            count = n[3];
        } else if (n[1] != -1) {
            float psize = box_size(boxes + (long)n[1] * 8, viewpoint);
            if (psize >= tau) {
                count = n[3];
                if (n[0] != 0) count += n[4];
            }
        }
        int parentgaussian = -1;
        if (n[1] != -1) parentgaussian = nodes[(long)n[1] * 7 + 2];
        for (int k = 0; k < count; k++) {
            render_indices[off] = n[2] + k;
            if (parent_indices) parent_indices[off] = parentgaussian;
            if (node_ids) node_ids[off] = i;
            off++;
        }
    }
    return off;
}

/* The same cut with the reference's three-stage structure spread over the host cores (markNodesForSize :495-529 -> inclusive scan
 * :1021-1023 -> putRenderIndices :68-93): counts in parallel, a two-level scan over contiguous node chunks, scatter in parallel.
 * Identical output to orc_expand_to_size; `counts` is caller scratch of N ints. Returns K; *threads_used reports the team size. */
#ifdef _OPENMP
#include <omp.h>
#endif
int orc_expand_to_size_omp(const int32_t* nodes, const float* boxes, int N, float tau, const float* viewpoint, int32_t* render_indices,
                           int32_t* parent_indices, int32_t* node_ids, int32_t* counts, int* threads_used) {
    int T = 1;
#ifdef _OPENMP
    T = omp_get_max_threads();
#endif
    if (T > 256) T = 256;
    long chunk_sum[257];
    int total = 0;
#pragma omp parallel num_threads(T)
    {
        int t = 0, nt = 1;
#ifdef _OPENMP
        t = omp_get_thread_num();
        nt = omp_get_num_threads();
#endif
        const long lo = (long)N * t / nt, hi = (long)N * (t + 1) / nt;
        long sum = 0;
        for (long i = lo; i < hi; i++) {
            const int32_t* n = nodes + i * 7;
            float size = box_size(boxes + i * 8, viewpoint);
            int count = 0;
            if (size >= tau) {
                count = n[3];
            } else if (n[1] != -1) {
                float psize = box_size(boxes + (long)n[1] * 8, viewpoint);
                if (psize >= tau) {
                    count = n[3];
                    if (n[0] != 0) count += n[4];
                }
            }
            counts[i] = count;
            sum += count;
        }
        chunk_sum[t + 1] = sum;
#pragma omp barrier
#pragma omp single
        {
            chunk_sum[0] = 0;
            for (int k = 0; k < nt; k++) chunk_sum[k + 1] += chunk_sum[k];
            total = (int)chunk_sum[nt];
            if (threads_used) *threads_used = nt;
        }
        long off = chunk_sum[t];
        for (long i = lo; i < hi; i++) {
            const int32_t* n = nodes + i * 7;
            const int count = counts[i];
            if (!count) continue;
            int parentgaussian = -1;
            if (n[1] != -1) parentgaussian = nodes[(long)n[1] * 7 + 2];
            for (int k = 0; k < count; k++) {
                render_indices[off] = n[2] + k;
                if (parent_indices) parent_indices[off] = parentgaussian;
                if (node_ids) node_ids[off] = (int32_t)i;
                off++;
            }
        }
    }
    return total;
}

void orc_get_interpolation_weights(const int32_t* indices, int K, float tau, const int32_t* nodes, const float* boxes, const float* viewpoint,
                                   float* ts, int32_t* kids) {
    for (int idx = 0; idx < K; idx++) {
        int node_id = indices[idx];
        const int32_t* n = nodes + (long)node_id * 7;
        float t;
        if (n[1] == -1) {
            t = 1.0f;
        } else {
            float parentsize = box_size(boxes + (long)n[1] * 8, viewpoint);
            if (parentsize > 2.0f * tau) {
                t = 1.0f;
            } else {
                float size = box_size(boxes + (long)node_id * 8, viewpoint);
                float start = fmaxf(0.5f * parentsize, size);
                float diff = parentsize - start;
                if (diff <= 0) {
                    t = 1.0f;
                } else {
                    float tdiff = fmaxf(0.0f, tau - start);
                    t = fmaxf(1.0f - (tdiff / diff), 0.0f);
                }
            }
        }
        ts[idx] = t;
        kids[idx] = (n[1] == -1) ? 1 : nodes[(long)n[1] * 7 + 6];
    }
}

/* ---- dynamic hierarchy format (HierarchyNode = 6 x i32: depth parent child_count first_child next_sibling max_side_length,
 * GH/types.h:60-79): markNodesForSizeDynamic + InclusiveSum + putRenderIndicesDynamic (GH/runtime_switching.cu:533-582, :95-108),
 * is_in_frustum (:165-186), computeSizeGPUDynamic / pointgaussiandistCUDA (:146-163, :222-232), computeTsIndexedDynamic (:637-684).
 * Distance with the association of the reference SASS, which differs between its two kernels: the mark kernel compiles to
 * fma(dz,dz, fma(dy,dy, dx*dx)), the weights kernel to fma(dz,dz, fma(dx,dx, dy*dy)); dot product x*zx -> fma(y,zy,.) -> fma(z,zz,.). */
static float dyn_norm_assoc(const float* pos, const float* v, float* d, int y_first) {
    d[0] = v[0] - pos[0];
    d[1] = v[1] - pos[1];
    d[2] = v[2] - pos[2];
    if (y_first) return sqrtf(fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1])));
    return sqrtf(fmaf(d[2], d[2], fmaf(d[1], d[1], d[0] * d[0])));
}
static float dyn_norm(const float* pos, const float* v, float* d) { return dyn_norm_assoc(pos, v, d, 0); }
static float dyn_size(const float* scale, float norm) {
    if (norm < 0.0f) return 0.0f;
    return fmaxf(scale[0], fmaxf(scale[1], scale[2])) / norm;
}

int orc_expand_to_size_dynamic(const int32_t* nodes, const float* positions, const float* scales, int N, float tau, const float* viewpoint,
                               const float* zdir, int32_t* render_indices, int32_t* parent_indices, int32_t* node_ids) {
    int off = 0;
    for (int i = 0; i < N; i++) {
        const int32_t* n = nodes + (long)i * 6;
        float d[3];
        float norm = dyn_norm(positions + (long)i * 3, viewpoint, d);
        float c = fmaf(d[2] / norm, zdir[2], fmaf(d[1] / norm, zdir[1], (d[0] / norm) * zdir[0]));
        if (!(c < -0.5f)) continue;
        if (n[0] < 0) continue;
        float size = dyn_size(scales + (long)i * 3, norm);
        int count = 0;
        if (size >= tau && n[2] == 0) {
            count = 1;
        } else if (n[1] >= 0) {
            float pd[3];
            float pnorm = dyn_norm(positions + (long)n[1] * 3, viewpoint, pd);
            float psize = dyn_size(scales + (long)n[1] * 3, pnorm);
            if (psize >= tau && size < tau) count = 1;
        }
        if (count) {
            render_indices[off] = i;
            if (node_ids) node_ids[off] = i;
            if (parent_indices && n[1] != -1) parent_indices[off] = n[1];
            off++;
        }
    }
    return off;
}

void orc_get_interpolation_weights_dynamic(const int32_t* indices, int K, float tau, const int32_t* nodes, const float* positions,
                                           const float* scales, const float* viewpoint, float* ts, int32_t* kids) {
    for (int idx = 0; idx < K; idx++) {
        int node_id = indices[idx];
        const int32_t* n = nodes + (long)node_id * 6;
        float t = 1.0f, d[3];
        if (n[1] >= 0) {
            float parentsize = dyn_size(scales + (long)n[1] * 3, dyn_norm_assoc(positions + (long)n[1] * 3, viewpoint, d, 1));
            if (!(parentsize > 2.0f * tau)) {
                float size = dyn_size(scales + (long)node_id * 3, dyn_norm_assoc(positions + (long)node_id * 3, viewpoint, d, 1));
                float start = fmaxf(0.5f * parentsize, size);
                float diff = parentsize - start;
                if (!(diff <= 0)) {
                    float tdiff = fmaxf(0.0f, tau - start);
                    t = fmaxf(1.0f - (tdiff / diff), 0.0f);
                }
            }
        }
        ts[idx] = t;
        kids[idx] = n[1] < 0 ? 1 : nodes[(long)n[1] * 6 + 2];
    }
}

/* ---- SPT cut: Switching::getSPTCut (GH/runtime_switching.cu:881-994), thread loops of binary_search_SPTs (:784-810) and
 * populate_SPT (:812-852) restated one to one, then DeviceSelect::If(!= 0) and the exclusive sum of the per-slot counts.
 * cut must hold the concatenation length (sum of the interval sizes; call with cut = NULL to get it through *positions).
 * n_entries guards the one-past-the-end probe (the reference reads out of bounds there); such a probe counts as not kept. */
int orc_get_spt_cut(int S, const int32_t* gaussian_indices, int n_entries, const int32_t* starts, const float* spt_max, const float* spt_min,
                    const int32_t* spt_indices, const float* spt_distances, int32_t* cut, int32_t* counts_prefix, int32_t* positions) {
    if (S <= 0) { if (positions) *positions = 0; return 0; }
    int32_t* sizes = (int32_t*)malloc(sizeof(int32_t) * (size_t)S);
    int32_t* prefix = (int32_t*)malloc(sizeof(int32_t) * (size_t)S);
    int32_t* counts = (int32_t*)calloc((size_t)S, sizeof(int32_t));
    for (int idx = 0; idx < S; idx++) {
        int index = spt_indices[idx];
        float distance = spt_distances[idx];
        int low = starts[index], high = starts[index + 1];
        int pivot = (low + high) / 2;
        while ((high - low) > 1) {
            if (spt_max[pivot] > distance) low = pivot; else high = pivot;
            pivot = (low + high) / 2;
        }
        sizes[idx] = high - starts[index];
    }
    int sum = 0;
    for (int i = 0; i < S; i++) { prefix[i] = sum; sum += sizes[i]; }
    if (positions) *positions = sum;
    int K = 0;
    if (cut) {
        for (int idx = 0; idx < sum; idx++) {
            int low = 0, high = S, interval = S / 2;
            while (high - low > 1) {
                if (prefix[interval] >= idx) high = interval; else low = interval;
                interval = (high + low) / 2;
            }
            int offset = idx - prefix[interval];
            long g = (long)starts[spt_indices[interval]] + offset;
            int result = 0;
            if (g >= 0 && g < n_entries && spt_min[g] < spt_distances[interval]) {
                counts[interval] += 1;
                result = gaussian_indices[g];
            }
            if (result != 0) cut[K++] = result;
        }
    }
    int run = 0;
    for (int i = 0; i < S; i++) { if (counts_prefix) counts_prefix[i] = run; run += counts[i]; }
    free(sizes); free(prefix); free(counts);
    return K;
}

/* ---- Morton codes: getMortonCodeCUDA (GH/morton.cu:9-42), thread body restated; the float -> int64 conversion follows the
 * device's cvt.rzi.s64.f32 (NaN -> 0, saturating) instead of C's undefined behaviour out of range. */
static int64_t f2ll_rz(float v) {
    if (v != v) return 0;
    if (v >= 9223372036854775808.0f) return INT64_MAX;
    if (v <= -9223372036854775808.0f) return INT64_MIN;
    return (int64_t)v;
}

void orc_get_morton_codes(const float* xyz, const float* mn, const float* mx, int64_t* codes, int P) {
    float box[3] = {mx[0] - mn[0], mx[1] - mn[1], mx[2] - mn[2]};
    for (int idx = 0; idx < P; idx++) {
        int64_t q[3];
        for (int c = 0; c < 3; c++) {
            float v = (xyz[3 * (long)idx + c] - mn[c]) / box[c];
            v = v * (float)(1 << 21);
            q[c] = f2ll_rz(v);
        }
        int64_t code = 0;
        for (int i = 0; i < 21; ++i) {
            code = code | (q[0] >> i & 1) << (3 * i);
            code = code | (q[1] >> i & 1) << (3 * i + 1);
            code = code | (q[2] >> i & 1) << (3 * i + 2);
        }
        codes[idx] = code;
    }
}

/* ---- depth-target cut: recExpand / Traversal::expandToTarget (GH/traversal.cpp:15-39) restated with the same recursion. */
static void orc_rec_expand(int node_id, const int32_t* nodes, int32_t* out, long* n, int target) {
    const int32_t* node = nodes + 7 * (long)node_id;
    for (int i = 0; i < node[3]; i++) { if (out) out[*n] = node[2] + i; (*n)++; }
    if (node[0] <= target) {
        for (int i = 0; i < node[4]; i++) { if (out) out[*n] = node[2] + node[3] + i; (*n)++; }
    } else {
        for (int i = 0; i < node[6]; i++) orc_rec_expand(node[5] + i, nodes, out, n, target);
    }
}

long orc_expand_to_target(const int32_t* nodes, int target, int32_t* out) {
    long n = 0;
    orc_rec_expand(0, nodes, out, &n, target);
    return n;
}
