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
