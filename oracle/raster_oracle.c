/* raster_oracle.c — TEST INFRASTRUCTURE ONLY. Plain-C (OpenMP) restatement of the reference rasterizer:
 *   preprocessCUDA        HR/cuda_rasterizer/forward.cu:218-445  (computeCov3D :181-215, computeCov2D :141-176,
 *                         computeColorFromSH[Interp] :25-138, getRect/ndc2Pix auxiliary.h:53-80)
 *   InclusiveSum, duplicateWithKeys, SortPairs, identifyTileRanges   rasterizer_impl.cu:322, :70-115, :358-363, :120-142
 *   renderCUDA fwd        forward.cu:450-596
 *   renderCUDA bwd        backward.cu:498-721
 *   computeCov2DCUDA, preprocessCUDA bwd, computeCov3D bwd, computeColorFromSH bwd   backward.cu:147-495
 * Float results agree with the GPU to rounding (expf/powf are libm here, CUDA's there; FMA contraction is off), so
 * images/gradients are compared with the tolerances of BASELINE.json and integer outputs derived from floats may differ
 * in isolated entries; bit-exactness of integers is tested on the GPU against the reference's own CUDA build. */
#include "oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define BX 16
#define BY 16

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

static const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* 3x3 matrices in glm's column-major convention: m[c][r] */
typedef struct { float m[3][3]; } M3;
static M3 m3mul(M3 a, M3 b) {
    M3 r;
    for (int c = 0; c < 3; c++)
        for (int k = 0; k < 3; k++) r.m[c][k] = a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1] + a.m[2][k] * b.m[c][2];
    return r;
}
static M3 m3t(M3 a) {
    M3 r;
    for (int c = 0; c < 3; c++)
        for (int k = 0; k < 3; k++) r.m[c][k] = a.m[k][c];
    return r;
}
static M3 m3cols(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1, float c2) {
    M3 r = {{{a0, a1, a2}, {b0, b1, b2}, {c0, c1, c2}}};
    return r;
}
static M3 quat_rot(const float* q) {
    float r = q[0], x = q[1], y = q[2], z = q[3];
    return m3cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y), 2.f * (x * y + r * z),
                  1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x), 2.f * (x * z - r * y), 2.f * (y * z + r * x),
                  1.f - 2.f * (x * x + y * y));
}
static void cov3d(const float* s, float mod, const float* q, float* c6) {
    M3 S = m3cols(mod * s[0], 0, 0, 0, mod * s[1], 0, 0, 0, mod * s[2]);
    M3 Mx = m3mul(S, quat_rot(q));
    M3 Sg = m3mul(m3t(Mx), Mx);
    c6[0] = Sg.m[0][0]; c6[1] = Sg.m[0][1]; c6[2] = Sg.m[0][2]; c6[3] = Sg.m[1][1]; c6[4] = Sg.m[1][2]; c6[5] = Sg.m[2][2];
}
static void xf43(const float* p, const float* m, float* o) {
    o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
    o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
    o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

typedef struct { float t[3], txtz, tytz; M3 T; } Cov2Ctx;
static void cov2d(const float* mean, float fx, float fy, float tanx, float tany, const float* c6, const float* view, float* out3,
                  Cov2Ctx* ctx) {
    float t[3];
    xf43(mean, view, t);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    M3 J = m3cols(fx / t[2], 0.0f, -(fx * t[0]) / (t[2] * t[2]), 0.0f, fy / t[2], -(fy * t[1]) / (t[2] * t[2]), 0, 0, 0);
    M3 Wm = m3cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    M3 T = m3mul(Wm, J);
    M3 V = m3cols(c6[0], c6[1], c6[2], c6[1], c6[3], c6[4], c6[2], c6[4], c6[5]);
    M3 cov = m3mul(m3mul(m3t(T), m3t(V)), T);
    out3[0] = cov.m[0][0]; out3[1] = cov.m[0][1]; out3[2] = cov.m[1][1];
    if (ctx) { ctx->t[0] = t[0]; ctx->t[1] = t[1]; ctx->t[2] = t[2]; ctx->txtz = txtz; ctx->tytz = tytz; ctx->T = T; }
}
static void get_rect(const float* p, int ex, int ey, int gx, int gy, int* rmin, int* rmax) {
    int a;
    a = (int)((p[0] - ex) / BX); if (a < 0) a = 0; rmin[0] = a < gx ? a : gx;
    a = (int)((p[1] - ey) / BY); if (a < 0) a = 0; rmin[1] = a < gy ? a : gy;
    a = (int)((p[0] + ex + BX - 1) / BX); if (a < 0) a = 0; rmax[0] = a < gx ? a : gx;
    a = (int)((p[1] + ey + BY - 1) / BY); if (a < 0) a = 0; rmax[1] = a < gy ? a : gy;
}

/* SH -> RGB with coefficients sh[k][c] (already lerped for LOD entries), unit direction d */
static void sh_rgb(int deg, float sh[16][3], const float* d, float* out) {
    float x = d[0], y = d[1], z = d[2];
    for (int c = 0; c < 3; c++) {
        float r = SH_C0 * sh[0][c];
        if (deg > 0) {
            r = r - SH_C1 * y * sh[1][c] + SH_C1 * z * sh[2][c] - SH_C1 * x * sh[3][c];
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + SH_C2[0] * xy * sh[4][c] + SH_C2[1] * yz * sh[5][c] + SH_C2[2] * (2.0f * zz - xx - yy) * sh[6][c] +
                    SH_C2[3] * xz * sh[7][c] + SH_C2[4] * (xx - yy) * sh[8][c];
                if (deg > 2)
                    r = r + SH_C3[0] * y * (3.0f * xx - yy) * sh[9][c] + SH_C3[1] * xy * z * sh[10][c] +
                        SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11][c] + SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12][c] +
                        SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13][c] + SH_C3[5] * z * (xx - yy) * sh[14][c] +
                        SH_C3[6] * x * (xx - 3.0f * yy) * sh[15][c];
            }
        }
        out[c] = r + 0.5f;
    }
}

static uint32_t preprocess_one(const orc_scene* sc, orc_state* st, int t_idx, int gx, int gy, float fx, float fy) {
    const int r_idx = sc->indices ? sc->indices[t_idx] : t_idx;
    st->radii[t_idx] = 0;
    int p_idx = -1;
    float t = 0;
    if (sc->parent_indices) { p_idx = sc->parent_indices[t_idx]; if (p_idx != -1) t = sc->ts[t_idx]; }
    const int lod = p_idx != -1;
    const float* own = sc->means3D + 3 * (long)r_idx;
    float p[3] = {own[0], own[1], own[2]};
    if (lod) {
        const float* pa = sc->means3D + 3 * (long)p_idx;
        for (int i = 0; i < 3; i++) p[i] = fmaf(t, p[i], (1.0f - t) * pa[i]); /* nvcc contracts the lerp this way */
    }
    const float* pm = sc->projmatrix;
    float hom[4] = {pm[0] * p[0] + pm[4] * p[1] + pm[8] * p[2] + pm[12], pm[1] * p[0] + pm[5] * p[1] + pm[9] * p[2] + pm[13],
                    pm[2] * p[0] + pm[6] * p[1] + pm[10] * p[2] + pm[14], pm[3] * p[0] + pm[7] * p[1] + pm[11] * p[2] + pm[15]};
    float p_w = 1.0f / (hom[3] + 0.0000001f);
    float proj[3] = {hom[0] * p_w, hom[1] * p_w, hom[2] * p_w};
    float pv[3];
    xf43(p, sc->viewmatrix, pv);
    if (pv[2] <= 0.2f) return 0;
    for (int i = 0; i < 3; i++) if (p[i] < -FLT_MAX || p[i] > FLT_MAX) return 0;
    float s[3], q[4];
    for (int i = 0; i < 3; i++) s[i] = sc->scales[3 * (long)r_idx + i];
    for (int i = 0; i < 4; i++) q[i] = sc->rotations[4 * (long)r_idx + i];
    if (lod) {
        float o[4];
        for (int i = 0; i < 3; i++) s[i] = fmaf(t, s[i], (1.0f - t) * sc->scales[3 * (long)p_idx + i]);
        for (int i = 0; i < 4; i++) o[i] = sc->rotations[4 * (long)p_idx + i];
        float dot = (q[0] * o[0] + q[1] * o[1]) + (q[2] * o[2] + q[3] * o[3]);
        if (dot < 0.0) for (int i = 0; i < 4; i++) o[i] = -o[i];
        for (int i = 0; i < 4; i++) q[i] = fmaf(t, q[i], (1.0f - t) * o[i]);
    }
    float* c6 = st->cov3D + 6 * (long)t_idx;
    cov3d(s, sc->scale_modifier, q, c6);
    float cov[3];
    cov2d(p, fx, fy, sc->tan_fovx, sc->tan_fovy, c6, sc->viewmatrix, cov, NULL);
    const float h_var = 0.3f;
    const float det_cov = cov[0] * cov[2] - cov[1] * cov[1];
    cov[0] += h_var; cov[2] += h_var;
    const float det = cov[0] * cov[2] - cov[1] * cov[1];
    const float h_conv = sqrtf(fmaxf(0.000025f, det_cov / det));
    if (det == 0.0f) return 0;
    float det_inv = 1.f / det;
    float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    float mid = 0.5f * (cov[0] + cov[2]);
    float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det)), l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
    float pt[2] = {ndc2pix(proj[0], sc->W), ndc2pix(proj[1], sc->H)};
    int ex = (int)ceilf(3.f * sqrtf(cov[0])), ey = (int)ceilf(3.f * sqrtf(cov[2]));
    st->rects[2 * (long)t_idx] = ex; st->rects[2 * (long)t_idx + 1] = ey;
    int rmin[2], rmax[2];
    get_rect(pt, ex, ey, gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) return 0;
    if (!sc->colors_precomp) {
        float d[3] = {own[0] - sc->campos[0], own[1] - sc->campos[1], own[2] - sc->campos[2]};
        float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int i = 0; i < 3; i++) d[i] = d[i] / len;
        float sh[16][3];
        int nco = (sc->D + 1) * (sc->D + 1);
        const float* a = sc->shs + (long)r_idx * sc->M * 3;
        const float* b = lod ? sc->shs + (long)p_idx * sc->M * 3 : NULL;
        for (int k = 0; k < nco; k++)
            for (int c = 0; c < 3; c++) sh[k][c] = lod ? fmaf(t, a[3 * k + c], (1.0f - t) * b[3 * k + c]) : a[3 * k + c];
        float rgb[3];
        sh_rgb(sc->D, sh, d, rgb);
        for (int c = 0; c < 3; c++) {
            st->clamped[3 * (long)t_idx + c] = rgb[c] < 0;
            st->rgb[3 * (long)t_idx + c] = fmaxf(rgb[c], 0.0f);
        }
    }
    st->depths[t_idx] = pv[2];
    st->radii[t_idx] = (int)my_radius;
    st->means2D[2 * (long)t_idx] = pt[0]; st->means2D[2 * (long)t_idx + 1] = pt[1];
    float op = sc->opacities[r_idx];
    if (lod) op = fmaf(t, op, (1.0f - t) * sc->opacities[p_idx]);
    float* co = st->conic_opacity + 4 * (long)t_idx;
    co[0] = conic[0]; co[1] = conic[1]; co[2] = conic[2]; co[3] = op * h_conv;
    return (uint32_t)((rmax[1] - rmin[1]) * (rmax[0] - rmin[0]));
}

/* stable LSD radix sort of (key,value) pairs on the low `bits` bits (cub::DeviceRadixSort::SortPairs semantics) */
static void sort_pairs(uint64_t* k, uint32_t* v, uint64_t* k2, uint32_t* v2, int n, int bits) {
    uint64_t *ki = k, *ko = k2;
    uint32_t *vi = v, *vo = v2;
    int passes = (bits + 15) / 16;
    size_t* hist = (size_t*)malloc(65536 * sizeof(size_t));
    for (int p = 0; p < passes; p++) {
        int sh = 16 * p;
        memset(hist, 0, 65536 * sizeof(size_t));
        for (int i = 0; i < n; i++) hist[(ki[i] >> sh) & 0xffff]++;
        size_t run = 0;
        for (int d = 0; d < 65536; d++) { size_t c = hist[d]; hist[d] = run; run += c; }
        for (int i = 0; i < n; i++) { size_t o = hist[(ki[i] >> sh) & 0xffff]++; ko[o] = ki[i]; vo[o] = vi[i]; }
        uint64_t* tk = ki; ki = ko; ko = tk;
        uint32_t* tv = vi; vi = vo; vo = tv;
    }
    free(hist);
    if (ki != k2) { memcpy(k2, ki, (size_t)n * 8); memcpy(v2, vi, (size_t)n * 4); }
}

static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

void orc_free_instances(orc_state* st) {
    free(st->keys_unsorted); free(st->vals_unsorted); free(st->keys); free(st->point_list);
    st->keys_unsorted = st->keys = NULL; st->vals_unsorted = st->point_list = NULL; st->R = 0;
}

int orc_forward(const orc_scene* sc, orc_state* st) {
    const int P = sc->P, W = sc->W, H = sc->H;
    const int gx = (W + BX - 1) / BX, gy = (H + BY - 1) / BY;
    const float fy = H / (2.0f * sc->tan_fovy), fx = W / (2.0f * sc->tan_fovx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) st->tiles_touched[i] = preprocess_one(sc, st, i, gx, gy, fx, fy);
    uint32_t run = 0;
    for (int i = 0; i < P; i++) { run += st->tiles_touched[i]; st->point_offsets[i] = run; }
    const int R = (int)run;
    st->R = R;
    st->keys_unsorted = st->keys = NULL; st->vals_unsorted = st->point_list = NULL;
    memset(st->ranges, 0, (size_t)gx * gy * 8);
    if (R == 0) return 0;
    st->keys_unsorted = (uint64_t*)malloc((size_t)R * 8); st->keys = (uint64_t*)malloc((size_t)R * 8);
    st->vals_unsorted = (uint32_t*)malloc((size_t)R * 4); st->point_list = (uint32_t*)malloc((size_t)R * 4);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < P; i++) {
        if (st->radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0 : st->point_offsets[i - 1];
        int rmin[2], rmax[2];
        get_rect(st->means2D + 2 * (long)i, st->rects[2 * (long)i], st->rects[2 * (long)i + 1], gx, gy, rmin, rmax);
        uint32_t dbits;
        memcpy(&dbits, st->depths + i, 4);
        for (int y = rmin[1]; y < rmax[1]; y++)
            for (int x = rmin[0]; x < rmax[0]; x++) {
                uint64_t key = (uint64_t)(y * gx + x);
                key <<= 32; key |= dbits;
                st->keys_unsorted[off] = key; st->vals_unsorted[off] = (uint32_t)i; off++;
            }
    }
    {
        uint64_t* tk = (uint64_t*)malloc((size_t)R * 8);
        uint32_t* tv = (uint32_t*)malloc((size_t)R * 4);
        memcpy(tk, st->keys_unsorted, (size_t)R * 8); memcpy(tv, st->vals_unsorted, (size_t)R * 4);
        sort_pairs(tk, tv, st->keys, st->point_list, R, 32 + (int)higher_msb((uint32_t)(gx * gy)));
        free(tk); free(tv);
    }
    for (int idx = 0; idx < R; idx++) { /* identifyTileRanges, including its L==1 behaviour */
        uint32_t cur = (uint32_t)(st->keys[idx] >> 32);
        if (idx == 0) st->ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(st->keys[idx - 1] >> 32);
            if (cur != prev) { st->ranges[2 * prev + 1] = idx; st->ranges[2 * cur] = idx; }
            if (idx == R - 1) st->ranges[2 * cur + 1] = R;
        }
    }
    const float* feat = sc->colors_precomp ? sc->colors_precomp : st->rgb;
    const int lodb = sc->ts != NULL && sc->kids != NULL;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = st->ranges[2 * (ty * gx + tx)], r1 = st->ranges[2 * (ty * gx + tx) + 1];
            for (int py = ty * BY; py < ty * BY + BY && py < H; py++)
                for (int px = tx * BX; px < tx * BX + BX && px < W; px++) {
                    float T = 1.0f, C[3] = {0, 0, 0}, inv = 0.0f;
                    uint32_t contributor = 0, last = 0;
                    for (uint32_t i = r0; i < r1; i++) {
                        contributor++;
                        const uint32_t id = st->point_list[i];
                        const float* xy = st->means2D + 2 * (long)id;
                        const float* co = st->conic_opacity + 4 * (long)id;
                        float dx = xy[0] - (float)px, dy = xy[1] - (float)py;
                        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.0f) continue;
                        float my_alpha = fminf(0.99f, co[3] * expf(power));
                        float alpha = my_alpha;
                        if (lodb) {
                            float tt = sc->ts[id], frac = 1.0f / (float)sc->kids[id];
                            alpha = tt * my_alpha + (1.0f - tt) * (1.0f - powf(1.0f - my_alpha, frac));
                        }
                        if (alpha < 1.0f / 255.0f) continue;
                        float test_T = T * (1 - alpha);
                        if (test_T < 0.0001f) break;
                        st->seen[id] = 1;
                        for (int c = 0; c < 3; c++) C[c] += feat[3 * (long)id + c] * alpha * T;
                        if (sc->do_depth) inv += (1 / st->depths[id]) * alpha * T;
                        T = test_T;
                        last = contributor;
                    }
                    const long pid = (long)py * W + px;
                    st->final_T[pid] = T; st->n_contrib[pid] = last;
                    for (int c = 0; c < 3; c++) st->color[(long)c * H * W + pid] = C[c] + T * sc->background[c];
                    if (sc->do_depth) st->invdepth[pid] = inv;
                }
        }
    return R;
}

static inline void atomic_addf(float* p, float v) {
#pragma omp atomic
    *p += v;
}

static void blend_backward(const orc_scene* sc, const orc_state* st, const float* dL_dpix, const float* dL_dinv_pix, float* dL_dmean2D,
                           float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth) {
    const int W = sc->W, H = sc->H, gx = (W + BX - 1) / BX, gy = (H + BY - 1) / BY;
    const float* feat = sc->colors_precomp ? sc->colors_precomp : st->rgb;
    const int lodb = sc->ts != NULL && sc->kids != NULL;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = st->ranges[2 * (ty * gx + tx)], r1 = st->ranges[2 * (ty * gx + tx) + 1];
            for (int py = ty * BY; py < ty * BY + BY && py < H; py++)
                for (int px = tx * BX; px < tx * BX + BX && px < W; px++) {
                    const long pid = (long)py * W + px;
                    const float T_final = st->final_T[pid];
                    float T = T_final;
                    const uint32_t last_contributor = st->n_contrib[pid];
                    float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, dL_dpixel[3];
                    float last_alpha = 0, accum_inv = 0, last_inv = 0, dL_inv = dL_dinv_pix ? dL_dinv_pix[pid] : 0;
                    for (int c = 0; c < 3; c++) dL_dpixel[c] = dL_dpix[(long)c * H * W + pid];
                    float bg_dot = 0;
                    for (int c = 0; c < 3; c++) bg_dot += sc->background[c] * dL_dpixel[c];
                    for (uint32_t pos = last_contributor; pos-- > 0;) {
                        if (r0 + pos >= r1) continue;
                        const uint32_t id = st->point_list[r0 + pos];
                        const float* xy = st->means2D + 2 * (long)id;
                        const float* co = st->conic_opacity + 4 * (long)id;
                        const float dx = xy[0] - (float)px, dy = xy[1] - (float)py;
                        const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.0f) continue;
                        const float G = expf(power);
                        const float test_alpha = co[3] * G;
                        const int nullalpha = test_alpha > 0.99f;
                        const float my_alpha = fminf(0.99f, test_alpha);
                        float alpha = my_alpha, t = 1.f, frac = 1.f;
                        if (lodb) {
                            t = sc->ts[id]; frac = 1.0f / sc->kids[id];
                            alpha = t * my_alpha + (1.0f - t) * (1.0f - powf(1.0f - my_alpha, frac));
                        }
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.f - alpha);
                        const float weight = alpha * T;
                        const long gid = sc->indices ? sc->indices[id] : (long)id;
                        float dL_dalpha = 0.0f;
                        for (int c = 0; c < 3; c++) {
                            const float col = feat[3 * (long)id + c];
                            accum_rec[c] = last_alpha * last_color[c] + (1.f - last_alpha) * accum_rec[c];
                            last_color[c] = col;
                            dL_dalpha += (col - accum_rec[c]) * dL_dpixel[c];
                            atomic_addf(&dL_dcolor[3 * gid + c], weight * dL_dpixel[c]);
                        }
                        if (dL_dinvdepth) {
                            const float invd = 1.f / st->depths[id];
                            accum_inv = last_alpha * last_inv + (1.f - last_alpha) * accum_inv;
                            last_inv = invd;
                            dL_dalpha += (invd - accum_inv) * dL_inv;
                            atomic_addf(&dL_dinvdepth[gid], weight * dL_inv);
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                        dL_dalpha = nullalpha ? 0 : dL_dalpha;
                        const float dL_dG = co[3] * dL_dalpha, gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co[0] - gdy * co[1], dG_ddely = -gdy * co[2] - gdx * co[1];
                        atomic_addf(&dL_dmean2D[3 * gid + 0], dL_dG * dG_ddelx * ddelx_dx);
                        atomic_addf(&dL_dmean2D[3 * gid + 1], dL_dG * dG_ddely * ddely_dy);
                        atomic_addf(&dL_dconic[4 * gid + 0], -0.5f * gdx * dx * dL_dG);
                        atomic_addf(&dL_dconic[4 * gid + 1], -0.5f * gdx * dy * dL_dG);
                        atomic_addf(&dL_dconic[4 * gid + 3], -0.5f * gdy * dy * dL_dG);
                        float mult = 1.0f;
                        if (lodb) mult = t - powf(1.0f - my_alpha, frac - 1.0f) * (t - 1.0f) * frac;
                        atomic_addf(&dL_dopacity[gid], mult * G * dL_dalpha);
                    }
                }
        }
}

static float sqf(float x) { return x * x; }

/* computeCov2DCUDA + preprocessCUDA bwd (+ computeCov3D / computeColorFromSH bwd) for one list entry. LOD children:
 * own opacity/scale/rot/SH gradients are discarded and (1-t) * dL_dmean goes to the parent (backward.cu:458-494); the
 * clamp flags used are the ones the forward computed for the entry (the reference reads an uninitialised array there,
 * SURVEY §2.3-6). */
static void gaussian_backward(const orc_scene* sc, const orc_state* st, int t_idx, float fx, float fy, const float* dL_dmean2D,
                              const float* dL_dconic, float* dL_dopacity, const float* dL_dcolor, const float* dL_dinvdepth,
                              float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot) {
    if (!(st->radii[t_idx] > 0)) return;
    const long idx = sc->indices ? sc->indices[t_idx] : t_idx;
    const float* view = sc->viewmatrix;
    const float* proj = sc->projmatrix;
    const float* c6 = st->cov3D + 6 * (long)t_idx;
    const float* mean = sc->means3D + 3 * idx;
    const float dcon[3] = {dL_dconic[4 * idx], dL_dconic[4 * idx + 1], dL_dconic[4 * idx + 3]};
    Cov2Ctx ctx;
    float cov2[3];
    cov2d(mean, fx, fy, sc->tan_fovx, sc->tan_fovy, c6, view, cov2, &ctx);
    const float limx = 1.3f * sc->tan_fovx, limy = 1.3f * sc->tan_fovy;
    const float xg = (ctx.txtz < -limx || ctx.txtz > limx) ? 0 : 1, yg = (ctx.tytz < -limy || ctx.tytz > limy) ? 0 : 1;
    M3 T = ctx.T;
    M3 V = m3cols(c6[0], c6[1], c6[2], c6[1], c6[3], c6[4], c6[2], c6[4], c6[5]);
    M3 Wm = m3cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    float c_xx = cov2[0], c_xy = cov2[1], c_yy = cov2[2];
    const float h_var = 0.3f;
    const float det_cov = c_xx * c_yy - c_xy * c_xy;
    c_xx += h_var; c_yy += h_var;
    const float det_h = c_xx * c_yy - c_xy * c_xy;
    const float h_conv = sqrtf(fmaxf(0.000025f, det_cov / det_h));
    const float dop_v = dL_dopacity[idx];
    const float d_h = dop_v * sc->opacities[idx];
    const float dop_out = dop_v * h_conv;
    const float d_inside = (det_cov / det_h) <= 0.000025f ? 0.f : d_h / (2 * h_conv);
    float dxx, dxy, dyy;
    {
        const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
        const float denom_f = d_inside / sqf(w * w + w * (x + y) + x * y - z * z);
        dxx = w * (w * y + y * y + z * z) * denom_f;
        dyy = w * (w * x + x * x + z * z) * denom_f;
        dxy = -2.f * w * z * (w + x + y) * denom_f;
    }
    const float denom = c_xx * c_yy - c_xy * c_xy;
    const float d2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6] = {0, 0, 0, 0, 0, 0};
    if (d2inv != 0) {
        dxx += d2inv * (-c_yy * c_yy * dcon[0] + 2 * c_xy * c_yy * dcon[1] + (denom - c_xx * c_yy) * dcon[2]);
        dyy += d2inv * (-c_xx * c_xx * dcon[2] + 2 * c_xx * c_xy * dcon[1] + (denom - c_xx * c_yy) * dcon[0]);
        dxy += d2inv * 2 * (c_xy * c_yy * dcon[0] - (denom + 2 * c_xy * c_xy) * dcon[1] + c_xx * c_xy * dcon[2]);
        dcov[0] = T.m[0][0] * T.m[0][0] * dxx + T.m[0][0] * T.m[1][0] * dxy + T.m[1][0] * T.m[1][0] * dyy;
        dcov[3] = T.m[0][1] * T.m[0][1] * dxx + T.m[0][1] * T.m[1][1] * dxy + T.m[1][1] * T.m[1][1] * dyy;
        dcov[5] = T.m[0][2] * T.m[0][2] * dxx + T.m[0][2] * T.m[1][2] * dxy + T.m[1][2] * T.m[1][2] * dyy;
        dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dxx + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dxy + 2 * T.m[1][0] * T.m[1][1] * dyy;
        dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dxx + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dxy + 2 * T.m[1][0] * T.m[1][2] * dyy;
        dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dxx + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dxy + 2 * T.m[1][1] * T.m[1][2] * dyy;
    }
    for (int i = 0; i < 6; i++) dL_dcov3D[6 * idx + i] = dcov[i];
    float dT[2][3];
    for (int k = 0; k < 3; k++) {
        float a = T.m[0][0] * V.m[k][0] + T.m[0][1] * V.m[k][1] + T.m[0][2] * V.m[k][2];
        float b = T.m[1][0] * V.m[k][0] + T.m[1][1] * V.m[k][1] + T.m[1][2] * V.m[k][2];
        dT[0][k] = 2 * a * dxx + b * dxy;
        dT[1][k] = 2 * b * dyy + a * dxy;
    }
    const float dJ00 = Wm.m[0][0] * dT[0][0] + Wm.m[0][1] * dT[0][1] + Wm.m[0][2] * dT[0][2];
    const float dJ02 = Wm.m[2][0] * dT[0][0] + Wm.m[2][1] * dT[0][1] + Wm.m[2][2] * dT[0][2];
    const float dJ11 = Wm.m[1][0] * dT[1][0] + Wm.m[1][1] * dT[1][1] + Wm.m[1][2] * dT[1][2];
    const float dJ12 = Wm.m[2][0] * dT[1][0] + Wm.m[2][1] * dT[1][1] + Wm.m[2][2] * dT[1][2];
    const float tz = 1.f / ctx.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = xg * -fx * tz2 * dJ02, dty = yg * -fy * tz2 * dJ12;
    float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * ctx.t[0]) * tz3 * dJ02 + (2 * fy * ctx.t[1]) * tz3 * dJ12;
    if (dL_dinvdepth) dtz -= dL_dinvdepth[idx] / (ctx.t[2] * ctx.t[2]);
    float dmean[3] = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
                      view[8] * dtx + view[9] * dty + view[10] * dtz};
    {
        const float mw = proj[3] * mean[0] + proj[7] * mean[1] + proj[11] * mean[2] + proj[15];
        const float m_w = 1.0f / (mw + 0.0000001f);
        const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * m_w * m_w;
        const float gx2 = dL_dmean2D[3 * idx], gy2 = dL_dmean2D[3 * idx + 1];
        dmean[0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
        dmean[1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
        dmean[2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;
    }
    int parent = -1;
    if (sc->parent_indices) parent = sc->parent_indices[t_idx];
    const int child = parent != -1;
    if (sc->shs) {
        const float dir0[3] = {mean[0] - sc->campos[0], mean[1] - sc->campos[1], mean[2] - sc->campos[2]};
        const float len = sqrtf(dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2]);
        const float x = dir0[0] / len, y = dir0[1] / len, z = dir0[2] / len;
        float dRGB[3];
        for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[3 * idx + c] * (st->clamped[3 * (long)t_idx + c] ? 0 : 1);
        const float* sh = sc->shs + idx * sc->M * 3;
        const int deg = sc->D, nco = (deg + 1) * (deg + 1);
        float w[16], gxv[3] = {0, 0, 0}, gyv[3] = {0, 0, 0}, gzv[3] = {0, 0, 0};
        w[0] = SH_C0;
#define SHC(k, c) sh[3 * (k) + (c)]
        if (deg > 0) {
            w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x;
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            if (deg > 1) { w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2.f * zz - xx - yy); w[7] = SH_C2[3] * xz; w[8] = SH_C2[4] * (xx - yy); }
            if (deg > 2) {
                w[9] = SH_C3[0] * y * (3.f * xx - yy); w[10] = SH_C3[1] * xy * z; w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
                w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
                w[14] = SH_C3[5] * z * (xx - yy); w[15] = SH_C3[6] * x * (xx - 3.f * yy);
            }
            for (int c = 0; c < 3; c++) {
                gxv[c] = -SH_C1 * SHC(3, c); gyv[c] = -SH_C1 * SHC(1, c); gzv[c] = SH_C1 * SHC(2, c);
                if (deg > 1) {
                    gxv[c] += SH_C2[0] * y * SHC(4, c) + SH_C2[2] * 2.f * -x * SHC(6, c) + SH_C2[3] * z * SHC(7, c) + SH_C2[4] * 2.f * x * SHC(8, c);
                    gyv[c] += SH_C2[0] * x * SHC(4, c) + SH_C2[1] * z * SHC(5, c) + SH_C2[2] * 2.f * -y * SHC(6, c) + SH_C2[4] * 2.f * -y * SHC(8, c);
                    gzv[c] += SH_C2[1] * y * SHC(5, c) + SH_C2[2] * 2.f * 2.f * z * SHC(6, c) + SH_C2[3] * x * SHC(7, c);
                }
                if (deg > 2) {
                    gxv[c] += SH_C3[0] * SHC(9, c) * 3.f * 2.f * xy + SH_C3[1] * SHC(10, c) * yz + SH_C3[2] * SHC(11, c) * -2.f * xy +
                              SH_C3[3] * SHC(12, c) * -3.f * 2.f * xz + SH_C3[4] * SHC(13, c) * (-3.f * xx + 4.f * zz - yy) +
                              SH_C3[5] * SHC(14, c) * 2.f * xz + SH_C3[6] * SHC(15, c) * 3.f * (xx - yy);
                    gyv[c] += SH_C3[0] * SHC(9, c) * 3.f * (xx - yy) + SH_C3[1] * SHC(10, c) * xz + SH_C3[2] * SHC(11, c) * (-3.f * yy + 4.f * zz - xx) +
                              SH_C3[3] * SHC(12, c) * -3.f * 2.f * yz + SH_C3[4] * SHC(13, c) * -2.f * xy + SH_C3[5] * SHC(14, c) * -2.f * yz +
                              SH_C3[6] * SHC(15, c) * -3.f * 2.f * xy;
                    gzv[c] += SH_C3[1] * SHC(10, c) * xy + SH_C3[2] * SHC(11, c) * 4.f * 2.f * yz + SH_C3[3] * SHC(12, c) * 3.f * (2.f * zz - xx - yy) +
                              SH_C3[4] * SHC(13, c) * 4.f * 2.f * xz + SH_C3[5] * SHC(14, c) * (xx - yy);
                }
            }
        }
#undef SHC
        if (!child)
            for (int k = 0; k < nco; k++)
                for (int c = 0; c < 3; c++) dL_dsh[(idx * sc->M + k) * 3 + c] = w[k] * dRGB[c];
        const float dd[3] = {gxv[0] * dRGB[0] + gxv[1] * dRGB[1] + gxv[2] * dRGB[2], gyv[0] * dRGB[0] + gyv[1] * dRGB[1] + gyv[2] * dRGB[2],
                             gzv[0] * dRGB[0] + gzv[1] * dRGB[1] + gzv[2] * dRGB[2]};
        const float sum2 = dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2];
        const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmean[0] += ((+sum2 - dir0[0] * dir0[0]) * dd[0] - dir0[1] * dir0[0] * dd[1] - dir0[2] * dir0[0] * dd[2]) * inv32;
        dmean[1] += (-dir0[0] * dir0[1] * dd[0] + (sum2 - dir0[1] * dir0[1]) * dd[1] - dir0[2] * dir0[1] * dd[2]) * inv32;
        dmean[2] += (-dir0[0] * dir0[2] * dd[0] - dir0[1] * dir0[2] * dd[1] + (sum2 - dir0[2] * dir0[2]) * dd[2]) * inv32;
    }
    if (sc->scales && !child) {
        const float* q = sc->rotations + 4 * idx;
        const float r = q[0], x = q[1], y = q[2], z = q[3];
        M3 R = quat_rot(q);
        const float s[3] = {sc->scale_modifier * sc->scales[3 * idx], sc->scale_modifier * sc->scales[3 * idx + 1], sc->scale_modifier * sc->scales[3 * idx + 2]};
        M3 S = m3cols(s[0], 0, 0, 0, s[1], 0, 0, 0, s[2]);
        M3 Mx = m3mul(S, R);
        M3 dSig = m3cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4], 0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
        M3 M2;
        for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) M2.m[c][k] = 2.0f * Mx.m[c][k];
        M3 dM = m3mul(M2, dSig);
        M3 Rt = m3t(R), dMt = m3t(dM);
        for (int k = 0; k < 3; k++) dL_dscale[3 * idx + k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
        for (int k = 0; k < 3; k++) { dMt.m[0][k] *= s[0]; dMt.m[1][k] *= s[1]; dMt.m[2][k] *= s[2]; }
#define D(a, b) dMt.m[a][b]
        dL_drot[4 * idx + 0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
        dL_drot[4 * idx + 1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
        dL_drot[4 * idx + 2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
        dL_drot[4 * idx + 3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
    }
    if (child) {
        const float tt = sc->ts[t_idx];
        dL_dopacity[idx] = 0.f;
        for (int i = 0; i < 3; i++) atomic_addf(&dL_dmean3D[3 * (long)parent + i], (1.0f - tt) * dmean[i]);
    } else {
        dL_dopacity[idx] = dop_out;
        for (int i = 0; i < 3; i++) atomic_addf(&dL_dmean3D[3 * idx + i], dmean[i]);
    }
}

void orc_backward(const orc_scene* sc, const orc_state* st, const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dmean2D,
                  float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D,
                  float* dL_dsh, float* dL_dscale, float* dL_drot) {
    if (st->R > 0) blend_backward(sc, st, dL_dpix, dL_dinvdepth_pix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth);
    const float fy = sc->H / (2.0f * sc->tan_fovy), fx = sc->W / (2.0f * sc->tan_fovx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < sc->P; i++)
        gaussian_backward(sc, st, i, fx, fy, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh,
                          dL_dscale, dL_drot);
}
