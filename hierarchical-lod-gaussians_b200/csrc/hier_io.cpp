// hier_io.cpp — .hier reader/writer (host only).
// File format of HierarchyLoader::load (GH/hierarchy_loader.cpp:26-128) / HierarchyWriter::write
// (GH/hierarchy_writer.cpp:27-119): int P (negative => half-compressed), pos[P,3] f32, rot[P,4], log-scales[P,3],
// opacities[P], SH[P,48] (f32 or IEEE half, round-to-nearest), int N, nodes (7 x i32, or HalfNode = 3 x i32 +
// 4 x i16 {depth,count_children,count_leafs,count_merged}), boxes (8 x f32 or 8 x half).
#include "common.cuh"
#include <cstdio>
#include <cstring>
#include <vector>

namespace {

float half_to_float(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else {  // subnormal half -> normal float
            int e = -1;
            do { e++; man <<= 1; } while ((man & 0x400u) == 0);
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint16_t float_to_half(float f) {  // round to nearest even
    uint32_t x;
    memcpy(&x, &f, 4);
    uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    uint32_t abs = x & 0x7fffffffu;
    if (abs >= 0x7f800000u) return sign | 0x7c00u | (abs > 0x7f800000u ? 0x200u | ((abs >> 13) & 0x3ffu) : 0);
    if (abs >= 0x477ff000u) return sign | 0x7c00u;  // rounds to >= 65520 -> inf
    if (abs < 0x33000001u) return sign;             // < 2^-25 (or exactly 2^-25, ties to even 0)
    int e = (int)(abs >> 23) - 127;
    uint32_t m = (abs & 0x7fffffu) | 0x800000u;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;
    uint32_t hm = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1u);
    uint32_t half_ulp = 1u << (shift - 1);
    uint32_t base = (e < -14) ? hm : (((uint32_t)(e + 15) << 10) + (hm & 0x3ffu));
    if (rem > half_ulp || (rem == half_ulp && (base & 1u))) base++;
    return sign | (uint16_t)base;
}

struct File {
    FILE* f;
    explicit File(const char* name, const char* mode) : f(fopen(name, mode)) {}
    ~File() { if (f) fclose(f); }
    bool rd(void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
    bool wr(const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
};

struct HalfNodeRec {
    int32_t parent, start, start_children;
    int16_t dccc[4];
};
static_assert(sizeof(HalfNodeRec) == 20, "HalfNode layout");

}  // namespace

extern "C" int lodgs_hier_probe(const char* filename, int32_t* P, int32_t* N) {
    if (!filename || !P || !N) { lodgs::set_error("lodgs_hier_probe: null argument"); return LODGS_E_ARG; }
    File in(filename, "rb");
    if (!in.f) { lodgs::set_error("File not found!"); return LODGS_E_IO; }
    int32_t p;
    if (!in.rd(&p, 4)) { lodgs::set_error("lodgs_hier_probe: short read"); return LODGS_E_IO; }
    const bool half = p < 0;
    const long long ap = half ? -(long long)p : p;
    const long long per = half ? (12 + 8 + 6 + 2 + 96) : (12 + 16 + 12 + 4 + 192);
    if (fseek(in.f, (long)(4 + ap * per), SEEK_SET) != 0) { lodgs::set_error("lodgs_hier_probe: seek failed"); return LODGS_E_IO; }
    int32_t n;
    if (!in.rd(&n, 4)) { lodgs::set_error("lodgs_hier_probe: short read"); return LODGS_E_IO; }
    *P = (int32_t)ap;
    *N = n;
    return LODGS_OK;
}

extern "C" int lodgs_hier_load(const char* filename, float* pos, float* shs, float* alpha, float* scales, float* rot, int32_t* nodes,
                               float* boxes) {
    if (!filename) { lodgs::set_error("lodgs_hier_load: null filename"); return LODGS_E_ARG; }
    File in(filename, "rb");
    if (!in.f) { lodgs::set_error("File not found!"); return LODGS_E_IO; }
    int32_t p;
    bool ok = in.rd(&p, 4);
    if (!ok) { lodgs::set_error("lodgs_hier_load: short read"); return LODGS_E_IO; }
    if (p >= 0) {
        size_t P = (size_t)p;
        ok = in.rd(pos, P * 12) && in.rd(rot, P * 16) && in.rd(scales, P * 12) && in.rd(alpha, P * 4) && in.rd(shs, P * 192);
        int32_t n = 0;
        ok = ok && in.rd(&n, 4);
        ok = ok && in.rd(nodes, (size_t)n * 28) && in.rd(boxes, (size_t)n * 32);
    } else {
        size_t P = (size_t)(-(long long)p);
        std::vector<uint16_t> h(P * 48);
        ok = in.rd(pos, P * 12);
        ok = ok && in.rd(h.data(), P * 8);
        if (ok) for (size_t i = 0; i < P * 4; i++) rot[i] = half_to_float(h[i]);
        ok = ok && in.rd(h.data(), P * 6);
        if (ok) for (size_t i = 0; i < P * 3; i++) scales[i] = half_to_float(h[i]);
        ok = ok && in.rd(h.data(), P * 2);
        if (ok) for (size_t i = 0; i < P; i++) alpha[i] = half_to_float(h[i]);
        ok = ok && in.rd(h.data(), P * 96);
        if (ok) for (size_t i = 0; i < P * 48; i++) shs[i] = half_to_float(h[i]);
        int32_t n = 0;
        ok = ok && in.rd(&n, 4);
        if (ok) {
            size_t N = (size_t)n;
            std::vector<HalfNodeRec> hn(N);
            std::vector<uint16_t> hb(N * 8);
            ok = in.rd(hn.data(), N * sizeof(HalfNodeRec)) && in.rd(hb.data(), N * 16);
            if (ok) {
                for (size_t i = 0; i < N; i++) {
                    int32_t* o = nodes + i * 7;
                    o[0] = hn[i].dccc[0]; o[1] = hn[i].parent; o[2] = hn[i].start; o[3] = hn[i].dccc[2];
                    o[4] = hn[i].dccc[3]; o[5] = hn[i].start_children; o[6] = hn[i].dccc[1];
                    for (int j = 0; j < 8; j++) boxes[i * 8 + j] = half_to_float(hb[i * 8 + j]);
                }
            }
        }
    }
    if (!ok) { lodgs::set_error("lodgs_hier_load: short read"); return LODGS_E_IO; }
    return LODGS_OK;
}

extern "C" int lodgs_hier_write(const char* filename, int P, int N, const float* pos, const float* shs, const float* alpha,
                                const float* scales, const float* rot, const int32_t* nodes, const float* boxes, int compressed) {
    if (!filename || P < 0 || N < 0) { lodgs::set_error("lodgs_hier_write: bad argument"); return LODGS_E_ARG; }
    File out(filename, "wb");
    if (!out.f) { lodgs::set_error("File not created!"); return LODGS_E_IO; }
    bool ok;
    const size_t sp = (size_t)P, sn = (size_t)N;
    if (!compressed) {
        int32_t p = P, n = N;
        ok = out.wr(&p, 4) && out.wr(pos, sp * 12) && out.wr(rot, sp * 16) && out.wr(scales, sp * 12) && out.wr(alpha, sp * 4) &&
             out.wr(shs, sp * 192) && out.wr(&n, 4) && out.wr(nodes, sn * 28) && out.wr(boxes, sn * 32);
    } else {
        int32_t p = -P, n = N;
        std::vector<uint16_t> h(sp * 48);
        ok = out.wr(&p, 4) && out.wr(pos, sp * 12);
        for (size_t i = 0; i < sp * 4; i++) h[i] = float_to_half(rot[i]);
        ok = ok && out.wr(h.data(), sp * 8);
        for (size_t i = 0; i < sp * 3; i++) h[i] = float_to_half(scales[i]);
        ok = ok && out.wr(h.data(), sp * 6);
        for (size_t i = 0; i < sp; i++) h[i] = float_to_half(alpha[i]);
        ok = ok && out.wr(h.data(), sp * 2);
        for (size_t i = 0; i < sp * 48; i++) h[i] = float_to_half(shs[i]);
        ok = ok && out.wr(h.data(), sp * 96);
        std::vector<HalfNodeRec> hn(sn);
        std::vector<uint16_t> hb(sn * 8);
        for (size_t i = 0; i < sn; i++) {
            const int32_t* o = nodes + i * 7;
            if (o[0] > 32000 || o[6] > 32000 || o[3] > 32000 || o[4] > 32000) { lodgs::set_error("Would lose information!"); return LODGS_E_ARG; }
            hn[i].parent = o[1]; hn[i].start = o[2]; hn[i].start_children = o[5];
            hn[i].dccc[0] = (int16_t)o[0]; hn[i].dccc[1] = (int16_t)o[6]; hn[i].dccc[2] = (int16_t)o[3]; hn[i].dccc[3] = (int16_t)o[4];
            for (int j = 0; j < 8; j++) hb[i * 8 + j] = float_to_half(boxes[i * 8 + j]);
        }
        ok = ok && out.wr(&n, 4) && out.wr(hn.data(), sn * sizeof(HalfNodeRec)) && out.wr(hb.data(), sn * 16);
    }
    if (!ok) { lodgs::set_error("lodgs_hier_write: short write"); return LODGS_E_IO; }
    return LODGS_OK;
}

// ---- dynamic hierarchy files (the format this fork's creator writes and scene/gaussian_model.py create_from_hier loads).
// HierarchyLoader::loadDynamic (GH/hierarchy_loader.cpp:129-189) / HierarchyWriter::writeDynamic (GH/hierarchy_writer.cpp:122-168):
// int P, int sh_degree, pos[P,3], rot[P,4], log-scales[P,3], opacities[P], SH[P, 3*(deg+1)^2] packed, int N, HierarchyNode[N]
// (6 x i32). The reference loader reads the node count but then loads exactly P nodes (one node per Gaussian); so does this one,
// and the probe reports both numbers.
namespace {
inline int sh_floats(int degree) { return 3 * (degree + 1) * (degree + 1); }
}  // namespace

extern "C" int lodgs_dhier_probe(const char* filename, int32_t* P, int32_t* sh_degree, int32_t* N_in_file) {
    if (!filename || !P || !sh_degree) { lodgs::set_error("lodgs_dhier_probe: null argument"); return LODGS_E_ARG; }
    File in(filename, "rb");
    if (!in.f) { lodgs::set_error("File not found!"); return LODGS_E_IO; }
    int32_t p = 0, deg = 0, n = 0;
    if (!in.rd(&p, 4) || !in.rd(&deg, 4)) { lodgs::set_error("lodgs_dhier_probe: truncated header in %s", filename); return LODGS_E_IO; }
    if (p < 0 || deg < 0 || deg > 3) { lodgs::set_error("lodgs_dhier_probe: bad header in %s (P %d, SH degree %d)", filename, p, deg); return LODGS_E_IO; }
    const long long body = (long long)p * (12 + 16 + 12 + 4 + 4ll * sh_floats(deg));
    if (fseeko(in.f, (off_t)body, SEEK_CUR) != 0 || !in.rd(&n, 4)) {
        lodgs::set_error("lodgs_dhier_probe: truncated file %s", filename);
        return LODGS_E_IO;
    }
    *P = p;
    *sh_degree = deg;
    if (N_in_file) *N_in_file = n;
    return LODGS_OK;
}

extern "C" int lodgs_dhier_load(const char* filename, float* pos, float* shs, float* alpha, float* scales, float* rot, int32_t* nodes) {
    if (!filename) { lodgs::set_error("lodgs_dhier_load: null argument"); return LODGS_E_ARG; }
    File in(filename, "rb");
    if (!in.f) { lodgs::set_error("File not found!"); return LODGS_E_IO; }
    int32_t p = 0, deg = 0, n = 0;
    if (!in.rd(&p, 4) || !in.rd(&deg, 4) || p < 0 || deg < 0 || deg > 3) { lodgs::set_error("lodgs_dhier_load: bad header in %s", filename); return LODGS_E_IO; }
    const size_t P = (size_t)p;
    if (P && (!pos || !shs || !alpha || !scales || !rot || !nodes)) { lodgs::set_error("lodgs_dhier_load: null output"); return LODGS_E_ARG; }
    if (!in.rd(pos, P * 12) || !in.rd(rot, P * 16) || !in.rd(scales, P * 12) || !in.rd(alpha, P * 4) ||
        !in.rd(shs, P * 4 * (size_t)sh_floats(deg)) || !in.rd(&n, 4) || !in.rd(nodes, P * 24)) {
        lodgs::set_error("lodgs_dhier_load: truncated file %s", filename);
        return LODGS_E_IO;
    }
    return LODGS_OK;
}

extern "C" int lodgs_dhier_write(const char* filename, int P, int N, int sh_degree, const float* pos, const float* shs, const float* alpha,
                                 const float* scales, const float* rot, const int32_t* nodes) {
    if (!filename || P < 0 || N < 0 || sh_degree < 0 || sh_degree > 3) { lodgs::set_error("lodgs_dhier_write: bad argument"); return LODGS_E_ARG; }
    if ((P && (!pos || !shs || !alpha || !scales || !rot)) || (N && !nodes)) { lodgs::set_error("lodgs_dhier_write: null input"); return LODGS_E_ARG; }
    File out(filename, "wb");
    if (!out.f) { lodgs::set_error("File not created!"); return LODGS_E_IO; }
    const int32_t p = P, n = N, deg = sh_degree;
    const size_t Ps = (size_t)P;
    if (!out.wr(&p, 4) || !out.wr(&deg, 4) || !out.wr(pos, Ps * 12) || !out.wr(rot, Ps * 16) || !out.wr(scales, Ps * 12) ||
        !out.wr(alpha, Ps * 4) || !out.wr(shs, Ps * 4 * (size_t)sh_floats(deg)) || !out.wr(&n, 4) || !out.wr(nodes, (size_t)N * 24)) {
        lodgs::set_error("lodgs_dhier_write: write failed for %s", filename);
        return LODGS_E_IO;
    }
    return LODGS_OK;
}

// ---- depth-target cut on the host. Replaces Traversal::expandToTarget (GH/traversal.cpp:15-39, bound as expand_to_target,
// torch_interface.cpp:136-142): pre-order walk from node 0; every visited node contributes its leaf Gaussians, a node at
// depth <= target also its merged Gaussians and is not descended. The recursion is unrolled on an explicit stack (children
// pushed in reverse so they pop in id order). Two-call use: out = NULL returns the count only.
extern "C" int lodgs_expand_to_target(const int32_t* nodes, int N, int target, int32_t* out, int64_t capacity, int64_t* count) {
    if (!nodes || N <= 0 || !count) { lodgs::set_error("lodgs_expand_to_target: bad argument"); return LODGS_E_ARG; }
    std::vector<int32_t> stack(1, 0);
    int64_t n = 0;
    while (!stack.empty()) {
        const int32_t id = stack.back();
        stack.pop_back();
        if (id < 0 || id >= N) { lodgs::set_error("lodgs_expand_to_target: node id %d outside [0, %d)", id, N); return LODGS_E_ARG; }
        const int32_t* nd = nodes + 7 * (size_t)id;  // depth parent start count_leafs count_merged start_children count_children
        int64_t take = nd[3];
        const bool stop = nd[0] <= target;
        if (stop) take += nd[4];
        for (int64_t i = 0; i < take; i++, n++)
            if (out && n < capacity) out[n] = nd[2] + (int32_t)i;
        if (!stop)
            for (int32_t c = nd[6] - 1; c >= 0; c--) stack.push_back(nd[5] + c);
    }
    *count = n;
    if (out && n > capacity) { lodgs::set_error("lodgs_expand_to_target: %lld entries, capacity %lld", (long long)n, (long long)capacity); return LODGS_E_WORKSPACE; }
    return LODGS_OK;
}


// ---- host side of the hierarchy creator around lodgs_hier_build: what the reference does with a PLY record before and after the
// tree stages, with the same libm calls, so that a .dhier written from the GPU build carries the creator's bits.
#include <cmath>

// Loader::loadPly<LessRichPoint> (GH/loader.cpp:105-124; record GH/common.h:39-47 = position 3, normal 3, shs 12, opacity 1, scale 3,
// rotation 4) followed by the validity filter of GH/mainHierarchyCreator.cpp:87-149. keep[i] = 1 for the Gaussians the creator keeps.
extern "C" int lodgs_creator_activate(const float* records, int64_t count, float* pos, float* scale, float* rot, float* opacity, float* shs48,
                                      uint8_t* keep) {
    if (count < 0 || (count > 0 && (!records || !pos || !scale || !rot || !opacity || !shs48 || !keep))) {
        lodgs::set_error("lodgs_creator_activate: bad argument");
        return LODGS_E_ARG;
    }
    for (int64_t i = 0; i < count; i++) {
        const float* r = records + i * 26;
        const float o = 1.0f / (1.0f + expf(-r[18]));                       // common.h:20-23 sigmoid
        opacity[i] = o;
        bool ok = !std::isinf(o) && !std::isnan(o) && o != 0.f;
        for (int c = 0; c < 3; c++) {
            pos[3 * i + c] = r[c];
            const float sc = expf(r[19 + c]);
            scale[3 * i + c] = sc;
            ok = ok && !(sc > 10.f) && !std::isnan(sc) && !std::isnan(r[c]);
        }
        float n2 = r[22] * r[22];
        for (int c = 1; c < 4; c++) n2 += r[22 + c] * r[22 + c];
        const float n = sqrtf(n2);
        for (int c = 0; c < 4; c++) {
            const float q = n > 0.f ? r[22 + c] / n : r[22 + c];            // Eigen normalized(): unchanged when the norm is 0
            rot[4 * i + c] = q;
            ok = ok && !std::isnan(q);
        }
        float* sh = shs48 + i * 48;
        for (int c = 0; c < 48; c++) sh[c] = 0.f;
        for (int c = 0; c < 3; c++) { sh[c] = r[6 + c]; ok = ok && !std::isnan(r[6 + c]); }   // a LessRichPoint only carries its DC over
        keep[i] = ok ? 1 : 0;
    }
    return LODGS_OK;
}

// log of the scales as the writer stores them (GH/writer.cpp:119,130 `g.scale.array().log()`): libm logf on the host.
extern "C" int lodgs_logf_host(const float* in, int64_t n, float* out) {
    if (n < 0 || (n > 0 && (!in || !out))) { lodgs::set_error("lodgs_logf_host: bad argument"); return LODGS_E_ARG; }
    for (int64_t i = 0; i < n; i++) out[i] = logf(in[i]);
    return LODGS_OK;
}
