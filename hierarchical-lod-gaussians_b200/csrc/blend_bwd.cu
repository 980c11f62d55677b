// blend_bwd.cu — back-to-front replay of the tile blend, producing per-Gaussian screen-space gradients.
//
// Replaces renderCUDA<3> backward (HR/cuda_rasterizer/backward.cu:498-721), which issues 9-10 global float
// atomics per blended (pixel,Gaussian) pair. Here the ten partial gradients of a Gaussian are first reduced
// across the 32 pixels of a warp with a value-splitting butterfly (12 shuffles instead of 50: every exchange
// halves the number of values a lane still owns), the ten owning lanes add them to a per-batch shared-memory
// accumulator, and after the batch one thread per Gaussian flushes non-zero totals with one global RED per value:
// at most 10 atomics per (tile,Gaussian) instance instead of per pair. Records are gathered from the per-Gaussian
// array through the same async-copy ring as the forward, and the replay starts at the tile's last contributing
// splat instead of the end of its range.
#include "state.cuh"
#include <cstdlib>
#ifndef LODGS_LOD_CUT_DEFAULT
#define LODGS_LOD_CUT_DEFAULT 1
#endif

namespace lodgs {

// LOD-mode alpha cut-off switch (see blend_pair2); LODGS_LOD_CUT=0 in the environment sends every pair through the exponentials (A/B)
__constant__ int c_lod_cut = LODGS_LOD_CUT_DEFAULT;
static int set_lod_cut() {
    const char* e = getenv("LODGS_LOD_CUT");
    const int v = e ? (e[0] != '0') : LODGS_LOD_CUT_DEFAULT;
    return (int)cudaMemcpyToSymbol(c_lod_cut, &v, sizeof(v));
}


constexpr int NV = 10;  // dmean.x dmean.y dconic.xx dconic.xy dconic.yy dopacity dcol.r dcol.g dcol.b dinvdepth

// Q = vertically adjacent pixel PAIRS per thread: Q = 2 is a column of four pixels per thread (a warp covers 16x8 pixels), Q = 1 two pixels
// (a warp covers 16x4). Everything that is paid once per (warp, splat) — the three record loads, the loop, the two votes, the 12-shuffle
// reduction and the accumulate — is shared by twice the pixels with Q = 2, which is what a throughput-bound launch wants. A tile with a very
// long list is LATENCY-bound instead (its CTA walks the list alone while the SMs run dry): there the shorter per-splat chain of Q = 1 wins.
// The default kernel (NT = 64 threads) therefore takes a tile either as ONE CTA with Q = 2 or, for the tiles binning.cu's tile_order_kernel
// marked as long, as TWO CTAs with Q = 1 that each own half of the tile's rows and walk the same list. NT = 128 is the round-1 shape
// (one CTA of 16 x 8 threads per tile, Q = 1), kept for A/B (LODGS_BWD_PX=2). The batch (splats gathered per round, one per thread)
// follows the CTA size.
constexpr int LONG_TILES_MAX = 2048;  // upper bound of the number of long tiles (tile_order_kernel), i.e. of the extra CTAs of a launch
template <int NT>
struct BwdCfg {
    static constexpr int THREADS = NT;
    static constexpr int BATCH = NT;
    // accumulator copies per batch: with 64 threads every warp owns one (plain read-modify-write, summed when the batch is flushed); 128
    // threads keep one copy for their four warps and add to it with shared-memory atomics (a compare-and-swap loop for float)
    static constexpr int COPIES = NT == 64 ? 2 : 1;
    static constexpr int MIN_CTAS = NT == 64 ? 10 : 7;  // register budget asked of ptxas: 96 / 72 per thread
};

template <int NT>
struct BwdSmemT {
    SplatRec recs[2][BwdCfg<NT>::BATCH];
    float acc[2][BwdCfg<NT>::COPIES][BwdCfg<NT>::BATCH * NV];
    uint64_t bar[2];
    unsigned warp_max[BwdCfg<NT>::THREADS / 32];
};

struct BwdArgs {
    const uint2* ranges;
    const uint32_t* tile_order;   // [tiles] + the number of long tiles at [tiles]; NULL = tile-id order
    const SplatRec* recs;
    const uint32_t* point_list;
    int W, H;
    const float* bg_color;
    float bg_mult, null_thr;
    const int32_t* indices;
    const float* final_Ts;
    const uint32_t* n_contrib;
    const float *dL_dpixels, *dL_invdepths;
    float *dL_dmean2D, *dL_dconic2D, *dL_dopacity, *dL_dcolors, *dL_dinvdepths;
};

__device__ __forceinline__ float ex2_ftz(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Which of the ten sums lane `lane` holds after warp_reduce10 (-1: none). A function of the lane only: computed once per thread.
__device__ __forceinline__ int reduce_slot(unsigned lane) {
    const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2;
    const int sub = (b4 ? 2 : 0) + (b2 ? 1 : 0);
    const bool valid = !(lane & 1) && (b8 ? sub <= 1 : sub <= 2);
    return valid ? ((b16 ? 5 : 0) + (b8 ? 3 : 0) + sub) : -1;
}

// Sum v[0..9] over the warp with a value-splitting butterfly: every exchange halves the number of values a lane still owns
// (10 -> 5 -> 3 -> 2 -> 1 -> 1), 12 shuffles in all. On return lane l holds the total of value reduce_slot(l).
__device__ __forceinline__ float warp_reduce10(const float (&v)[NV], unsigned lane) {
    const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2;
    float w[6], u[4], t[2];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const float send = b16 ? v[k] : v[k + 5];
        const float keep = b16 ? v[k + 5] : v[k];
        w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    w[5] = 0.f;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float send = b8 ? w[k] : w[k + 3];
        const float keep = b8 ? w[k + 3] : w[k];
        u[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    u[3] = 0.f;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const float send = b4 ? u[k] : u[k + 2];
        const float keep = b4 ? u[k + 2] : u[k];
        t[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    {
        const float send = b2 ? t[0] : t[1];
        const float keep = b2 ? t[1] : t[0];
        t[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    t[0] += __shfl_xor_sync(0xffffffffu, t[0], 1);
    return t[0];
}

// ---- packed FP32 (sm_100 FFMA2 / FMUL2 / FADD2): the thread's two pixels ride in the halves of one 64-bit register pair, so
// the arithmetic of a (splat, pixel pair) costs one issue slot instead of two (the kernel is issue-bound, not FP-pipe-bound).
// Splat constants enter as broadcast operands (mov.b64 {x,x} folds into the instruction). ptxas contracts mul+add pairs of
// packed operations even with .rn; harmless here (gradients are held to 1e-3, not bit-exact) — the FORWARD kernel, whose
// rounding is pinned, restricts itself to the packed forms ptxas cannot re-contract (see blend_fwd.cu).
// Replay state of the thread's Q pixel pairs (pair h: lo = row 2h of the thread's column, hi = the one below it).
template <int Q>
struct BwdPix {
    // Tfn = T_final * (-bg_mult * bg . dL_dpixel).  S = dL_dpixel . (colour accumulated behind the current splat, inverse depth included
    // with DEPTH), lcg = dL_dpixel . (colour of the previously processed splat): backward.cu:640-668 keeps the accumulated colour per
    // channel and forms sum_c (c_c - accum_c) * dL_dpixel_c; the recurrence accum' = la * last_c + (1 - la) * accum is linear, so its dot
    // product with the (per-pixel constant) dL_dpixel obeys the same recurrence and one value per pixel replaces four.
    f2 T[Q], Tfn[Q], last_alpha[Q], S[Q], lcg[Q], g0[Q], g1[Q], g2[Q], gd[Q];
    f2 npy;  // -(y) of pair 0; pair h lies 2h rows further down
    float px;
    unsigned last[2 * Q];
};

// One splat against the thread's 2Q pixels (backward.cu:600-716). A pixel the splat does not contribute to (behind its last contributor,
// power > 0, below the cut-off or alpha < 1/255) takes part as a ZERO-ALPHA splat: T / (1 - 0) = T, the running colour
// recurrence ar' = last_alpha*lc + (1-last_alpha)*ar passes through it unchanged (the next splat sees last_alpha = 0), and
// its k is forced to 0 — so no per-pixel branches and no per-pixel copies of the previous colour are needed. The same holds
// for a whole LANE: the two early-outs are warp votes (the return value is warp-uniform), a lane whose pixels all skip the
// splat produces ten exact zeros, so there is no divergence inside the loop and nothing to clear before the reduction.
template <bool LOD, bool DEPTH, int Q>
__device__ __forceinline__ bool bwd_splat(BwdPix<Q>& p, const unsigned pos, const float4 q0, const float4 q1, const float4 q2,
                                          const float null_thr, float (&v)[NV]) {
    // log2(e) * power = dy * (dy * A + B) + C with A = -0.5*l2e*con.yy, B = -l2e*con.xy*dx, C = -0.5*l2e*con.xx*dx*dx: two packed FMAs per
    // pixel pair; the exponential is ex2 of it, the cut-offs are compared in the same units
    constexpr float L2E = 1.4426950408889634f;
    const float dx = q0.x - p.px;
    const float cA = (-0.5f * L2E) * q1.x, cB = (-L2E * q0.w) * dx, cC = ((-0.5f * L2E) * q0.z * dx) * dx;
    float cut;
    if (!LOD) cut = q2.z * L2E;  // certainly below the 1/255 cut (cut-off stored by the preprocess kernel)
    else      // the same cut rebuilt per splat (LOD alpha <= opacity*exp(power) for t in [0,1], kids >= 1; see blend_fwd.cu)
        cut = (c_lod_cut && q2.z >= 0.0f && q2.z <= 1.0f && q2.w > 0.0f && q2.w <= 1.0f) ? -__log2f(255.0f * q1.y) - 0.03f : -__builtin_huge_valf();
    f2 dy[Q], power[Q];
    bool ok[2 * Q];
    bool any = false;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        dy[h] = add2(bc(q0.y - 2.0f * h), p.npy);
        power[h] = fma2(dy[h], fma2(dy[h], bc(cA), bc(cB)), bc(cC));
        const float pw0 = lo(power[h]), pw1 = hi(power[h]);
        ok[2 * h] = pos < p.last[2 * h] && !(pw0 > 0.0f) && !(pw0 < cut);
        ok[2 * h + 1] = pos < p.last[2 * h + 1] && !(pw1 > 0.0f) && !(pw1 < cut);
        any = any || ok[2 * h] || ok[2 * h + 1];
    }
    if (!__any_sync(0xffffffffu, any)) return false;  // warp-uniform: lanes without a contribution ride along with alpha 0
    // ex2.approx.ftz on log2(e) * power instead of the forward's full-range expf: alpha moves by ~1e-7 relative, far inside the
    // gradient budget; flushing results below 2^-126 is immaterial (alpha < 1/255 is skipped a few lines further down)
    f2 G[Q];
    float ta[2 * Q], ma[2 * Q], al[2 * Q];
    float t = 1.f, frac = 1.f;
    if (LOD) { t = q2.z; frac = q2.w; }
    any = false;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        G[h] = pk(ex2_ftz(lo(power[h])), ex2_ftz(hi(power[h])));
        const f2 ta2 = mul2(bc(q1.y), G[h]);
        ta[2 * h] = lo(ta2);
        ta[2 * h + 1] = hi(ta2);
#pragma unroll
        for (int e = 2 * h; e < 2 * h + 2; e++) {
            ma[e] = fminf(0.99f, ta[e]);
            al[e] = ma[e];
            if (LOD) al[e] = fmaf(ma[e], t, (1.0f - t) * (1.0f - __powf(1.0f - ma[e], frac)));
            ok[e] = ok[e] && !(al[e] < 1.0f / 255.0f);
            any = any || ok[e];
        }
    }
    if (!__any_sync(0xffffffffu, any)) return false;
    f2 A[NV], M[3];
#pragma unroll
    for (int h = 0; h < Q; h++) {
        const bool ok0 = ok[2 * h], ok1 = ok[2 * h + 1];
        const f2 alpha = pk(ok0 ? al[2 * h] : 0.f, ok1 ? al[2 * h + 1] : 0.f);
        const f2 oma = fma2(alpha, bc(-1.0f), bc(1.0f));
        const f2 rcp1ma = pk(rcp_approx(lo(oma)), rcp_approx(hi(oma)));  // 1-alpha in [0.01,1]; rcp.approx(1) == 1 exactly
        p.T[h] = mul2(p.T[h], rcp1ma);
        const f2 weight = mul2(alpha, p.T[h]);
        f2 cg = mul2(p.g0[h], bc(q1.z));
        cg = fma2(p.g1[h], bc(q1.w), cg);
        cg = fma2(p.g2[h], bc(q2.x), cg);
        A[6] = h ? fma2(weight, p.g0[h], A[6]) : mul2(weight, p.g0[h]);
        A[7] = h ? fma2(weight, p.g1[h], A[7]) : mul2(weight, p.g1[h]);
        A[8] = h ? fma2(weight, p.g2[h], A[8]) : mul2(weight, p.g2[h]);
        if (DEPTH) {
            cg = fma2(p.gd[h], bc(q2.y), cg);
            A[9] = h ? fma2(weight, p.gd[h], A[9]) : mul2(weight, p.gd[h]);
        }
        p.S[h] = fma2(p.last_alpha[h], fma2(p.S[h], bc(-1.0f), p.lcg[h]), p.S[h]);
        p.lcg[h] = cg;
        f2 dL_dalpha = fma2(p.S[h], bc(-1.0f), cg);
        p.last_alpha[h] = alpha;
        dL_dalpha = fma2(p.Tfn[h], rcp1ma, mul2(dL_dalpha, p.T[h]));
        f2 k = mul2(G[h], dL_dalpha);
        const float k0 = (ok0 && !(ta[2 * h] > null_thr)) ? lo(k) : 0.f;  // 0.99 clamp (backward.cu:620,687); +inf threshold with AR's convention
        const float k1 = (ok1 && !(ta[2 * h + 1] > null_thr)) ? hi(k) : 0.f;
        k = pk(k0, k1);
        // Per pair only the moments of k = G * dL_dalpha over (dx, dy) are accumulated; the splat-constant factors (conic,
        // opacity, 0.5*W, 0.5*H, -0.5) of backward.cu:676-702 are applied once per (tile, splat) when the sums are flushed.
        // dx is the same for all pixels of the thread (one column): only sum k, sum k*dy and sum k*dy*dy are accumulated per pixel
        const f2 kdy = mul2(k, dy[h]);
        M[0] = h ? add2(M[0], k) : k;
        M[1] = h ? add2(M[1], kdy) : kdy;
        M[2] = h ? fma2(kdy, dy[h], M[2]) : mul2(kdy, dy[h]);
        if (LOD) {
            const float w0 = t - __powf(1.0f - ma[2 * h], frac - 1.0f) * (t - 1.0f) * frac;
            const float w1 = t - __powf(1.0f - ma[2 * h + 1], frac - 1.0f) * (t - 1.0f) * frac;
            A[5] = h ? fma2(pk(w0, w1), k, A[5]) : mul2(pk(w0, w1), k);
        }
    }
    // the pairs were added lane-wise first: one horizontal add per value
    const float s0 = hsum(M[0]), s1 = hsum(M[1]);
    v[0] = dx * s0;
    v[1] = s1;
    v[2] = dx * v[0];
    v[3] = dx * s1;
    v[4] = hsum(M[2]);
    v[5] = LOD ? hsum(A[5]) : s0;
    v[6] = hsum(A[6]);
    v[7] = hsum(A[7]);
    v[8] = hsum(A[8]);
    v[9] = DEPTH ? hsum(A[9]) : 0.f;
    return true;
}

// One tile (rows y_off .. y_off + 2*Q*NT/16 - 1 of it) replayed by the NT threads of this CTA.
template <bool LOD, bool DEPTH, int Q, int NT>
__device__ __forceinline__ void bwd_tile(const BwdArgs& a, BwdSmemT<NT>& sm, const unsigned tile, const unsigned y_off) {
    constexpr int BWD_THREADS = BwdCfg<NT>::THREADS, BWD_BATCH = BwdCfg<NT>::BATCH, COPIES = BwdCfg<NT>::COPIES;
    const uint2* __restrict__ ranges = a.ranges;
    const SplatRec* __restrict__ recs = a.recs;
    const uint32_t* __restrict__ point_list = a.point_list;
    const int W = a.W, H = a.H;
    const float* __restrict__ bg_color = a.bg_color;
    const float bg_mult = a.bg_mult, null_thr = a.null_thr;
    const int32_t* __restrict__ indices = a.indices;
    const float* __restrict__ final_Ts = a.final_Ts;
    const uint32_t* __restrict__ n_contrib = a.n_contrib;
    const float* __restrict__ dL_dpixels = a.dL_dpixels;
    const float* __restrict__ dL_invdepths = a.dL_invdepths;
    float* __restrict__ dL_dmean2D = a.dL_dmean2D;
    float* __restrict__ dL_dconic2D = a.dL_dconic2D;
    float* __restrict__ dL_dopacity = a.dL_dopacity;
    float* __restrict__ dL_dcolors = a.dL_dcolors;
    float* __restrict__ dL_dinvdepths = a.dL_dinvdepths;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, tx = tid & 15u, ty = tid >> 4;
    const unsigned horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const unsigned tile_x = tile % horizontal_blocks, tile_y = tile / horizontal_blocks;
    const unsigned pix_x = tile_x * TILE_X + tx, pix_y0 = tile_y * TILE_Y + y_off + 2u * Q * ty;
    const uint2 range = ranges[tile];
    const size_t HW = (size_t)H * W;
    const float bg0 = bg_color[0], bg1 = bg_color[1], bg2 = bg_color[2];

    BwdPix<Q> p2;
    unsigned tmax = 0;
    {
        p2.px = (float)pix_x;
        p2.npy = pk(-(float)pix_y0, -(float)(pix_y0 + 1u));
#pragma unroll
        for (int q = 0; q < Q; q++) {
            float Tf[2], g0[2], g1[2], g2[2], gd[2];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const unsigned y = pix_y0 + 2u * q + h;
                const bool inside = pix_x < (unsigned)W && y < (unsigned)H;
                const uint32_t pix_id = W * y + pix_x;
                Tf[h] = inside ? final_Ts[pix_id] : 0.f;
                p2.last[2 * q + h] = inside ? n_contrib[pix_id] : 0u;
                tmax = max(tmax, p2.last[2 * q + h]);
                g0[h] = inside ? dL_dpixels[0 * HW + pix_id] : 0.f;
                g1[h] = inside ? dL_dpixels[1 * HW + pix_id] : 0.f;
                g2[h] = inside ? dL_dpixels[2 * HW + pix_id] : 0.f;
                gd[h] = (DEPTH && inside) ? dL_invdepths[pix_id] : 0.f;
            }
            p2.T[q] = pk(Tf[0], Tf[1]);
            p2.last_alpha[q] = p2.S[q] = p2.lcg[q] = bc(0.f);
            p2.g0[q] = pk(g0[0], g0[1]);
            p2.g1[q] = pk(g1[0], g1[1]);
            p2.g2[q] = pk(g2[0], g2[1]);
            p2.gd[q] = pk(gd[0], gd[1]);
            p2.Tfn[q] = pk(Tf[0] * (-bg_mult * (bg0 * g0[0] + bg1 * g1[0] + bg2 * g2[0])), Tf[1] * (-bg_mult * (bg0 * g0[1] + bg1 * g1[1] + bg2 * g2[1])));
        }
    }

    // the replay only has to cover splats up to the tile's deepest contributor
    const unsigned wmax = __reduce_max_sync(0xffffffffu, tmax);
    if (lane == 0) sm.warp_max[warp] = wmax;
    if (tid == 0) {
        mbar_init(&sm.bar[0], BWD_THREADS);
        mbar_init(&sm.bar[1], BWD_THREADS);
        mbar_fence_init();
    }
    for (int i = tid; i < 2 * COPIES * BWD_BATCH * NV; i += BWD_THREADS) (&sm.acc[0][0][0])[i] = 0.f;
    __syncthreads();
    unsigned L = 0;
#pragma unroll
    for (int w = 0; w < BWD_THREADS / 32; w++) L = max(L, sm.warp_max[w]);
    L = min(L, range.y - range.x);
    const int rounds = (int)((L + BWD_BATCH - 1) / BWD_BATCH);

    // batch b covers sorted positions [lo_b, hi_b) with hi_b = L - b*BATCH, walked from hi_b-1 down to lo_b; thread tid
    // gathers the record of position lo_b + tid. The ids are fetched one batch ahead of the copies.
    constexpr uint32_t NO_ID = 0xffffffffu;
    auto load_id = [&](int b) -> uint32_t {
        const unsigned hi = L - (unsigned)b * BWD_BATCH;
        const unsigned lo = hi > BWD_BATCH ? hi - BWD_BATCH : 0u;
        return lo + tid < hi ? __ldg(point_list + range.x + lo + tid) : NO_ID;
    };
    auto issue = [&](int b, uint32_t id) {
        const int s = b & 1;
        if (id != NO_ID) {
            const float4* src = reinterpret_cast<const float4*>(recs + id);
            float4* dst = reinterpret_cast<float4*>(&sm.recs[s][tid]);
            cp_async16(dst, src);
            cp_async16(dst + 1, src + 1);
            cp_async16(dst + 2, src + 2);
        }
        cp_async_arrive(&sm.bar[s]);  // counts this thread in once its copies have landed
    };
    uint32_t id_cur = NO_ID, id_next = NO_ID;
    if (rounds > 0) {
        id_cur = load_id(0);
        issue(0, id_cur);
    }
    if (rounds > 1) id_next = load_id(1);

    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;

    // the lane id and the lane's slot in the reduction are pinned to registers: ptxas otherwise re-derives both from %tid inside the loop
    unsigned lane_r = lane;
    int my_slot = reduce_slot(lane);
    if (!DEPTH && my_slot == 9) my_slot = -1;
    asm volatile("" : "+r"(lane_r), "+r"(my_slot));

    int pending = -1;
    int my_gid = -1;
    float4 my_con = {0.f, 0.f, 0.f, 0.f};  // conic xx, xy, yy and opacity of the splat this thread flushes
    auto flush = [&]() {
        if (pending < 0) return;
        if (my_gid >= 0) {
            float v[NV];
            bool any = false;
#pragma unroll
            for (int k = 0; k < NV; k += 2) {
                float2 t2 = make_float2(0.f, 0.f);
#pragma unroll
                for (int c = 0; c < COPIES; c++) {
                    float2* a2 = reinterpret_cast<float2*>(&sm.acc[pending & 1][c][tid * NV + k]);
                    const float2 x = *a2;
                    if (x.x != 0.f || x.y != 0.f) *a2 = make_float2(0.f, 0.f);
                    t2.x += x.x;
                    t2.y += x.y;
                }
                v[k] = t2.x;
                v[k + 1] = t2.y;
                any |= (t2.x != 0.f) | (t2.y != 0.f);
            }
            if (any) {
                const float mo = -my_con.w, ho = -0.5f * my_con.w;
                atomicAdd(&dL_dmean2D[3 * (size_t)my_gid + 0], mo * ddelx_dx * (my_con.x * v[0] + my_con.y * v[1]));
                atomicAdd(&dL_dmean2D[3 * (size_t)my_gid + 1], mo * ddely_dy * (my_con.z * v[1] + my_con.y * v[0]));
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 0], ho * v[2]);
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 1], ho * v[3]);
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 3], ho * v[4]);
                atomicAdd(&dL_dopacity[my_gid], v[5]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 0], v[6]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 1], v[7]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 2], v[8]);
                if (DEPTH) atomicAdd(&dL_dinvdepths[my_gid], v[9]);
            }
        }
        pending = -1;
    };

    for (int b = 0; b < rounds; b++) {
        __syncthreads();  // all warps have left batch b-1: its accumulator is complete and its ring slot is free
        flush();
        if (b + 1 < rounds) issue(b + 1, id_next);
        const uint32_t id = id_cur;
        id_cur = id_next;
        if (b + 2 < rounds) id_next = load_id(b + 2);
        const int s = b & 1;
        mbar_wait(&sm.bar[s], (unsigned)(b >> 1) & 1u);
        const unsigned hi = L - (unsigned)b * BWD_BATCH;
        const unsigned lo = hi > BWD_BATCH ? hi - BWD_BATCH : 0u;
        const int cnt = (int)(hi - lo);
        {
            my_gid = -1;
            if ((int)tid < cnt) {
                my_gid = indices ? __ldg(indices + id) : (int)id;
                const SplatRec& r = sm.recs[s][tid];  // read now: the ring slot is refilled before this batch is flushed
                my_con = make_float4(r.conx, r.cony, r.conz, r.opac);
            }
            pending = b;
        }
        if (lo >= wmax) continue;  // nothing in this batch reaches any pixel of this warp
        const int jtop = (int)min((unsigned)cnt, wmax - lo) - 1;
        const float4* __restrict__ rp = reinterpret_cast<const float4*>(&sm.recs[s][0]) + 3 * jtop;
        float* accp = &sm.acc[s][COPIES == 1 ? 0 : warp][jtop * NV];
        unsigned pos = lo + (unsigned)jtop;  // 0-based position in the tile's list == contributor index - 1
        _Pragma("unroll 2") for (int j = jtop; j >= 0; j--, rp -= 3, accp -= NV, pos--) {
            float v[NV];
            const float4 q0 = rp[0], q1 = rp[1], q2 = rp[2];
            if (!bwd_splat<LOD, DEPTH, Q>(p2, pos, q0, q1, q2, null_thr, v)) continue;  // the same answer in every lane
            const float total = warp_reduce10(v, lane_r);
            if (my_slot >= 0 && total != 0.f) {
                if (COPIES == 1) atomicAdd(accp + my_slot, total);
                else accp[my_slot] += total;  // this warp's copy: lanes own distinct slots, splats are taken one after the other
            }
        }
    }
    __syncthreads();
    flush();
}

template <bool LOD, bool DEPTH, int NT>
__global__ void __launch_bounds__(BwdCfg<NT>::THREADS, BwdCfg<NT>::MIN_CTAS) blend_bwd_kernel(const BwdArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BwdSmemT<NT>& sm = *reinterpret_cast<BwdSmemT<NT>*>(smem_raw);
    const unsigned ntiles = ((a.W + TILE_X - 1) / TILE_X) * ((a.H + TILE_Y - 1) / TILE_Y);
    const unsigned cta = blockIdx.x;
    if (NT == 64) {
        // CTAs [0, 2 n_long): the long tiles, two CTAs each (upper / lower eight rows, two pixels per thread); then one CTA per remaining tile
        const unsigned n_long = a.tile_order ? __ldg(a.tile_order + ntiles) : 0u;
        if (cta < 2u * n_long) {
            bwd_tile<LOD, DEPTH, 1, NT>(a, sm, __ldg(a.tile_order + (cta >> 1)), 8u * (cta & 1u));
            return;
        }
        const unsigned i = cta - n_long;
        if (i >= ntiles) return;
        bwd_tile<LOD, DEPTH, 2, NT>(a, sm, a.tile_order ? __ldg(a.tile_order + i) : i, 0u);   // longest lists first (binning.cu)
    } else {
        if (cta >= ntiles) return;
        bwd_tile<LOD, DEPTH, 1, NT>(a, sm, a.tile_order ? __ldg(a.tile_order + cta) : cta, 0u);
    }
}

template <bool LOD, bool DEPTH, int NT>
static int launch_bwd_variant(unsigned grid, cudaStream_t s, const BwdArgs& a) {
    static std::atomic<bool> attr_set[MAX_DEVICES];
    const int dslot = device_slot();
    if (!attr_set[dslot].load(std::memory_order_acquire)) {
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<LOD, DEPTH, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemT<NT>)));
        attr_set[dslot].store(true, std::memory_order_release);
    }
    blend_bwd_kernel<LOD, DEPTH, NT><<<grid, NT, sizeof(BwdSmemT<NT>), s>>>(a);
    return LODGS_OK;
}

int launch_blend_bwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                     const float* bg, float bg_mult, float null_thr, const int32_t* indices, const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, cudaStream_t s) {
    const unsigned ntiles = (unsigned)((W + TILE_X - 1) / TILE_X) * (unsigned)((H + TILE_Y - 1) / TILE_Y);
    static const bool tile_id_order = getenv("LODGS_BLEND_TILE_ID_ORDER") != nullptr;  // A/B: CTAs in tile-id order
    static const int px = getenv("LODGS_BWD_PX") ? atoi(getenv("LODGS_BWD_PX")) : 4;   // A/B: 2 = 128-thread CTAs with two pixels per thread
    static const int lod_cut_rc = set_lod_cut();
    if (lod_cut_rc) { set_error("blend: setting the LOD cut-off switch failed"); return LODGS_E_CUDA; }
    BwdArgs a;
    a.ranges = img.ranges; a.tile_order = tile_id_order ? nullptr : img.tile_order; a.recs = recs; a.point_list = bin.point_list; a.W = W; a.H = H;
    a.bg_color = bg; a.bg_mult = bg_mult; a.null_thr = null_thr; a.indices = indices; a.final_Ts = img.final_T; a.n_contrib = img.n_contrib;
    a.dL_dpixels = dL_dpix; a.dL_invdepths = dL_dinvdepth_pix; a.dL_dmean2D = dL_dmean2D; a.dL_dconic2D = dL_dconic; a.dL_dopacity = dL_dopacity;
    a.dL_dcolors = dL_dcolor; a.dL_dinvdepths = dL_dinvdepth;
    int rc;
    if (px == 2) {
        if (lod && depth) rc = launch_bwd_variant<true, true, 128>(ntiles, s, a);
        else if (lod) rc = launch_bwd_variant<true, false, 128>(ntiles, s, a);
        else if (depth) rc = launch_bwd_variant<false, true, 128>(ntiles, s, a);
        else rc = launch_bwd_variant<false, false, 128>(ntiles, s, a);
    } else {
        const unsigned grid = ntiles + (a.tile_order ? (unsigned)LONG_TILES_MAX : 0u);   // the CTAs beyond tiles + n_long exit at once
        if (lod && depth) rc = launch_bwd_variant<true, true, 64>(grid, s, a);
        else if (lod) rc = launch_bwd_variant<true, false, 64>(grid, s, a);
        else if (depth) rc = launch_bwd_variant<false, true, 64>(grid, s, a);
        else rc = launch_bwd_variant<false, false, 64>(grid, s, a);
    }
    if (rc) return rc;
    LODGS_LAUNCHED("blend_bwd_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
