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


constexpr int BWD_BATCH = 128;
constexpr int BWD_THREADS = 128;  // 16 x 8 threads, two vertically adjacent pixels each (a warp covers a 16x4 block)
constexpr int NV = 10;  // dmean.x dmean.y dconic.xx dconic.xy dconic.yy dopacity dcol.r dcol.g dcol.b dinvdepth

struct BwdSmem {
    SplatRec recs[2][BWD_BATCH];
    float acc[2][BWD_BATCH * NV];
    uint64_t bar[2];
    unsigned warp_max[BWD_THREADS / 32];
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

// Sum v[0..9] over the warp. On return lanes with even lane id and a valid slot hold the total of value
// `slot` (returned, -1 otherwise) in the return value `out`.
__device__ __forceinline__ float warp_reduce10(const float (&v)[NV], unsigned lane, int& slot) {
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
    const int sub = (b4 ? 2 : 0) + (b2 ? 1 : 0);
    const bool valid = !(lane & 1) && (b8 ? sub <= 1 : sub <= 2);
    slot = valid ? ((b16 ? 5 : 0) + (b8 ? 3 : 0) + sub) : -1;
    return t[0];
}

// ---- packed FP32 (sm_100 FFMA2 / FMUL2 / FADD2): the thread's two pixels ride in the halves of one 64-bit register pair, so
// the arithmetic of a (splat, pixel pair) costs one issue slot instead of two (the kernel is issue-bound, not FP-pipe-bound).
// Splat constants enter as broadcast operands (mov.b64 {x,x} folds into the instruction). ptxas contracts mul+add pairs of
// packed operations even with .rn; harmless here (gradients are held to 1e-3, not bit-exact) — the FORWARD kernel, whose
// rounding is pinned, restricts itself to the packed forms ptxas cannot re-contract (see blend_fwd.cu).
// Replay state of the thread's two pixels (lo = upper pixel, hi = the one below it).
struct BwdPix2 {
    f2 npy, T, T_final, last_alpha, ar0, ar1, ar2, ar_invd, g0, g1, g2, gd, nbg_dot;
    float px, lc0, lc1, lc2, last_invd;  // the colour / inverse depth of the previously processed splat is the same for both pixels
    unsigned last0, last1;
};

// One splat against both pixels (backward.cu:600-716). A pixel the splat does not contribute to (behind its last contributor,
// power > 0, below the cut-off or alpha < 1/255) takes part as a ZERO-ALPHA splat: T / (1 - 0) = T, the running colour
// recurrence ar' = last_alpha*lc + (1-last_alpha)*ar passes through it unchanged (the next splat sees last_alpha = 0), and
// its k is forced to 0 — so no per-pixel branches and no per-pixel copies of the previous colour are needed. The same holds
// for a whole LANE: the two early-outs are warp votes (the return value is warp-uniform), a lane whose pixels both skip the
// splat produces ten exact zeros, so there is no divergence inside the loop and nothing to clear before the reduction.
template <bool LOD, bool DEPTH>
__device__ __forceinline__ bool bwd_pair2(BwdPix2& p, const unsigned pos, const float4 q0, const float4 q1, const float4 q2,
                                          const float null_thr, float (&v)[NV]) {
    const float dx = q0.x - p.px;
    const f2 dy = add2(bc(q0.y), p.npy);
    const f2 quad = fma2(dy, mul2(dy, bc(q1.x)), bc(dx * (dx * q0.z)));
    const f2 power = fma2(dy, bc(-(dx * q0.w)), mul2(quad, bc(-0.5f)));
    const float pw0 = lo(power), pw1 = hi(power);
    bool ok0 = pos < p.last0 && !(pw0 > 0.0f), ok1 = pos < p.last1 && !(pw1 > 0.0f);
    if (!LOD) {  // certainly below the 1/255 cut (cut-off stored by the preprocess kernel)
        ok0 = ok0 && !(pw0 < q2.z);
        ok1 = ok1 && !(pw1 < q2.z);
    } else {     // the same cut rebuilt per splat (LOD alpha <= opacity*exp(power) for t in [0,1], kids >= 1; see blend_fwd.cu)
        const float cut = (c_lod_cut && q2.z >= 0.0f && q2.z <= 1.0f && q2.w > 0.0f && q2.w <= 1.0f) ? -__logf(255.0f * q1.y) - 0.02f : -__builtin_huge_valf();
        ok0 = ok0 && !(pw0 < cut);
        ok1 = ok1 && !(pw1 < cut);
    }
    if (!__any_sync(0xffffffffu, ok0 || ok1)) return false;  // warp-uniform: lanes without a contribution ride along with alpha 0
    // ex2.approx.ftz on power * log2(e) instead of the forward's full-range expf: alpha moves by ~1e-7 relative, far inside the
    // gradient budget; flushing results below 2^-126 is immaterial (alpha < 1/255 is skipped two lines further down)
    const f2 pl = mul2(power, bc(1.4426950408889634f));
    const f2 G = pk(ex2_ftz(lo(pl)), ex2_ftz(hi(pl)));
    const f2 ta = mul2(bc(q1.y), G);
    const float ta0 = lo(ta), ta1 = hi(ta);
    const float ma0 = fminf(0.99f, ta0), ma1 = fminf(0.99f, ta1);
    float a0 = ma0, a1 = ma1, t = 1.f, frac = 1.f;
    if (LOD) {
        t = q2.z;
        frac = q2.w;
        a0 = fmaf(ma0, t, (1.0f - t) * (1.0f - __powf(1.0f - ma0, frac)));
        a1 = fmaf(ma1, t, (1.0f - t) * (1.0f - __powf(1.0f - ma1, frac)));
    }
    ok0 = ok0 && !(a0 < 1.0f / 255.0f);
    ok1 = ok1 && !(a1 < 1.0f / 255.0f);
    if (!__any_sync(0xffffffffu, ok0 || ok1)) return false;
    const f2 alpha = pk(ok0 ? a0 : 0.f, ok1 ? a1 : 0.f);
    const f2 oma = fma2(alpha, bc(-1.0f), bc(1.0f));
    const f2 rcp1ma = pk(rcp_approx(lo(oma)), rcp_approx(hi(oma)));  // 1-alpha in [0.01,1]; rcp.approx(1) == 1 exactly
    p.T = mul2(p.T, rcp1ma);
    const f2 weight = mul2(alpha, p.T);
    const f2 oml = fma2(p.last_alpha, bc(-1.0f), bc(1.0f));
    p.ar0 = fma2(p.last_alpha, bc(p.lc0), mul2(oml, p.ar0));
    p.ar1 = fma2(p.last_alpha, bc(p.lc1), mul2(oml, p.ar1));
    p.ar2 = fma2(p.last_alpha, bc(p.lc2), mul2(oml, p.ar2));
    p.lc0 = q1.z; p.lc1 = q1.w; p.lc2 = q2.x;
    f2 dL_dalpha = mul2(fma2(p.ar0, bc(-1.0f), bc(q1.z)), p.g0);
    dL_dalpha = fma2(fma2(p.ar1, bc(-1.0f), bc(q1.w)), p.g1, dL_dalpha);
    dL_dalpha = fma2(fma2(p.ar2, bc(-1.0f), bc(q2.x)), p.g2, dL_dalpha);
    v[6] = hsum(mul2(weight, p.g0));
    v[7] = hsum(mul2(weight, p.g1));
    v[8] = hsum(mul2(weight, p.g2));
    if (DEPTH) {
        p.ar_invd = fma2(p.last_alpha, bc(p.last_invd), mul2(oml, p.ar_invd));
        p.last_invd = q2.y;
        dL_dalpha = fma2(fma2(p.ar_invd, bc(-1.0f), bc(q2.y)), p.gd, dL_dalpha);
        v[9] = hsum(mul2(weight, p.gd));
    }
    p.last_alpha = alpha;
    dL_dalpha = fma2(mul2(p.T_final, rcp1ma), p.nbg_dot, mul2(dL_dalpha, p.T));
    f2 k = mul2(G, dL_dalpha);
    const float k0 = (ok0 && !(ta0 > null_thr)) ? lo(k) : 0.f;  // 0.99 clamp (backward.cu:620,687); +inf threshold with AR's convention
    const float k1 = (ok1 && !(ta1 > null_thr)) ? hi(k) : 0.f;
    k = pk(k0, k1);
    // Per pair only the moments of k = G * dL_dalpha over (dx, dy) are accumulated; the splat-constant factors (conic,
    // opacity, 0.5*W, 0.5*H, -0.5) of backward.cu:676-702 are applied once per (tile, splat) when the sums are flushed.
    const f2 kdx = mul2(k, bc(dx)), kdy = mul2(k, dy);
    v[0] = hsum(kdx);
    v[1] = hsum(kdy);
    v[2] = hsum(mul2(kdx, bc(dx)));
    v[3] = hsum(mul2(kdx, dy));
    v[4] = hsum(mul2(kdy, dy));
    if (LOD) {
        const float m0 = t - __powf(1.0f - ma0, frac - 1.0f) * (t - 1.0f) * frac;
        const float m1 = t - __powf(1.0f - ma1, frac - 1.0f) * (t - 1.0f) * frac;
        v[5] = fmaf(m0, k0, m1 * k1);
    } else {
        v[5] = k0 + k1;
    }
    return true;
}

template <bool LOD, bool DEPTH>
__global__ void __launch_bounds__(BWD_THREADS) blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const SplatRec* __restrict__ recs,
                                                               const uint32_t* __restrict__ point_list, int W, int H,
                                                               const float* __restrict__ bg_color, const float bg_mult,
                                                               const float null_thr, const int32_t* __restrict__ indices,
                                                               const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                                                               const float* __restrict__ dL_dpixels, const float* __restrict__ dL_invdepths,
                                                               float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic2D,
                                                               float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors,
                                                               float* __restrict__ dL_dinvdepths) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, tx = tid & 15u, ty = tid >> 4;
    const unsigned horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const unsigned cta = blockIdx.y * horizontal_blocks + blockIdx.x;
    const unsigned tile = tile_order ? __ldg(tile_order + cta) : cta;   // longest lists first (binning.cu)
    const unsigned tile_x = tile % horizontal_blocks, tile_y = tile / horizontal_blocks;
    const unsigned pix_x = tile_x * TILE_X + tx, pix_y0 = tile_y * TILE_Y + 2u * ty;
    const uint2 range = ranges[tile];
    const size_t HW = (size_t)H * W;
    const float bg0 = bg_color[0], bg1 = bg_color[1], bg2 = bg_color[2];

    BwdPix2 p2;
    {
        float Tf[2], g0[2], g1[2], g2[2], gd[2];
        unsigned last[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const unsigned y = pix_y0 + h;
            const bool inside = pix_x < (unsigned)W && y < (unsigned)H;
            const uint32_t pix_id = W * y + pix_x;
            Tf[h] = inside ? final_Ts[pix_id] : 0.f;
            last[h] = inside ? n_contrib[pix_id] : 0u;
            g0[h] = inside ? dL_dpixels[0 * HW + pix_id] : 0.f;
            g1[h] = inside ? dL_dpixels[1 * HW + pix_id] : 0.f;
            g2[h] = inside ? dL_dpixels[2 * HW + pix_id] : 0.f;
            gd[h] = (DEPTH && inside) ? dL_invdepths[pix_id] : 0.f;
        }
        p2.px = (float)pix_x;
        p2.npy = pk(-(float)pix_y0, -(float)(pix_y0 + 1u));
        p2.T_final = pk(Tf[0], Tf[1]);
        p2.T = p2.T_final;
        p2.last0 = last[0];
        p2.last1 = last[1];
        p2.last_alpha = p2.ar0 = p2.ar1 = p2.ar2 = p2.ar_invd = bc(0.f);
        p2.lc0 = p2.lc1 = p2.lc2 = p2.last_invd = 0.f;
        p2.g0 = pk(g0[0], g0[1]);
        p2.g1 = pk(g1[0], g1[1]);
        p2.g2 = pk(g2[0], g2[1]);
        p2.gd = pk(gd[0], gd[1]);
        p2.nbg_dot = pk(-bg_mult * (bg0 * g0[0] + bg1 * g1[0] + bg2 * g2[0]), -bg_mult * (bg0 * g0[1] + bg1 * g1[1] + bg2 * g2[1]));
    }

    // the replay only has to cover splats up to the tile's deepest contributor
    const unsigned wmax = __reduce_max_sync(0xffffffffu, max(p2.last0, p2.last1));
    if (lane == 0) sm.warp_max[warp] = wmax;
    if (tid == 0) {
        mbar_init(&sm.bar[0], BWD_THREADS);
        mbar_init(&sm.bar[1], BWD_THREADS);
        mbar_fence_init();
    }
    for (int i = tid; i < 2 * BWD_BATCH * NV; i += BWD_THREADS) (&sm.acc[0][0])[i] = 0.f;
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

    int pending = -1;
    int my_gid = -1;
    float4 my_con = {0.f, 0.f, 0.f, 0.f};  // conic xx, xy, yy and opacity of the splat this thread flushes
    auto flush = [&]() {
        if (pending < 0) return;
        float* a = &sm.acc[pending & 1][tid * NV];
        if (my_gid >= 0) {
            float v[NV];
            bool any = false;
#pragma unroll
            for (int k = 0; k < NV; k += 2) {
                const float2 t2 = *reinterpret_cast<const float2*>(a + k);
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
#pragma unroll
                for (int k = 0; k < NV; k += 2) *reinterpret_cast<float2*>(a + k) = make_float2(0.f, 0.f);
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
        float* accp = &sm.acc[s][jtop * NV];
        unsigned pos = lo + (unsigned)jtop;  // 0-based position in the tile's list == contributor index - 1
        _Pragma("unroll 2") for (int j = jtop; j >= 0; j--, rp -= 3, accp -= NV, pos--) {
            float v[NV];
            const float4 q0 = rp[0], q1 = rp[1], q2 = rp[2];
            if (!bwd_pair2<LOD, DEPTH>(p2, pos, q0, q1, q2, null_thr, v)) continue;  // the same answer in every lane
            int slot;
            const float total = warp_reduce10(v, lane, slot);
            if (slot >= 0 && (DEPTH || slot < 9) && total != 0.f) atomicAdd(accp + slot, total);
        }
    }
    __syncthreads();
    flush();
}

int launch_blend_bwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                     const float* bg, float bg_mult, float null_thr, const int32_t* indices, const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, cudaStream_t s) {
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1), block(BWD_THREADS, 1, 1);
    static std::atomic<bool> attr_set[MAX_DEVICES];
    const int dslot = device_slot();
    if (!attr_set[dslot].load(std::memory_order_acquire)) {
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        attr_set[dslot].store(true, std::memory_order_release);
    }
    const size_t smem = sizeof(BwdSmem);
    static const bool tile_id_order = getenv("LODGS_BLEND_TILE_ID_ORDER") != nullptr;  // A/B: CTAs in tile-id order
    static const int lod_cut_rc = set_lod_cut();
    if (lod_cut_rc) { set_error("blend: setting the LOD cut-off switch failed"); return LODGS_E_CUDA; }
    const uint32_t* order = tile_id_order ? nullptr : img.tile_order;
#define LODGS_BWD_ARGS img.ranges, order, recs, bin.point_list, W, H, bg, bg_mult, null_thr, indices, img.final_T, img.n_contrib, dL_dpix, dL_dinvdepth_pix, \
                       dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth
    if (lod && depth) blend_bwd_kernel<true, true><<<grid, block, smem, s>>>(LODGS_BWD_ARGS);
    else if (lod) blend_bwd_kernel<true, false><<<grid, block, smem, s>>>(LODGS_BWD_ARGS);
    else if (depth) blend_bwd_kernel<false, true><<<grid, block, smem, s>>>(LODGS_BWD_ARGS);
    else blend_bwd_kernel<false, false><<<grid, block, smem, s>>>(LODGS_BWD_ARGS);
#undef LODGS_BWD_ARGS
    LODGS_LAUNCHED("blend_bwd_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
