// blend_fwd.cu — per-tile front-to-back alpha blending with hierarchical-LOD alpha.
//
// Replaces renderCUDA<3> forward (HR/cuda_rasterizer/forward.cu:450-596). One CTA per 16x16 tile; the 48-byte
// blend records of the tile's sorted Gaussians are gathered (through the sorted id list) straight from the
// per-Gaussian array the preprocess kernel wrote — L2-resident, 48 MB per million Gaussians — into a two-stage
// shared-memory ring by 16-byte async copies (cp.async -> LDGSTS, completion on an mbarrier), so the gather of
// batch b+1 overlaps the blending of batch b, no per-instance record array is ever built, and colours/depths never
// come from global memory inside the pixel loop (the reference fetches them per blended pair). A warp leaves a
// batch as soon as all of its 64 pixels are saturated (ballot), the CTA leaves when all warps have.
// The arithmetic keeps the reference's expression order (quadratic form, expf, __powf, blend FMAs) so that
// n_contrib — an integer output decided by the alpha<1/255 and T<1e-4 tests — and the image are bit-identical.
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


// Q = vertically adjacent pixel PAIRS per thread (Q = 2: a column of four pixels, a warp covers 16x8; Q = 1: two pixels, a warp covers 16x4).
// What is paid per (warp, splat) — three record loads, the loop with its exit vote, the column-constant dx terms — is shared by twice the
// pixels with Q = 2, which a throughput-bound launch wants; a tile with a very long list is latency-bound and wants the shorter per-splat
// chain of Q = 1. The default kernel (NT = 64 threads) takes a tile as ONE CTA with Q = 2 or, for the tiles tile_order_kernel (binning.cu)
// marked as long, as TWO CTAs with Q = 1 that own the upper / lower eight rows and walk the same list; per pixel the arithmetic is the same
// in every mapping, so the outputs stay bit-identical. NT = 128 is the round-1 shape (LODGS_FWD_PX=2). Every thread gathers two records
// per batch, so the batch follows the CTA size.
constexpr int LONG_TILES_MAX = 2048;  // upper bound of the number of long tiles, i.e. of the extra CTAs of a launch
template <int NT>
struct BlendCfg {
    static constexpr int THREADS = NT;
    static constexpr int BATCH = 2 * NT;
};

template <int NT>
struct BlendSmemT {
    SplatRec recs[2][BlendCfg<NT>::BATCH];
    uint8_t flag[2][BlendCfg<NT>::BATCH];
    uint64_t bar[2];
};

struct FwdArgs {
    const uint2* ranges;
    const uint32_t* tile_order;   // [tiles] + the number of long tiles at [tiles]; NULL = tile-id order
    const SplatRec* recs;
    const uint32_t* point_list;
    int W, H;
    const float* bg_color;
    float* final_T;
    uint32_t* n_contrib;
    float *out_color, *out_invdepth;
    int32_t* seen;
};

// Blend state of the thread's 2Q pixels in packed FP32 pairs (pair h: lo = row 2h of the thread's column, hi = the one below it).
template <int Q>
struct PixState {
    f2 npy[Q], T[Q], C0[Q], C1[Q], C2[Q], D[Q];
    float px;
    uint32_t last[2 * Q];
};

// One splat against the thread's pixels, rounding pinned to the reference build's SASS for renderCUDA (forward.cu:529-585):
//   power = (dx*(dx*con.x) + dy*(dy*con.z)) * -0.5 - dy*(dx*con.y),   C += T * (alpha * c),   T *= (1 - alpha)
// evaluated on packed pairs: every half performs exactly the IEEE operations of the scalar code —
//   dy = y + (-py);  quad = fma(dx, dx*con.x, dy*(dy*con.z));  power = fma(quad, -0.5, dy*(-(dx*con.y)))  [quad * -0.5 is exact,
//   so the fused form equals the reference's rounded product minus rounded product];  1 - alpha = fma(alpha, -1, 1);
//   C = fma(T, alpha*c, C) — and only sequences ptxas cannot re-contract are used (see common.cuh).
// A pixel that does not take the splat (saturated, power > 0, below the cut-off, alpha < 1/255, or saturating on this splat)
// rides along with alpha 0: fma(T, 0*c, C) == C and its T is kept by a select, so its results are untouched bit for bit.
template <bool LOD, bool DEPTH, bool SEEN, int Q>
__device__ __forceinline__ void blend_splat(PixState<Q>& p, unsigned (&done)[2 * Q], const float4 q0, const float4 q1,
                                            const float4 q2, uint32_t contributor, uint8_t* flagp) {
    const float dx = __fsub_rn(q0.x, p.px), dxa = __fmul_rn(dx, q0.z), ndxb = -__fmul_rn(dx, q0.w);
    float cut;
    if (!LOD) {  // alpha is certainly below 1/255 (cut-off precomputed per splat, see preprocess_one)
        cut = q2.z;
    } else {
        // Both spare record slots carry t and 1/kids here, so the cut-off is rebuilt per splat: the LOD alpha
        // t*a + (1-t)*(1 - (1-a)^(1/kids)) never exceeds a = opacity*exp(power) for t in [0,1] and kids >= 1, and
        // power < -log(255*opacity) - 0.02 puts a 2 % below 1/255 (the approximations of __logf/__powf move it by ~1e-6), so
        // exactly pairs the alpha test below would drop are dropped here — before two exponentials and two powers.
        cut = (c_lod_cut && q2.z >= 0.0f && q2.z <= 1.0f && q2.w > 0.0f && q2.w <= 1.0f) ? -__logf(255.0f * q1.y) - 0.02f : -__builtin_huge_valf();
    }
    float pw[2 * Q];
    bool ok[2 * Q];
    bool any = false;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        const f2 dy = add2(bc(q0.y), p.npy[h]);
        const f2 quad = fma2(bc(dx), bc(dxa), mul2(dy, mul2(dy, bc(q1.x))));
        const f2 power = fma2(quad, bc(-0.5f), mul2(dy, bc(ndxb)));
        pw[2 * h] = lo(power);
        pw[2 * h + 1] = hi(power);
#pragma unroll
        for (int e = 2 * h; e < 2 * h + 2; e++) {
            ok[e] = !done[e] && !(pw[e] > 0.0f) && !(pw[e] < cut);
            any = any || ok[e];
        }
    }
    if (!any) return;
    float a[2 * Q];
    any = false;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        const f2 al = mul2(bc(q1.y), pk(expf(pw[2 * h]), expf(pw[2 * h + 1])));
        a[2 * h] = fminf(0.99f, lo(al));
        a[2 * h + 1] = fminf(0.99f, hi(al));
#pragma unroll
        for (int e = 2 * h; e < 2 * h + 2; e++) {
            if (LOD) {
                const float k = __fsub_rn(1.0f, __powf(__fsub_rn(1.0f, a[e]), q2.w));
                const float omt = __fsub_rn(1.0f, q2.z);
                a[e] = __fmaf_rn(a[e], q2.z, __fmul_rn(omt, k));
            }
            ok[e] = ok[e] && !(a[e] < 1.0f / 255.0f);
            any = any || ok[e];
        }
    }
    if (!any) return;
    bool blended = false;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        const f2 test_T = mul2(p.T[h], fma2(pk(a[2 * h], a[2 * h + 1]), bc(-1.0f), bc(1.0f)));
        const float t0 = lo(test_T), t1 = hi(test_T);
        const bool sat0 = ok[2 * h] && t0 < 0.0001f, sat1 = ok[2 * h + 1] && t1 < 0.0001f;
        if (sat0) done[2 * h] = 1u;
        if (sat1) done[2 * h + 1] = 1u;
        const bool bl0 = ok[2 * h] && !sat0, bl1 = ok[2 * h + 1] && !sat1;
        blended = blended || bl0 || bl1;
        const f2 ae = pk(bl0 ? a[2 * h] : 0.f, bl1 ? a[2 * h + 1] : 0.f);
        fma2_acc(p.C0[h], p.T[h], mul2(ae, bc(q1.z)));
        fma2_acc(p.C1[h], p.T[h], mul2(ae, bc(q1.w)));
        fma2_acc(p.C2[h], p.T[h], mul2(ae, bc(q2.x)));
        if (DEPTH) fma2_acc(p.D[h], p.T[h], mul2(ae, bc(q2.y)));  // q2.y = rcp.rn(depth), precomputed per Gaussian
        p.T[h] = pk(bl0 ? t0 : lo(p.T[h]), bl1 ? t1 : hi(p.T[h]));
        if (bl0) p.last[2 * h] = contributor;
        if (bl1) p.last[2 * h + 1] = contributor;
    }
    if (SEEN && blended) *flagp = 1;
}

// One tile (rows y_off .. y_off + 2*Q*NT/16 - 1 of it) blended by the NT threads of this CTA.
template <bool LOD, bool DEPTH, bool SEEN, int Q, int NT>
__device__ __forceinline__ void fwd_tile(const FwdArgs& a, BlendSmemT<NT>& sm, const unsigned tile, const unsigned y_off) {
    constexpr int BLEND_THREADS = BlendCfg<NT>::THREADS, BLEND_BATCH = BlendCfg<NT>::BATCH;
    const uint2* __restrict__ ranges = a.ranges;
    const SplatRec* __restrict__ recs = a.recs;
    const uint32_t* __restrict__ point_list = a.point_list;
    const int W = a.W, H = a.H;
    const float* __restrict__ bg_color = a.bg_color;
    float* __restrict__ final_T = a.final_T;
    uint32_t* __restrict__ n_contrib = a.n_contrib;
    float* __restrict__ out_color = a.out_color;
    float* __restrict__ out_invdepth = a.out_invdepth;
    int32_t* __restrict__ seen = a.seen;
    const unsigned tid = threadIdx.x, tx = tid & 15u, ty = tid >> 4;
    const unsigned horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const unsigned tile_x = tile % horizontal_blocks, tile_y = tile / horizontal_blocks;
    const unsigned pix_x = tile_x * TILE_X + tx, pix_y0 = tile_y * TILE_Y + y_off + 2u * Q * ty;
    unsigned done[2 * Q];
    PixState<Q> p;
    p.px = (float)pix_x;
#pragma unroll
    for (int h = 0; h < Q; h++) {
        p.npy[h] = pk(-(float)(pix_y0 + 2u * h), -(float)(pix_y0 + 2u * h + 1u));
        p.T[h] = bc(1.0f);
        p.C0[h] = p.C1[h] = p.C2[h] = p.D[h] = bc(0.f);
    }
    unsigned all_done = 1u;
#pragma unroll
    for (int e = 0; e < 2 * Q; e++) {
        done[e] = (pix_x < (unsigned)W && pix_y0 + e < (unsigned)H) ? 0u : 1u;
        all_done &= done[e];
        p.last[e] = 0u;
    }

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BLEND_BATCH - 1) / BLEND_BATCH;

    if (tid == 0) {
        mbar_init(&sm.bar[0], BLEND_THREADS);
        mbar_init(&sm.bar[1], BLEND_THREADS);
        mbar_fence_init();
    }
    if (SEEN) {
        for (int i = tid; i < 2 * BLEND_BATCH; i += BLEND_THREADS) (&sm.flag[0][0])[i] = 0;
    }
    __syncthreads();

    constexpr uint32_t NO_ID = 0xffffffffu;
    // every thread gathers the records of slots tid and tid + BLEND_THREADS of a batch; their ids are fetched one batch ahead
    auto load_ids = [&](int b, uint32_t(&id)[2]) {
        const unsigned start = range.x + (unsigned)b * BLEND_BATCH;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const unsigned pos = start + tid + h * BLEND_THREADS;
            id[h] = pos < range.y ? __ldg(point_list + pos) : NO_ID;
        }
    };
    auto issue = [&](int b, const uint32_t(&id)[2]) {
        const int s = b & 1;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (id[h] != NO_ID) {
                const float4* src = reinterpret_cast<const float4*>(recs + id[h]);
                float4* dst = reinterpret_cast<float4*>(&sm.recs[s][tid + h * BLEND_THREADS]);
                cp_async16(dst, src);
                cp_async16(dst + 1, src + 1);
                cp_async16(dst + 2, src + 2);
            }
        }
        cp_async_arrive(&sm.bar[s]);  // counts this thread in once its copies have landed
    };
    uint32_t id_cur[2] = {NO_ID, NO_ID}, id_next[2] = {NO_ID, NO_ID};
    if (rounds > 0) {
        load_ids(0, id_cur);
        issue(0, id_cur);
    }
    if (rounds > 1) load_ids(1, id_next);

    uint32_t my_id[2] = {0, 0};  // ids of the two splats this thread flushes `seen` for (previous batch)
    int pending = -1;            // batch whose seen flags still have to be flushed

    auto flush_seen = [&]() {
        if (SEEN && pending >= 0) {
            const int ps = pending & 1;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const unsigned slot = tid + h * BLEND_THREADS;
                if (sm.flag[ps][slot]) {
                    seen[my_id[h]] = 1;
                    sm.flag[ps][slot] = 0;
                }
            }
        }
    };

    int toDo = total;
    for (int b = 0; b < rounds; b++, toDo -= BLEND_BATCH) {
        // barrier: everyone has left batch b-1, so its ring slot may be refilled; also the CTA-wide exit vote
        all_done = 1u;
#pragma unroll
        for (int e = 0; e < 2 * Q; e++) all_done &= done[e];
        const int num_done = __syncthreads_count(all_done != 0u);
        flush_seen();
        pending = -1;
        const int s = b & 1;
        if (num_done == BLEND_THREADS) {
            // this thread's copies of batch b are still in flight: they must land before its CTA (and the smem) goes away
            cp_async_wait_all();
            break;
        }
        if (b + 1 < rounds) issue(b + 1, id_next);
        if (SEEN) {
            my_id[0] = id_cur[0];
            my_id[1] = id_cur[1];
            pending = b;
        }
        id_cur[0] = id_next[0];
        id_cur[1] = id_next[1];
        if (b + 2 < rounds) load_ids(b + 2, id_next);
        mbar_wait(&sm.bar[s], (unsigned)(b >> 1) & 1u);
        const int cnt = min(BLEND_BATCH, toDo);
        const float4* __restrict__ rp = reinterpret_cast<const float4*>(&sm.recs[s][0]);
        uint8_t* fl = &sm.flag[s][0];
        uint32_t contributor = (uint32_t)b * BLEND_BATCH;
        _Pragma("unroll 2") for (int j = 0; j < cnt; j++, rp += 3, fl++) {
            contributor++;
            // ballot-based early termination: the warp stops scanning once all of its 64 pixels are saturated
            all_done = 1u;
#pragma unroll
            for (int e = 0; e < 2 * Q; e++) all_done &= done[e];
            if (__all_sync(0xffffffffu, all_done != 0u)) break;
            const float4 q0 = rp[0];  // x y conx cony
            const float4 q1 = rp[1];  // conz opac r g
            const float4 q2 = rp[2];  // b 1/depth t 1/kids
            blend_splat<LOD, DEPTH, SEEN, Q>(p, done, q0, q1, q2, contributor, fl);
        }
    }
    if (SEEN) {
        __syncthreads();
        flush_seen();
    }
    const size_t HW = (size_t)H * W;
    const float bg0 = bg_color[0], bg1 = bg_color[1], bg2 = bg_color[2];
#pragma unroll
    for (int e = 0; e < 2 * Q; e++) {
        const unsigned y = pix_y0 + e;
        if (pix_x < (unsigned)W && y < (unsigned)H) {
            const int h = e >> 1;
            const uint32_t pix_id = W * y + pix_x;
            const float T = (e & 1) ? hi(p.T[h]) : lo(p.T[h]);
            final_T[pix_id] = T;
            n_contrib[pix_id] = p.last[e];
            out_color[0 * HW + pix_id] = __fmaf_rn(T, bg0, (e & 1) ? hi(p.C0[h]) : lo(p.C0[h]));
            out_color[1 * HW + pix_id] = __fmaf_rn(T, bg1, (e & 1) ? hi(p.C1[h]) : lo(p.C1[h]));
            out_color[2 * HW + pix_id] = __fmaf_rn(T, bg2, (e & 1) ? hi(p.C2[h]) : lo(p.C2[h]));
            if (DEPTH) out_invdepth[pix_id] = (e & 1) ? hi(p.D[h]) : lo(p.D[h]);
        }
    }
}

template <bool LOD, bool DEPTH, bool SEEN, int NT>
__global__ void __launch_bounds__(BlendCfg<NT>::THREADS) blend_fwd_kernel(const FwdArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BlendSmemT<NT>& sm = *reinterpret_cast<BlendSmemT<NT>*>(smem_raw);
    const unsigned ntiles = ((a.W + TILE_X - 1) / TILE_X) * ((a.H + TILE_Y - 1) / TILE_Y);
    const unsigned cta = blockIdx.x;
    if (NT == 64) {
        // CTAs [0, 2 n_long): the long tiles, two CTAs each (upper / lower eight rows, two pixels per thread); then one CTA per remaining tile
        const unsigned n_long = a.tile_order ? __ldg(a.tile_order + ntiles) : 0u;
        if (cta < 2u * n_long) {
            fwd_tile<LOD, DEPTH, SEEN, 1, NT>(a, sm, __ldg(a.tile_order + (cta >> 1)), 8u * (cta & 1u));
            return;
        }
        const unsigned i = cta - n_long;
        if (i >= ntiles) return;
        fwd_tile<LOD, DEPTH, SEEN, 2, NT>(a, sm, a.tile_order ? __ldg(a.tile_order + i) : i, 0u);   // longest lists first (binning.cu)
    } else {
        if (cta >= ntiles) return;
        fwd_tile<LOD, DEPTH, SEEN, 1, NT>(a, sm, a.tile_order ? __ldg(a.tile_order + cta) : cta, 0u);
    }
}

template <bool LOD, bool DEPTH, bool SEEN, int NT>
static int launch_blend_q(unsigned grid, cudaStream_t s, const FwdArgs& a) {
    static std::atomic<bool> attr_set[MAX_DEVICES];
    const int dslot = device_slot();
    if (!attr_set[dslot].load(std::memory_order_acquire)) {
        LODGS_CUDA(cudaFuncSetAttribute(blend_fwd_kernel<LOD, DEPTH, SEEN, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BlendSmemT<NT>)));
        attr_set[dslot].store(true, std::memory_order_release);
    }
    blend_fwd_kernel<LOD, DEPTH, SEEN, NT><<<grid, NT, sizeof(BlendSmemT<NT>), s>>>(a);
    LODGS_LAUNCHED("blend_fwd_kernel");
    return LODGS_OK;
}

template <bool LOD, bool DEPTH, bool SEEN>
static int launch_blend_one(dim3 grid3, cudaStream_t s, const uint2* ranges, const uint32_t* order, const SplatRec* recs, const uint32_t* point_list, int W, int H,
                            const float* bg, float* final_T, uint32_t* n_contrib, float* out_color, float* out_invdepth, int32_t* seen) {
    static const int px = getenv("LODGS_FWD_PX") ? atoi(getenv("LODGS_FWD_PX")) : 4;   // A/B: 2 = 128-thread CTAs with two pixels per thread
    const unsigned ntiles = grid3.x * grid3.y;
    FwdArgs a;
    a.ranges = ranges; a.tile_order = order; a.recs = recs; a.point_list = point_list; a.W = W; a.H = H; a.bg_color = bg; a.final_T = final_T;
    a.n_contrib = n_contrib; a.out_color = out_color; a.out_invdepth = out_invdepth; a.seen = seen;
    if (px == 2) return launch_blend_q<LOD, DEPTH, SEEN, 128>(ntiles, s, a);
    return launch_blend_q<LOD, DEPTH, SEEN, 64>(ntiles + (order ? (unsigned)LONG_TILES_MAX : 0u), s, a);   // the CTAs beyond tiles + n_long exit at once
}

template <bool LOD, bool DEPTH>
static int launch_blend_seen(bool has_seen, dim3 grid, dim3 block, cudaStream_t s, const uint2* ranges, const uint32_t* order, const SplatRec* recs,
                             const uint32_t* point_list, int W, int H, const float* bg, float* final_T, uint32_t* n_contrib,
                             float* out_color, float* out_invdepth, int32_t* seen) {
    (void)block;
    if (has_seen) return launch_blend_one<LOD, DEPTH, true>(grid, s, ranges, order, recs, point_list, W, H, bg, final_T, n_contrib, out_color, out_invdepth, seen);
    return launch_blend_one<LOD, DEPTH, false>(grid, s, ranges, order, recs, point_list, W, H, bg, final_T, n_contrib, out_color, out_invdepth, seen);
}

int launch_blend_fwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                     const float* bg, float* out_color, float* out_invdepth, int32_t* seen, cudaStream_t s) {
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1), block(TILE_X, TILE_Y, 1);
    const bool hs = seen != nullptr;
    static const bool tile_id_order = getenv("LODGS_BLEND_TILE_ID_ORDER") != nullptr;  // A/B: CTAs in tile-id order
    static const int lod_cut_rc = set_lod_cut();
    if (lod_cut_rc) { set_error("blend: setting the LOD cut-off switch failed"); return LODGS_E_CUDA; }
    const uint32_t* order = tile_id_order ? nullptr : img.tile_order;
    if (lod && depth) return launch_blend_seen<true, true>(hs, grid, block, s, img.ranges, order, recs, bin.point_list, W, H, bg, img.final_T, img.n_contrib, out_color, out_invdepth, seen);
    if (lod && !depth) return launch_blend_seen<true, false>(hs, grid, block, s, img.ranges, order, recs, bin.point_list, W, H, bg, img.final_T, img.n_contrib, out_color, out_invdepth, seen);
    if (!lod && depth) return launch_blend_seen<false, true>(hs, grid, block, s, img.ranges, order, recs, bin.point_list, W, H, bg, img.final_T, img.n_contrib, out_color, out_invdepth, seen);
    return launch_blend_seen<false, false>(hs, grid, block, s, img.ranges, order, recs, bin.point_list, W, H, bg, img.final_T, img.n_contrib, out_color, out_invdepth, seen);
}

// Work counters of the blend (bench.py roofline): replays the forward's per-pixel decisions and counts
// N_test (pairs whose quadratic form is evaluated), N_alpha (pairs that reach the exponential) and N_blend
// (pairs that are blended). Not on any product path.
template <bool LOD>
__global__ void __launch_bounds__(TILE_PIX) blend_stats_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                              const SplatRec* __restrict__ grecs, int W, int H,
                                                              unsigned long long* __restrict__ counters) {
    const unsigned horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const uint2 pix = {blockIdx.x * TILE_X + threadIdx.x, blockIdx.y * TILE_Y + threadIdx.y};
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    const uint2 range = ranges[blockIdx.y * horizontal_blocks + blockIdx.x];
    unsigned n_test = 0, n_alpha = 0, n_blend = 0;
    float T = 1.0f;
    if (inside) {
        for (unsigned i = range.x; i < range.y; i++) {
            const float4* r4 = reinterpret_cast<const float4*>(grecs + __ldg(point_list + i));
            const float4 q0 = __ldg(r4);
            const float4 q1 = __ldg(r4 + 1);
            n_test++;
            const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
            const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
            if (power > 0.0f) continue;
            n_alpha++;
            const float my_alpha = min(0.99f, q1.y * expf(power));
            float alpha = my_alpha;
            if (LOD) {
                const float4 q2 = __ldg(r4 + 2);
                alpha = q2.z * my_alpha + (1.0f - q2.z) * (1.0f - __powf(1.0f - my_alpha, q2.w));
            }
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break;
            n_blend++;
            T = test_T;
        }
    }
    n_test = __reduce_add_sync(0xffffffffu, n_test);
    n_alpha = __reduce_add_sync(0xffffffffu, n_alpha);
    n_blend = __reduce_add_sync(0xffffffffu, n_blend);
    if (((threadIdx.y * TILE_X + threadIdx.x) & 31) == 0) {
        atomicAdd(counters + 0, (unsigned long long)n_test);
        atomicAdd(counters + 1, (unsigned long long)n_alpha);
        atomicAdd(counters + 2, (unsigned long long)n_blend);
    }
}

int launch_blend_stats(bool lod, const ImageState& img, const BinningState& bin, const SplatRec* recs, int W, int H,
                       unsigned long long* counters, cudaStream_t s) {
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1), block(TILE_X, TILE_Y, 1);
    LODGS_CUDA(cudaMemsetAsync(counters, 0, 3 * sizeof(unsigned long long), s));
    if (lod) blend_stats_kernel<true><<<grid, block, 0, s>>>(img.ranges, bin.point_list, recs, W, H, counters);
    else blend_stats_kernel<false><<<grid, block, 0, s>>>(img.ranges, bin.point_list, recs, W, H, counters);
    LODGS_LAUNCHED("blend_stats_kernel");
    return LODGS_OK;
}

}  // namespace lodgs
