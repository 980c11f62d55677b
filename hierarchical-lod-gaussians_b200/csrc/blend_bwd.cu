// blend_bwd.cu — back-to-front replay of the tile blend, producing per-Gaussian screen-space gradients.
//
// Replaces renderCUDA<3> backward (HR/cuda_rasterizer/backward.cu:498-721), which issues 9-10 global float
// atomics per blended (pixel,Gaussian) pair. Here the ten partial gradients of a Gaussian are first reduced
// across the 32 pixels of a warp with a value-splitting butterfly (12 shuffles instead of 50: every exchange
// halves the number of values a lane still owns), the ten owning lanes add them to a per-batch shared-memory
// accumulator, and after the batch one thread per Gaussian flushes non-zero totals with one global RED per value:
// at most 10 atomics per (tile,Gaussian) instance instead of per pair. Records arrive through the same
// bulk-copy ring as the forward, and the replay starts at the tile's last contributing splat instead of the
// end of its range.
#include "state.cuh"

namespace lodgs {

constexpr int BWD_BATCH = 256;
constexpr int NV = 10;  // dmean.x dmean.y dconic.xx dconic.xy dconic.yy dopacity dcol.r dcol.g dcol.b dinvdepth

struct BwdSmem {
    SplatRec recs[2][BWD_BATCH];
    float acc[2][BWD_BATCH * NV];
    uint32_t ids[2][BWD_BATCH + 8];
    uint64_t bar[2];
    unsigned warp_max[TILE_PIX / 32];
};

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

template <bool LOD, bool DEPTH>
__global__ void __launch_bounds__(TILE_PIX) blend_bwd_kernel(const uint2* __restrict__ ranges, const SplatRec* __restrict__ recs,
                                                            const uint32_t* __restrict__ point_list, int W, int H,
                                                            const float* __restrict__ bg_color, const int32_t* __restrict__ indices,
                                                            const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                                                            const float* __restrict__ dL_dpixels, const float* __restrict__ dL_invdepths,
                                                            float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic2D,
                                                            float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors,
                                                            float* __restrict__ dL_dinvdepths) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);
    const unsigned tid = threadIdx.y * TILE_X + threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const uint2 pix = {blockIdx.x * TILE_X + threadIdx.x, blockIdx.y * TILE_Y + threadIdx.y};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    const uint2 range = ranges[blockIdx.y * horizontal_blocks + blockIdx.x];

    const float T_final = inside ? final_Ts[pix_id] : 0;
    float T = T_final;
    const unsigned last_contributor = inside ? n_contrib[pix_id] : 0;

    // the replay only has to cover splats up to the tile's deepest contributor
    const unsigned wmax = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) sm.warp_max[warp] = wmax;
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        mbar_fence_init();
    }
    for (int i = tid; i < 2 * BWD_BATCH * NV; i += TILE_PIX) (&sm.acc[0][0])[i] = 0.f;
    __syncthreads();
    unsigned L = 0;
#pragma unroll
    for (int w = 0; w < TILE_PIX / 32; w++) L = max(L, sm.warp_max[w]);
    L = min(L, range.y - range.x);
    const int rounds = (int)((L + BWD_BATCH - 1) / BWD_BATCH);

    // batch b covers sorted positions [lo_b, hi_b) with hi_b = L - b*BATCH, walked from hi_b-1 down to lo_b
    auto issue = [&](int b) {
        const int s = b & 1;
        const unsigned hi = L - (unsigned)b * BWD_BATCH;
        const unsigned lo = hi > BWD_BATCH ? hi - BWD_BATCH : 0u;
        const unsigned cnt = hi - lo;
        const unsigned start = range.x + lo;
        const unsigned rec_bytes = cnt * (unsigned)sizeof(SplatRec);
        const unsigned a0 = start & ~3u;
        const unsigned id_bytes = (((start - a0) + cnt) * 4u + 15u) & ~15u;
        mbar_expect_tx(&sm.bar[s], rec_bytes + id_bytes);
        bulk_g2s(&sm.recs[s][0], recs + start, rec_bytes, &sm.bar[s]);
        bulk_g2s(&sm.ids[s][0], point_list + a0, id_bytes, &sm.bar[s]);
    };
    if (tid == 0 && rounds > 0) issue(0);

    float accum_rec[3] = {0.f, 0.f, 0.f};
    float dL_dpixel[3] = {0.f, 0.f, 0.f};
    float dL_invdepth = 0.f;
    float accum_invdepth_rec = 0.f;
    if (inside) {
        const size_t HW = (size_t)H * W;
#pragma unroll
        for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
        if (DEPTH) dL_invdepth = dL_invdepths[pix_id];
    }
    float last_alpha = 0.f;
    float last_color[3] = {0.f, 0.f, 0.f};
    float last_invdepth = 0.f;
    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;
    float bg_dot_dpixel = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) bg_dot_dpixel += bg_color[i] * dL_dpixel[i];

    int pending = -1;
    int my_gid = -1;
    auto flush = [&]() {
        if (pending < 0) return;
        float* a = &sm.acc[pending & 1][tid * NV];
        if (my_gid >= 0) {
            float v[NV];
            bool any = false;
#pragma unroll
            for (int k = 0; k < NV; k++) {
                v[k] = a[k];
                any |= (v[k] != 0.f);
            }
            if (any) {
                atomicAdd(&dL_dmean2D[3 * (size_t)my_gid + 0], v[0]);
                atomicAdd(&dL_dmean2D[3 * (size_t)my_gid + 1], v[1]);
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 0], v[2]);
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 1], v[3]);
                atomicAdd(&dL_dconic2D[4 * (size_t)my_gid + 3], v[4]);
                atomicAdd(&dL_dopacity[my_gid], v[5]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 0], v[6]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 1], v[7]);
                atomicAdd(&dL_dcolors[3 * (size_t)my_gid + 2], v[8]);
                if (DEPTH) atomicAdd(&dL_dinvdepths[my_gid], v[9]);
#pragma unroll
                for (int k = 0; k < NV; k++) a[k] = 0.f;
            }
        }
        pending = -1;
    };

    for (int b = 0; b < rounds; b++) {
        __syncthreads();  // all warps have left batch b-1: its accumulator is complete and its ring slot is free
        flush();
        if (tid == 0 && b + 1 < rounds) issue(b + 1);
        const int s = b & 1;
        mbar_wait(&sm.bar[s], (unsigned)(b >> 1) & 1u);
        const unsigned hi = L - (unsigned)b * BWD_BATCH;
        const unsigned lo = hi > BWD_BATCH ? hi - BWD_BATCH : 0u;
        const int cnt = (int)(hi - lo);
        {
            const unsigned mis = (range.x + lo) & 3u;
            my_gid = -1;
            if ((int)tid < cnt) {
                const uint32_t id = sm.ids[s][mis + tid];
                my_gid = indices ? __ldg(indices + id) : (int)id;
            }
            pending = b;
        }
        const SplatRec* __restrict__ rb = sm.recs[s];
        float* acc = sm.acc[s];
        if (lo >= wmax) continue;  // nothing in this batch reaches any pixel of this warp
        for (int j = cnt - 1; j >= 0; j--) {
            const unsigned pos = lo + (unsigned)j;  // 0-based position in the tile's list == contributor index
            if (pos >= wmax) continue;
            float v[NV];
#pragma unroll
            for (int k = 0; k < NV; k++) v[k] = 0.f;
            bool contrib = false;
            if (pos < last_contributor) {
                const float4 q0 = *reinterpret_cast<const float4*>(&rb[j].x);
                const float4 q1 = *reinterpret_cast<const float4*>(&rb[j].conz);
                const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                if (power <= 0.0f) {
                    const float G = expf(power);
                    const float test_alpha = q1.y * G;
                    const bool nullalpha = test_alpha > 0.99f;
                    const float my_alpha = min(0.99f, test_alpha);
                    const float4 q2 = *reinterpret_cast<const float4*>(&rb[j].b);
                    float alpha, t = 1.f, frac = 1.f;
                    if (LOD) {
                        t = q2.z;
                        frac = q2.w;
                        const float kidsqrt_alpha = 1.0f - powf(1.0f - my_alpha, frac);
                        alpha = t * my_alpha + (1.0f - t) * kidsqrt_alpha;
                    } else {
                        alpha = my_alpha;
                    }
                    if (alpha >= 1.0f / 255.0f) {
                        contrib = true;
                        T = T / (1.f - alpha);
                        const float weight = alpha * T;
                        float dL_dalpha = 0.0f;
                        const float col[3] = {q1.z, q1.w, q2.x};
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            const float c = col[ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = c;
                            const float dL_dchannel = dL_dpixel[ch];
                            dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
                            v[6 + ch] = weight * dL_dchannel;
                        }
                        if (DEPTH) {
                            const float invd = 1.f / q2.y;
                            accum_invdepth_rec = last_alpha * last_invdepth + (1.f - last_alpha) * accum_invdepth_rec;
                            last_invdepth = invd;
                            dL_dalpha += (invd - accum_invdepth_rec) * dL_invdepth;
                            v[9] = weight * dL_invdepth;
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                        dL_dalpha = nullalpha ? 0 : dL_dalpha;
                        const float dL_dG = q1.y * dL_dalpha;
                        const float gdx = G * d.x;
                        const float gdy = G * d.y;
                        const float dG_ddelx = -gdx * q0.z - gdy * q0.w;
                        const float dG_ddely = -gdy * q1.x - gdx * q0.w;
                        v[0] = dL_dG * dG_ddelx * ddelx_dx;
                        v[1] = dL_dG * dG_ddely * ddely_dy;
                        v[2] = -0.5f * gdx * d.x * dL_dG;
                        v[3] = -0.5f * gdx * d.y * dL_dG;
                        v[4] = -0.5f * gdy * d.y * dL_dG;
                        float opac_grad_mult = 1.0f;
                        if (LOD) opac_grad_mult = t - powf(1.0f - my_alpha, frac - 1.0f) * (t - 1.0f) * frac;
                        v[5] = opac_grad_mult * G * dL_dalpha;
                    }
                }
            }
            if (!__any_sync(0xffffffffu, contrib)) continue;
            int slot;
            const float total = warp_reduce10(v, lane, slot);
            if (slot >= 0 && (DEPTH || slot < 9) && total != 0.f) atomicAdd(&acc[j * NV + slot], total);
        }
    }
    __syncthreads();
    flush();
}

int launch_blend_bwd(bool lod, bool depth, const ImageState& img, const BinningState& bin, int W, int H, const float* bg,
                     const int32_t* indices, const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, cudaStream_t s) {
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1), block(TILE_X, TILE_Y, 1);
    static bool attr_set = false;
    if (!attr_set) {
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        LODGS_CUDA(cudaFuncSetAttribute(blend_bwd_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
        attr_set = true;
    }
    const size_t smem = sizeof(BwdSmem);
#define LODGS_BWD_ARGS img.ranges, bin.recs, bin.point_list, W, H, bg, indices, img.final_T, img.n_contrib, dL_dpix, dL_dinvdepth_pix, \
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
