// ssim_loss.cu — fused L1 + SSIM image loss, forward and backward.
//
// Replaces the training loss of the reference, train_post.py:558-559 / train_single.py:106-110:
//     Ll1 = l1_loss(image, gt)                                       utils/loss_utils.py:17-18
//     loss = (1 - lambda) * Ll1 + lambda * (1 - fused_ssim(image[None], gt[None]))
// where fused_ssim is an external package (not vendored; absent here) computing the SSIM of utils/loss_utils.py:33-63: 11x11 Gaussian
// window (sigma 1.5, zero "same" padding), mu / sigma / covariance maps, C1 = 0.01^2, C2 = 0.03^2, mean over channels and pixels.
// The eager form is 5 grouped conv2d launches + ~15 element-wise kernels forward and as many backward over 5 full-resolution maps.
// Here: ONE forward kernel per 16x16 tile and channel (26x26 halo tiles of both images in shared memory, separable 11-tap passes for
// the five moments, the SSIM value, its three partial derivatives and the L1 term, block partial sums) and ONE backward kernel that
// convolves the three derivative maps back (the window is symmetric) and adds the L1 sign term. Sums are reduced in a fixed order.
#include "common.cuh"
#include <cmath>

namespace lodgs {

constexpr int SS_T = 16, SS_R = 5, SS_H = SS_T + 2 * SS_R;   // tile, radius, halo tile side (26)
__constant__ float c_gauss[11];

struct SsimParams {
    const float* img; const float* gt; int C, H, W;
    float* dmu; float* ds; float* dc;      // derivative maps [C,H,W]
    double* partial;                       // [blocks][2]: sum |x-y|, sum ssim
};

__global__ void __launch_bounds__(SS_T * SS_T) ssim_fwd_kernel(const SsimParams p) {
    __shared__ float sx[SS_H][SS_H + 1], sy[SS_H][SS_H + 1];
    __shared__ float hz[5][SS_H][SS_T + 1];
    __shared__ double red[2][SS_T * SS_T / 32];
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
    const int c = blockIdx.z, x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t plane = (size_t)c * p.H * p.W;
    for (int i = tid; i < SS_H * SS_H; i += SS_T * SS_T) {
        const int ly = i / SS_H, lx = i - ly * SS_H;
        const int gy = y0 + ly - SS_R, gx = x0 + lx - SS_R;
        const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
        sx[ly][lx] = in ? __ldg(p.img + plane + (size_t)gy * p.W + gx) : 0.f;
        sy[ly][lx] = in ? __ldg(p.gt + plane + (size_t)gy * p.W + gx) : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < SS_H * SS_T; i += SS_T * SS_T) {   // horizontal pass: 26 rows x 16 columns
        const int ly = i / SS_T, lx = i - ly * SS_T;
        float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_gauss[k], u = sx[ly][lx + k], v = sy[ly][lx + k];
            a += w * u; b += w * v; aa += w * u * u; bb += w * v * v; ab += w * u * v;
        }
        hz[0][ly][lx] = a; hz[1][ly][lx] = b; hz[2][ly][lx] = aa; hz[3][ly][lx] = bb; hz[4][ly][lx] = ab;
    }
    __syncthreads();
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float w = c_gauss[k];
        mu1 += w * hz[0][ty + k][tx]; mu2 += w * hz[1][ty + k][tx]; e11 += w * hz[2][ty + k][tx]; e22 += w * hz[3][ty + k][tx];
        e12 += w * hz[4][ty + k][tx];
    }
    const int gx = x0 + tx, gy = y0 + ty;
    const bool in = gx < p.W && gy < p.H;
    double l1 = 0.0, ss = 0.0;
    if (in) {
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float mu1s = mu1 * mu1, mu2s = mu2 * mu2, mu12 = mu1 * mu2;
        const float s1 = e11 - mu1s, s2 = e22 - mu2s, s12 = e12 - mu12;
        const float A1 = 2.f * mu12 + C1, A2 = 2.f * s12 + C2, B1 = mu1s + mu2s + C1, B2 = s1 + s2 + C2;
        const float inv = 1.f / (B1 * B2);
        const float m = A1 * A2 * inv;
        // partial derivatives of the map value with respect to mu1 (sigma terms held), sigma1^2 and sigma12; then the total in mu1
        const float d_s1 = -m / B2;
        const float d_s12 = 2.f * A1 * inv;
        const float d_mu1 = 2.f * (mu2 * A2 - mu1 * m * B2) * inv;
        const size_t o = plane + (size_t)gy * p.W + gx;
        p.dmu[o] = d_mu1 - 2.f * mu1 * d_s1 - mu2 * d_s12;
        p.ds[o] = d_s1;
        p.dc[o] = d_s12;
        ss = m;
        l1 = fabsf(sx[ty + SS_R][tx + SS_R] - sy[ty + SS_R][tx + SS_R]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { l1 += __shfl_xor_sync(0xffffffffu, l1, o); ss += __shfl_xor_sync(0xffffffffu, ss, o); }
    if ((tid & 31) == 0) { red[0][tid >> 5] = l1; red[1][tid >> 5] = ss; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < SS_T * SS_T / 32; w++) { a += red[0][w]; b += red[1][w]; }
        const size_t blk = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.partial[2 * blk] = a; p.partial[2 * blk + 1] = b;
    }
}

// fixed-order final sum: out[0] = a_l1 * mean|x-y| + a_ssim * mean(ssim) + bias, out[1] = mean |x-y|, out[2] = mean ssim
__global__ void ssim_sum_kernel(const double* __restrict__ partial, int blocks, double n, float a_l1, float a_ssim, float bias, float* out) {
    __shared__ double s[2][256];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < blocks; i += 256) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    s[0][threadIdx.x] = a; s[1][threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s[0][threadIdx.x] += s[0][threadIdx.x + o]; s[1][threadIdx.x] += s[1][threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double l1 = s[0][0] / n, ss = s[1][0] / n;
        out[0] = (float)(a_l1 * l1 + a_ssim * ss + bias); out[1] = (float)l1; out[2] = (float)ss;
    }
}

struct SsimBwdParams {
    const float* img; const float* gt; const float* dmu; const float* ds; const float* dc; int C, H, W;
    float g_l1, g_ssim;   // a_l1 / n and a_ssim / n
    float* dL_dimg;
};

__global__ void __launch_bounds__(SS_T * SS_T) ssim_bwd_kernel(const SsimBwdParams p) {
    __shared__ float sm[3][SS_H][SS_H + 1];
    __shared__ float hz[3][SS_H][SS_T + 1];
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
    const int c = blockIdx.z, x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t plane = (size_t)c * p.H * p.W;
    for (int i = tid; i < SS_H * SS_H; i += SS_T * SS_T) {
        const int ly = i / SS_H, lx = i - ly * SS_H;
        const int gy = y0 + ly - SS_R, gx = x0 + lx - SS_R;
        const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
        const size_t o = plane + (size_t)gy * p.W + gx;
        sm[0][ly][lx] = in ? __ldg(p.dmu + o) : 0.f;
        sm[1][ly][lx] = in ? __ldg(p.ds + o) : 0.f;
        sm[2][ly][lx] = in ? __ldg(p.dc + o) : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < SS_H * SS_T; i += SS_T * SS_T) {
        const int ly = i / SS_T, lx = i - ly * SS_T;
        float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_gauss[k];
            a += w * sm[0][ly][lx + k]; b += w * sm[1][ly][lx + k]; d += w * sm[2][ly][lx + k];
        }
        hz[0][ly][lx] = a; hz[1][ly][lx] = b; hz[2][ly][lx] = d;
    }
    __syncthreads();
    const int gx = x0 + tx, gy = y0 + ty;
    if (gx >= p.W || gy >= p.H) return;
    float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float w = c_gauss[k];
        a += w * hz[0][ty + k][tx]; b += w * hz[1][ty + k][tx]; d += w * hz[2][ty + k][tx];
    }
    const size_t o = plane + (size_t)gy * p.W + gx;
    const float x = __ldg(p.img + o), y = __ldg(p.gt + o);
    const float diff = x - y;
    const float sgn = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);   // torch.abs backward: sign, 0 at 0
    p.dL_dimg[o] = p.g_ssim * (a + 2.f * x * b + y * d) + p.g_l1 * sgn;
}

static int set_gauss() {
    // utils/loss_utils.py:23-25: float32 tensor of exp(-(x-5)^2 / (2*1.5^2)) divided by its float32 sum
    float g[11], sum = 0.f;
    for (int i = 0; i < 11; i++) { g[i] = (float)std::exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); sum += g[i]; }
    for (int i = 0; i < 11; i++) g[i] = g[i] / sum;
    return (int)cudaMemcpyToSymbol(c_gauss, g, sizeof(g));
}

}  // namespace lodgs

using namespace lodgs;

static size_t ssim_blocks(int C, int H, int W) { return (size_t)C * ((H + SS_T - 1) / SS_T) * ((W + SS_T - 1) / SS_T); }

extern "C" size_t lodgs_l1_ssim_workspace_bytes(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return 256;
    return 3 * align_up((size_t)C * H * W * 4, 128) + align_up(ssim_blocks(C, H, W) * 16, 128) + 256;
}

extern "C" int lodgs_l1_ssim_loss(const float* img, const float* gt, int C, int H, int W, float a_l1, float a_ssim, float bias, float* out3,
                                  float* dL_dimg, void* ws, size_t ws_bytes, void* stream) {
    if (!img || !gt || C <= 0 || H <= 0 || W <= 0 || !out3 || !ws) { set_error("lodgs_l1_ssim_loss: bad argument"); return LODGS_E_ARG; }
    if (((uintptr_t)ws & 127) != 0) { set_error("lodgs_l1_ssim_loss: workspace must be 128-byte aligned"); return LODGS_E_ARG; }
    if (ws_bytes < lodgs_l1_ssim_workspace_bytes(C, H, W) - 256) { set_error("lodgs_l1_ssim_loss: workspace too small"); return LODGS_E_WORKSPACE; }
    static std::atomic<bool> ready[MAX_DEVICES];
    const int dslot = device_slot();
    if (!ready[dslot].load(std::memory_order_acquire)) {
        if (set_gauss()) { set_error("lodgs_l1_ssim_loss: uploading the window failed"); return LODGS_E_CUDA; }
        ready[dslot].store(true, std::memory_order_release);
    }
    cudaStream_t s = (cudaStream_t)stream;
    prof_stage_begin(s);
    const size_t map = align_up((size_t)C * H * W * 4, 128);
    SsimParams p;
    p.img = img; p.gt = gt; p.C = C; p.H = H; p.W = W;
    p.dmu = (float*)ws; p.ds = (float*)((char*)ws + map); p.dc = (float*)((char*)ws + 2 * map); p.partial = (double*)((char*)ws + 3 * map);
    const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C), block(SS_T, SS_T, 1);
    ssim_fwd_kernel<<<grid, block, 0, s>>>(p);
    LODGS_LAUNCHED("ssim_fwd_kernel");
    const double n = (double)C * H * W;
    ssim_sum_kernel<<<1, 256, 0, s>>>(p.partial, (int)ssim_blocks(C, H, W), n, a_l1, a_ssim, bias, out3);
    LODGS_LAUNCHED("ssim_sum_kernel");
    if (dL_dimg) {
        SsimBwdParams b;
        b.img = img; b.gt = gt; b.dmu = p.dmu; b.ds = p.ds; b.dc = p.dc; b.C = C; b.H = H; b.W = W;
        b.g_l1 = (float)(a_l1 / n); b.g_ssim = (float)(a_ssim / n); b.dL_dimg = dL_dimg;
        ssim_bwd_kernel<<<grid, block, 0, s>>>(b);
        LODGS_LAUNCHED("ssim_bwd_kernel");
    }
    return LODGS_OK;
}
