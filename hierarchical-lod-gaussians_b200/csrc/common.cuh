// common.cuh — shared helpers for the lodgs kernels (sm_100a only).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/lodgs.h"

#define TILE_X 16
#define TILE_Y 16
#define TILE_PIX 256

namespace lodgs {

void set_error(const char* fmt, ...);
void count_launch();
int check_cuda(cudaError_t e, const char* what);
void prof_stage_begin(cudaStream_t s);
void prof_mark(const char* name);
// Small device->host results (counts the caller sizes buffers from) are published by a store into a mapped pinned word,
// not by a cudaMemcpy: a device-to-host copy queues on the D2H copy engine behind whatever bulk transfer the application has
// in flight. publish_words enqueues a one-thread kernel copying n <= 8 ints from `src_dev` and records an event;
// wait_words blocks the host on that event and returns the host copy (thread-local, valid until the next publish).
int publish_words(const int32_t* src_dev, int n, cudaStream_t s);
// One-time per-DEVICE setup (function attributes, occupancy queries and events belong to a device, and a process may drive several):
// slot of the current device for the small per-device tables below; the tables are atomics, so concurrent host threads are safe
// (setting an attribute twice is harmless).
constexpr int MAX_DEVICES = 64;
static inline int device_slot() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= MAX_DEVICES) d = 0;
    return d;
}
const int32_t* wait_words();

#define LODGS_CUDA(call)                                              \
    do {                                                              \
        cudaError_t e__ = (call);                                     \
        if (e__ != cudaSuccess) return ::lodgs::check_cuda(e__, #call); \
    } while (0)

#define LODGS_LAUNCHED(name)                                          \
    do {                                                              \
        ::lodgs::count_launch();                                      \
        ::lodgs::prof_mark(name);                                     \
        cudaError_t e__ = cudaGetLastError();                         \
        if (e__ != cudaSuccess) return ::lodgs::check_cuda(e__, name); \
    } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- 3x3 column-major matrix with glm-compatible semantics: m[c][r], product
// R[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2] summed left to right. Keeping the same
// expression trees as the reference's glm calls lets nvcc contract them into the same FMA chains, which
// is what makes radii / rects / tile counts bit-identical (SURVEY §7 hard part 1).
struct mat3 {
    float m[3][3];
    __host__ __device__ mat3() {}
    __host__ __device__ mat3(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1, float c2) {
        m[0][0] = a0; m[0][1] = a1; m[0][2] = a2;
        m[1][0] = b0; m[1][1] = b1; m[1][2] = b2;
        m[2][0] = c0; m[2][1] = c1; m[2][2] = c2;
    }
    __host__ __device__ float* operator[](int c) { return m[c]; }
    __host__ __device__ const float* operator[](int c) const { return m[c]; }
};

__host__ __device__ inline mat3 mul(const mat3& a, const mat3& b) {
    mat3 r;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) r.m[c][k] = a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1] + a.m[2][k] * b.m[c][2];
    return r;
}
__host__ __device__ inline mat3 transpose(const mat3& a) {
    mat3 r;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) r.m[c][k] = a.m[k][c];
    return r;
}

// ---- decoupled look-back (single-pass chained scan) state word: {flag:2 | value:62} packed in u64.
// flag 0 = empty, 1 = aggregate available, 2 = inclusive prefix available.
__device__ __forceinline__ unsigned long long pack_state(unsigned flag, unsigned long long v) {
    return (static_cast<unsigned long long>(flag) << 62) | v;
}
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Look back from tile `tile` (>0) for the exclusive prefix; called by one full warp. Returns the exclusive prefix to all
// lanes. Every lane inspects LB_PER_LANE consecutive predecessors per step (1: a wider window was measured and only adds
// polling traffic on the few L2 lines that hold the youngest states).
constexpr int LB_PER_LANE = 1;
__device__ __forceinline__ unsigned long long lookback_warp(const unsigned long long* state, int tile) {
    const unsigned lane = threadIdx.x & 31;
    unsigned long long exclusive = 0;
    int base = tile - 1;
    while (true) {
        // predecessor number p = lane*LB_PER_LANE + k of this window is tile base - p (p = 0 is the nearest)
        unsigned long long local = 0;
        bool found = false;
#pragma unroll
        for (int k = 0; k < LB_PER_LANE; k++) {
            const int idx = base - ((int)lane * LB_PER_LANE + k);
            unsigned long long s = pack_state(2, 0);  // virtual tile before 0: inclusive prefix 0
            if (idx >= 0) {
                do { s = ld_relaxed(state + idx); } while ((s >> 62) == 0);
            }
            if (!found) {
                local += s & ((1ull << 62) - 1);
                found = (s >> 62) == 2;
            }
        }
        const unsigned incl_mask = __ballot_sync(0xffffffffu, found);
        // lanes before the first one that met an inclusive prefix contribute all their aggregates, that lane its partial sum
        const int first = incl_mask ? (__ffs(incl_mask) - 1) : 32;
        unsigned long long contrib = ((int)lane <= first) ? local : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
        exclusive += contrib;
        if (incl_mask) break;
        base -= 32 * LB_PER_LANE;
    }
    return exclusive;
}

// ---- packed FP32 pairs (sm_100: FFMA2 / FMUL2 / FADD2 take one issue slot for two lanes of work). mov.b64 {x,x} folds into
// the instruction as a broadcast operand. NOTE: ptxas contracts a packed mul feeding a packed add into one FFMA2 even when both
// carry .rn — code whose rounding is pinned must only use sequences where that cannot happen (explicit fma2, mul feeding mul/fma).
typedef unsigned long long f2;
__device__ __forceinline__ f2 pk(float lo, float hi) {
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f2 bc(float x) { return pk(x, x); }
__device__ __forceinline__ float lo(f2 v) {
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
}
__device__ __forceinline__ float hi(f2 v) {
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
}
__device__ __forceinline__ float hsum(f2 v) { return lo(v) + hi(v); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    f2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
// c = a * b + c with the accumulator tied to one register pair (keeps ptxas from computing into a temporary and copying back).
__device__ __forceinline__ void fma2_acc(f2& c, f2 a, f2 b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(a), "l"(b)); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    f2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    f2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier + 1-D bulk async copy (TMA unit; SASS: UBLKCP)
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
// shared -> global bulk store (TMA unit), tracked by the per-thread bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_addr(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit_wait_all() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// make generic-proxy writes to shared memory visible to the async proxy (required before a bulk store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// ---- 16-byte async copies (LDGSTS) for gathers; completion is reported to an mbarrier: every participating thread
// calls cp_async_arrive once per phase (the barrier is initialised with the number of participating threads).
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src_gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dst_smem)), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

}  // namespace lodgs
