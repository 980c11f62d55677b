// micro-probe: issue cost of scalar FFMA pairs vs packed FFMA2 mixed with integer work (not part of the product)
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long f2;
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
template <int MODE>
__global__ void k(float* out, int iters, float seed) {
    float a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float m = 1.0001f, c = 0.5f;
    int acc = threadIdx.x;
    if (MODE == 0) {
        for (int i = 0; i < iters; i++) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
            acc = acc * 3 + i; acc ^= acc >> 3; acc += i; acc ^= acc << 2;
        }
    } else {
        f2 p0, p1, p2, p3, mm, cc;
        asm("mov.b64 %0, {%1,%2};" : "=l"(p0) : "f"(a0), "f"(a1));
        asm("mov.b64 %0, {%1,%2};" : "=l"(p1) : "f"(a2), "f"(a3));
        asm("mov.b64 %0, {%1,%2};" : "=l"(p2) : "f"(a4), "f"(a5));
        asm("mov.b64 %0, {%1,%2};" : "=l"(p3) : "f"(a6), "f"(a7));
        asm("mov.b64 %0, {%1,%2};" : "=l"(mm) : "f"(m), "f"(m));
        asm("mov.b64 %0, {%1,%2};" : "=l"(cc) : "f"(c), "f"(c));
        for (int i = 0; i < iters; i++) {
            p0 = fma2(p0, mm, cc); p1 = fma2(p1, mm, cc); p2 = fma2(p2, mm, cc); p3 = fma2(p3, mm, cc);
            acc = acc * 3 + i; acc ^= acc >> 3; acc += i; acc ^= acc << 2;
        }
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a0), "=f"(a1) : "l"(p0));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a2), "=f"(a3) : "l"(p1));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a4), "=f"(a5) : "l"(p2));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a6), "=f"(a7) : "l"(p3));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + acc;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(out, 20000, 1.0f); else k<1><<<148 * 8, 256>>>(out, 20000, 1.0f);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep == 2) printf("mode %d (%s): %.3f ms\n", mode, mode ? "4 FFMA2 + 6 int" : "8 FFMA + 6 int", ms);
        }
    }
    return 0;
}
