// Probe: issue throughput of FFMA vs FFMA2 (fma.rn.f32x2) per SM, 8 warps per SMSP-friendly block.
#include <cuda_runtime.h>
#include <cstdio>
template <int MODE>
__global__ void __launch_bounds__(512) k(float* out, int iters) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float m = 0.999f, c = 0.001f;
    unsigned long long p0, p1, p2, p3, pm, pc;
    asm("mov.b64 %0, {%1, %2};" : "=l"(p0) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(p1) : "f"(a2), "f"(a3));
    asm("mov.b64 %0, {%1, %2};" : "=l"(p2) : "f"(a4), "f"(a5));
    asm("mov.b64 %0, {%1, %2};" : "=l"(p3) : "f"(a6), "f"(a7));
    asm("mov.b64 %0, {%1, %1};" : "=l"(pm) : "f"(m));
    asm("mov.b64 %0, {%1, %1};" : "=l"(pc) : "f"(c));
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            if (MODE == 0) {
                a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
                a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
            } else {
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p0) : "l"(pm), "l"(pc));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p1) : "l"(pm), "l"(pc));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2) : "l"(pm), "l"(pc));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p3) : "l"(pm), "l"(pc));
            }
        }
    }
    float r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    float x, y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(p0)); r += x + y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(p1)); r += x + y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(p2)); r += x + y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(p3)); r += x + y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run() {
    float* d; cudaMalloc(&d, 148 * 2 * 512 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    k<MODE><<<148 * 2, 512>>>(d, iters); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<MODE><<<148 * 2, 512>>>(d, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fma = 148.0 * 2 * 512 * iters * 16 * 8;
    printf("%s: %.3f ms, %.1f TFLOP/s fp32 (%.1f FMA/clk/SM at 1.9 GHz)\n", MODE ? "FFMA2" : "FFMA ", ms, 2 * fma / (ms * 1e-3) / 1e12,
           fma / (ms * 1e-3) / 148 / 1.9e9);
}
int main() { run<0>(); run<1>(); run<0>(); run<1>(); return 0; }
