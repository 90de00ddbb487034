// Probe: tensor-memory read bandwidth per SM. W warps (W = 4, 8, 12; warp w reads lane quadrant w % 4) issue
// tcgen05.ld.32x32b.x32 (4 KB per warp instruction) back to back; prints bytes per clock per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o _bin/ldtm_rate ldtm_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters, long long* cycles) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t r[32];
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                : "r"(base + c * 32 + (it & 3) * 128)
                : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            acc += __uint_as_float(r[(it + c) & 31] & 0x3fffffffu);
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}
int main() {
    float* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
    const int iters = 2048;
    for (int warps : {4, 8, 12, 16}) {
        for (int rep = 0; rep < 2; ++rep) { k<<<148, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize(); }
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        double bytes = double(iters) * 4 * warps * 4096;
        printf("warps/SM %2d: %.1f B/clk/SM (%.1f cycles per x32 load per warp)  %s\n", warps, bytes / c, double(c) / (iters * 4),
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
