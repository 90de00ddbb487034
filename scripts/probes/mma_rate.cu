// Probe (not part of the library): issue rate of tcgen05.mma kind::f16 for several N, cta_group::1 (M=128) and
// cta_group::2 (M=256 over a CTA pair). Operands are whatever is in shared memory; only time matters.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/mma_rate scripts/probes/mma_rate.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "../../codename-symbiont_b200/csrc/common.cuh"
using namespace b200;

__device__ __forceinline__ void umma2_bf16(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma2_commit(uint32_t bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void csync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int CG, int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(long long* out, int iters, int nacc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t base = smem_u32(smem);
    // A: 128 rows x 64 k (16 KB) x 4 stages; B: N/CG rows x 64 k x 4 stages
    const uint32_t sa = base, sb = base + 4 * 16384, bar = sb + 4 * 32768, slot = bar + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (4 * 16384 + 4 * 32768) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 0) {
        if (CG == 1) tmem_alloc(slot, 512);
        else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    if (CG == 2) csync();
    tc_fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + (slot - base));
    const bool leader = CG == 1 || ctarank() == 0;
    long long t0 = 0, t1 = 0;
    if (warp == 1 && lane == 0 && leader) {
        constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, N);
        t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            const int st = i & 3;
            const uint64_t da = umma_desc_sw128(sa + st * 16384), db = umma_desc_sw128(sb + st * 32768);
            const uint32_t d = tmem + (nacc > 1 ? (i % nacc) * N : 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) { if (CG == 1) umma_bf16(d, da + 2 * k, db + 2 * k, idesc, 1u); else umma2_bf16(d, da + 2 * k, db + 2 * k, idesc, 1u); }
        }
        if (CG == 1) umma_commit(bar); else umma2_commit(bar, 3);
    }
    if (CG == 2 || true) {
        if (warp == 1 && lane == 0) {
            if (leader) { mbar_wait(bar, 0); t1 = clock64(); out[blockIdx.x] = t1 - t0; }
            else if (CG == 2) mbar_wait(bar, 0);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (CG == 2) csync();
    if (warp == 0) {
        if (CG == 1) tmem_dealloc(tmem, 512);
        else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

template <int CG, int N>
void run(int grid, int iters, int nacc) {
    long long* d; cudaMalloc(&d, sizeof(long long) * 296); cudaMemset(d, 0, sizeof(long long) * 296);
    auto kern = rate_kernel<CG, N>;
    const int smem = 4 * 16384 + 4 * 32768 + 64;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        cudaError_t err = cudaLaunchKernelEx(&cfg, kern, d, iters, nacc);
        cudaEventRecord(e1);
        if (err != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { printf("CG %d N %d: error %s\n", CG, N, cudaGetErrorString(cudaGetLastError())); return; }
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[296]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    const double per = double(mx) / (iters * 4.0);
    const double flops = 2.0 * 128 * CG * N * 16 * 4.0 * iters * (grid / CG);
    printf("cta_group %d M %3d N %3d grid %3d nacc %d: %7.1f cyc/MMA (ideal %5.1f) -> %5.1f%% of rate; %7.1f TF/s by events\n", CG, 128 * CG, N, grid, nacc,
           per, N / 2.0, 100.0 * (N / 2.0) / per, flops / (ms * 1e-3) / 1e12);
    cudaFree(d);
}
int main() {
    for (int grid : {2, 148}) {
        run<1, 32>(grid, 4096, 1); run<1, 64>(grid, 4096, 1); run<1, 128>(grid, 4096, 1); run<1, 192>(grid, 4096, 1); run<1, 256>(grid, 4096, 1);
        run<1, 128>(grid, 4096, 2); run<1, 256>(grid, 4096, 2);
        run<2, 64>(grid, 4096, 1); run<2, 128>(grid, 4096, 1); run<2, 192>(grid, 4096, 1); run<2, 256>(grid, 4096, 1);
        run<2, 128>(grid, 4096, 2); run<2, 256>(grid, 4096, 2);
    }
    return 0;
}
