// K2: bf16 tensor-core GEMM for sm_100a:  C[M,N] = epilogue(A[M,K] . W[N,K]^T + bias)
//
// Replaces the four candle `Linear` layers of one BERT layer (candle-transformers 0.9.1 bert.rs:
// BertSelfAttention query/key/value, BertSelfOutput.dense, BertIntermediate.dense, BertOutput.dense;
// called from /root/reference/services/preprocessing_service/src/embedding_generator.rs:198),
// which the reference runs as f32 `gemm`-crate matmuls followed by separate bias / GELU /
// residual / LayerNorm passes.
//
// Design (one persistent CTA per SM, 12 warps):
//   warp 0      TMA producer: A tile 128x64 and W tile BLOCK_Nx64 (bf16, 128B swizzle) per stage
//   warp 1      one thread issues tcgen05.mma (128 x UMMA_N x 16, fp32 accumulators in TMEM)
//   warp 2      TMEM allocator
//   warps 4-11  epilogue: tcgen05.ld -> bias / activation / residual + LayerNorm -> global
// Accumulators are double-buffered in TMEM when 2*BLOCK_N <= 512 columns so the epilogue of tile i
// overlaps the MMAs of tile i+1. With BLOCK_N == N (<= 384) a CTA owns whole output rows and the
// residual add + LayerNorm run in the epilogue (EPI_RES_LN), so the pre-LN activations never
// reach HBM. Both operands are K-major ("TN"): activations [M,K] and HF weights [N,K] as stored.
#include "gemm_tc.h"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace b200 {

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARP0 = 4;
constexpr int EPI_THREADS = 256;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;

__host__ __device__ constexpr int b_stage_bytes(int block_n) { return block_n * BLOCK_K * 2; }
__host__ __device__ constexpr int num_stages(int block_n) {
    int s = (196 * 1024) / (A_STAGE_BYTES + b_stage_bytes(block_n));
    return s > 8 ? 8 : s;
}
// barriers + tmem pointer + LayerNorm partial statistics
constexpr int AUX_BYTES = 256 + 2 * BLOCK_M * 8;
__host__ __device__ constexpr int smem_bytes(int block_n) {
    return 1024 + num_stages(block_n) * (A_STAGE_BYTES + b_stage_bytes(block_n)) + AUX_BYTES;
}

struct GemmDev {
    int m;              // rows covered by the tensor map
    const int* d_rows;  // optional device-side row count (<= m)
    int n, k;
    const float* bias;
    const __nv_bfloat16* residual;     // [m, n] (EPI_RES_LN): high half of the residual stream
    const __nv_bfloat16* residual_lo;  // low half (residual = hi + lo) or nullptr
    const float* gamma;
    const float* beta;
    float eps;
    void* out;              // [m, n] bf16 or fp32
    __nv_bfloat16* out_lo;  // EPI_RES_LN: bf16(y - bf16(y)) so that the next residual add sees ~16 mantissa bits
};

template <typename OutT>
__device__ __forceinline__ void store_chunk32(OutT* dst, const float (&f)[32]);
template <>
__device__ __forceinline__ void store_chunk32<__nv_bfloat16>(__nv_bfloat16* dst, const float (&f)[32]) {
    uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(f[q * 8 + 0], f[q * 8 + 1]);
        u.y = pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
        u.z = pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]);
        u.w = pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
        d4[q] = u;
    }
}
template <>
__device__ __forceinline__ void store_chunk32<float>(float* dst, const float (&f)[32]) {
    float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
    for (int q = 0; q < 8; ++q) d4[q] = make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
}

// lo[j] = bf16(f[j] - bf16(f[j])): the part of the fp32 value the bf16 "hi" tensor dropped
__device__ __forceinline__ void store_chunk32_lo(__nv_bfloat16* dst, const float (&f)[32]) {
    uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float a = f[q * 8 + 2 * j], b = f[q * 8 + 2 * j + 1];
            const float2 h = unpack_bf16x2(pack_bf16x2(a, b));
            w[j] = pack_bf16x2(a - h.x, b - h.y);
        }
        d4[q] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

template <int BLOCK_N, int EPI, typename OutT>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmDev p) {
    constexpr int NSPLIT = BLOCK_N > 256 ? 2 : 1;
    constexpr int UMMA_N = BLOCK_N / NSPLIT;
    constexpr int ACC_STAGES = (2 * BLOCK_N <= 512) ? 2 : 1;
    constexpr int B_STAGE_BYTES = b_stage_bytes(BLOCK_N);
    constexpr int STAGES = num_stages(BLOCK_N);
    constexpr int HALF_N = BLOCK_N / 2;  // columns per epilogue warp group
    static_assert(UMMA_N % 16 == 0 && UMMA_N <= 256, "invalid UMMA N");
    static_assert(HALF_N % 32 == 0, "epilogue works in 32-column chunks");

    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    const uint32_t base = (raw_addr + 1023u) & ~1023u;
    uint8_t* base_ptr = smem_raw + (base - raw_addr);
    const uint32_t smem_a = base;
    const uint32_t smem_b = base + STAGES * A_STAGE_BYTES;
    const uint32_t aux = base + STAGES * (A_STAGE_BYTES + B_STAGE_BYTES);
    uint8_t* aux_ptr = base_ptr + STAGES * (A_STAGE_BYTES + B_STAGE_BYTES);
    const uint32_t bar_full = aux;                // STAGES x 8 B
    const uint32_t bar_empty = aux + 64;          // STAGES x 8 B
    const uint32_t bar_tfull = aux + 128;         // 2 x 8 B
    const uint32_t bar_tempty = aux + 144;        // 2 x 8 B
    const uint32_t tmem_slot = aux + 160;         // 4 B
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(aux_ptr + 160);
    float2* ln_part = reinterpret_cast<float2*>(aux_ptr + 256);  // [2][BLOCK_M]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    int rows = p.m;
    if (p.d_rows != nullptr) {
        const int r = *p.d_rows;
        rows = r < rows ? r : rows;
    }
    const int num_m_tiles = (rows + BLOCK_M - 1) / BLOCK_M;
    const int num_n_tiles = p.n / BLOCK_N;
    const int num_tiles = num_m_tiles * num_n_tiles;
    const int num_kb = (p.k + BLOCK_K - 1) / BLOCK_K;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_tfull + 8 * s, 1);
            mbar_init(bar_tempty + 8 * s, EPI_THREADS);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m_tile = t / num_n_tiles, n_tile = t % num_n_tiles;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    mbar_arrive_expect_tx(bar_full + 8 * stage, A_STAGE_BYTES + B_STAGE_BYTES);
                    tma_load_2d(smem_a + stage * A_STAGE_BYTES, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M,
                                bar_full + 8 * stage);
#pragma unroll
                    for (int s = 0; s < NSPLIT; ++s)
                        tma_load_2d(smem_b + stage * B_STAGE_BYTES + s * (UMMA_N * BLOCK_K * 2), &tmap_b,
                                    kb * BLOCK_K, n_tile * BLOCK_N + s * UMMA_N, bar_full + 8 * stage);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, UMMA_N);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const int acc = it % ACC_STAGES;
                const uint32_t acc_phase = (it / ACC_STAGES) & 1;
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after_sync();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after_sync();
                    const uint64_t desc_a = umma_desc_sw128(smem_a + stage * A_STAGE_BYTES);
                    const uint64_t desc_b = umma_desc_sw128(smem_b + stage * B_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
#pragma unroll
                        for (int s = 0; s < NSPLIT; ++s) {
                            // +2 per 16-element K step (32 B >> 4); second N split starts UMMA_N rows later
                            umma_bf16(tmem_d + s * UMMA_N, desc_a + 2 * k,
                                      desc_b + 2 * k + ((s * UMMA_N * BLOCK_K * 2) >> 4), idesc,
                                      (kb > 0 || k > 0) ? 1u : 0u);
                        }
                    }
                    umma_commit(bar_empty + 8 * stage);  // frees the smem stage when these MMAs finish
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(bar_tfull + 8 * acc);  // accumulator complete
            }
        }
    } else if (warp >= EPI_WARP0) {
        // ------------------------------ epilogue ----------------------------------
        const int quad = warp & 3;               // TMEM lane quadrant this warp may access
        const int half = (warp - EPI_WARP0) >> 2;  // which half of the tile's columns
        const int row_in_tile = quad * 32 + lane;
        OutT* out = reinterpret_cast<OutT*>(p.out);
        int it = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            const int m_tile = t / num_n_tiles, n_tile = t % num_n_tiles;
            const int acc = it % ACC_STAGES;
            const uint32_t acc_phase = (it / ACC_STAGES) & 1;
            const int row = m_tile * BLOCK_M + row_in_tile;
            const bool row_ok = row < rows;
            const int col_base = n_tile * BLOCK_N + half * HALF_N;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BLOCK_N + half * HALF_N;
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after_sync();

            if constexpr (EPI != EPI_RES_LN) {
#pragma unroll 1
                for (int c = 0; c < HALF_N / 32; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c * 32, v);
                    tmem_ld_wait();
                    const int col0 = col_base + c * 32;
                    float f[32];
                    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 b = __ldg(b4 + q);
                        f[q * 4 + 0] = __uint_as_float(v[q * 4 + 0]) + b.x;
                        f[q * 4 + 1] = __uint_as_float(v[q * 4 + 1]) + b.y;
                        f[q * 4 + 2] = __uint_as_float(v[q * 4 + 2]) + b.z;
                        f[q * 4 + 3] = __uint_as_float(v[q * 4 + 3]) + b.w;
                    }
                    if constexpr (EPI == EPI_GELU_ERF || EPI == EPI_GELU_TANH || EPI == EPI_RELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = activate(f[j], EPI - EPI_GELU_ERF);
                    }
                    if (row_ok) store_chunk32<OutT>(out + static_cast<size_t>(row) * p.n + col0, f);
                }
            } else {
                // pass 1: v = acc + bias + residual, kept in TMEM; row sum and sum of squares
                float sum = 0.f, sumsq = 0.f;
                const __nv_bfloat16* res = p.residual + static_cast<size_t>(row_ok ? row : 0) * p.n;
                const __nv_bfloat16* res_lo =
                    p.residual_lo ? p.residual_lo + static_cast<size_t>(row_ok ? row : 0) * p.n : nullptr;
#pragma unroll 1
                for (int c = 0; c < HALF_N / 32; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c * 32, v);
                    tmem_ld_wait();
                    const int col0 = col_base + c * 32;
                    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
                    const uint4* r4 = reinterpret_cast<const uint4*>(res + col0);
                    const uint4* l4 = reinterpret_cast<const uint4*>(res_lo + col0);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 r = __ldg(r4 + q);
                        const float4 b0 = __ldg(b4 + 2 * q), b1 = __ldg(b4 + 2 * q + 1);
                        float2 r0 = unpack_bf16x2(r.x), r1 = unpack_bf16x2(r.y), r2 = unpack_bf16x2(r.z),
                               r3 = unpack_bf16x2(r.w);
                        if (res_lo != nullptr) {
                            const uint4 l = __ldg(l4 + q);
                            const float2 l0 = unpack_bf16x2(l.x), l1 = unpack_bf16x2(l.y), l2 = unpack_bf16x2(l.z),
                                         l3 = unpack_bf16x2(l.w);
                            r0.x += l0.x; r0.y += l0.y; r1.x += l1.x; r1.y += l1.y;
                            r2.x += l2.x; r2.y += l2.y; r3.x += l3.x; r3.y += l3.y;
                        }
                        float x[8];
                        x[0] = __uint_as_float(v[q * 8 + 0]) + b0.x + r0.x;
                        x[1] = __uint_as_float(v[q * 8 + 1]) + b0.y + r0.y;
                        x[2] = __uint_as_float(v[q * 8 + 2]) + b0.z + r1.x;
                        x[3] = __uint_as_float(v[q * 8 + 3]) + b0.w + r1.y;
                        x[4] = __uint_as_float(v[q * 8 + 4]) + b1.x + r2.x;
                        x[5] = __uint_as_float(v[q * 8 + 5]) + b1.y + r2.y;
                        x[6] = __uint_as_float(v[q * 8 + 6]) + b1.z + r3.x;
                        x[7] = __uint_as_float(v[q * 8 + 7]) + b1.w + r3.y;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            sum += x[j];
                            sumsq = fmaf(x[j], x[j], sumsq);
                            v[q * 8 + j] = __float_as_uint(x[j]);
                        }
                    }
                    tmem_st_32x32(taddr + c * 32, v);
                }
                tmem_st_wait();
                ln_part[half * BLOCK_M + row_in_tile] = make_float2(sum, sumsq);
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
                const float2 other = ln_part[(half ^ 1) * BLOCK_M + row_in_tile];
                const float inv_n = 1.0f / static_cast<float>(p.n);
                const float mean = (sum + other.x) * inv_n;
                const float var = fmaxf((sumsq + other.y) * inv_n - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.eps);
                // pass 2: normalise, scale, shift, store
#pragma unroll 1
                for (int c = 0; c < HALF_N / 32; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c * 32, v);
                    tmem_ld_wait();
                    const int col0 = col_base + c * 32;
                    const float4* g4 = reinterpret_cast<const float4*>(p.gamma + col0);
                    const float4* be4 = reinterpret_cast<const float4*>(p.beta + col0);
                    float f[32];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 g = __ldg(g4 + q), be = __ldg(be4 + q);
                        f[q * 4 + 0] = (__uint_as_float(v[q * 4 + 0]) - mean) * rstd * g.x + be.x;
                        f[q * 4 + 1] = (__uint_as_float(v[q * 4 + 1]) - mean) * rstd * g.y + be.y;
                        f[q * 4 + 2] = (__uint_as_float(v[q * 4 + 2]) - mean) * rstd * g.z + be.z;
                        f[q * 4 + 3] = (__uint_as_float(v[q * 4 + 3]) - mean) * rstd * g.w + be.w;
                    }
                    if (row_ok) {
                        store_chunk32<OutT>(out + static_cast<size_t>(row) * p.n + col0, f);
                        if (p.out_lo != nullptr) store_chunk32_lo(p.out_lo + static_cast<size_t>(row) * p.n + col0, f);
                    }
                }
                // the partial-stat slots are reused by the next tile
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
            }
            tc_fence_before_sync();
            mbar_arrive(bar_tempty + 8 * acc);
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
    });
    return fn;
}

// bf16 row-major [rows, cols] matrix, box = 64 columns x box_rows rows, 128-byte swizzle.
bool make_tmap(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    auto fn = get_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {BLOCK_K, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int BLOCK_N, int EPI, typename OutT>
cudaError_t launch(const GemmTcArgs& a, cudaStream_t stream, int num_sms) {
    constexpr int NSPLIT = BLOCK_N > 256 ? 2 : 1;
    alignas(64) CUtensorMap ta, tb;
    if (!make_tmap(&ta, a.a, a.m, a.k, BLOCK_M)) return cudaErrorInvalidValue;
    if (!make_tmap(&tb, a.b, a.n, a.k, BLOCK_N / NSPLIT)) return cudaErrorInvalidValue;
    auto kern = gemm_tc_kernel<BLOCK_N, EPI, OutT>;
    constexpr int smem = smem_bytes(BLOCK_N);
    static bool attr_set = false;  // per instantiation; benign race (idempotent)
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    GemmDev p;
    p.m = a.m; p.d_rows = a.d_rows; p.n = a.n; p.k = a.k; p.bias = a.bias;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(a.residual);
    p.residual_lo = reinterpret_cast<const __nv_bfloat16*>(a.residual_lo);
    p.out_lo = reinterpret_cast<__nv_bfloat16*>(a.out_lo);
    p.gamma = a.gamma; p.beta = a.beta; p.eps = a.eps; p.out = a.out;
    const int tiles = ((a.m + BLOCK_M - 1) / BLOCK_M) * (a.n / BLOCK_N);
    const int grid = tiles < num_sms ? tiles : num_sms;
    if (grid <= 0) return cudaSuccess;
    kern<<<grid, NUM_THREADS, smem, stream>>>(ta, tb, p);
    return cudaGetLastError();
}

}  // namespace

int gemm_tc_pick_block_n(int n, int epi) {
    if (epi == EPI_RES_LN) return (n == 128 || n == 256 || n == 384) ? n : 0;
    if (n % 256 == 0) return 256;
    if (epi == EPI_BIAS && n % 192 == 0) return 192;
    if (n % 128 == 0) return 128;
    return 0;
}

cudaError_t gemm_tc_launch(const GemmTcArgs& a, cudaStream_t stream, int num_sms) {
    if (a.m <= 0) return cudaSuccess;
    if (a.k % 8 != 0 || a.n % a.block_n != 0) return cudaErrorInvalidValue;
#define B200_CASE(BN, EP, T) \
    if (a.block_n == BN && a.epi == EP && a.out_f32 == (sizeof(T) == 4)) return launch<BN, EP, T>(a, stream, num_sms);
    B200_CASE(128, EPI_BIAS, __nv_bfloat16)
    B200_CASE(192, EPI_BIAS, __nv_bfloat16)
    B200_CASE(256, EPI_BIAS, __nv_bfloat16)
    B200_CASE(128, EPI_BIAS, float)
    B200_CASE(256, EPI_BIAS, float)
    B200_CASE(128, EPI_GELU_ERF, __nv_bfloat16)
    B200_CASE(256, EPI_GELU_ERF, __nv_bfloat16)
    B200_CASE(128, EPI_GELU_TANH, __nv_bfloat16)
    B200_CASE(256, EPI_GELU_TANH, __nv_bfloat16)
    B200_CASE(128, EPI_RELU, __nv_bfloat16)
    B200_CASE(256, EPI_RELU, __nv_bfloat16)
    B200_CASE(128, EPI_RES_LN, __nv_bfloat16)
    B200_CASE(256, EPI_RES_LN, __nv_bfloat16)
    B200_CASE(384, EPI_RES_LN, __nv_bfloat16)
#undef B200_CASE
    return cudaErrorInvalidValue;
}

}  // namespace b200
