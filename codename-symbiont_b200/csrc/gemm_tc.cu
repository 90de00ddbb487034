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
//   warps 4-11  epilogue. Warp w owns TMEM lane quadrant w%4 (32 rows) and one half of the tile's
//               columns, and walks them in 32-column chunks: tcgen05.ld -> math in registers ->
//               swizzled shared-memory staging -> one TMA store per chunk. A thread holds one row,
//               so writing global memory directly would touch 32 rows x 16 B per instruction
//               (ncu r01a: 32 sectors per request); the TMA store writes full 64 B row segments.
// Accumulators are double-buffered in TMEM when 2*BLOCK_N <= 512 columns so the epilogue of tile i
// overlaps the MMAs of tile i+1. With BLOCK_N == N (<= 384) a CTA owns whole output rows and the
// residual add + LayerNorm run in the epilogue (EPI_RES_LN): pass 1 adds bias and the residual
// (TMA-loaded per chunk into a per-warp ring, prefetched one tile ahead), parks the sum back in TMEM
// and accumulates the row statistics; pass 2 normalises and stores. The pre-LN activations never
// reach HBM. For N = 384 / 768 the row is split over a thread-block CLUSTER of 2 / 4 CTAs (192 columns each, so
// the TMEM accumulators stay double-buffered and the smem stages small); the CTAs of a cluster walk the same row
// blocks and exchange their per-row partial (sum, sum of squares) through distributed shared memory.
// Shared operand tiles are loaded once per cluster with TMA multicast: the CTAs of a LayerNorm cluster share the A
// tile (each loads 1/CL of its rows and multicasts), and the plain GEMMs run as pairs of CTAs on adjacent row
// blocks that share the W tile (each loads half of it). Both halve that operand's L2 -> SM traffic, which is what
// bounds 128-row tiles at K = 384 (12 TB/s of L2 reads at 930 TFLOP/s, profiles/r01b). A stage is released to the
// producers of all sharing CTAs by a multicast tcgen05.commit.
// Both operands are K-major ("TN"): activations [M,K] and HF weights [N,K] as stored.
#include "gemm_tc.h"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tc_helpers.cuh"

namespace b200 {

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARP0 = 4;
constexpr int EPI_WARPS = 8;
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;

__host__ __device__ constexpr int b_stage_bytes(int block_n) { return block_n * BLOCK_K * 2; }
// per-warp epilogue shared memory: 2 output staging buffers (+ a `ring`-stage (hi, lo) residual ring for LN)
__host__ __device__ constexpr int epi_warp_bytes(int epi, bool out_f32, int ring, bool wide = false) {
    return (epi == EPI_RES_LN ? ring * 2 * CHUNK_BF16 : 0) + 2 * ((out_f32 || wide) ? CHUNK_F32 : CHUNK_BF16);
}
// mbarriers (<= 8 + 8 stage, 4 TMEM, 16 residual, 2 cluster exchange) + TMEM pointer, then the LayerNorm partial
// statistics of the two column halves, then (clusters) the slots peers write their row partials into
constexpr int MAX_CLUSTER = 4;
constexpr int AUX_BARRIERS = 512;
constexpr int AUX_LN_PART = 2 * BLOCK_M * 8;
constexpr int AUX_LN_PEER = 2 * MAX_CLUSTER * BLOCK_M * 8;  // [tile parity][sender rank][row] float2
// plain epilogues instead keep the bias of the current and the next tile's columns there ([tile parity][BLOCK_N] fp32)
__host__ __device__ constexpr int aux_bytes(int epi, int block_n) {
    return AUX_BARRIERS + (epi == EPI_RES_LN ? AUX_LN_PART + AUX_LN_PEER : 2 * block_n * 4);
}
// kslots == 0: A and W stream through the same ring of stages. kslots > 0 ("resident A"): the whole 128 x K
// activation tile sits in `kslots` fixed 16 KB slots while the CTA walks every column tile of that row block, and
// only W streams through the ring.
// pair == 2 (one MMA over a CTA pair): each CTA stages only its half of the W tile's rows
__host__ __device__ constexpr int num_stages(int block_n, int epi, bool out_f32, int kslots, int ring, bool wide = false,
                                             int pair = 0) {
    int s = (SMEM_LIMIT - aux_bytes(epi, block_n) - EPI_WARPS * epi_warp_bytes(epi, out_f32, ring, wide) - kslots * A_STAGE_BYTES) /
            ((kslots ? 0 : A_STAGE_BYTES) + b_stage_bytes(pair == 2 ? block_n / 2 : block_n));
    return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int smem_bytes(int block_n, int epi, bool out_f32, int kslots, int ring, bool wide = false,
                                             int pair = 0) {
    return num_stages(block_n, epi, out_f32, kslots, ring, wide, pair) *
               ((kslots ? 0 : A_STAGE_BYTES) + b_stage_bytes(pair == 2 ? block_n / 2 : block_n)) +
           kslots * A_STAGE_BYTES + EPI_WARPS * epi_warp_bytes(epi, out_f32, ring, wide) + aux_bytes(epi, block_n);
}

// One 8-byte store into a peer CTA's shared memory that also counts 8 bytes on the peer's mbarrier
// (complete_tx): the data is visible to whoever observes that barrier phase complete. No fence, no remote arrive.
__device__ __forceinline__ void st_async_cluster_b64(uint32_t remote_addr, float a, float b, uint32_t remote_bar) {
    const unsigned long long v =
        static_cast<unsigned long long>(__float_as_uint(a)) | (static_cast<unsigned long long>(__float_as_uint(b)) << 32);
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr), "l"(v),
                 "r"(remote_bar)
                 : "memory");
}

struct GemmDev {
    int m;              // rows covered by the tensor maps
    const int* d_rows;  // optional device-side row count (<= m)
    int n, k;
    const float* bias;
    __nv_bfloat16* out_bf16;  // direct-store epilogue only
    const float* gamma;
    const float* beta;
    float eps;
    int has_res_lo, has_out_lo;
    int dbg;  // timing experiments only (B200_EMBED_GEMM_DBG): 1 no TMA stores, 2 no cluster exchange, 4 no pass-2 math
};

template <int BLOCK_N, int EPI, typename OutT, int CL, int PAIR, int KSLOTS, int RING, int NSP, int WST>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_c_lo,
               const __grid_constant__ CUtensorMap tmap_r, const __grid_constant__ CUtensorMap tmap_r_lo,
               const GemmDev p) {
    constexpr bool OUT_F32 = sizeof(OutT) == 4;
    constexpr int NSPLIT = NSP ? NSP : (BLOCK_N > 256 ? 2 : 1);  // independent accumulator column ranges per k-step
    constexpr int UMMA_N = BLOCK_N / NSPLIT;
    constexpr int ACC_STAGES = (2 * BLOCK_N <= 512) ? 2 : 1;
    constexpr bool CG2 = PAIR == 2;  // one cta_group::2 MMA per k-step over a pair of CTAs (256 x BLOCK_N per pair)
    constexpr int B_STAGE_BYTES = b_stage_bytes(CG2 ? BLOCK_N / 2 : BLOCK_N);
    constexpr bool WIDE = WST == 1;    // stage two chunks side by side and store 128-byte row segments
    constexpr bool DIRECT = WST == 2;  // no staging: every lane stores its 64 bytes of the chunk straight from registers
    constexpr int STAGES = num_stages(BLOCK_N, EPI, OUT_F32, KSLOTS, RING, WIDE, PAIR);
    static_assert(!CG2 || (EPI != EPI_RES_LN && !OUT_F32 && NSPLIT == 1 && KSLOTS == 0), "CTA-pair MMA: plain bf16 epilogues");
    constexpr bool RES_A = KSLOTS > 0;
    static_assert(!RES_A || (CL == 1 && PAIR == 0 && KSLOTS <= 8), "resident A is for the plain single-CTA GEMMs");
    constexpr int HALF_N = BLOCK_N / 2;  // columns per epilogue warp group
    constexpr int NCH = HALF_N / CHUNK;  // chunks per warp per tile
    constexpr int EPI_WARP_BYTES = epi_warp_bytes(EPI, OUT_F32, RING, WIDE);
    static_assert(!WIDE || (EPI != EPI_RES_LN && !OUT_F32 && NCH % 2 == 0), "wide stores: plain bf16 epilogues, even chunk count");
    constexpr bool KEEP_IN_REGS = NCH <= 3;  // LN: hold acc + bias + residual in registers between the two passes
    static_assert(RING >= 2 && RING <= 3, "residual ring depth");
    constexpr int OUT_CHUNK_BYTES = OUT_F32 ? CHUNK_F32 : CHUNK_BF16;
    static_assert(UMMA_N % 16 == 0 && UMMA_N <= 256, "invalid UMMA N");
    static_assert(HALF_N % CHUNK == 0, "epilogue works in 32-column chunks");
    static_assert(STAGES >= 2, "not enough shared memory for a pipeline");
    static_assert(EPI != EPI_RES_LN || !OUT_F32, "LayerNorm epilogue writes bf16 (hi, lo)");
    static_assert(CL == 1 || EPI == EPI_RES_LN, "row-splitting clusters are only used by the LayerNorm epilogue");
    static_assert(CL <= MAX_CLUSTER, "cluster too large");
    static_assert(PAIR == 0 || (CL == 1 && NSPLIT == 1), "W-sharing pairs are for the plain epilogues");
    constexpr int CSIZE = CL > 1 ? CL : (PAIR ? 2 : 1);       // CTAs per cluster
    constexpr uint16_t CMASK = static_cast<uint16_t>((1u << CSIZE) - 1);
    constexpr int A_SLICE_ROWS = BLOCK_M / (CL > 1 ? CL : 1);  // rows of the A tile this CTA loads (and multicasts)
    constexpr int B_SLICE_ROWS = BLOCK_N / (PAIR ? 2 : 1);     // rows of the W tile this CTA loads (and multicasts)

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    if ((base & 1023u) != 0) __trap();  // swizzled TMA / UMMA tiles need 1024-byte alignment
    const uint32_t smem_a = base;  // STAGES ring slots, or KSLOTS resident k-blocks
    const uint32_t smem_b = base + (RES_A ? KSLOTS : STAGES) * A_STAGE_BYTES;
    const uint32_t smem_epi = smem_b + STAGES * B_STAGE_BYTES;
    const uint32_t aux = smem_epi + EPI_WARPS * EPI_WARP_BYTES;
    uint8_t* aux_ptr = smem_raw + (aux - base);
    const uint32_t bar_full = aux;          // STAGES x 8 B
    const uint32_t bar_empty = aux + 64;    // STAGES x 8 B
    const uint32_t bar_tfull = aux + 128;   // 2 x 8 B
    const uint32_t bar_tempty = aux + 144;  // 2 x 8 B
    const uint32_t tmem_slot = aux + 160;   // 4 B
    const uint32_t bar_ln = aux + 176;      // 2 x 8 B: peers' row partials have arrived (per tile parity)
    const uint32_t bar_a_full = aux + 192;  // 8 x 8 B: resident A k-block landed
    const uint32_t bar_a_empty = aux + 256; // 8 x 8 B: resident A k-block no longer needed by this row block
    const uint32_t bar_res = aux + 320;     // EPI_WARPS x RING x 8 B (residual ring of each epilogue warp)
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(aux_ptr + 160);
    float2* ln_part = reinterpret_cast<float2*>(aux_ptr + AUX_BARRIERS);                // [2][BLOCK_M]
    float2* ln_peer = reinterpret_cast<float2*>(aux_ptr + AUX_BARRIERS + AUX_LN_PART);  // [2][MAX_CLUSTER][BLOCK_M]
    const uint32_t ln_peer_addr = aux + AUX_BARRIERS + AUX_LN_PART;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    int rows = p.m;
    if (p.d_rows != nullptr) {
        const int r = *p.d_rows;
        rows = r < rows ? r : rows;
    }
    const int num_m_tiles = (rows + BLOCK_M - 1) / BLOCK_M;
    const int num_n_tiles = p.n / BLOCK_N;  // == CL when the row is split over a cluster
    const int num_kb = (p.k + BLOCK_K - 1) / BLOCK_K;
    // Work units, walked identically by all CTAs of a cluster:
    //   plain:      unit = (m_tile, n_tile), one CTA each;
    //   CL > 1:     unit = m_tile, CTA `crank` takes columns [crank * BLOCK_N, +BLOCK_N) of it;
    //   PAIR:       unit = (pair of adjacent m_tiles, n_tile), CTA `crank` takes m_tile 2 * pair + crank. With an odd
    //               number of row blocks the last pair's second CTA works on an out-of-range block: its loads are
    //               zero-filled and its stores clipped, but it keeps its side of the shared-stage protocol.
    const int crank = CSIZE > 1 ? static_cast<int>(cluster_ctarank()) : 0;
    //   RES_A:      unit = m_tile; the CTA walks all n_tiles of it back to back with the A tile resident.
    const int num_units =
        (CL > 1 || RES_A) ? num_m_tiles : (PAIR ? ((num_m_tiles + 1) / 2) * num_n_tiles : num_m_tiles * num_n_tiles);
    const int unit0 = static_cast<int>(blockIdx.x) / CSIZE;
    const int unit_step = static_cast<int>(gridDim.x) / CSIZE;
    const int my_units = unit0 < num_units ? (num_units - unit0 + unit_step - 1) / unit_step : 0;
    const int my_tiles = RES_A ? my_units * num_n_tiles : my_units;  // output tiles this CTA produces, in order
    auto tile_m = [&](int i) {
        if constexpr (RES_A) return unit0 + (i / num_n_tiles) * unit_step;
        const int t = unit0 + i * unit_step;
        return CL > 1 ? t : (PAIR ? 2 * (t / num_n_tiles) + crank : t / num_n_tiles);
    };
    auto tile_n = [&](int i) {
        if constexpr (RES_A) return i % num_n_tiles;
        return CL > 1 ? crank : (unit0 + i * unit_step) % num_n_tiles;
    };

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        prefetch_tmap(&tmap_c);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, CG2 ? 1 : CSIZE);  // one multicast commit from every CTA that issues MMAs on the stage
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_tfull + 8 * s, 1);
            mbar_init(bar_tempty + 8 * s, CG2 ? 2 * EPI_WARPS : EPI_THREADS);  // pair: one arrive per epilogue warp of both CTAs
        }
        for (int s = 0; s < RING * EPI_WARPS; ++s) mbar_init(bar_res + 8 * s, 1);
        for (int s = 0; s < 2; ++s) mbar_init(bar_ln + 8 * s, 1);
        for (int s = 0; s < 8; ++s) {
            mbar_init(bar_a_full + 8 * s, 1);
            mbar_init(bar_a_empty + 8 * s, 1);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        if constexpr (CG2) tmem_alloc_cg2(tmem_slot, 512);
        else tmem_alloc(tmem_slot, 512);
    }
    tc_fence_before_sync();
    __syncthreads();
    if constexpr (CSIZE > 1) cluster_sync_all();  // peers' barriers exist before anyone arrives on them remotely
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            const uint64_t pol_w = l2_policy_evict_last();
            for (int i = 0; i < my_tiles; ++i) {
                const int m_tile = tile_m(i), n_tile = tile_n(i);
                if ((p.dbg & 16) && i + 2 < my_tiles) {  // A tile of the tile after next -> L2
                    const int pm = tile_m(i + 2);
                    if (CL > 1 || tile_n(i + 2) == 0 || (p.dbg & 128))
                        for (int kb = 0; kb < num_kb; ++kb) tma_prefetch_2d(&tmap_a, kb * BLOCK_K, pm * BLOCK_M);
                }
                for (int kb = 0; kb < num_kb; ++kb) {
                    if constexpr (RES_A) {
                        if (n_tile == 0) {  // new row block: (re)fill resident slot kb once the previous block let go of it
                            const uint32_t au = static_cast<uint32_t>(i / num_n_tiles);
                            mbar_wait(bar_a_empty + 8 * kb, (au & 1) ^ 1);
                            mbar_arrive_expect_tx(bar_a_full + 8 * kb, A_STAGE_BYTES);
                            tma_load_2d(smem_a + kb * A_STAGE_BYTES, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M,
                                        bar_a_full + 8 * kb);
                        }
                    }
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    if constexpr (CG2) {  // both CTAs' boxes count on the leader's barrier
                        if (crank == 0) mbar_arrive_expect_tx(bar_full + 8 * stage, 2 * (A_STAGE_BYTES + B_STAGE_BYTES));
                        const uint32_t fb = map_to_cta(bar_full + 8 * stage, 0);
                        tma_load_2d_cg2(smem_a + stage * A_STAGE_BYTES, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M, fb);
                        tma_load_2d_cg2(smem_b + stage * B_STAGE_BYTES, &tmap_b, kb * BLOCK_K,
                                        n_tile * BLOCK_N + crank * B_SLICE_ROWS, fb);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    mbar_arrive_expect_tx(bar_full + 8 * stage, (RES_A ? 0 : A_STAGE_BYTES) + B_STAGE_BYTES);
                    if constexpr (RES_A) {
                    } else if constexpr (CL > 1) {  // my slice of the A tile, delivered to every CTA of the cluster
                        tma_load_2d_multicast(smem_a + stage * A_STAGE_BYTES + crank * (A_SLICE_ROWS * BLOCK_K * 2), &tmap_a,
                                              kb * BLOCK_K, m_tile * BLOCK_M + crank * A_SLICE_ROWS, bar_full + 8 * stage,
                                              CMASK);
                    } else {
                        tma_load_2d(smem_a + stage * A_STAGE_BYTES, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M,
                                    bar_full + 8 * stage);
                    }
                    if constexpr (PAIR) {  // my half of the W tile, delivered to both CTAs of the pair
                        tma_load_2d_multicast(smem_b + stage * B_STAGE_BYTES + crank * (B_SLICE_ROWS * BLOCK_K * 2), &tmap_b,
                                              kb * BLOCK_K, n_tile * BLOCK_N + crank * B_SLICE_ROWS, bar_full + 8 * stage,
                                              CMASK);
                    } else {
#pragma unroll
                        for (int s = 0; s < NSPLIT; ++s) {
                            if (p.dbg & 64)
                                tma_load_2d_hint(smem_b + stage * B_STAGE_BYTES + s * (UMMA_N * BLOCK_K * 2), &tmap_b,
                                                 kb * BLOCK_K, n_tile * BLOCK_N + s * UMMA_N, bar_full + 8 * stage, pol_w);
                            else
                                tma_load_2d(smem_b + stage * B_STAGE_BYTES + s * (UMMA_N * BLOCK_K * 2), &tmap_b,
                                            kb * BLOCK_K, n_tile * BLOCK_N + s * UMMA_N, bar_full + 8 * stage);
                        }
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        // The whole warp walks the loop (waits included) so that control flow stays uniform; one elected lane issues.
        if (!CG2 || crank == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(CG2 ? 2 * BLOCK_M : BLOCK_M, UMMA_N);
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const int acc = it % ACC_STAGES;
                const uint32_t acc_phase = (it / ACC_STAGES) & 1;
                const int n_tile = tile_n(it);
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after_sync();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb) {
                    if constexpr (RES_A) {
                        if (n_tile == 0) mbar_wait(bar_a_full + 8 * kb, static_cast<uint32_t>(it / num_n_tiles) & 1);
                    }
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t desc_a = umma_desc_sw128(smem_a + (RES_A ? kb : stage) * A_STAGE_BYTES);
                        const uint64_t desc_b = umma_desc_sw128(smem_b + stage * B_STAGE_BYTES);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
#pragma unroll
                            for (int s = 0; s < NSPLIT; ++s) {
                                // +2 per 16-element K step (32 B >> 4); second N split starts UMMA_N rows later
                                if constexpr (CG2)
                                    umma_bf16_cg2(tmem_d, desc_a + 2 * k, desc_b + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                                else
                                    umma_bf16(tmem_d + s * UMMA_N, desc_a + 2 * k,
                                              desc_b + 2 * k + ((s * UMMA_N * BLOCK_K * 2) >> 4), idesc,
                                              (kb > 0 || k > 0) ? 1u : 0u);
                            }
                        }
                        // frees the smem stage (in every CTA that shares it) when these MMAs finish
                        if constexpr (CG2) umma_commit_cg2(bar_empty + 8 * stage, CMASK);
                        else if constexpr (CSIZE > 1) umma_commit_multicast(bar_empty + 8 * stage, CMASK);
                        else umma_commit(bar_empty + 8 * stage);
                        if constexpr (RES_A) {  // last column tile of this row block: slot kb may take the next block's A
                            if (n_tile == num_n_tiles - 1) umma_commit(bar_a_empty + 8 * kb);
                        }
                        if (kb == num_kb - 1) {
                            if constexpr (CG2) umma_commit_cg2(bar_tfull + 8 * acc, CMASK);  // accumulator halves complete in both CTAs
                            else umma_commit(bar_tfull + 8 * acc);                             // accumulator complete
                        }
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= EPI_WARP0) {
        // ------------------------------ epilogue ----------------------------------
        const int ew = warp - EPI_WARP0;
        const int quad = warp & 3;  // TMEM lane quadrant this warp may access
        const int half = ew >> 2;   // which half of the tile's columns
        const int row_in_tile = quad * 32 + lane;
        const uint32_t my_smem = smem_epi + ew * EPI_WARP_BYTES;
        const uint32_t out_buf = my_smem + (EPI == EPI_RES_LN ? RING * 2 * CHUNK_BF16 : 0);  // 2 x OUT_CHUNK_BYTES
        uint32_t st_it = 0;  // output staging buffers used so far (alternating)
        const uint64_t pol_out = l2_policy_evict_first();

        // stages one chunk and hands it to the TMA store engine; at most one older store may still be reading
        auto store_chunk = [&](const CUtensorMap* map, int col0, int row0, const float (&f)[32], bool lo) {
            const uint32_t buf = out_buf + (st_it & 1) * OUT_CHUNK_BYTES;
            if (p.dbg & 1) return;
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
            if constexpr (OUT_F32) stage_row_f32(buf, lane, f);
            else if (lo) stage_row_bf16_lo(buf, lane, f);
            else stage_row_bf16(buf, lane, f);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                if (p.dbg & 32) tma_store_2d_hint(map, buf, col0, row0, pol_out);
                else tma_store_2d(map, buf, col0, row0);
                tma_store_commit();
            }
            ++st_it;
        };

        // residual ring (EPI_RES_LN): stage s of this warp = my_smem + s * 4 KB: [hi 2 KB][lo 2 KB]. Chunks are numbered
        // g = tile * NCH + c over the CTA's whole tile sequence and loaded RING chunks ahead of their use.
        uint32_t ld_used = 0;
        auto issue_residual = [&](uint32_t g) {
            const int tile = static_cast<int>(g) / NCH, c = static_cast<int>(g) % NCH;
            if (lane == 0 && tile < my_tiles) {
                const int m_tile = tile_m(tile), n_tile = tile_n(tile);
                const uint32_t s = g % RING;
                const uint32_t bar = bar_res + 8 * (RING * ew + s);
                const uint32_t dst = my_smem + s * (2 * CHUNK_BF16);
                const int col0 = n_tile * BLOCK_N + half * HALF_N + c * CHUNK, row0 = m_tile * BLOCK_M + quad * 32;
                mbar_arrive_expect_tx(bar, p.has_res_lo ? 2 * CHUNK_BF16 : CHUNK_BF16);
                tma_load_2d(dst, &tmap_r, col0, row0, bar);
                if (p.has_res_lo) tma_load_2d(dst + CHUNK_BF16, &tmap_r_lo, col0, row0, bar);
            }
        };
        if constexpr (EPI == EPI_RES_LN) {
#pragma unroll
            for (int g = 0; g < RING; ++g) issue_residual(g);
        }

        for (int it = 0; it < my_tiles; ++it) {
            const int t = it;
            const int m_tile = tile_m(t), n_tile = tile_n(t);
            const int acc = it % ACC_STAGES;
            const uint32_t acc_phase = (it / ACC_STAGES) & 1;
            const int row0 = m_tile * BLOCK_M + quad * 32;
            const int col_base = n_tile * BLOCK_N + half * HALF_N;
            const uint32_t taddr =
                tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BLOCK_N + half * HALF_N;
            if (p.dbg & 8) {  // mainloop only
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_before_sync();
                if constexpr (CG2) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(map_to_cta(bar_tempty + 8 * acc, 0));
                } else {
                    mbar_arrive(bar_tempty + 8 * acc);
                }
                continue;
            }
            if constexpr (EPI != EPI_RES_LN && !OUT_F32) {
                // Bias of this tile's columns -> shared memory (one warp per column half loads, the half's four warps
                // meet on a named barrier); the accumulator chunks are then read one ahead of the math on them.
                const uint32_t bias_s = aux + AUX_BARRIERS + ((it & 1) * BLOCK_N + half * HALF_N) * 4;
                if (quad == 0 && lane * 4 < HALF_N) {
                    const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col_base) + lane);
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + lane * 16), "f"(b.x), "f"(b.y),
                                 "f"(b.z), "f"(b.w)
                                 : "memory");
                }
                asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after_sync();
                uint32_t v[2][32];
                tmem_ld_32x32(taddr, v[0]);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    tmem_ld_wait();
                    if (c + 1 < NCH) {
                        tmem_ld_32x32(taddr + (c + 1) * CHUNK, v[(c + 1) & 1]);
                    } else {  // accumulator drained: the MMA warp may start the tile after next
                        tc_fence_before_sync();
                        if constexpr (CG2) {
                            __syncwarp();
                            if (lane == 0) mbar_arrive_cluster(map_to_cta(bar_tempty + 8 * acc, 0));
                        } else {
                            mbar_arrive(bar_tempty + 8 * acc);
                        }
                    }
                    const int col0 = col_base + c * CHUNK;
                    uint32_t w[16];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const uint4 bq = lds_v4(bias_s + (c * CHUNK + q * 4) * 4);
                        w[q * 2 + 0] = epi_pair_bf16<EPI>(v[c & 1][q * 4 + 0], v[c & 1][q * 4 + 1], __uint_as_float(bq.x),
                                                          __uint_as_float(bq.y));
                        w[q * 2 + 1] = epi_pair_bf16<EPI>(v[c & 1][q * 4 + 2], v[c & 1][q * 4 + 3], __uint_as_float(bq.z),
                                                          __uint_as_float(bq.w));
                    }
                    if (p.dbg & 1) continue;
                    if constexpr (DIRECT) {
                        const int row = row0 + lane;
                        if (row < p.m) {
                            __nv_bfloat16* dst = p.out_bf16 + static_cast<size_t>(row) * p.n + col0;
                            stg_v8(dst, w);
                            stg_v8(dst + 16, w + 8);
                        }
                    } else if constexpr (WIDE) {
                        const uint32_t buf = out_buf + ((st_it >> 1) & 1) * CHUNK_F32;
                        if ((c & 1) == 0) {
                            if (lane == 0) tma_store_wait_read<1>();
                            __syncwarp();
                        }
                        const uint32_t row = buf + lane * 128;
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            sts_v4(row + ((((c & 1) * 4 + q) ^ (lane & 7)) << 4), w[q * 4], w[q * 4 + 1], w[q * 4 + 2], w[q * 4 + 3]);
                        if (c & 1) {
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0 && !(p.dbg & 256)) {
                                tma_store_2d(&tmap_c, buf, col0 - CHUNK, (p.dbg & 512) ? (row0 & 4095) : row0);
                                tma_store_commit();
                            }
                        }
                    } else {
                        const uint32_t buf = out_buf + (st_it & 1) * CHUNK_BF16;
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                        const uint32_t row = buf + lane * 64;
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            sts_v4(row + ((q ^ ((lane >> 1) & 3)) << 4), w[q * 4], w[q * 4 + 1], w[q * 4 + 2], w[q * 4 + 3]);
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0 && !(p.dbg & 256)) {
                            tma_store_2d(&tmap_c, buf, col0, (p.dbg & 512) ? (row0 & 4095) : row0);
                            tma_store_commit();
                        }
                    }
                    ++st_it;
                }
            } else if constexpr (EPI != EPI_RES_LN) {
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after_sync();
#pragma unroll 1
                for (int c = 0; c < NCH; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c * CHUNK, v);
                    tmem_ld_wait();
                    if (c == NCH - 1) {  // accumulator drained: the MMA warp may start the tile after next
                        tc_fence_before_sync();
                        mbar_arrive(bar_tempty + 8 * acc);
                    }
                    const int col0 = col_base + c * CHUNK;
                    float f[32];
                    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 b = __ldg(b4 + q);
                        f[q * 4 + 0] = epi_act<EPI>(__uint_as_float(v[q * 4 + 0]) + b.x);
                        f[q * 4 + 1] = epi_act<EPI>(__uint_as_float(v[q * 4 + 1]) + b.y);
                        f[q * 4 + 2] = epi_act<EPI>(__uint_as_float(v[q * 4 + 2]) + b.z);
                        f[q * 4 + 3] = epi_act<EPI>(__uint_as_float(v[q * 4 + 3]) + b.w);
                    }
                    store_chunk(&tmap_c, col0, row0, f, false);
                }
            } else {
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after_sync();
                // pass 1: x = acc + bias + residual; row sum and sum of squares. x stays in registers when the warp's
                // share of the row is at most 3 chunks (the accumulator is then free after this pass), else it is
                // parked back in TMEM.
                float sum = 0.f, sumsq = 0.f;
                float xs[KEEP_IN_REGS ? NCH : 1][32];
                auto pass1_chunk = [&](int c, uint32_t (&v)[32]) {
                    tmem_ld_32x32(taddr + c * CHUNK, v);
                    const uint32_t s = ld_used % RING;
                    mbar_wait(bar_res + 8 * (RING * ew + s), (ld_used / RING) & 1);
                    const uint32_t rbuf = my_smem + s * (2 * CHUNK_BF16) + lane * 64;
                    const int sw = (lane >> 1) & 3;
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col_base + c * CHUNK);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 r = lds_v4(rbuf + ((q ^ sw) << 4));
                        const float4 b0 = __ldg(b4 + 2 * q), b1 = __ldg(b4 + 2 * q + 1);
                        float2 r0 = unpack_bf16x2(r.x), r1 = unpack_bf16x2(r.y), r2 = unpack_bf16x2(r.z),
                               r3 = unpack_bf16x2(r.w);
                        if (p.has_res_lo) {
                            const uint4 l = lds_v4(rbuf + CHUNK_BF16 + ((q ^ sw) << 4));
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
                    __syncwarp();  // every lane has read this ring stage: refill it with the chunk RING ahead
                    issue_residual(ld_used + RING);
                    ++ld_used;
                };
                if constexpr (KEEP_IN_REGS) {
#pragma unroll
                    for (int c = 0; c < NCH; ++c) pass1_chunk(c, *reinterpret_cast<uint32_t(*)[32]>(&xs[c][0]));
                    tc_fence_before_sync();
                    mbar_arrive(bar_tempty + 8 * acc);  // accumulator fully read: the MMA warp may reuse it
                } else {
#pragma unroll 1
                    for (int c = 0; c < NCH; ++c) {
                        uint32_t v[32];
                        pass1_chunk(c, v);
                        tmem_st_32x32(taddr + c * CHUNK, v);
                    }
                    tmem_st_wait();
                }
                ln_part[half * BLOCK_M + row_in_tile] = make_float2(sum, sumsq);
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
                const float2 other = ln_part[(half ^ 1) * BLOCK_M + row_in_tile];
                sum += other.x;
                sumsq += other.y;
                if (CL > 1 && !(p.dbg & 2)) {
                    // this CTA's 2 x 128 threads now hold the partials of its BLOCK_N columns. One thread arms the
                    // barrier of this tile parity with the bytes the peers will deliver; one thread per row stores
                    // its (sum, sumsq) into every peer's slot [parity][my rank][row] with st.async, which counts the
                    // bytes on the peer's barrier. A peer can be at most one tile ahead, so two parities suffice.
                    const uint32_t par = it & 1;
                    if (ew == 0 && lane == 0) mbar_arrive_expect_tx(bar_ln + 8 * par, (CL - 1) * BLOCK_M * 8);
                    if (half == 0) {
                        const uint32_t slot = ln_peer_addr + ((par * MAX_CLUSTER + crank) * BLOCK_M + row_in_tile) * 8;
#pragma unroll
                        for (int pr = 0; pr < CL; ++pr) {
                            if (pr == crank) continue;
                            st_async_cluster_b64(map_to_cta(slot, pr), sum, sumsq, map_to_cta(bar_ln + 8 * par, pr));
                        }
                    }
                    mbar_wait(bar_ln + 8 * par, (it >> 1) & 1);
#pragma unroll
                    for (int pr = 0; pr < CL; ++pr) {
                        if (pr == crank) continue;
                        const float2 o = ln_peer[(par * MAX_CLUSTER + pr) * BLOCK_M + row_in_tile];
                        sum += o.x;
                        sumsq += o.y;
                    }
                }
                const float inv_n = 1.0f / static_cast<float>(p.n);
                const float mean = sum * inv_n;
                const float var = fmaxf(sumsq * inv_n - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.eps);
                // pass 2: normalise, scale, shift, store (hi, lo)
                auto pass2_chunk = [&](int c, float (&f)[32]) {
                    const int col0 = col_base + c * CHUNK;
                    const float4* g4 = reinterpret_cast<const float4*>(p.gamma + col0);
                    const float4* be4 = reinterpret_cast<const float4*>(p.beta + col0);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 g = __ldg(g4 + q), be = __ldg(be4 + q);
                        f[q * 4 + 0] = fmaf((f[q * 4 + 0] - mean) * rstd, g.x, be.x);
                        f[q * 4 + 1] = fmaf((f[q * 4 + 1] - mean) * rstd, g.y, be.y);
                        f[q * 4 + 2] = fmaf((f[q * 4 + 2] - mean) * rstd, g.z, be.z);
                        f[q * 4 + 3] = fmaf((f[q * 4 + 3] - mean) * rstd, g.w, be.w);
                    }
                    store_chunk(&tmap_c, col0, row0, f, false);
                    if (p.has_out_lo) store_chunk(&tmap_c_lo, col0, row0, f, true);
                };
                if constexpr (KEEP_IN_REGS) {
#pragma unroll
                    for (int c = 0; c < NCH; ++c) pass2_chunk(c, xs[c]);
                } else {
#pragma unroll 1
                    for (int c = 0; c < NCH; ++c) {
                        float f[32];
                        tmem_ld_32x32(taddr + c * CHUNK, *reinterpret_cast<uint32_t(*)[32]>(&f[0]));
                        tmem_ld_wait();
                        if (c == NCH - 1) {
                            tc_fence_before_sync();
                            mbar_arrive(bar_tempty + 8 * acc);
                        }
                        pass2_chunk(c, f);
                    }
                }
                // the partial-stat slots are reused by the next tile
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
            }
        }
        if (lane == 0) tma_store_wait<0>();  // all stores of this warp are complete before the CTA exits
    }

    tc_fence_before_sync();
    __syncthreads();
    if constexpr (CSIZE > 1) cluster_sync_all();  // no CTA leaves while a peer may still write into its shared memory
    if (warp == 2) {
        tc_fence_after_sync();
        if constexpr (CG2) tmem_dealloc_cg2(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
template <int BLOCK_N, int EPI, typename OutT, int CL = 1, int PAIR = 0, int KSLOTS = 0, int RING = 2, int NSP = 0, int WST = 0>
cudaError_t launch(const GemmTcArgs& a, cudaStream_t stream, int num_sms) {
    constexpr int NSPLIT = NSP ? NSP : (BLOCK_N > 256 ? 2 : 1);
    constexpr bool OUT_F32 = sizeof(OutT) == 4;
    constexpr int CSIZE = CL > 1 ? CL : (PAIR ? 2 : 1);
    alignas(64) CUtensorMap ta, tb, tc, tcl, tr, trl;
    if (!make_tmap(&ta, a.a, a.m, a.k, BLOCK_K, BLOCK_M / (CL > 1 ? CL : 1), false, CU_TENSOR_MAP_SWIZZLE_128B))
        return cudaErrorInvalidValue;
    if (!make_tmap(&tb, a.b, a.n, a.k, BLOCK_K, BLOCK_N / NSPLIT / (PAIR ? 2 : 1), false, CU_TENSOR_MAP_SWIZZLE_128B))
        return cudaErrorInvalidValue;
    const CUtensorMapSwizzle csw = (OUT_F32 || WST == 1) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    if (!make_tmap(&tc, a.out, a.m, a.n, WST == 1 ? 2 * CHUNK : CHUNK, 32, OUT_F32, csw)) return cudaErrorInvalidValue;
    tcl = tc; tr = tc; trl = tc;
    if (a.out_lo && !make_tmap(&tcl, a.out_lo, a.m, a.n, CHUNK, 32, false, csw)) return cudaErrorInvalidValue;
    if (EPI == EPI_RES_LN) {
        if (!a.residual) return cudaErrorInvalidValue;
        if (!make_tmap(&tr, a.residual, a.m, a.n, CHUNK, 32, false, csw)) return cudaErrorInvalidValue;
        trl = tr;
        if (a.residual_lo && !make_tmap(&trl, a.residual_lo, a.m, a.n, CHUNK, 32, false, csw))
            return cudaErrorInvalidValue;
    }
    auto kern = gemm_tc_kernel<BLOCK_N, EPI, OutT, CL, PAIR, KSLOTS, RING, NSP, WST>;
    constexpr int smem = smem_bytes(BLOCK_N, EPI, OUT_F32, KSLOTS, RING, WST == 1, PAIR);
    static_assert(num_stages(BLOCK_N, EPI, OUT_F32, KSLOTS, RING, WST == 1, PAIR) >= 2, "not enough shared memory for a pipeline");
    static_assert(smem <= SMEM_LIMIT, "shared memory budget exceeded");
    static bool attr_set = false;  // per instantiation; benign race (idempotent)
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    GemmDev p;
    p.m = a.m; p.d_rows = a.d_rows; p.n = a.n; p.k = a.k; p.bias = a.bias;
    p.out_bf16 = static_cast<__nv_bfloat16*>(a.out);
    p.gamma = a.gamma; p.beta = a.beta; p.eps = a.eps;
    p.has_res_lo = a.residual_lo != nullptr; p.has_out_lo = a.out_lo != nullptr;
    p.dbg = 0;
    if (const char* dv = getenv("B200_EMBED_GEMM_DBG")) p.dbg = atoi(dv);
    const int m_tiles = (a.m + BLOCK_M - 1) / BLOCK_M;
    const int units = (CL > 1 || KSLOTS > 0) ? m_tiles : (PAIR ? ((m_tiles + 1) / 2) : m_tiles) * (a.n / BLOCK_N);
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CSIZE;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = CSIZE > 1 ? 1 : 0;
    // persistent grid: as many clusters as the device can hold at once (size-4 clusters do not tile all 148 SMs)
    static int max_units = 0;
    if (max_units == 0) {
        max_units = num_sms / CSIZE;
        if (CSIZE > 1) {
            cfg.gridDim = dim3((num_sms / CSIZE) * CSIZE);
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0 && n < max_units) max_units = n;
        }
    }
    const int grid = (units < max_units ? units : max_units) * CSIZE;
    if (grid <= 0) return cudaSuccess;
    cfg.gridDim = dim3(grid);
    return cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tcl, tr, trl, p);
}

}  // namespace

int gemm_tc_pick_cluster(int n, int epi) {
    if (epi != EPI_RES_LN) return 1;
    if (n == 128 || n == 256) return 1;
    if (n == 384 || n == 512) return 2;
    if (n == 768 || n == 1024) return 4;
    return 0;
}

int gemm_tc_pick_block_n(int n, int epi) {
    if (epi == EPI_RES_LN) {
        const int cl = gemm_tc_pick_cluster(n, epi);
        return cl ? n / cl : 0;
    }
    if (n % 256 == 0) return 256;
    if (epi == EPI_BIAS && n % 192 == 0) return 192;
    if (n % 128 == 0) return 128;
    return 0;
}

cudaError_t gemm_tc_launch(const GemmTcArgs& a, cudaStream_t stream, int num_sms) {
    if (a.m <= 0) return cudaSuccess;
    // TMA needs 16-byte row pitches: k % 8 (bf16 operands) and n % 8 (bf16 output rows)
    if (a.k % 8 != 0 || a.n % a.block_n != 0) return cudaErrorInvalidValue;
    if (a.epi == EPI_RES_LN) {  // a CTA, or a cluster of 2 / 4 CTAs, owns whole rows
        const int cl = a.n / a.block_n;
        if (a.block_n == 128 && cl == 1) return launch<128, EPI_RES_LN, __nv_bfloat16, 1>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 1) return launch<256, EPI_RES_LN, __nv_bfloat16, 1>(a, stream, num_sms);
        // short K: the epilogue is the critical path -> deeper residual prefetch, fewer mainloop stages
        if (a.block_n == 192 && cl == 2 && a.k <= 768) return launch<192, EPI_RES_LN, __nv_bfloat16, 2, 0, 0, 3>(a, stream, num_sms);
        if (a.block_n == 192 && cl == 2) return launch<192, EPI_RES_LN, __nv_bfloat16, 2>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 2) return launch<256, EPI_RES_LN, __nv_bfloat16, 2>(a, stream, num_sms);
        if (a.block_n == 192 && cl == 4) return launch<192, EPI_RES_LN, __nv_bfloat16, 4>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 4) return launch<256, EPI_RES_LN, __nv_bfloat16, 4>(a, stream, num_sms);
        return cudaErrorInvalidValue;
    }
    if (a.pair == 2 && !a.out_f32) {  // one MMA over a CTA pair; each CTA stages half of the W tile
#define B200_CG2(BN, EP)                                                                                     \
    if (a.block_n == BN && a.epi == EP)                                                                      \
        return a.wide_store == 1 ? launch<BN, EP, __nv_bfloat16, 1, 2, 0, 2, 0, ((BN / 64) % 2 == 0 ? 1 : 0)>(a, stream, num_sms) \
                            : launch<BN, EP, __nv_bfloat16, 1, 2, 0, 2, 0, 0>(a, stream, num_sms);
        B200_CG2(256, EPI_BIAS)
        B200_CG2(192, EPI_BIAS)
        B200_CG2(128, EPI_BIAS)
        B200_CG2(256, EPI_GELU_ERF)
        B200_CG2(128, EPI_GELU_ERF)
#undef B200_CG2
    }
    const bool nsplit2 = a.block_n == 256 && a.nsplit == 2;
    if (a.wide_store == 1 && !a.out_f32 && !a.pair && a.epi != EPI_RES_LN) {  // 128-byte row segments per TMA store
        if (a.block_n == 256 && a.epi == EPI_BIAS) return launch<256, EPI_BIAS, __nv_bfloat16, 1, 0, 0, 2, 0, 1>(a, stream, num_sms);
        if (a.block_n == 256 && a.epi == EPI_GELU_ERF) return launch<256, EPI_GELU_ERF, __nv_bfloat16, 1, 0, 0, 2, 0, 1>(a, stream, num_sms);
        if (a.block_n == 128 && a.epi == EPI_BIAS) return launch<128, EPI_BIAS, __nv_bfloat16, 1, 0, 0, 2, 0, 1>(a, stream, num_sms);
    }
    if (a.wide_store == 2 && !a.out_f32 && !a.pair && a.epi != EPI_RES_LN) {  // direct stores from registers
        if (a.block_n == 256 && a.epi == EPI_BIAS) return launch<256, EPI_BIAS, __nv_bfloat16, 1, 0, 0, 2, 0, 2>(a, stream, num_sms);
        if (a.block_n == 256 && a.epi == EPI_GELU_ERF) return launch<256, EPI_GELU_ERF, __nv_bfloat16, 1, 0, 0, 2, 0, 2>(a, stream, num_sms);
        if (a.block_n == 192 && a.epi == EPI_BIAS) return launch<192, EPI_BIAS, __nv_bfloat16, 1, 0, 0, 2, 0, 2>(a, stream, num_sms);
        if (a.block_n == 128 && a.epi == EPI_BIAS) return launch<128, EPI_BIAS, __nv_bfloat16, 1, 0, 0, 2, 0, 2>(a, stream, num_sms);
    }
    // K <= 384 with several column tiles per row block: keep the activation tile resident (6 x 64 columns)
    const bool res_a = !a.pair && !a.no_resident_a && a.k <= 6 * BLOCK_K && a.n / a.block_n >= 2;
    // nsplit == 2 (B200_EMBED_GEMM_NSPLIT=2): 256-wide tiles issued as two independent 128-column accumulator chains per
    // k-step. Measured 62 % -> 71 % of the MMA rate at K = 4096 but slower at K = 384 with the GELU epilogue
    // (FFN1 0.78 vs 0.68 ms), so it is off by default.
#define B200_CASE(BN, EP, T)                                                                   \
    if (a.block_n == BN && a.epi == EP && a.out_f32 == (sizeof(T) == 4))                       \
        return a.pair == 1 ? launch<BN, EP, T, 1, 1, 0, 2, 0>(a, stream, num_sms)                \
               : res_a   ? launch<BN, EP, T, 1, 0, 6, 2, 0>(a, stream, num_sms)                \
               : nsplit2 ? launch<BN, EP, T, 1, 0, 0, 2, (BN == 256 ? 2 : 0)>(a, stream, num_sms) \
                         : launch<BN, EP, T, 1, 0, 0, 2, 0>(a, stream, num_sms);
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
#undef B200_CASE
    return cudaErrorInvalidValue;
}

}  // namespace b200
