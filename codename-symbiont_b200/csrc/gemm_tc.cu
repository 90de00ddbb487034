// K2: bf16 tensor-core GEMM for sm_100a:  C[M,N] = epilogue(A[M,K] . W[N,K]^T + bias)
//
// Replaces the four candle `Linear` layers of one BERT layer (candle-transformers 0.9.1 bert.rs:
// BertSelfAttention query/key/value, BertSelfOutput.dense, BertIntermediate.dense, BertOutput.dense;
// called from /root/reference/services/preprocessing_service/src/embedding_generator.rs:198),
// which the reference runs as f32 `gemm`-crate matmuls followed by separate bias / GELU /
// residual / LayerNorm passes.
//
// Design (persistent CTAs, one per SM, 12 warps):
//   warp 0      TMA producer: A tile 128x64 and W tile BLOCK_Nx64 (bf16, 128B swizzle) per stage
//   warp 1      MMA issuer: the warp walks the pipeline together, one elected lane issues tcgen05.mma
//               (128 x BLOCK_N x 16, fp32 accumulators in TMEM). Keeping the control flow warp-uniform matters: under
//               `if (lane == 0)` the compiler wraps every MMA in a divergence-safe ELECT loop and the issuing thread,
//               not the tensor pipe, bounds the kernel.
//   warp 2      TMEM allocator
//   warps 4-11  epilogue. Warp w owns TMEM lane quadrant w%4 (32 rows) and one half of the tile's columns, and walks
//               them in 32-column chunks: tcgen05.ld (one chunk ahead) -> packed-pair fp32 math in registers ->
//               swizzled shared-memory staging -> one TMA store per chunk. A thread holds one row, so writing global
//               memory directly would touch 32 rows per instruction; measured, direct 32-byte stores, 64-byte and
//               128-byte TMA rows all cost the same (the SM -> L2 write path is the limit), so the simplest is kept.
// Accumulators are double-buffered in TMEM (2 * BLOCK_N <= 512 columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1.
//
// PAIR (plain bf16 epilogues): two CTAs of a cluster work on adjacent row blocks as one cta_group::2 MMA (256 x
// BLOCK_N per pair). Each CTA stages its own A tile but only half of the W tile, so a stage is 1/3 smaller (more
// stages in flight for the same shared memory) and the W traffic from L2 halves. The leader CTA's "full" barrier
// collects the bytes of both CTAs' loads, its elected lane issues the pair's MMAs, and its commits arrive on the
// barriers of both CTAs.
//
// EPI_RES_LN: with BLOCK_N * CL == N a CTA, or a thread-block CLUSTER of CL = 2 / 4 CTAs, owns whole output rows and
// the residual add + LayerNorm run in the epilogue: pass 1 adds bias and the residual ((hi, lo) pairs, TMA-loaded per
// chunk into a per-warp ring ahead of use) and accumulates the row statistics; pass 2 normalises and stores (hi, lo).
// The pre-LN activations never reach HBM. For N = 384 / 768 the row is split over 2 / 4 CTAs (192 columns each, so
// the TMEM accumulators stay double-buffered and the values stay in registers between the passes); the CTAs of a
// cluster walk the same row blocks, share the A tile (each loads 1/CL of its rows and multicasts) and exchange their
// per-row partial (sum, sum of squares) through distributed shared memory.
// Both operands are K-major ("TN"): activations [M,K] and HF weights [N,K] as stored.
#include "gemm_tc.h"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tc_helpers.cuh"

namespace b200 {
B200_NS_THIS {

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARP0 = 4;
constexpr int EPI_WARPS = 8;
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;

// W bytes per stage in one CTA (a pair stages half of the tile's rows in each CTA)
__host__ __device__ constexpr int b_stage_bytes(int block_n, bool pair) { return (pair ? block_n / 2 : block_n) * BLOCK_K * 2; }
// per-warp epilogue shared memory: 2 output staging buffers (+ a `ring`-stage (hi, lo) residual ring for LN).
// ring == 1 is the lean LayerNorm epilogue for long K (one residual slot, one staging buffer): with 48 k-blocks per
// tile the epilogue hides behind the next tile's mainloop anyway, and the 40 KB it gives back are a 4th stage.
__host__ __device__ constexpr int staging_buffers(int ring) { return ring == 1 ? 1 : 2; }
__host__ __device__ constexpr int epi_warp_bytes(int epi, bool out_f32, int ring) {
    return (epi == EPI_RES_LN ? ring * 2 * CHUNK_BF16 : 0) + staging_buffers(ring) * (out_f32 ? CHUNK_F32 : CHUNK_BF16);
}
// mbarriers (<= 8 + 8 stage, 4 TMEM, 24 residual, 2 cluster exchange) + TMEM pointer; then
//   LayerNorm: partial statistics of the two column halves, the slots peers write their row partials into, and
//              bias | gamma | beta of this CTA's columns;
//   plain:     the bias of the current and the next tile's columns ([tile parity][BLOCK_N] fp32)
constexpr int MAX_CLUSTER = 4;
constexpr int AUX_BARRIERS = 512;
constexpr int AUX_LN_PART = 2 * BLOCK_M * 8;
__host__ __device__ constexpr int aux_ln_peer(int cl) { return 2 * (cl - 1) * BLOCK_M * 8; }  // [tile parity][peer][row] float2
__host__ __device__ constexpr int aux_bytes(int epi, int block_n, int cl) {
    return AUX_BARRIERS + (epi == EPI_RES_LN ? AUX_LN_PART + aux_ln_peer(cl) + 3 * block_n * 4 : 2 * block_n * 4);
}
__host__ __device__ constexpr int num_stages(int block_n, int epi, bool out_f32, int ring, bool pair, int cl) {
    int s = (SMEM_LIMIT - aux_bytes(epi, block_n, cl) - EPI_WARPS * epi_warp_bytes(epi, out_f32, ring)) /
            (A_STAGE_BYTES + b_stage_bytes(block_n, pair));
    return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int smem_bytes(int block_n, int epi, bool out_f32, int ring, bool pair, int cl) {
    return num_stages(block_n, epi, out_f32, ring, pair, cl) * (A_STAGE_BYTES + b_stage_bytes(block_n, pair)) +
           EPI_WARPS * epi_warp_bytes(epi, out_f32, ring) + aux_bytes(epi, block_n, cl);
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
    const float* gamma;
    const float* beta;
    float eps;
    int has_res_lo, has_out_lo;
    int dbg;  // timing experiments only (B200_EMBED_GEMM_DBG): 1 no output stores, 2 no cluster exchange, 8 mainloop only
};

template <int BLOCK_N, int EPI, typename OutT, int CL, bool PAIR, int RING>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_c_lo,
               const __grid_constant__ CUtensorMap tmap_r, const __grid_constant__ CUtensorMap tmap_r_lo,
               const GemmDev p) {
    constexpr bool OUT_F32 = sizeof(OutT) == 4;
    constexpr int ACC_STAGES = 2;
    constexpr int B_STAGE_BYTES = b_stage_bytes(BLOCK_N, PAIR);
    constexpr int STAGES = num_stages(BLOCK_N, EPI, OUT_F32, RING, PAIR, CL);
    constexpr int HALF_N = BLOCK_N / 2;  // columns per epilogue warp group
    constexpr int NCH = HALF_N / CHUNK;  // chunks per warp per tile
    constexpr int EPI_WARP_BYTES = epi_warp_bytes(EPI, OUT_F32, RING);
    constexpr bool KEEP_IN_REGS = NCH <= 3;  // LN: hold acc + bias + residual in registers between the two passes
    constexpr int OUT_CHUNK_BYTES = OUT_F32 ? CHUNK_F32 : CHUNK_BF16;
    static_assert(RING >= 1 && RING <= 3, "residual ring depth");
    constexpr int STG = staging_buffers(RING);
    static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "invalid UMMA N / TMEM double buffering needs 2 * BLOCK_N <= 512");
    static_assert(HALF_N % CHUNK == 0, "epilogue works in 32-column chunks");
    static_assert(STAGES >= 2, "not enough shared memory for a pipeline");
    static_assert(EPI != EPI_RES_LN || !OUT_F32, "LayerNorm epilogue writes bf16 (hi, lo)");
    static_assert(CL == 1 || EPI == EPI_RES_LN, "row-splitting clusters are only used by the LayerNorm epilogue");
    static_assert(CL <= MAX_CLUSTER, "cluster too large");
    static_assert(!PAIR || (CL == 1 && EPI != EPI_RES_LN && !OUT_F32), "CTA-pair MMA: plain bf16 epilogues");
    // Issue style of the producer and MMA warps. Warp-uniform (whole warp walks the loop, one elected lane issues) gives
    // the leanest instruction stream; a lone lane reacts faster to a barrier flip, which is what the LayerNorm
    // variant with its 2 shallow stages needs (measured: FFN2 1.03 ms uniform vs 0.76 ms lone lane).
    constexpr bool UNIFORM = PAIR;
    constexpr int CSIZE = CL > 1 ? CL : (PAIR ? 2 : 1);        // CTAs per cluster
    constexpr uint16_t CMASK = static_cast<uint16_t>((1u << CSIZE) - 1);
    // rows of the A tile this CTA loads (and multicasts); a cluster of 3 splits the tile between its first two CTAs
    constexpr int A_LOADERS = CL == 3 ? 2 : (CL > 1 ? CL : 1);
    constexpr int A_SLICE_ROWS = BLOCK_M / A_LOADERS;
    constexpr int B_SLICE_ROWS = BLOCK_N / (PAIR ? 2 : 1);     // rows of the W tile this CTA loads

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    if ((base & 1023u) != 0) __trap();  // swizzled TMA / UMMA tiles need 1024-byte alignment
    const uint32_t smem_a = base;
    const uint32_t smem_b = base + STAGES * A_STAGE_BYTES;
    const uint32_t smem_epi = smem_b + STAGES * B_STAGE_BYTES;
    const uint32_t aux = smem_epi + EPI_WARPS * EPI_WARP_BYTES;
    uint8_t* aux_ptr = smem_raw + (aux - base);
    const uint32_t bar_full = aux;          // STAGES x 8 B
    const uint32_t bar_empty = aux + 64;    // STAGES x 8 B
    const uint32_t bar_tfull = aux + 128;   // 2 x 8 B
    const uint32_t bar_tempty = aux + 144;  // 2 x 8 B
    const uint32_t tmem_slot = aux + 160;   // 4 B
    const uint32_t bar_ln = aux + 176;      // 2 x 8 B: peers' row partials have arrived (per tile parity)
    const uint32_t bar_res = aux + 192;     // EPI_WARPS x RING x 8 B (residual ring of each epilogue warp)
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(aux_ptr + 160);
    float2* ln_part = reinterpret_cast<float2*>(aux_ptr + AUX_BARRIERS);                // [2][BLOCK_M]
    float2* ln_peer = reinterpret_cast<float2*>(aux_ptr + AUX_BARRIERS + AUX_LN_PART);  // [2][CL - 1][BLOCK_M]
    const uint32_t ln_peer_addr = aux + AUX_BARRIERS + AUX_LN_PART;
    const uint32_t ln_const = aux + AUX_BARRIERS + AUX_LN_PART + aux_ln_peer(CL);  // bias | gamma | beta, BLOCK_N floats each

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    pdl_trigger();
    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        prefetch_tmap(&tmap_c);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, PAIR ? 1 : CSIZE);  // one (multicast) commit from every CTA that issues MMAs on the stage
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_tfull + 8 * s, 1);
            mbar_init(bar_tempty + 8 * s, PAIR ? 2 * EPI_WARPS : EPI_THREADS);  // pair: one arrive per epilogue warp of both CTAs
        }
        for (int s = 0; s < RING * EPI_WARPS; ++s) mbar_init(bar_res + 8 * s, 1);
        for (int s = 0; s < 2; ++s) mbar_init(bar_ln + 8 * s, 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        if constexpr (PAIR) tmem_alloc_cg2(tmem_slot, 512);
        else tmem_alloc(tmem_slot, 512);
    }
    tc_fence_before_sync();
    __syncthreads();
    if constexpr (CSIZE > 1) cluster_sync_all();  // peers' barriers exist before anyone arrives on them remotely
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();  // nothing above reads or writes global memory

    int rows = p.m;
    if (p.d_rows != nullptr) {
        const int r = *p.d_rows;
        rows = r < rows ? r : rows;
    }
    const int num_m_tiles = (rows + BLOCK_M - 1) / BLOCK_M;
    const int num_n_tiles = p.n / BLOCK_N;  // == CL when the row is split over a cluster
    const int num_kb = (p.k + BLOCK_K - 1) / BLOCK_K;
    // Work units, walked identically by all CTAs of a cluster:
    //   plain:  unit = (m_tile, n_tile), one CTA each;
    //   CL > 1: unit = m_tile, CTA `crank` takes columns [crank * BLOCK_N, +BLOCK_N) of it;
    //   PAIR:   unit = (pair of adjacent m_tiles, n_tile), CTA `crank` takes m_tile 2 * pair + crank. With an odd number
    //           of row blocks the last pair's second CTA works on an out-of-range block: its loads are zero-filled and
    //           its stores clipped, but it keeps its side of every handshake.
    const int crank = CSIZE > 1 ? static_cast<int>(cluster_ctarank()) : 0;
    const int num_units = CL > 1 ? num_m_tiles : (PAIR ? ((num_m_tiles + 1) / 2) * num_n_tiles : num_m_tiles * num_n_tiles);
    const int unit0 = static_cast<int>(blockIdx.x) / CSIZE;
    const int unit_step = static_cast<int>(gridDim.x) / CSIZE;
    const int my_tiles = unit0 < num_units ? (num_units - unit0 + unit_step - 1) / unit_step : 0;
    auto tile_m = [&](int i) {
        const int t = unit0 + i * unit_step;
        return CL > 1 ? t : (PAIR ? 2 * (t / num_n_tiles) + crank : t / num_n_tiles);
    };
    auto tile_n = [&](int i) { return CL > 1 ? crank : (unit0 + i * unit_step) % num_n_tiles; };


    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        // UNIFORM: the warp walks the loop together, one elected lane issues; else lane 0 walks it alone
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t leader_full = PAIR ? map_to_cta(bar_full, 0) : bar_full;
        for (int i = 0; i < ((UNIFORM || lane == 0) ? my_tiles : 0); ++i) {
            const int m_tile = tile_m(i), n_tile = tile_n(i);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                if (!UNIFORM || elect_one()) {
                    const uint32_t sa = smem_a + stage * A_STAGE_BYTES, sb = smem_b + stage * B_STAGE_BYTES;
                    if constexpr (PAIR) {  // both CTAs' boxes count on the leader's barrier
                        if (crank == 0) mbar_arrive_expect_tx(bar_full + 8 * stage, 2 * (A_STAGE_BYTES + B_STAGE_BYTES));
                        tma_load_2d_cg2(sa, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M, leader_full + 8 * stage);
                        tma_load_2d_cg2(sb, &tmap_b, kb * BLOCK_K, n_tile * BLOCK_N + crank * B_SLICE_ROWS,
                                        leader_full + 8 * stage);
                    } else {
                        mbar_arrive_expect_tx(bar_full + 8 * stage, A_STAGE_BYTES + B_STAGE_BYTES);
                        if constexpr (CL > 1) {  // my slice of the A tile, delivered to every CTA of the cluster
                            if (crank < A_LOADERS)
                                tma_load_2d_multicast(sa + crank * (A_SLICE_ROWS * BLOCK_K * 2), &tmap_a, kb * BLOCK_K,
                                                      m_tile * BLOCK_M + crank * A_SLICE_ROWS, bar_full + 8 * stage, CMASK);
                        } else
                            tma_load_2d(sa, &tmap_a, kb * BLOCK_K, m_tile * BLOCK_M, bar_full + 8 * stage);
                        tma_load_2d(sb, &tmap_b, kb * BLOCK_K, n_tile * BLOCK_N, bar_full + 8 * stage);
                    }
                }
                if constexpr (UNIFORM) __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        if ((!PAIR || crank == 0) && (UNIFORM || lane == 0)) {
            constexpr uint32_t idesc = umma_idesc_op(PAIR ? 2 * BLOCK_M : BLOCK_M, BLOCK_N);
            const uint64_t da0 = umma_desc_sw128(smem_a), db0 = umma_desc_sw128(smem_b);
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const int acc = it % ACC_STAGES;
                const uint32_t acc_phase = (it / ACC_STAGES) & 1;
                if constexpr (PAIR) mbar_wait_cluster(bar_tempty + 8 * acc, acc_phase ^ 1);
                else mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after_sync();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after_sync();
                    if (!UNIFORM || elect_one()) {
                        const uint64_t desc_a = da0 + ((stage * A_STAGE_BYTES) >> 4);
                        const uint64_t desc_b = db0 + ((stage * B_STAGE_BYTES) >> 4);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {  // +2 per 16-element K step (32 B >> 4)
                            if constexpr (PAIR)
                                umma_bf16_cg2(tmem_d, desc_a + 2 * k, desc_b + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                            else
                                umma_bf16(tmem_d, desc_a + 2 * k, desc_b + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        }
                        // frees the smem stage (in every CTA that shares it) when these MMAs finish
                        if constexpr (PAIR) umma_commit_cg2(bar_empty + 8 * stage, CMASK);
                        else if constexpr (CSIZE > 1) umma_commit_multicast(bar_empty + 8 * stage, CMASK);
                        else umma_commit(bar_empty + 8 * stage);
                        if (kb == num_kb - 1) {  // accumulator complete (pair: both CTAs' halves)
                            if constexpr (PAIR) umma_commit_cg2(bar_tfull + 8 * acc, CMASK);
                            else umma_commit(bar_tfull + 8 * acc);
                        }
                    }
                    if constexpr (UNIFORM) __syncwarp();
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
        const uint32_t leader_tempty = PAIR ? map_to_cta(bar_tempty, 0) : bar_tempty;

        // accumulator `acc` has been read by this warp: the MMA warp may start the tile after next in it
        auto release_acc = [&](int acc) {
            tc_fence_before_sync();
            if constexpr (PAIR) {
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(leader_tempty + 8 * acc);
            } else {
                mbar_arrive(bar_tempty + 8 * acc);
            }
        };
        // hands a staged chunk to the TMA store engine; at most one older store may still be reading its buffer
        auto stage_begin = [&]() {
            if (lane == 0) tma_store_wait_read<STG - 1>();
            __syncwarp();
            return out_buf + (st_it % STG) * OUT_CHUNK_BYTES;
        };
        auto stage_end = [&](const CUtensorMap* map, uint32_t buf, int col0, int row0) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(map, buf, col0, row0);
                tma_store_commit();
            }
            ++st_it;
        };

        // residual ring (EPI_RES_LN): stage s of this warp = my_smem + s * 4 KB: [hi 2 KB][lo 1 KB, e5m2 bytes, then 1 KB unused]. Chunks are numbered
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
                mbar_arrive_expect_tx(bar, p.has_res_lo ? CHUNK_BF16 + CHUNK_BF16 / 2 : CHUNK_BF16);  // lo: one byte per element
                tma_load_2d(dst, &tmap_r, col0, row0, bar);
                if (p.has_res_lo) tma_load_2d(dst + CHUNK_BF16, &tmap_r_lo, col0, row0, bar);
            }
        };
        if constexpr (EPI == EPI_RES_LN) {
#pragma unroll
            for (int g = 0; g < RING; ++g) issue_residual(g);
            // bias | gamma | beta of this CTA's columns (the same for every tile it produces) -> shared memory
            const int n0 = crank * BLOCK_N;
            for (int i = threadIdx.x - EPI_WARP0 * 32; i < 3 * BLOCK_N; i += EPI_THREADS) {
                const float* src = i < BLOCK_N ? p.bias + n0 + i
                                               : (i < 2 * BLOCK_N ? p.gamma + n0 + (i - BLOCK_N) : p.beta + n0 + (i - 2 * BLOCK_N));
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(ln_const + i * 4), "f"(__ldg(src)) : "memory");
            }
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
        }

        for (int it = 0; it < my_tiles; ++it) {
            const int m_tile = tile_m(it), n_tile = tile_n(it);
            const int acc = it % ACC_STAGES;
            const uint32_t acc_phase = (it / ACC_STAGES) & 1;
            const int row0 = m_tile * BLOCK_M + quad * 32;
            const int col_base = n_tile * BLOCK_N + half * HALF_N;
            const uint32_t taddr =
                tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BLOCK_N + half * HALF_N;
            if (p.dbg & 8) {  // mainloop only
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                release_acc(acc);
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
                    if (c + 1 < NCH) tmem_ld_32x32(taddr + (c + 1) * CHUNK, v[(c + 1) & 1]);
                    else release_acc(acc);
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
                    const uint32_t buf = stage_begin();
                    stage_words_bf16(buf, lane, w);
                    stage_end(&tmap_c, buf, col_base + c * CHUNK, row0);
                }
            } else if constexpr (EPI != EPI_RES_LN) {  // fp32 output (pre-LayerNorm rows of the unfused route)
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after_sync();
#pragma unroll 1
                for (int c = 0; c < NCH; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(taddr + c * CHUNK, v);
                    tmem_ld_wait();
                    if (c == NCH - 1) release_acc(acc);
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
                    if (p.dbg & 1) continue;
                    const uint32_t buf = stage_begin();
                    stage_row_f32(buf, lane, f);
                    stage_end(&tmap_c, buf, col0, row0);
                }
            } else {
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after_sync();
                // pass 1: x = acc + bias + residual; row sum and sum of squares (packed pairs). x stays in registers when
                // the warp's share of the row is at most 3 chunks (the accumulator is then free after this pass), else
                // it is parked back in TMEM.
                const int lcol = half * HALF_N;  // this warp's first column within the CTA's BLOCK_N
                uint64_t sum2 = 0, sq2 = 0;      // (0.f, 0.f)
                uint32_t xs[KEEP_IN_REGS ? NCH : 1][32];
                auto pass1_chunk = [&](int c, uint32_t (&v)[32]) {
                    tmem_ld_32x32(taddr + c * CHUNK, v);
                    const uint32_t s = ld_used % RING;
                    mbar_wait(bar_res + 8 * (RING * ew + s), (ld_used / RING) & 1);
                    const uint32_t rbuf = my_smem + s * (2 * CHUNK_BF16) + lane * 64;
                    uint32_t hi_addr[4], lo_addr[2];
#pragma unroll
                    for (int q = 0; q < 4; ++q) hi_addr[q] = rbuf + ((q ^ ((lane >> 1) & 3)) << 4);
                    lo8_piece_addrs(my_smem + s * (2 * CHUNK_BF16) + CHUNK_BF16, lane, lo_addr);
                    tmem_ld_wait();
                    ln_accumulate_chunk(v, ln_const + (lcol + c * CHUNK) * 4, hi_addr, lo_addr, p.has_res_lo != 0, sum2, sq2);
                    __syncwarp();  // every lane has read this ring stage: refill it with the chunk RING ahead
                    issue_residual(ld_used + RING);
                    ++ld_used;
                };
                if constexpr (KEEP_IN_REGS) {
#pragma unroll
                    for (int c = 0; c < NCH; ++c) pass1_chunk(c, xs[c]);
                    release_acc(acc);  // accumulator fully read: the MMA warp may reuse it
                } else {
#pragma unroll 1
                    for (int c = 0; c < NCH; ++c) {
                        uint32_t v[32];
                        pass1_chunk(c, v);
                        tmem_st_32x32(taddr + c * CHUNK, v);
                    }
                    tmem_st_wait();
                }
                float sum, sumsq;
                {
                    float a, b;
                    upk2(sum2, a, b); sum = a + b;
                    upk2(sq2, a, b); sumsq = a + b;
                }
                ln_part[half * BLOCK_M + row_in_tile] = make_float2(sum, sumsq);
                asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
                const float2 other = ln_part[(half ^ 1) * BLOCK_M + row_in_tile];
                sum += other.x;
                sumsq += other.y;
                if (CL > 1 && !(p.dbg & 2)) {
                    // this CTA's 2 x 128 threads now hold the partials of its BLOCK_N columns. One thread arms the
                    // barrier of this tile parity with the bytes the peers will deliver; one thread per row stores
                    // its (sum, sumsq) into every peer's slot [parity][my index among that peer's peers][row] with
                    // st.async, which counts the bytes on the peer's barrier. A peer can be at most one tile ahead, so
                    // two parities suffice.
                    const uint32_t par = it & 1;
                    if (ew == 0 && lane == 0) mbar_arrive_expect_tx(bar_ln + 8 * par, (CL - 1) * BLOCK_M * 8);
                    if (half == 0) {
#pragma unroll
                        for (int pr = 0; pr < CL; ++pr) {
                            if (pr == crank) continue;
                            const int me = crank < pr ? crank : crank - 1;  // my slot in peer pr's table
                            const uint32_t slot = ln_peer_addr + ((par * (CL - 1) + me) * BLOCK_M + row_in_tile) * 8;
                            st_async_cluster_b64(map_to_cta(slot, pr), sum, sumsq, map_to_cta(bar_ln + 8 * par, pr));
                        }
                    }
                    mbar_wait(bar_ln + 8 * par, (it >> 1) & 1);
#pragma unroll
                    for (int sl = 0; sl < CL - 1; ++sl) {
                        const float2 o = ln_peer[(par * (CL - 1) + sl) * BLOCK_M + row_in_tile];
                        sum += o.x;
                        sumsq += o.y;
                    }
                }
                const float inv_n = 1.0f / static_cast<float>(p.n);
                const float mean = sum * inv_n;
                const float var = fmaxf(sumsq * inv_n - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.eps);
                // pass 2: normalise, scale, shift, store (hi, lo)
                auto pass2_chunk = [&](int c, const uint32_t (&x)[32]) {
                    const int col0 = col_base + c * CHUNK;
                    uint32_t hi[16], lo[8];
                    ln_normalize_chunk(x, mean, rstd, ln_const + (BLOCK_N + lcol + c * CHUNK) * 4,
                                       ln_const + (2 * BLOCK_N + lcol + c * CHUNK) * 4, hi, lo);
                    if (p.dbg & 1) return;
                    uint32_t buf = stage_begin();
                    stage_words_bf16(buf, lane, hi);
                    stage_end(&tmap_c, buf, col0, row0);
                    if (p.has_out_lo) {
                        buf = stage_begin();
                        stage_words_lo8(buf, lane, lo);
                        stage_end(&tmap_c_lo, buf, col0, row0);
                    }
                };
                if constexpr (KEEP_IN_REGS) {
#pragma unroll
                    for (int c = 0; c < NCH; ++c) pass2_chunk(c, xs[c]);
                } else {
#pragma unroll 1
                    for (int c = 0; c < NCH; ++c) {
                        uint32_t x[32];
                        tmem_ld_32x32(taddr + c * CHUNK, x);
                        tmem_ld_wait();
                        if (c == NCH - 1) release_acc(acc);
                        pass2_chunk(c, x);
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
        if constexpr (PAIR) tmem_dealloc_cg2(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

// True when every byte of [ptr, ptr + bytes) is plain device memory of the CURRENT device. The range is walked
// allocation by allocation (cuMemGetAddressRange gives the extent of the allocation or mapping that holds an address), so
// a virtual range stitched together from several mappings, one of them another GPU's memory, is refused as well; so is
// anything the driver cannot describe (IPC / managed / host mappings, ranges with holes).
bool operand_is_local_device_memory(const void* ptr, size_t bytes) {
    using range_fn_t = CUresult (*)(CUdeviceptr*, size_t*, CUdeviceptr);
    static range_fn_t range_fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<range_fn_t>(f);
    }();
    int dev = -1;
    if (!ptr || bytes == 0 || !range_fn || cudaGetDevice(&dev) != cudaSuccess) return false;
    const uintptr_t end = reinterpret_cast<uintptr_t>(ptr) + bytes;
    uintptr_t cur = reinterpret_cast<uintptr_t>(ptr);
    for (int pieces = 0; cur < end; ++pieces) {
        cudaPointerAttributes pa{};
        CUdeviceptr base = 0;
        size_t size = 0;
        if (pieces >= 64 || cudaPointerGetAttributes(&pa, reinterpret_cast<const void*>(cur)) != cudaSuccess ||
            pa.type != cudaMemoryTypeDevice || pa.device != dev ||
            range_fn(&base, &size, static_cast<CUdeviceptr>(cur)) != CUDA_SUCCESS || size == 0 ||
            static_cast<uintptr_t>(base) > cur) {
            cudaGetLastError();
            return false;
        }
        cur = static_cast<uintptr_t>(base) + size;
    }
    return true;
}
template <int BLOCK_N, int EPI, typename OutT, int CL = 1, bool PAIR = false, int RING = 2>
cudaError_t launch(const GemmTcArgs& a, cudaStream_t stream, int num_sms) {
    constexpr bool OUT_F32 = sizeof(OutT) == 4;
    constexpr int CSIZE = CL > 1 ? CL : (PAIR ? 2 : 1);
    alignas(64) CUtensorMap ta, tb, tc, tcl, tr, trl;
    if constexpr (CL > 1) {
        // The A tile is multicast inside the cluster. TMA multicast loads here read only this GPU's memory, since on B200 a
        // TMA multicast load from another GPU's memory hung the GPU. The tensor map of a multicast load must therefore
        // cover this GPU's memory alone: every byte of the operand is checked, not only its first one.
        if (!operand_is_local_device_memory(a.a, static_cast<size_t>(a.m) * a.k * 2)) return cudaErrorInvalidDevicePointer;
    }
    if (!make_tmap(&ta, a.a, a.m, a.k, BLOCK_K, BLOCK_M / (CL == 3 ? 2 : (CL > 1 ? CL : 1)), false, CU_TENSOR_MAP_SWIZZLE_128B))
        return cudaErrorInvalidValue;
    if (!make_tmap(&tb, a.b, a.n, a.k, BLOCK_K, BLOCK_N / (PAIR ? 2 : 1), false, CU_TENSOR_MAP_SWIZZLE_128B))
        return cudaErrorInvalidValue;
    const CUtensorMapSwizzle csw = OUT_F32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    if (!make_tmap(&tc, a.out, a.m, a.n, CHUNK, 32, OUT_F32, csw)) return cudaErrorInvalidValue;
    tcl = tc; tr = tc; trl = tc;
    if (a.out_lo && !make_tmap_lo8(&tcl, a.out_lo, a.m, a.n, CHUNK, 32)) return cudaErrorInvalidValue;
    if (EPI == EPI_RES_LN) {
        if (!a.residual) return cudaErrorInvalidValue;
        if (!make_tmap(&tr, a.residual, a.m, a.n, CHUNK, 32, false, csw)) return cudaErrorInvalidValue;
        trl = tr;
        if (a.residual_lo && !make_tmap_lo8(&trl, a.residual_lo, a.m, a.n, CHUNK, 32))
            return cudaErrorInvalidValue;
    }
    auto kern = gemm_tc_kernel<BLOCK_N, EPI, OutT, CL, PAIR, RING>;
    constexpr int smem = smem_bytes(BLOCK_N, EPI, OUT_F32, RING, PAIR, CL);
    static_assert(num_stages(BLOCK_N, EPI, OUT_F32, RING, PAIR, CL) >= 2, "not enough shared memory for a pipeline");
    static_assert(smem <= SMEM_LIMIT, "shared memory budget exceeded");
    static PerDeviceOnce attr_set;  // per instantiation AND per device: function attributes belong to a device
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_set.mark();
    }
    GemmDev p;
    p.m = a.m; p.d_rows = a.d_rows; p.n = a.n; p.k = a.k; p.bias = a.bias;
    p.gamma = a.gamma; p.beta = a.beta; p.eps = a.eps;
    p.has_res_lo = a.residual_lo != nullptr; p.has_out_lo = a.out_lo != nullptr;
#if defined(B200_GEMM_DEBUG)  // timing experiments (results wrong by construction): never in the shipped library
    static const int dbg = [] { const char* dv = getenv("B200_EMBED_GEMM_DBG"); return dv ? atoi(dv) : 0; }();
    p.dbg = dbg;
#else
    p.dbg = 0;
#endif
    const int m_tiles = (a.m + BLOCK_M - 1) / BLOCK_M;
    const int units = CL > 1 ? m_tiles : (PAIR ? ((m_tiles + 1) / 2) : m_tiles) * (a.n / BLOCK_N);
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    pdl_attribute(&attr[0]);
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = CSIZE;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = CSIZE > 1 ? 2 : 1;
    // persistent grid: as many clusters as the device can hold at once (size-4 clusters do not tile all 148 SMs)
    static PerDeviceValue max_units_dev;
    int max_units = max_units_dev.get();
    if (max_units == 0) {
        max_units = num_sms / CSIZE;
        if (CSIZE > 1) {
            cfg.gridDim = dim3((num_sms / CSIZE) * CSIZE);
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0 && n < max_units) max_units = n;
        }
        max_units_dev.set(max_units);
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
        // N = 768 with a long K (BERT-base FFN2): 3 x 256 columns instead of 4 x 192 (experiment, B200_EMBED_LN_CL3=1)
        static const bool cl3 = getenv("B200_EMBED_LN_CL3") != nullptr;
        if (cl3 && a.n == 768 && a.k >= 1536) return launch<256, EPI_RES_LN, __nv_bfloat16, 3, false, 1>(a, stream, num_sms);
        const int cl = a.n / a.block_n;
        if (a.block_n == 128 && cl == 1) return launch<128, EPI_RES_LN, __nv_bfloat16, 1>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 1) return launch<256, EPI_RES_LN, __nv_bfloat16, 1>(a, stream, num_sms);
        // short K: the epilogue is the critical path -> deeper residual prefetch, fewer mainloop stages
        if (a.block_n == 192 && cl == 2 && a.k <= 768) return launch<192, EPI_RES_LN, __nv_bfloat16, 2, false, 3>(a, stream, num_sms);
        // long K: the mainloop is the critical path -> lean epilogue, one more mainloop stage
        const bool lean = a.k >= 1536 && getenv("B200_EMBED_LN_LEAN_OFF") == nullptr;
        if (a.block_n == 192 && cl == 2 && lean) return launch<192, EPI_RES_LN, __nv_bfloat16, 2, false, 1>(a, stream, num_sms);
        if (a.block_n == 192 && cl == 4 && lean) return launch<192, EPI_RES_LN, __nv_bfloat16, 4, false, 1>(a, stream, num_sms);
        if (a.block_n == 192 && cl == 2) return launch<192, EPI_RES_LN, __nv_bfloat16, 2>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 2) return launch<256, EPI_RES_LN, __nv_bfloat16, 2>(a, stream, num_sms);
        if (a.block_n == 192 && cl == 4) return launch<192, EPI_RES_LN, __nv_bfloat16, 4>(a, stream, num_sms);
        if (a.block_n == 256 && cl == 4) return launch<256, EPI_RES_LN, __nv_bfloat16, 4>(a, stream, num_sms);
        return cudaErrorInvalidValue;
    }
    if (a.out_f32) {
        if (a.epi != EPI_BIAS) return cudaErrorInvalidValue;
        if (a.block_n == 128) return launch<128, EPI_BIAS, float>(a, stream, num_sms);
        if (a.block_n == 256) return launch<256, EPI_BIAS, float>(a, stream, num_sms);
        return cudaErrorInvalidValue;
    }
#define B200_CASE(BN, EP)                                                                              \
    if (a.block_n == BN && a.epi == EP)                                                                \
        return a.pair ? launch<BN, EP, __nv_bfloat16, 1, true>(a, stream, num_sms)                     \
                      : launch<BN, EP, __nv_bfloat16, 1, false>(a, stream, num_sms);
    B200_CASE(128, EPI_BIAS)
    B200_CASE(192, EPI_BIAS)
    B200_CASE(256, EPI_BIAS)
    B200_CASE(128, EPI_GELU_ERF)
    B200_CASE(256, EPI_GELU_ERF)
    B200_CASE(128, EPI_GELU_TANH)
    B200_CASE(256, EPI_GELU_TANH)
    B200_CASE(128, EPI_RELU)
    B200_CASE(256, EPI_RELU)
#undef B200_CASE
    return cudaErrorInvalidValue;
}

}  // operand-format namespace
}  // namespace b200
