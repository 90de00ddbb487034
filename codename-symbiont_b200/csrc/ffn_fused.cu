// K2f: the whole feed-forward block of one BERT layer in one kernel (hidden size <= 384):
//
//     x' = LayerNorm( gelu(a . W1^T + b1) . W2^T + b2 + a )
//
// Replaces candle bert.rs BertIntermediate::forward + BertOutput::forward (called from
// /root/reference/services/preprocessing_service/src/embedding_generator.rs:198), which the reference runs as
// two f32 matmuls with separate bias / GELU / residual / LayerNorm passes. Run as two GEMM kernels the
// [tokens, I] intermediate costs an HBM write and a read (3.2 GB per layer at c2), and both GEMMs are bound by
// that traffic, not by the tensor pipe (HBM write bandwidth is 3.9 TB/s; profiles/). Here it never leaves the SM.
//
// A persistent pair of CTAs (one thread-block cluster on two SMs) owns 256 token rows at a time, 128 per CTA, and
// runs every GEMM step as one cta_group::2 MMA: each CTA stages its own rows of a and h but only HALF of every W
// block, so the W stream (the same for all rows, 2.4 MB per 128 rows from L2) is shared by 256 rows and the rings
// cover twice the lookahead. (A single-CTA version measured 1.6 ms per layer at c2 against 1.4 ms for the two
// separate GEMMs: it could not keep 192 KB of W per chunk in flight.) The activation tile a[128,H] of each CTA is
// loaded once and stays in shared memory. The intermediate is produced and consumed in chunks of 128 columns:
//     S1  = a . W1[chunk]^T                (tcgen05.mma 256x128xH, fp32 in TMEM columns 0..127 of both CTAs)
//     h   = gelu(S1 + b1[chunk]) -> bf16   (epilogue warps: tcgen05.ld -> registers -> 128B-swizzled K-major smem)
//     OUT += h . W2[:, chunk]^T            (tcgen05.mma 256xHx128, fp32 in TMEM columns 128..128+H)
// software-pipelined so that the tensor pipe runs S1 of chunk c+1 and OUT of chunk c-1 while the epilogue warps work
// on chunk c. After the last chunk the epilogue warps add b2 and the residual (hi, lo), LayerNorm the rows and
// store (hi, lo) exactly as the LayerNorm epilogue of gemm_tc.cu does.
//   warp 0      TMA producer: own a tile, own halves of the W1 blocks [64 x 64] and W2 blocks [H/NS/2 x 64] through
//               two rings; all loads of the pair complete on the leader's (rank 0) "full" barriers
//   warp 1      leader only: one thread issues all tcgen05.mma for the pair; its commits arrive in both CTAs
//   warp 2      TMEM allocator (512 columns)
//   warps 4-11  epilogue (warp w: TMEM lane quadrant w % 4, column half (w - 4) / 4)
#include "kernels.h"
#include "tc_helpers.cuh"

namespace b200 {
B200_NS_THIS {

namespace {

constexpr int FF_THREADS = 384;
constexpr int FF_EPI_WARP0 = 4;
constexpr int FF_EPI_WARPS = 8;
constexpr int FF_EPI_THREADS = FF_EPI_WARPS * 32;
constexpr int BM = 128;            // token rows per tile
constexpr int BK = 64;             // bf16 per 128-byte swizzle row
constexpr int IC = 128;            // intermediate columns per chunk
constexpr int BLK16 = BM * BK * 2;  // a [128 x 64] bf16 block: 16 KB

template <int H>
struct FfnCfg {
    static constexpr int KB1 = H / BK;                // k-blocks of GEMM1 (and of the resident a tile)
    static constexpr int NS = H > 256 ? 2 : 1;        // GEMM2 N split (UMMA N <= 256)
    static constexpr int UN2 = H / NS;                // GEMM2 UMMA N
    static constexpr int W1_BLK = (IC / 2) * BK * 2;  // this CTA's half of a W1 block: [64 x 64]
    static constexpr int W2_BLK = (UN2 / 2) * BK * 2; // this CTA's half of a W2 block: [UN2/2 x 64]
    // Both rings hold a whole number of GEMM steps, so a block's ring slot is a compile-time function of its place
    // in the step (the MMA-issuing warp has no index arithmetic to do: its issue rate bounds the kernel otherwise).
    static constexpr int R1 = H == 384 ? 1 : (H == 256 ? 2 : 3);  // GEMM1 steps in the W1 ring
    static constexpr int R2 = H == 384 ? 1 : 2;       // GEMM2 steps in the W2 ring
    static constexpr int NB2 = NS * (IC / BK);        // W2 blocks per GEMM2 step
    static constexpr int S1N = R1 * KB1;              // W1 ring stages
    static constexpr int S2N = R2 * NB2;              // W2 ring stages
    static_assert(S1N <= 8 && S2N <= 8, "barrier slots");
    static constexpr int OFF_A = 0;
    static constexpr int OFF_W1 = KB1 * BLK16;
    static constexpr int OFF_H = OFF_W1 + S1N * W1_BLK;  // h chunk: 2 k-blocks [128 x 64]
    static constexpr int OFF_W2 = OFF_H + 2 * BLK16;
    static constexpr int OFF_AUX = OFF_W2 + S2N * W2_BLK;
    static constexpr int AUX = 1024 + 2 * BM * 8;     // barriers + tmem slot, LayerNorm partial statistics
    static constexpr int BYTES = OFF_AUX + AUX;
    // h + the W2 ring: per-warp scratch of the final epilogue. The W1 ring is NOT part of it, so that the first GEMM1 of
    // the next tile is loaded and issued while the epilogue warps are still normalising this one.
    static constexpr int SCRATCH = OFF_AUX - OFF_H;
    static constexpr int LO_SLOTS = H == 128 ? 3 : 4;  // residual-lo ring of a warp; pass 2 reuses it as output staging
    static constexpr int HALF = H / 2;                // output columns per epilogue warp
    static constexpr int NCH = HALF / CHUNK;
    static_assert(H % 128 == 0 && H <= 384, "fused FFN needs the OUT accumulator (H columns) next to S1 (128) in TMEM");
    static_assert(SCRATCH >= FF_EPI_WARPS * LO_SLOTS * CHUNK_BF16 + 3 * H * 4, "final epilogue scratch does not fit");
    static_assert(BYTES <= SMEM_LIMIT, "shared memory budget exceeded");
    static_assert(W1_BLK % 1024 == 0 && W2_BLK % 1024 == 0 && (UN2 / 2) % 8 == 0, "swizzle atoms must stay 1024-byte aligned");
};

struct FfnDev {
    int m;
    const int* d_rows;
    int inter;  // I
    const float *b1, *b2, *gamma, *beta;
    float eps;
    int has_lo;
};

template <int H>
__global__ void __launch_bounds__(FF_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w1,
                 const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_r,
                 const __grid_constant__ CUtensorMap tmap_r_lo, const __grid_constant__ CUtensorMap tmap_lo_pf,
                 const __grid_constant__ CUtensorMap tmap_c,
                 const __grid_constant__ CUtensorMap tmap_c_lo, const FfnDev p) {
    using C = FfnCfg<H>;
    constexpr int KB1 = C::KB1, UN2 = C::UN2, W1_BLK = C::W1_BLK, W2_BLK = C::W2_BLK, S1N = C::S1N, S2N = C::S2N;
    constexpr int R1 = C::R1, R2 = C::R2, NB2 = C::NB2;
    constexpr uint16_t BOTH = 3;  // multicast mask: both CTAs of the pair
    constexpr int HALF = C::HALF, NCH = C::NCH;
    constexpr int OUT_COL = IC;  // OUT accumulator starts after the 128 S1 columns

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    if ((base & 1023u) != 0) __trap();
    const uint32_t smem_a = base + C::OFF_A, smem_h = base + C::OFF_H, smem_w1 = base + C::OFF_W1,
                   smem_w2 = base + C::OFF_W2, aux = base + C::OFF_AUX;
    uint8_t* aux_ptr = smem_raw + C::OFF_AUX;
    const uint32_t bar_a_full = aux, bar_a_empty = aux + 8;
    const uint32_t bar_w1_full = aux + 16, bar_w1_empty = aux + 80;    // <= 8 each
    const uint32_t bar_w2_full = aux + 144, bar_w2_empty = aux + 208;  // <= 8 each
    const uint32_t bar_s1_full = aux + 272, bar_s1_empty = aux + 280, bar_h_full = aux + 288, bar_h_empty = aux + 296;
    const uint32_t bar_out_full = aux + 304, bar_out_empty = aux + 312, bar_epi_done = aux + 320;
    const uint32_t bar_res = aux + 384;   // FF_EPI_WARPS x 4
    const uint32_t tmem_slot = aux + 768;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(aux_ptr + 768);
    float2* ln_part = reinterpret_cast<float2*>(aux_ptr + 1024);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_trigger();
    const int NC = p.inter / IC;  // chunks per tile
    // Pair `pr` owns row tiles 2 pr (leader) and 2 pr + 1 (peer). With an odd tile count the last peer tile is out of
    // range: its loads are zero-filled and its stores clipped, and it keeps its side of every handshake.
    const uint32_t crank = cluster_ctarank();
    const bool leader = crank == 0;
    const int pair0 = static_cast<int>(blockIdx.x) / 2, pair_step = static_cast<int>(gridDim.x) / 2;
    // "full" barriers and the barriers the MMA thread waits on live in the leader CTA
    const uint32_t l_a_full = map_to_cta(bar_a_full, 0), l_w1_full = map_to_cta(bar_w1_full, 0),
                   l_w2_full = map_to_cta(bar_w2_full, 0), l_s1_empty = map_to_cta(bar_s1_empty, 0),
                   l_h_full = map_to_cta(bar_h_full, 0), l_out_empty = map_to_cta(bar_out_empty, 0);

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_w1); prefetch_tmap(&tmap_w2);
        mbar_init(bar_a_full, 1);
        mbar_init(bar_a_empty, 1 + FF_EPI_WARPS);  // last GEMM1 done + every epilogue warp has read its residual rows
        for (int s = 0; s < S1N; ++s) { mbar_init(bar_w1_full + 8 * s, 1); mbar_init(bar_w1_empty + 8 * s, 1); }
        for (int s = 0; s < S2N; ++s) { mbar_init(bar_w2_full + 8 * s, 1); mbar_init(bar_w2_empty + 8 * s, 1); }
        // one arrive per epilogue warp of both CTAs on the barriers the MMA thread waits on
        mbar_init(bar_s1_full, 1); mbar_init(bar_s1_empty, 2 * FF_EPI_WARPS);
        mbar_init(bar_h_full, 2 * FF_EPI_WARPS); mbar_init(bar_h_empty, 1);
        mbar_init(bar_out_full, 1); mbar_init(bar_out_empty, 2 * FF_EPI_WARPS);
        mbar_init(bar_epi_done, FF_EPI_THREADS);
        for (int s = 0; s < 4 * FF_EPI_WARPS; ++s) mbar_init(bar_res + 8 * s, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_cg2(tmem_slot, 512);
    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();  // the peer's barriers exist before anything arrives on them
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();  // nothing above reads or writes global memory; the predecessor's output (and the row count) from here on
    int rows = p.m;
    if (p.d_rows != nullptr) {
        const int r = *p.d_rows;
        rows = r < rows ? r : rows;
    }
    const int num_tiles = (rows + BM - 1) / BM;
    const int num_pairs = (num_tiles + 1) / 2;

    // Both the producer and the MMA thread walk the tensor-pipe schedule of a tile in the same order:
    //   G1(0), [G1(1), G2(0)], [G1(2), G2(1)], ..., [G1(NC-1), G2(NC-2)], G2(NC-1)
    if (warp == 0) {
        // The whole warp walks the schedule so that control flow stays uniform; one elected lane issues the copies.
        uint32_t g1_it = 0, g2_it = 0, tile_it = 0;
        for (int pr = pair0; pr < num_pairs; pr += pair_step, ++tile_it) {
            const int t = 2 * pr + static_cast<int>(crank);
            // the a tile: free once the previous tile's last GEMM1 has completed
            mbar_wait(bar_a_empty, (tile_it & 1) ^ 1);
            if (elect_one()) {
                if (leader) mbar_arrive_expect_tx(bar_a_full, 2 * KB1 * BLK16);
#pragma unroll
                for (int kb = 0; kb < KB1; ++kb) tma_load_2d_cg2(smem_a + kb * BLK16, &tmap_a, kb * BK, t * BM, l_a_full);
                if (p.has_lo) {  // the low halves of the same rows are needed by the final epilogue: pull them into L2 now
#pragma unroll
                    for (int kb = 0; kb < KB1; ++kb) tma_prefetch_2d(&tmap_lo_pf, kb * BK, t * BM);
                }
            }
            __syncwarp();
            for (int step = 0; step <= NC; ++step) {
                if (step < NC) {
                    const uint32_t s0 = (g1_it % R1) * KB1, ph = ((g1_it / R1) & 1) ^ 1;
#pragma unroll
                    for (int kb = 0; kb < KB1; ++kb) {
                        mbar_wait(bar_w1_empty + 8 * (s0 + kb), ph);
                        if (elect_one()) {
                            if (leader) mbar_arrive_expect_tx(bar_w1_full + 8 * (s0 + kb), 2 * W1_BLK);
                            tma_load_2d_cg2(smem_w1 + (s0 + kb) * W1_BLK, &tmap_w1, kb * BK, step * IC + crank * (IC / 2),
                                            l_w1_full + 8 * (s0 + kb));
                        }
                        __syncwarp();
                    }
                    ++g1_it;
                }
                if (step >= 1) {
                    const int ic = step - 1;
                    // the W2 ring (and h) double as scratch of the previous tile's final epilogue
                    if (ic == 0) mbar_wait(bar_epi_done, (tile_it & 1) ^ 1);
                    const uint32_t s0 = (g2_it % R2) * NB2, ph = ((g2_it / R2) & 1) ^ 1;
#pragma unroll
                    for (int blk = 0; blk < NB2; ++blk) {
                        const int ns = blk / (IC / BK), kb = blk % (IC / BK);
                        mbar_wait(bar_w2_empty + 8 * (s0 + blk), ph);
                        if (elect_one()) {
                            if (leader) mbar_arrive_expect_tx(bar_w2_full + 8 * (s0 + blk), 2 * W2_BLK);
                            tma_load_2d_cg2(smem_w2 + (s0 + blk) * W2_BLK, &tmap_w2, ic * IC + kb * BK,
                                            ns * UN2 + crank * (UN2 / 2), l_w2_full + 8 * (s0 + blk));
                        }
                        __syncwarp();
                    }
                    ++g2_it;
                }
            }
        }
    } else if (warp == 1) {
        // The whole warp walks the schedule (waits included) so that control flow stays uniform; one elected lane issues.
        if (leader) {
            constexpr uint32_t idesc1 = umma_idesc_op(2 * BM, IC);
            constexpr uint32_t idesc2 = umma_idesc_op(2 * BM, UN2);
            uint32_t tile_it = 0, g1_it = 0, g2_it = 0;  // g*_it: chunks issued so far
            const uint64_t da0 = umma_desc_sw128(smem_a), dh0 = umma_desc_sw128(smem_h);
            const uint64_t dw1 = umma_desc_sw128(smem_w1), dw2 = umma_desc_sw128(smem_w2);
            for (int pr = pair0; pr < num_pairs; pr += pair_step, ++tile_it) {
                mbar_wait(bar_a_full, tile_it & 1);
                for (int step = 0; step <= NC; ++step) {
                    if (step < NC) {  // G1(step): S1 = a . W1[chunk]^T
                        mbar_wait_cluster(bar_s1_empty, (g1_it & 1) ^ 1);  // both epilogues have drained the previous S1
                        tc_fence_after_sync();
                        const uint32_t s0 = (g1_it % R1) * KB1, ph = (g1_it / R1) & 1;
#pragma unroll
                        for (int kb = 0; kb < KB1; ++kb) {
                            mbar_wait(bar_w1_full + 8 * (s0 + kb), ph);
                            tc_fence_after_sync();
                            if (elect_one()) {
                                const uint64_t da = da0 + ((kb * BLK16) >> 4);
                                const uint64_t db = dw1 + (((s0 + kb) * W1_BLK) >> 4);
#pragma unroll
                                for (int k = 0; k < BK / 16; ++k)
                                    umma_bf16_cg2(tmem_base, da + 2 * k, db + 2 * k, idesc1, (kb > 0 || k > 0) ? 1u : 0u);
                                umma_commit_cg2(bar_w1_empty + 8 * (s0 + kb), BOTH);
                                if (kb == KB1 - 1) {
                                    umma_commit_cg2(bar_s1_full, BOTH);
                                    // the a tiles have been read for the last time
                                    if (step == NC - 1) umma_commit_cg2(bar_a_empty, BOTH);
                                }
                            }
                            __syncwarp();
                        }
                        ++g1_it;
                    }
                    if (step >= 1) {  // G2(step - 1): OUT += h . W2[:, chunk]^T
                        const int ic = step - 1;
                        mbar_wait_cluster(bar_h_full, g2_it & 1);  // h(ic) is in both CTAs' shared memory
                        if (ic == 0) mbar_wait_cluster(bar_out_empty, (tile_it & 1) ^ 1);  // previous tiles' OUT has been read
                        tc_fence_after_sync();
                        const uint32_t s0 = (g2_it % R2) * NB2, ph = (g2_it / R2) & 1;
#pragma unroll
                        for (int blk = 0; blk < NB2; ++blk) {
                            const int ns = blk / (IC / BK), kb = blk % (IC / BK);
                            mbar_wait(bar_w2_full + 8 * (s0 + blk), ph);
                            tc_fence_after_sync();
                            if (elect_one()) {
                                const uint64_t da = dh0 + ((kb * BLK16) >> 4);
                                const uint64_t db = dw2 + (((s0 + blk) * W2_BLK) >> 4);
#pragma unroll
                                for (int k = 0; k < BK / 16; ++k)
                                    umma_bf16_cg2(tmem_base + OUT_COL + ns * UN2, da + 2 * k, db + 2 * k, idesc2,
                                                  (ic > 0 || kb > 0 || k > 0) ? 1u : 0u);
                                umma_commit_cg2(bar_w2_empty + 8 * (s0 + blk), BOTH);
                                if (blk == NB2 - 1) {
                                    umma_commit_cg2(bar_h_empty, BOTH);  // h(ic) consumed
                                    if (ic == NC - 1) umma_commit_cg2(bar_out_full, BOTH);
                                }
                            }
                            __syncwarp();
                        }
                        ++g2_it;
                    }
                }
            }
        }
    } else if (warp >= FF_EPI_WARP0) {
        const int ew = warp - FF_EPI_WARP0;
        const int quad = warp & 3, half = ew >> 2;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        // final-epilogue scratch of this warp inside the (then idle) h / W2-ring region: a residual-lo ring for the first
        // pass, reused as (hi, lo) output staging by the second
        constexpr int LO_SLOTS = C::LO_SLOTS;
        const uint32_t my_smem = smem_h + ew * (LO_SLOTS * CHUNK_BF16);
        const uint32_t const_s = smem_h + FF_EPI_WARPS * (LO_SLOTS * CHUNK_BF16);  // b2 | gamma | beta, fp32 (final epilogue)
        uint32_t chunk_it = 0, tile_it = 0, ld_issued = 0, ld_used = 0;

        // residual: the hi half is the a tile itself (still in shared memory); the lo half comes through a 4-slot ring
        auto issue_lo = [&](int tile, int c) {
            if (lane == 0) {
                const uint32_t s = ld_issued % LO_SLOTS;
                const uint32_t bar = bar_res + 8 * (4 * ew + s);
                mbar_arrive_expect_tx(bar, CHUNK_BF16 / 2);  // 32 x 32 e5m2 bytes in the first half of the slot
                tma_load_2d(my_smem + s * CHUNK_BF16, &tmap_r_lo, half * HALF + c * CHUNK, tile * BM + quad * 32, bar);
            }
            ++ld_issued;
        };

        for (int pr = pair0; pr < num_pairs; pr += pair_step, ++tile_it) {
            const int t = 2 * pr + static_cast<int>(crank);
            // ---- per chunk: h = gelu(S1 + b1) as the bf16 A operand of the second GEMM ----
            for (int ic = 0; ic < NC; ++ic, ++chunk_it) {
                // this warp's 64 bias values -> its 256 B slot of the (now idle) LayerNorm scratch, fetched while S1 is
                // still being computed
                const uint32_t bias_s = aux + 1024 + ew * 256;
                {
                    const float2 b = __ldg(reinterpret_cast<const float2*>(p.b1 + ic * IC + half * 64) + lane);
                    __syncwarp();  // every lane has finished reading the previous chunk's values
                    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(bias_s + lane * 8), "f"(b.x), "f"(b.y) : "memory");
                    __syncwarp();
                }
                mbar_wait(bar_s1_full, chunk_it & 1);
                tc_fence_after_sync();
                uint32_t v[64];  // this thread's row, S1 columns half*64 .. +63
                tmem_ld_32x32(t_lane + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                tmem_ld_32x32(t_lane + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                tmem_ld_wait();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(l_s1_empty);  // S1 is in registers: the next chunk's GEMM1 may overwrite it
                uint32_t hp[32];
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const uint4 b = lds_v4(bias_s + q * 16);
                    hp[q * 2 + 0] = epi_pair_bf16<EPI_GELU_ERF>(v[q * 4 + 0], v[q * 4 + 1], __uint_as_float(b.x), __uint_as_float(b.y));
                    hp[q * 2 + 1] = epi_pair_bf16<EPI_GELU_ERF>(v[q * 4 + 2], v[q * 4 + 3], __uint_as_float(b.z), __uint_as_float(b.w));
                }
                // the h buffer is free once GEMM2 of the previous chunk has completed
                mbar_wait(bar_h_empty, (chunk_it & 1) ^ 1);
                // k-block `half` of the h chunk, K-major, 128-byte rows, SWIZZLE_128B
                const uint32_t hrow = smem_h + half * BLK16 + row_in_tile * 128;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const uint32_t a = hrow + ((static_cast<uint32_t>(q) ^ (row_in_tile & 7)) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(hp[q * 4 + 0]), "r"(hp[q * 4 + 1]),
                                 "r"(hp[q * 4 + 2]), "r"(hp[q * 4 + 3])
                                 : "memory");
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(l_h_full);
            }
            // ---- final: x' = LayerNorm(OUT + b2 + a) -> (hi, lo) ----
            mbar_wait(bar_out_full, tile_it & 1);  // every MMA of this tile has completed: h and the rings are idle
            tc_fence_after_sync();
            if (p.has_lo) {
#pragma unroll
                for (int c = 0; c < (NCH < LO_SLOTS ? NCH : LO_SLOTS); ++c) issue_lo(t, c);
            }
            // b2, gamma, beta -> scratch behind the per-warp areas, read back with broadcast LDS in the loops
            for (int i = threadIdx.x - FF_EPI_WARP0 * 32; i < 3 * H; i += FF_EPI_THREADS) {
                const float* src = i < H ? p.b2 + i : (i < 2 * H ? p.gamma + (i - H) : p.beta + (i - 2 * H));
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(const_s + i * 4), "f"(__ldg(src)) : "memory");
            }
            asm volatile("bar.sync 1, %0;" ::"n"(FF_EPI_THREADS) : "memory");
            const uint32_t taddr = t_lane + OUT_COL + half * HALF;
            const int row0 = t * BM + quad * 32;
            uint64_t sum2 = 0, sq2 = 0;  // (0.f, 0.f)
            {
                uint32_t v[2][32];
                tmem_ld_32x32(taddr, v[0]);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const int col0 = half * HALF + c * CHUNK;
                    uint32_t hi_addr[4], lo_addr[2];
                    // a tile: k-block col0 / 64, 128-byte rows, SWIZZLE_128B; lo slot: 32-byte rows, SWIZZLE_32B
                    const uint32_t arow = smem_a + (col0 / BK) * BLK16 + row_in_tile * 128;
#pragma unroll
                    for (int q = 0; q < 4; ++q) hi_addr[q] = arow + ((((col0 % BK) / 8 + q) ^ (row_in_tile & 7)) << 4);
                    lo8_piece_addrs(my_smem + (ld_used % LO_SLOTS) * CHUNK_BF16, lane, lo_addr);
                    if (p.has_lo) {
                        mbar_wait(bar_res + 8 * (4 * ew + (ld_used % LO_SLOTS)), (ld_used / LO_SLOTS) & 1);
                        ++ld_used;
                    }
                    tmem_ld_wait();
                    if (c + 1 < NCH) tmem_ld_32x32(taddr + (c + 1) * CHUNK, v[(c + 1) & 1]);
                    ln_accumulate_chunk(v[c & 1], const_s + col0 * 4, hi_addr, lo_addr, p.has_lo != 0, sum2, sq2);
                    tmem_st_32x32(taddr + c * CHUNK, v[c & 1]);
                    __syncwarp();
                    if (p.has_lo && c + LO_SLOTS < NCH) issue_lo(t, c + LO_SLOTS);
                }
            }
            if (lane == 0) mbar_arrive(bar_a_empty);  // this warp's residual rows have been read from the a tile
            float sum, sumsq;
            {
                float a, b;
                upk2(sum2, a, b); sum = a + b;
                upk2(sq2, a, b); sumsq = a + b;
            }
            tmem_st_wait();
            ln_part[half * BM + row_in_tile] = make_float2(sum, sumsq);
            asm volatile("bar.sync 1, %0;" ::"n"(FF_EPI_THREADS) : "memory");
            const float2 other = ln_part[(half ^ 1) * BM + row_in_tile];
            const float inv_n = 1.0f / static_cast<float>(H);
            const float mean = (sum + other.x) * inv_n;
            const float var = fmaxf((sumsq + other.y) * inv_n - mean * mean, 0.f);
            const float rstd = rsqrtf(var + p.eps);
            {
                uint32_t x[2][32];
                tmem_ld_32x32(taddr, x[0]);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    tmem_ld_wait();
                    if (c + 1 < NCH) {
                        tmem_ld_32x32(taddr + (c + 1) * CHUNK, x[(c + 1) & 1]);
                    } else {  // OUT fully read: the next tile's GEMM2 may overwrite it
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(l_out_empty);
                    }
                    const int col0 = half * HALF + c * CHUNK;
                    uint32_t hi[16], lo[8];
                    ln_normalize_chunk(x[c & 1], mean, rstd, const_s + (H + col0) * 4, const_s + (2 * H + col0) * 4, hi, lo);
                    const uint32_t buf_hi = my_smem, buf_lo = my_smem + CHUNK_BF16;
                    if (lane == 0) tma_store_wait_read<0>();
                    __syncwarp();
                    stage_words_bf16(buf_hi, lane, hi);
                    if (p.has_lo) stage_words_lo8(buf_lo, lane, lo);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&tmap_c, buf_hi, col0, row0);
                        if (p.has_lo) tma_store_2d(&tmap_c_lo, buf_lo, col0, row0);
                        tma_store_commit();
                    }
                }
            }
            // scratch (h + rings) may be refilled once the TMA stores have read their staging buffers
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
            asm volatile("bar.sync 1, %0;" ::"n"(FF_EPI_THREADS) : "memory");  // also protects ln_part
            mbar_arrive(bar_epi_done);
        }
        if (lane == 0) tma_store_wait<0>();
    }

    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();  // neither CTA leaves while the pair's MMAs or barrier arrivals may still touch it
    if (warp == 2) {
        tc_fence_after_sync();
        tmem_dealloc_cg2(tmem_base, 512);
    }
}

template <int H>
cudaError_t launch_ffn(const FfnFusedArgs& a, cudaStream_t stream, int num_sms) {
    using C = FfnCfg<H>;
    alignas(64) CUtensorMap ta, tw1, tw2, tr, trl, tlp, tc, tcl;
    if (!make_tmap(&ta, a.a, a.m, H, BK, BM, false, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
    if (!make_tmap(&tw1, a.w1, a.inter, H, BK, IC / 2, false, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
    if (!make_tmap(&tw2, a.w2, H, a.inter, BK, C::UN2 / 2, false, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
    if (!make_tmap(&tr, a.a, a.m, H, CHUNK, 32, false, CU_TENSOR_MAP_SWIZZLE_64B)) return cudaErrorInvalidValue;
    trl = tr;
    if (a.a_lo && !make_tmap_lo8(&trl, a.a_lo, a.m, H, CHUNK, 32)) return cudaErrorInvalidValue;
    tlp = ta;
    if (a.a_lo && !make_tmap_es(&tlp, a.a_lo, a.m, H, BK, BM, 1, CU_TENSOR_MAP_SWIZZLE_NONE)) return cudaErrorInvalidValue;  // L2 prefetch only
    if (!make_tmap(&tc, a.out, a.m, H, CHUNK, 32, false, CU_TENSOR_MAP_SWIZZLE_64B)) return cudaErrorInvalidValue;
    tcl = tc;
    if (a.out_lo && !make_tmap_lo8(&tcl, a.out_lo, a.m, H, CHUNK, 32)) return cudaErrorInvalidValue;
    auto kern = ffn_fused_kernel<H>;
    static PerDeviceOnce attr_set;  // function attributes belong to a (function, device) pair
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::BYTES);
        if (e != cudaSuccess) return e;
        attr_set.mark();
    }
    FfnDev p;
    p.m = a.m; p.d_rows = a.d_rows; p.inter = a.inter; p.b1 = a.b1; p.b2 = a.b2; p.gamma = a.gamma; p.beta = a.beta;
    p.eps = a.eps; p.has_lo = (a.a_lo != nullptr && a.out_lo != nullptr) ? 1 : 0;
    const int pairs = ((a.m + BM - 1) / BM + 1) / 2;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(FF_THREADS);
    cfg.dynamicSmemBytes = C::BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    pdl_attribute(&attr[1]);
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    static PerDeviceValue max_pairs_dev;  // CTA pairs the device holds at once
    int max_pairs = max_pairs_dev.get();
    if (max_pairs == 0) {
        max_pairs = num_sms / 2;
        cfg.gridDim = dim3(2 * (num_sms / 2));
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0 && n < max_pairs) max_pairs = n;
        max_pairs_dev.set(max_pairs);
    }
    const int grid = 2 * (pairs < max_pairs ? pairs : max_pairs);
    if (grid <= 0) return cudaSuccess;
    cfg.gridDim = dim3(grid);
    return cudaLaunchKernelEx(&cfg, kern, ta, tw1, tw2, tr, trl, tlp, tc, tcl, p);
}

}  // namespace

bool ffn_fused_supported(int hidden, int inter, int act) {
    return (hidden == 128 || hidden == 256 || hidden == 384) && inter % IC == 0 && act == 0;
}

cudaError_t ffn_fused_launch(const FfnFusedArgs& a, cudaStream_t stream, int num_sms) {
    if (a.m <= 0) return cudaSuccess;
    if (a.inter % IC != 0) return cudaErrorInvalidValue;
    if (a.hidden == 128) return launch_ffn<128>(a, stream, num_sms);
    if (a.hidden == 256) return launch_ffn<256>(a, stream, num_sms);
    if (a.hidden == 384) return launch_ffn<384>(a, stream, num_sms);
    return cudaErrorInvalidValue;
}

}  // operand-format namespace
}  // namespace b200
