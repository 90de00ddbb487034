// K23: Q/K/V projection + self-attention of one sentence in one kernel (hidden <= 384, head_dim 32, sentences of at
// most 128 tokens):
//
//     ctx[s, h*d:(h+1)*d] = softmax(q_h . k_h^T / sqrt(d) + mask) . v_h,   [q_h | k_h | v_h] = x_s . Wqkv_h^T + bqkv_h
//
// Replaces candle bert.rs BertSelfAttention::forward as a whole (query / key / value Linear, scores, softmax, .v, head
// merge; called from /root/reference/services/preprocessing_service/src/embedding_generator.rs:198). Run as the K2 QKV
// GEMM followed by K3 the packed [tokens, 3H] qkv tensor costs an HBM write and a read (2.4 GB per layer at c2) and the
// GEMM is bound by exactly that write; here q, k, v never leave the SM.
//
// Work item = a TILE of up to 128 token rows holding whole sentences (one sentence of 65..128 tokens, or several short
// ones packed by pack_tiles_kernel; attention is then block-diagonal: every row only sees the keys of its own
// sentence). A persistent CTA (one per SM) keeps the tile's activations x[128, H] resident in shared memory (rows past
// the tile's tokens belong to the next tile or are zero-filled: they are computed and never used) and walks its heads.
// Five warpgroups (640 threads; setmaxnreg moves registers from the control and output warps to the softmax groups):
//   warp 0      TMA producer: the x tile once per sentence; per head KB blocks of Wqkv rows {h*d.., H + h*d.., 2H + h*d..}
//               x 64 columns (three 4 KB boxes = one 96-row K-major B operand) through a small ring
//   warp 1      MMA issuer of the projections  ACC[h & 1] = x . Wqkv_h^T  (128 x 96 x H per head h, up to two heads ahead;
//               the whole warp walks the loop, one elected lane issues)
//   warps 2, 3  MMA issuers of group 0 / 1: S = Q . K_u^T (128 x 64 x d) and O_u = P_u . V_u (128 x d x 64) per unit u of
//               64 keys. Three issuers, each blocking on its own dependencies only: one issuer with a fixed schedule kept the
//               groups in lockstep (softmax warps spent 27 % of their samples waiting for S / O / ACC, profiles/r02a).
//               Warp 2 is also the TMEM allocator (512 columns: 2 x 96 ACC, 2 x 64 S, 2 x 2 x 32 O, 2 x 32 P)
//   warps 4-7 / 8-11  softmax group 0 / 1 (even / odd heads), thread = token row = TMEM lane. Per unit the softmax of
//               attention_tc.cu: S read once into registers, P written as 16-bit pairs into tensor memory (the P.V MMA reads its
//               A operand there). Nothing else: a unit's P.V goes to its own O buffer, and the head's rescaling factor and
//               row sum are handed to the output warps through shared memory
//   warps 12-15 convert: ACC + bias -> 16-bit Q / K / V tiles of the head's group in shared memory, in the layouts K3
//               gets from TMA (64-byte rows, SWIZZLE_64B; V consumed MN-major as stored), head after head
//   warps 16-19 output: ctx rows = (O_0 * corr + O_1) / l of finished heads, stored as whole 32-byte sectors
// With convert and output inside the softmax warps (12 warps, 168 registers) a head took a group 7.5 k cycles of which 1.8 k
// were these two; as separate warps under a 128-register launch bound they spilled (0.91 ms) - setmaxnreg is what makes
// the split pay: 13.83 vs 14.14 ms per c2 step in a same-box A/B.
#include <cuda.h>
#include <cudaTypedefs.h>
#ifdef QA_TIMELINE
#include <cstdio>
#endif

#include "common.cuh"
#include "kernels.h"
#include "tc_helpers.cuh"

namespace b200 {
B200_NS_THIS {

namespace {

constexpr int QA_THREADS = 640;    // five warpgroups: control, softmax 0, softmax 1, convert, output
constexpr int QA_GROUP_WARP0 = 4;  // group 0: warps 4-7, group 1: warps 8-11
constexpr int QA_CONVERT_WARP0 = 12, QA_OUTPUT_WARP0 = 16;
constexpr int QA_GROUP = 128;      // threads per group
// registers per thread after the roles have split (setmaxnreg; 640 threads start at 96, the sum over the five warpgroups
// must stay <= 5 * 96)
constexpr int QA_REGS_CONTROL = 80, QA_REGS_SOFTMAX = 112, QA_REGS_CONVERT = 104, QA_REGS_OUTPUT = 72;
static_assert(QA_REGS_CONTROL + 2 * QA_REGS_SOFTMAX + QA_REGS_CONVERT + QA_REGS_OUTPUT <= 5 * 96, "register budget");
constexpr int BLK = 128;           // token rows per item
constexpr int UNIT = 64;           // keys per softmax unit
constexpr int BK = 64;             // bf16 per 128-byte swizzle row
constexpr int X_BLK = BLK * BK * 2;

// Measured options, off (same-box A/B at c2, kernel alone: 0.626 ms as shipped): a second V tile per group (QA_VBUFS = 2)
// so that the next head's Q / K / V are staged before the current head's last P.V has finished: 0.636 ms, and with the next
// head's first S then issued ahead of that P.V 0.624-0.672 ms (that issue order is no longer in the source): the tensor pipe
// executes in issue order and the chain a group waits on is not shortened by queueing more behind the projections.
// QA_JOINT = 1 projects both groups' heads in one N = 192 chain into one accumulator (x read from shared memory once per
// head pair): 0.655 ms, 13.92 vs 13.77 ms per step.
#ifndef QA_JOINT
#define QA_JOINT 0
#endif
#ifndef QA_VBUFS
#define QA_VBUFS 1
#endif

template <int H, int D>
struct QaCfg {
    static constexpr int KB = H / BK;
    static constexpr int HEADS = H / D;
    static constexpr int NQ = 3 * D;                 // accumulator columns per head: q | k | v
    static constexpr int W_SUB = D * BK * 2;         // one [D x 64] box of Wqkv rows
    static constexpr int PASS_HEADS = QA_JOINT ? 2 : 1;   // heads per projection pass (one accumulator of PASS_HEADS * NQ columns)
    static constexpr int W_BLK = PASS_HEADS * 3 * W_SUB;  // [PASS_HEADS * 3D x 64]: one K-major B operand block
    // W ring slots (KB % WS == 0: slots are compile-time in the k loop). A ring holding the whole head (6 slots at H = 384)
    // was measured slower: 0.703 vs 0.644 ms per launch at c2
    static constexpr int WS = QA_JOINT ? 2 : (KB % 3 == 0 ? 3 : 2);
    static constexpr int T_TILE = BLK * D * 2;       // one Q / K / V tile
    static constexpr int OFF_X = 0;
    static constexpr int OFF_W = KB * X_BLK;
    static constexpr int OFF_T = OFF_W + WS * W_BLK;   // per group: Q, K, V of even heads, V of odd heads (of the group's heads)
    static constexpr int G_TILES = 2 + QA_VBUFS;
    static constexpr int OFF_AUX = OFF_T + 2 * G_TILES * T_TILE;
    static constexpr int OFF_BIAS = OFF_AUX + 512;     // bqkv [3H] fp32
    static constexpr int OFF_ROWS = OFF_BIAS + 3 * H * 4;  // per group: corr[128], l[128] of the head being finished
    // score-bias tiles of a packed tile (see the convert warps): [128 rows x 32 columns] 16-bit, Q side and K side, in the
    // layout of the Q / K tiles (64-byte rows, SWIZZLE_64B)
    static constexpr int OFF_BIASQ = (OFF_ROWS + 2 * 2 * BLK * 4 + 1023) / 1024 * 1024;
    static constexpr int OFF_BIASK = OFF_BIASQ + T_TILE;
    static constexpr int BYTES = OFF_BIASK + T_TILE;
    static constexpr int MAX_PACKED = 31;  // sentences per tile the score-bias columns can tell apart
    // tensor memory: ACC[2] (q | k | v of the head being projected), per group S (one unit of 64 keys), O (two units), and P:
    // the probabilities as 16-bit pairs, one 32-bit column per two keys, which the P.V MMA reads straight from tensor
    // memory (A operand in TMEM) - they never go through shared memory
    static constexpr int ACC_COL = 0, ACC_STRIDE = NQ, S_COL = 2 * NQ, O_COL = S_COL + 2 * UNIT, P_COL = O_COL + 4 * D;
    static_assert(P_COL + 2 * (UNIT / 2) <= 512, "tensor memory budget");
    static_assert(HEADS % 2 == 0, "the two groups take alternate heads");
    static_assert(D == 32, "the q / k / v tiles use 64-byte rows");
    static_assert(KB % WS == 0 && H % BK == 0 && H % D == 0, "shape");
    static_assert(W_SUB % 1024 == 0, "swizzle atoms must stay 1024-byte aligned");
    static_assert(NQ % 16 == 0 && NQ <= ACC_STRIDE, "UMMA N");
    static_assert(BYTES <= SMEM_LIMIT, "shared memory budget exceeded");
};

struct QaDev {
    const int* cu;
    const int* tile_first;  // [n_tiles + 1] first sentence of every tile
    const int* n_tiles;     // device scalar
    const float* bias;  // [3H]
    int batch, H;
    float scale_log2;
    __nv_bfloat16* ctx;
};

// -DQA_TIMELINE: cycles per phase of one thread per role (block 0), printed at kernel end - a measuring build only
#ifdef QA_TIMELINE
#define TL_DECL uint32_t tl[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; uint32_t tl_last = static_cast<uint32_t>(clock())
#define TL(i) do { const uint32_t now_ = static_cast<uint32_t>(clock()); tl[i] += now_ - tl_last; tl_last = now_; } while (0)
#define TL_PRINT(name) do { if (blockIdx.x == 0 && lane == 0) printf("TL %s w%d: %u %u %u %u %u %u %u %u %u %u %u %u\n", name, warp, \
    tl[0], tl[1], tl[2], tl[3], tl[4], tl[5], tl[6], tl[7], tl[8], tl[9], tl[10], tl[11]); } while (0)
#else
#define TL_DECL
#define TL(i)
#define TL_PRINT(name)
#endif

template <int H, int D>
__global__ void __launch_bounds__(QA_THREADS, 1)
qkv_attention_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w, const QaDev p) {
    using C = QaCfg<H, D>;
    constexpr int KB = C::KB, HEADS = C::HEADS, NQ = C::NQ, W_SUB = C::W_SUB, W_BLK = C::W_BLK, WS = C::WS;
    constexpr int T_TILE = C::T_TILE;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    if ((base & 1023u) != 0) __trap();
    const uint32_t smem_x = base + C::OFF_X, smem_w = base + C::OFF_W, smem_t = base + C::OFF_T;
    const uint32_t aux = base + C::OFF_AUX;
    const uint32_t bar_x_full = aux, bar_x_empty = aux + 8;
    const uint32_t bar_w_full = aux + 16, bar_w_empty = aux + 48;        // WS each
    // two of each from here on, index = ACC buffer / group
    const uint32_t bar_acc_full = aux + 80, bar_acc_empty = aux + 96;
    const uint32_t bar_t_full = aux + 112;                               // Q / K / V tiles of the group's head are staged
    const uint32_t bar_qk_empty = aux + 128;                             // [group]: the head's last S has read Q and K
    const uint32_t bar_v_empty = aux + 272;                              // [group][head & 1]: the head's last P.V has read V
    const uint32_t bar_s_full = aux + 144, bar_s_empty = aux + 160, bar_p_full = aux + 176;
    const uint32_t bar_o_full = aux + 192;                               // [group][unit]: P.V of that unit has completed
    const uint32_t bar_o_empty = aux + 224;                              // [group]: the output warps have read the head's O
    const uint32_t bar_c_full = aux + 240;                               // [group]: the head's corr / l are in shared memory
    const uint32_t bar_sb_full = aux + 304, bar_sb_empty = aux + 312;    // the tile's score-bias tiles are written / have been read
    const uint32_t tmem_slot = aux + 256;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + C::OFF_AUX + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_trigger();

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        mbar_init(bar_x_full, 1);
        mbar_init(bar_x_empty, 1);
        for (int s = 0; s < WS; ++s) { mbar_init(bar_w_full + 8 * s, 1); mbar_init(bar_w_empty + 8 * s, 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_acc_full + 8 * s, 1);
            mbar_init(bar_acc_empty + 8 * s, C::PASS_HEADS * QA_GROUP);
            mbar_init(bar_t_full + 8 * s, QA_GROUP);
            mbar_init(bar_qk_empty + 8 * s, 1);
            mbar_init(bar_s_full + 8 * s, 1);
            mbar_init(bar_s_empty + 8 * s, QA_GROUP);
            mbar_init(bar_p_full + 8 * s, QA_GROUP);
            mbar_init(bar_o_empty + 8 * s, QA_GROUP);
            mbar_init(bar_c_full + 8 * s, QA_GROUP);
        }
        for (int s = 0; s < 4; ++s) { mbar_init(bar_o_full + 8 * s, 1); mbar_init(bar_v_empty + 8 * s, 1); }
        mbar_init(bar_sb_full, QA_GROUP);
        mbar_init(bar_sb_empty, 2);  // one commit per group issuer
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    const uint32_t smem_bias = base + C::OFF_BIAS;
    for (int i = threadIdx.x; i < 3 * H; i += QA_THREADS)  // (weights: written at create time, not by the predecessor)
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_bias + i * 4), "f"(__ldg(p.bias + i)) : "memory");
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();  // the tile list, cu and x of the predecessor kernels from here on

    // The roles split the register file: 640 threads start with 96 registers each; the control warps and the output
    // warps hand registers to the two softmax groups (a whole warpgroup executes the same setmaxnreg).
    // (each setmaxnreg sits at the head of its role's branch below: ptxas budgets the code that follows it)

    // Work items are TILES: runs of whole sentences that together have at most 128 tokens (built on the device by
    // pack_tiles_kernel; a sentence of 65..128 tokens is a tile of its own, short ones share one). Every role walks the
    // same tile sequence b = blockIdx.x, + gridDim.x, ...; empty tiles are skipped by all. The bounds of the NEXT tile are
    // fetched while the current one is processed.
    const int n_items = min(__ldg(p.n_tiles), p.batch);
    struct Walk { int b, first, last, s0, s1; };
    auto fetch = [&](Walk& w) {
        if (w.b < n_items) {
            w.first = __ldg(p.tile_first + w.b);
            w.last = __ldg(p.tile_first + w.b + 1);
            w.s0 = __ldg(p.cu + w.first);
            w.s1 = __ldg(p.cu + w.last);
        }
    };
    auto walk_init = [&]() {
        Walk w{static_cast<int>(blockIdx.x), 0, 0, 0, 0};
        fetch(w);
        return w;
    };
    // current tile -> (first token row, rows); advances w
    auto take2 = [&](Walk& w, int& start, int& len, int& first, int& last) {
        start = w.s0;
        len = min(w.s1 - w.s0, BLK);
        first = w.first;
        last = w.last;
        w.b += gridDim.x;
        fetch(w);
        return len > 0;
    };
    auto take = [&](Walk& w, int& start, int& len) {
        int f, l;
        return take2(w, start, len, f, l);
    };

    if (warp < QA_GROUP_WARP0) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(QA_REGS_CONTROL));
    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        uint32_t it = 0, wb = 0;  // sentences taken, W blocks loaded (ring slot wb % WS)
        Walk w = walk_init();
        while (w.b < n_items) {
            int start, len;
            if (!take(w, start, len)) continue;
            mbar_wait(bar_x_empty, (it & 1) ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(bar_x_full, KB * X_BLK);
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) tma_load_2d(smem_x + kb * X_BLK, &tmap_x, kb * BK, start, bar_x_full);
            }
            __syncwarp();
            ++it;
            for (int h = 0; h < HEADS; h += C::PASS_HEADS) {
#pragma unroll
                for (int kb = 0; kb < KB; ++kb, ++wb) {
                    const uint32_t s = kb % WS;
                    mbar_wait(bar_w_empty + 8 * s, ((wb / WS) & 1) ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(bar_w_full + 8 * s, W_BLK);
#pragma unroll
                        for (int hh = 0; hh < C::PASS_HEADS; ++hh)
#pragma unroll
                            for (int j = 0; j < 3; ++j)  // q, k, v rows of this head
                                tma_load_2d(smem_w + s * W_BLK + (hh * 3 + j) * W_SUB, &tmap_w, kb * BK, j * H + (h + hh) * D,
                                            bar_w_full + 8 * s);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer 1 of 3: the projections ------------------------------
        // ACC[g & 1] = x . Wqkv_h^T (128 x 96 x H) for head after head, as far ahead as the two accumulators allow. The
        // three issuers (this warp and one per group) block on their own dependencies only: with a single issuing warp
        // walking a fixed schedule the two groups ran in lockstep and each waited for the other's softmax.
        constexpr uint32_t idesc_g = umma_idesc_op(BLK, C::PASS_HEADS * NQ);
        const uint64_t dx0 = umma_desc_sw128(smem_x), dw0 = umma_desc_sw128(smem_w);
        Walk wp = walk_init();
        uint32_t pit = 0, gg = 0, wb = 0;  // tiles entered, heads projected, W blocks consumed
        TL_DECL;
        while (wp.b < n_items) {
            int a, b;
            if (!take(wp, a, b)) continue;
            TL(0);
            mbar_wait(bar_x_full, pit & 1);
            TL(1);
            ++pit;
            for (int ph = 0; ph < HEADS; ph += C::PASS_HEADS, ++gg) {
                // one head per pass: accumulators alternate; two heads per pass: one accumulator of 2 * NQ columns
                const uint32_t ab = QA_JOINT ? 0u : (gg & 1);
                mbar_wait(bar_acc_empty + 8 * ab, ((QA_JOINT ? gg : gg >> 1) & 1) ^ 1);  // the convert warps have drained it
                TL(2);
                tc_fence_after_sync();
#pragma unroll
                for (int kb = 0; kb < KB; ++kb, ++wb) {
                    const uint32_t sl = kb % WS;
                    TL(0);
                    mbar_wait(bar_w_full + 8 * sl, (wb / WS) & 1);
                    TL(3);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t da = dx0 + ((kb * X_BLK) >> 4), db = dw0 + ((sl * W_BLK) >> 4);
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            umma_bf16(tmem_base + C::ACC_COL + ab * C::ACC_STRIDE, da + 2 * k, db + 2 * k, idesc_g,
                                      (kb > 0 || k > 0) ? 1u : 0u);
                        umma_commit(bar_w_empty + 8 * sl);
                        if (kb == KB - 1) {
                            umma_commit(bar_acc_full + 8 * ab);
                            if (ph == HEADS - C::PASS_HEADS) umma_commit(bar_x_empty);  // the x tile has been read for the last time
                        }
                    }
                    __syncwarp();
                }
            }
        }
        TL(0);
        TL_PRINT("proj: issue x_full acc_empty w_full");
    } else {
        // ------------------------------ MMA issuers 2, 3 of 3: one per group ------------------------------
        // Per head of the group: S = Q . K_u^T (128 x 64 x d) and O_u = P_u . V_u (128 x d x 64) per unit u of 64 keys,
        // in the order S(u0), S(u1), P.V(u0), P.V(u1): S(u1) only needs the softmax warps to have read S(u0).
        const int G = warp - 2;
        constexpr uint32_t idesc_s = umma_idesc_op(BLK, UNIT);
        constexpr uint32_t idesc_o = umma_idesc_op(BLK, D) | (1u << 16);  // B (= V) is MN-major
        const uint64_t dq0 = umma_desc_k_sw64(smem_t + G * C::G_TILES * T_TILE);
        const uint64_t dv0 = umma_desc_mn<D>(smem_t + G * C::G_TILES * T_TILE + 2 * T_TILE);
        const uint64_t dbq0 = umma_desc_k_sw64(base + C::OFF_BIASQ), dbk0 = umma_desc_k_sw64(base + C::OFF_BIASK);
        uint32_t hc = 0;          // heads of this group completed
        uint32_t us = 0, uo = 0;  // S / P.V units issued
        uint32_t nb = 0;          // packed tiles (with score-bias tiles) seen
        TL_DECL;
        // S of unit u. `packed`: the tile holds several sentences and the convert warps have written its score-bias tiles:
        // two more k-steps add 0 to the scores of a row's own sentence and -30000 to all others (keys past the tile's
        // tokens included), so the softmax warps need no masking at all.
        auto issue_s = [&](int u, bool first, bool last, bool packed, bool sb_wait, bool sb_release) {
            TL(0);
            if (first) mbar_wait(bar_t_full + 8 * G, hc & 1);       // this head's Q / K / V are staged
            if (sb_wait) mbar_wait(bar_sb_full, nb & 1);            // this tile's score-bias tiles are written
            TL(1);
            mbar_wait(bar_s_empty + 8 * G, (us & 1) ^ 1);           // the group has read the previous S
            TL(2);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint64_t dk = dq0 + ((T_TILE + u * (UNIT * D * 2)) >> 4);
#pragma unroll
                for (int k = 0; k < D / 16; ++k)
                    umma_bf16(tmem_base + C::S_COL + G * UNIT, dq0 + 2 * k, dk + 2 * k, idesc_s, k > 0 ? 1u : 0u);
                if (packed) {
                    const uint64_t dbk = dbk0 + ((u * (UNIT * 64)) >> 4);
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        umma_bf16(tmem_base + C::S_COL + G * UNIT, dbq0 + 2 * k, dbk + 2 * k, idesc_s, 1u);
                }
                umma_commit(bar_s_full + 8 * G);
                if (last) umma_commit(bar_qk_empty + 8 * G);  // Q and K may be restaged for the next head
                if (sb_release) umma_commit(bar_sb_empty);    // this group's last S of the tile
            }
            __syncwarp();
            ++us;
        };
        auto issue_pv = [&](int u, bool last) {
            TL(0);
            mbar_wait(bar_p_full + 8 * G, uo & 1);                          // P_u is in tensor memory
            TL(3);
            if (u == 0) mbar_wait(bar_o_empty + 8 * G, (hc & 1) ^ 1);       // the previous head's O has been read
            TL(4);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t tp = tmem_base + C::P_COL + G * (UNIT / 2);
                const uint64_t dv = dv0 + (((hc % QA_VBUFS) * T_TILE + u * (UNIT * D * 2)) >> 4);
#pragma unroll
                for (int k = 0; k < UNIT / 16; ++k)  // P: 8 columns per 16 keys; V: 16 rows of D*2 B
                    umma_ts(tmem_base + C::O_COL + G * 64 + u * D, tp + 8 * k, dv + ((k * 16 * D * 2) >> 4), idesc_o,
                            k > 0 ? 1u : 0u);
                umma_commit(bar_o_full + 8 * (2 * G + u));
                if (last) umma_commit(bar_v_empty + 8 * (2 * G + (hc % QA_VBUFS)));  // this head's V buffer is consumed
            }
            __syncwarp();
            ++uo;
        };
        Walk w = walk_init();
        while (w.b < n_items) {
            int start, len, first, last;
            if (!take2(w, start, len, first, last)) continue;
            const int nunits = (len + UNIT - 1) / UNIT;
            const bool packed = last - first > 1 && last - first <= C::MAX_PACKED;
            for (int h = G; h < HEADS; h += 2, ++hc) {
                const bool tfirst = h == G, tlast = h + 2 >= HEADS;
                issue_s(0, true, nunits == 1, packed, packed && tfirst, packed && tlast && nunits == 1);
                if (nunits > 1) {
                    issue_s(1, false, true, packed, false, packed && tlast);
                    issue_pv(0, false);
                }
                issue_pv(nunits - 1, true);
            }
            if (packed) ++nb;
        }
        TL(0);
        TL_PRINT("grp issuer: issue t_full s_empty p_full o_empty");
    }
    } else if (warp < QA_CONVERT_WARP0) {
        // ------------------------------ softmax groups: alternate heads ------------------------------
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(QA_REGS_SOFTMAX));
        const int G = (warp - QA_GROUP_WARP0) >> 2;
        const int quad = warp & 3;
        const int row = quad * 32 + lane;  // token row inside the item = TMEM lane
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const uint32_t my_rows = base + C::OFF_ROWS + G * (2 * BLK * 4) + row * 4;  // corr; l at + BLK * 4
        uint32_t hc = 0;   // heads done by this group
        uint32_t h2 = 0;   // ... of which had two units
        uint32_t ug = 0;   // units done by this group
        uint32_t prev_bar = 0, prev_par = 0;  // the barrier that says the previous unit's P.V has read P (0: no previous unit)
        TL_DECL;
        Walk w = walk_init();
        while (w.b < n_items) {
            int start, len, first, last;
            if (!take2(w, start, len, first, last)) continue;
            const int nunits = (len + UNIT - 1) / UNIT;
            // keys [klo, khi) of the tile are this row's own sentence (rows past the tile's tokens: none)
            // (tiles of up to 31 sentences are masked by the S MMA itself - see the convert warps - and need none of this)
            const bool packed = last - first > 1 && last - first <= C::MAX_PACKED;
            int klo = 0, khi = row < len ? len : 0;
            if (!packed && last - first > 1 && row < len) {  // more sentences than that share the tile: find mine
                int a = first, b = last;          // cu[a] - start <= row < cu[b] - start
                while (b - a > 1) {
                    const int mid = (a + b) >> 1;
                    if (__ldg(p.cu + mid) - start <= row) a = mid; else b = mid;
                }
                klo = __ldg(p.cu + a) - start;
                khi = __ldg(p.cu + b) - start;
            }
            // the row's keys as bit masks, one 64-bit mask per unit, made once per tile (not per head): bit j of unit u is
            // set when key u*64 + j belongs to the row's sentence
            uint32_t km[2][2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int jlo = min(max(klo - u * UNIT, 0), UNIT), jhi = min(max(khi - u * UNIT, 0), UNIT);
                const uint64_t m = (jhi >= 64 ? ~0ull : ((1ull << jhi) - 1)) & ~(jlo >= 64 ? ~0ull : ((1ull << jlo) - 1));
                km[u][0] = static_cast<uint32_t>(m);
                km[u][1] = static_cast<uint32_t>(m >> 32);
            }
            for (int h = G; h < HEADS; h += 2, ++hc) {
                float m = -INFINITY, l = 0.f;
                for (int u = 0; u < nunits; ++u, ++ug) {
                    TL(11);
                    mbar_wait(bar_s_full + 8 * G, ug & 1);
                    TL(4);
                    tc_fence_after_sync();
                    uint32_t v[UNIT];
                    tmem_ld_32x32(t_lane + C::S_COL + G * UNIT, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                    tmem_ld_32x32(t_lane + C::S_COL + G * UNIT + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                    tmem_ld_wait();
                    tc_fence_before_sync();
                    mbar_arrive(bar_s_empty + 8 * G);  // S is in registers: the MMA warp may overwrite the buffer
                    TL(5);
                    const uint32_t k0 = u == 0 ? km[0][0] : km[1][0], k1 = u == 0 ? km[0][1] : km[1][1];
                    // (a mask test is one instruction per key; comparing j with the range bounds took two: 0.412 -> 0.362 ms per
                    // launch on corpus-length tiles)
                    if (!packed && (k0 & k1) != 0xffffffffu) {  // keys of other sentences / past the end -> -inf (exp2 then gives exactly 0)
#pragma unroll
                        for (int j = 0; j < UNIT; ++j)
                            if (((j < 32 ? k0 : k1) & (1u << (j & 31))) == 0) v[j] = 0xff800000u;
                    }
                    const float mx = row_max(v);
                    const float m_new = fmaxf(m, mx * p.scale_log2);
                    // a row with no key yet (its sentence lies in the other unit, or the row is past the tile's tokens)
                    // keeps m = -inf; 0 stands in for it below so that exp2 sees -inf, not inf - inf
                    const float m_use = m_new == -INFINITY ? 0.f : m_new;
                    const float corr = ex2(m - m_use);  // unit 0: exp2(-inf) = 0
                    m = m_new;
                    TL(6);
                    const uint64_t sc2 = pk2(p.scale_log2, p.scale_log2), nm2 = pk2(-m_use, -m_use);
                    uint64_t lb2[4] = {0, 0, 0, 0};
                    uint32_t pw[UNIT / 2];  // this row's probabilities, two keys per word
#pragma unroll
                    for (int q = 0; q < UNIT / 8; ++q) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float a, bb;
                            upk2(fma2(pk2(__uint_as_float(v[q * 8 + 2 * j]), __uint_as_float(v[q * 8 + 2 * j + 1])), sc2, nm2), a, bb);
                            const float pa = ex2(a), pb = ex2(bb);
                            lb2[j] = add2(lb2[j], pk2(pa, pb));
                            pw[q * 4 + j] = pack_op2(pa, pb);
                        }
                    }
                    TL(7);
                    if (prev_bar != 0) mbar_wait(prev_bar, prev_par);  // the previous unit's P.V has read the P buffer
                    TL(8);
                    tc_fence_after_sync();
                    tmem_st_32x32(t_lane + C::P_COL + G * (UNIT / 2), pw);
                    float lb;
                    {
                        float a0, a1, a2, a3;
                        upk2(add2(lb2[0], lb2[1]), a0, a1);
                        upk2(add2(lb2[2], lb2[3]), a2, a3);
                        lb = (a0 + a1) + (a2 + a3);
                    }
                    l = fmaf(l, corr, lb);
                    tmem_st_wait();
                    tc_fence_before_sync();
                    mbar_arrive(bar_p_full + 8 * G);
                    prev_bar = bar_o_full + 8 * (2 * G + u);
                    prev_par = (u == 0 ? hc : h2) & 1;
                    TL(10);
                    if (u == nunits - 1) {
                        // hand the head to the output warps: out = (O_0 * corr + O_1) / l (one unit: O_0 / l)
                        mbar_wait(bar_o_empty + 8 * G, (hc & 1) ^ 1);  // they have read the previous head's pair
                        asm volatile("st.shared.f32 [%0], %1;" ::"r"(my_rows), "f"(corr) : "memory");
                        asm volatile("st.shared.f32 [%0], %1;" ::"r"(my_rows + BLK * 4), "f"(l) : "memory");
                        mbar_arrive(bar_c_full + 8 * G);
                        TL(9);
                    }
                }
                if (nunits > 1) ++h2;
            }
        }
        TL(11);
        TL_PRINT("softmax: - - - - s_full ldS max exp o_full(P free) handoff stP other");
    } else if (warp < QA_OUTPUT_WARP0) {
        // ------------------------------ convert warps: ACC + bias -> Q / K / V tiles, head after head ------------------------------
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(QA_REGS_CONVERT));
        const int quad = warp & 3;
        const int row = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        uint32_t gh = 0;  // heads converted; head gh belongs to group gh & 1 and is that group's head number gh >> 1
        uint32_t nb = 0;  // packed tiles seen
        TL_DECL;
        Walk w = walk_init();
        while (w.b < n_items) {
            int start, len, first, last;
            if (!take2(w, start, len, first, last)) continue;
            if (last - first > 1 && last - first <= C::MAX_PACKED) {
                // Several sentences share the tile: attention is block-diagonal. Instead of masking 64 scores per row and
                // unit in registers, the mask rides on the tensor pipe: 32 extra columns of the S contraction hold, on the Q
                // side, a one-hot row "my sentence is number c of this tile" and, on the K side, 0 in the column of the key's
                // own sentence and -30000 everywhere else (keys past the tile's tokens: -30000 in every column). Their
                // product adds 0 to the scores inside a sentence and -30000 (x log2 e / sqrt d, far below any score) to all
                // others, whose exponentials are then exactly 0. Written once per tile, read by the S MMAs of all heads.
                int sid = C::MAX_PACKED;  // rows past the tile's tokens: a sentence nobody has
                if (row < len) {
                    int a = first, b = last;  // cu[a] - start <= row < cu[b] - start
                    while (b - a > 1) {
                        const int mid = (a + b) >> 1;
                        if (__ldg(p.cu + mid) - start <= row) a = mid; else b = mid;
                    }
                    sid = a - first;
                }
                uint32_t wq[16], wk[16];
                const uint32_t neg2 = pack_op2(-30000.0f, -30000.0f);
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const bool mine = c == (sid >> 1);
                    wq[c] = mine ? ((sid & 1) ? pack_op2(0.0f, 1.0f) : pack_op2(1.0f, 0.0f)) : 0u;
                    wk[c] = (mine && row < len) ? ((sid & 1) ? pack_op2(-30000.0f, 0.0f) : pack_op2(0.0f, -30000.0f)) : neg2;
                }
                mbar_wait(bar_sb_empty, (nb & 1) ^ 1);  // the previous packed tile's last S MMAs (both groups) have read them
                stage_words_bf16(base + C::OFF_BIASQ, row, wq);
                stage_words_bf16(base + C::OFF_BIASK, row, wk);
                fence_proxy_async_smem();
                mbar_arrive(bar_sb_full);
                ++nb;
            }
            for (int h = 0; h < HEADS; ++h, ++gh) {
                const uint32_t G = gh & 1, hcg = gh >> 1;
                const uint32_t tile = smem_t + G * C::G_TILES * T_TILE;
                TL(11);
                const uint32_t ab = QA_JOINT ? 0u : G;
                mbar_wait(bar_acc_full + 8 * ab, hcg & 1);  // (two heads per pass: pass number = the group's head number)
                TL(0);
                tc_fence_after_sync();
                uint32_t words[3][16];
#pragma unroll
                for (int j = 0; j < 3; ++j) {  // q, k, v
                    uint32_t v[32];
                    tmem_ld_32x32(t_lane + C::ACC_COL + G * C::ACC_STRIDE + j * D, v);
                    const uint32_t bs = smem_bias + (j * H + h * D) * 4;
                    tmem_ld_wait();
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const uint4 b = lds_v4(bs + q * 16);
                        words[j][q * 2 + 0] = epi_pair_bf16<EPI_BIAS>(v[q * 4 + 0], v[q * 4 + 1], __uint_as_float(b.x), __uint_as_float(b.y));
                        words[j][q * 2 + 1] = epi_pair_bf16<EPI_BIAS>(v[q * 4 + 2], v[q * 4 + 3], __uint_as_float(b.z), __uint_as_float(b.w));
                    }
                }
                tc_fence_before_sync();
                mbar_arrive(bar_acc_empty + 8 * ab);  // the accumulator is in registers
                TL(1);
                // Q and K: the previous head's last S has read them; V alternates between two tiles, this one was read by the
                // head before the previous one - so the tiles are staged while the previous head's softmax is still running
                mbar_wait(bar_qk_empty + 8 * G, (hcg & 1) ^ 1);
                mbar_wait(bar_v_empty + 8 * (2 * G + (hcg % QA_VBUFS)), ((hcg / QA_VBUFS) & 1) ^ 1);
                TL(2);
                stage_words_bf16(tile, row, words[0]);
                stage_words_bf16(tile + T_TILE, row, words[1]);
                stage_words_bf16(tile + (2 + (hcg % QA_VBUFS)) * T_TILE, row, words[2]);
                fence_proxy_async_smem();
                mbar_arrive(bar_t_full + 8 * G);
                TL(3);
            }
        }
        TL(11);
        TL_PRINT("convert: acc_full cvt qkv_empty stage");
    } else {
        // ------------------------------ output warps: ctx rows of finished heads ------------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(QA_REGS_OUTPUT));
        const int quad = warp & 3;
        const int row = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        uint32_t gh = 0;   // heads written
        uint32_t h2 = 0;   // two-unit heads per group before this tile
        TL_DECL;
        Walk w = walk_init();
        while (w.b < n_items) {
            int start, len;
            if (!take(w, start, len)) continue;
            const int nunits = (len + UNIT - 1) / UNIT;
            for (int h = 0; h < HEADS; ++h, ++gh) {
                const uint32_t G = gh & 1, hcg = gh >> 1;
                const uint32_t rows_g = base + C::OFF_ROWS + G * (2 * BLK * 4) + row * 4;
                TL(11);
                mbar_wait(bar_c_full + 8 * G, hcg & 1);
                TL(0);
                float corr, l;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(corr) : "r"(rows_g) : "memory");
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(l) : "r"(rows_g + BLK * 4) : "memory");
                // the last unit's P.V is complete, and with it the first's (one issuing thread, in order)
                mbar_wait(bar_o_full + 8 * (2 * G + (nunits - 1)), (nunits == 1 ? hcg : h2 + (h >> 1)) & 1);
                TL(1);
                tc_fence_after_sync();
                const float inv = 1.0f / l;
                const float c0 = nunits > 1 ? corr * inv : inv;
                __nv_bfloat16* dst = p.ctx + static_cast<size_t>(start + row) * p.H + h * D;
                const uint32_t to = t_lane + C::O_COL + G * 64;
#pragma unroll
                for (int q = 0; q < D / 16; ++q) {
                    uint32_t a[16], b[16];
                    tmem_ld_32x16(to + q * 16, a);
                    if (nunits > 1) tmem_ld_32x16(to + D + q * 16, b);
                    tmem_ld_wait();
                    uint32_t wd[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float o0 = __uint_as_float(a[2 * j]) * c0, o1 = __uint_as_float(a[2 * j + 1]) * c0;
                        if (nunits > 1) {
                            o0 = fmaf(__uint_as_float(b[2 * j]), inv, o0);
                            o1 = fmaf(__uint_as_float(b[2 * j + 1]), inv, o1);
                        }
                        wd[j] = pack_op2(o0, o1);
                    }
                    if (row < len) stg_v8(dst + q * 16, wd);
                }
                tc_fence_before_sync();
                mbar_arrive(bar_o_empty + 8 * G);
                TL(2);
            }
            if (nunits > 1) h2 += HEADS / 2;
        }
        TL(11);
        TL_PRINT("output: c_full o_full write");
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
}

template <int H, int D>
cudaError_t launch_qa(const __nv_bfloat16* x, const __nv_bfloat16* wqkv, const float* bqkv, const int* cu,
                      const int* tile_first, const int* n_tiles, int rows_cap, int batch, __nv_bfloat16* ctx, int num_sms,
                      cudaStream_t s) {
    using C = QaCfg<H, D>;
    alignas(64) CUtensorMap tx, tw;
    if (!make_tmap(&tx, x, rows_cap, H, BK, BLK, false, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
    if (!make_tmap(&tw, wqkv, 3 * H, H, BK, D, false, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
    auto kern = qkv_attention_kernel<H, D>;
    static PerDeviceOnce attr_set;  // function attributes belong to a (function, device) pair
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::BYTES);
        if (e != cudaSuccess) return e;
        attr_set.mark();
    }
    QaDev p;
    p.cu = cu; p.tile_first = tile_first; p.n_tiles = n_tiles; p.bias = bqkv; p.batch = batch; p.H = H; p.ctx = ctx;
    p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    const int grid = batch < num_sms ? batch : num_sms;
    if (grid <= 0) return cudaSuccess;
    return launch_pdl(kern, dim3(grid), dim3(QA_THREADS), C::BYTES, s, tx, tw, p);
}

}  // namespace

bool qkv_attention_supported(int hidden, int heads, int max_seqlen) {
    return (hidden == 384 || hidden == 256 || hidden == 128) && heads * 32 == hidden && max_seqlen <= BLK;
}

cudaError_t launch_qkv_attention(const __nv_bfloat16* x, const __nv_bfloat16* wqkv, const float* bqkv, const int* cu,
                                 const int* tile_first, const int* n_tiles, int rows_cap, int batch, int max_seqlen,
                                 int hidden, int heads, __nv_bfloat16* ctx, int num_sms, cudaStream_t s) {
    if (batch <= 0 || max_seqlen <= 0) return cudaSuccess;
    if (!qkv_attention_supported(hidden, heads, max_seqlen)) return cudaErrorInvalidValue;
    if (hidden == 384) return launch_qa<384, 32>(x, wqkv, bqkv, cu, tile_first, n_tiles, rows_cap, batch, ctx, num_sms, s);
    if (hidden == 256) return launch_qa<256, 32>(x, wqkv, bqkv, cu, tile_first, n_tiles, rows_cap, batch, ctx, num_sms, s);
    return launch_qa<128, 32>(x, wqkv, bqkv, cu, tile_first, n_tiles, rows_cap, batch, ctx, num_sms, s);
}

}  // operand-format namespace
}  // namespace b200
