// K3 (tcgen05): varlen self-attention on the 5th-generation tensor cores.
//
// Replaces candle bert.rs BertSelfAttention::forward after the Q/K/V projections (scores = q.k^T/sqrt(d)
// + (1-mask)*f32::MIN, softmax, .v, head merge; called from embedding_generator.rs:198)
// (the mma.sync variant of round 1 is gone); this kernel is the one the forward uses when K23 does not apply.
//
// Work item = (sentence, head, block of 128 query rows). A persistent CTA (3 per SM for d = 32, 2 for d = 64) walks items; per
// item it walks the sentence's keys in UNITS of 64 keys, two units in flight:
//   warp 0     TMA producer: Q tile once per item, K and V tiles per 128-key block (2-stage rings), straight from
//              the packed qkv [tokens, 3H] activations (box D x 128 rows; D*2-byte swizzle)
//   warp 1     one thread issues tcgen05.mma: S_u = Q.K_u^T (128 x 64, fp32 in TMEM) one unit ahead of the softmax, then
//              O += P_u.V_u (128 x D, one accumulator per item) as soon as P_u is written
//   warp 2     TMEM allocator (S 64 + O 32 + P 32 = 128 columns at d = 32, 2 x 64 S + 64 O + 2 x 32 P = 256 at d = 64)
//   warps 4-7  softmax: thread r owns query row r (= TMEM lane r). One tcgen05.ld brings the unit's 64 scores
//              into registers (read once: row max, p = exp2(s*scale - max) and the row sum all work from there),
//              P goes as 16-bit pairs into tensor memory (the A operand of the second MMA is read there). The units of an
//              item accumulate in one tensor-memory accumulator: a unit's exponentials use the maximum the row had
//              before it unless the new one is more than 2^8 above (then the accumulator is rescaled in place), so
//              there is no per-unit read-back of O; running max / sum live in registers (fp32), the item's rows are
//              written out during the next item's first unit.
// Keys past the sentence end get p = 0, which is exactly what the reference's additive f32::MIN mask yields.
// Rows of the 128-row tiles that lie past the sentence end belong to the next sentence (or are zero-filled
// by TMA past the last token); they are computed and never stored.
#include "kernels.h"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"
#include "tc_helpers.cuh"

namespace b200 {
B200_NS_THIS {

namespace {

constexpr int BLK = 128;   // query rows per item; keys per K/V tile
constexpr int UNIT = 64;   // keys per softmax unit
constexpr int ATT_THREADS = 256;
constexpr int SM_WARP0 = 4;       // softmax warps 4-7; the single-thread roles keep the low warp ids: putting them
constexpr int PRODUCER_WARP = 0;  // after the softmax warps (ids 4-6) was measured 28 % slower
constexpr int MMA_WARP = 1;
constexpr int ALLOC_WARP = 2;
constexpr int SM_THREADS = 128;

struct AttnDev {
    const int* cu;
    int batch, heads, H, q_blocks;
    float scale_log2;
    __nv_bfloat16* ctx;
};

// CTAs per SM at head_dim 64 (2: one O accumulator per CTA and registers moved to the softmax warps by setmaxnreg)
#ifndef K3_CTAS_D64
#define K3_CTAS_D64 2
#endif
// CTAs per SM at head_dim 32 (3: one S, one P and one O buffer per CTA - 128 tensor-memory columns - and the register split)
#ifndef K3_CTAS_D32
#define K3_CTAS_D32 3
#endif

template <int D>
struct AttnSmem {
    static constexpr int TILE = BLK * D * 2;         // one Q / K / V tile
    static constexpr int OFF_Q = 0;                   // 2 stages
    static constexpr int OFF_KV = 2 * TILE;           // 2 stages x (K, V)
    static constexpr int OFF_AUX = OFF_KV + 4 * TILE;
    static constexpr int BYTES = OFF_AUX + 256;
};

template <int D>
__global__ void __launch_bounds__(ATT_THREADS, D == 32 ? K3_CTAS_D32 : K3_CTAS_D64)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const AttnDev p) {
    using L = AttnSmem<D>;
    constexpr int TILE = L::TILE;
    // tensor memory: 2 x 64 S | 2 x D O | 2 x 32 P. The probabilities are written there as 16-bit pairs (one 32-bit column per
    // two keys) and the P.V MMA reads its A operand from tensor memory: P never goes through shared memory, whose operand
    // fetch (128 B per clock) paces these small MMAs.
    // d = 64 with two CTAs per SM: one O accumulator (the next item's first P.V waits for this item's rows to be written,
    // which happens right after that item's first P is in place), so that each CTA fits 256 of the SM's 512 columns
    constexpr int OB = ((D == 64 && K3_CTAS_D64 == 2) || (D == 32 && K3_CTAS_D32 == 3)) ? 1 : 2;
    // d = 32 with three CTAs per SM: S and P single-buffered as well (S of the next unit is issued once the softmax warps
    // have the current one in registers, P is rewritten once the previous unit's P.V has completed - both long before they
    // are needed), 64 + 32 + 32 = 128 columns per CTA
    constexpr int NB = (D == 32 && K3_CTAS_D32 == 3) ? 1 : 2;
    constexpr int O_COL = NB * UNIT, P_COL = O_COL + OB * D;
    constexpr uint32_t TMEM_COLS = P_COL + NB * (UNIT / 2) <= 128 ? 128 : (P_COL + NB * (UNIT / 2) <= 256 ? 256 : 512);

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = smem_u32(smem_raw);
    if ((base & 1023u) != 0) __trap();
    const uint32_t smem_q = base + L::OFF_Q, smem_kv = base + L::OFF_KV;
    const uint32_t aux = base + L::OFF_AUX;
    const uint32_t bar_q_full = aux, bar_q_empty = aux + 16, bar_kv_full = aux + 32, bar_kv_empty = aux + 48;
    const uint32_t bar_s_full = aux + 64, bar_s_empty = aux + 80, bar_p_full = aux + 96, bar_o_full = aux + 112,
                   bar_o_empty = aux + 128;  // two of each: buffer u & 1
    const uint32_t tmem_slot = aux + 160;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + L::OFF_AUX + 160);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int items = p.batch * p.heads * p.q_blocks;

    pdl_trigger();
    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_qkv);
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_q_full + 8 * s, 1);
            mbar_init(bar_q_empty + 8 * s, 1);
            mbar_init(bar_kv_full + 8 * s, 1);
            mbar_init(bar_kv_empty + 8 * s, 1);
            mbar_init(bar_s_full + 8 * s, 1);
            mbar_init(bar_s_empty + 8 * s, SM_THREADS);
            mbar_init(bar_p_full + 8 * s, SM_THREADS);
            mbar_init(bar_o_full + 8 * s, 1);
            mbar_init(bar_o_empty + 8 * s, SM_THREADS);
        }
        fence_barrier_init();
    }
    if (warp == ALLOC_WARP) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();  // nothing above reads or writes global memory

    // item i -> (sentence b, query block qb, head h), i = (b * q_blocks + qb) * heads + h. Items advance by gridDim.x,
    // so (b, r) is carried along instead of dividing every time; items whose query block lies past the sentence
    // end are empty and skipped by every role alike.
    const int per_b = p.heads * p.q_blocks;
    const int step_b = static_cast<int>(gridDim.x) / per_b, step_r = static_cast<int>(gridDim.x) % per_b;
    // Every role walks the same item sequence. The sentence bounds (two loads from cu_seqlens) of the NEXT item are
    // fetched while the current one is processed, so their latency is never on an item's critical path.
    struct Walk {
        int i, b, r;   // item index, sentence, (query block, head) index inside the sentence
        int s0, s1;    // cu[b], cu[b + 1] of this item (prefetched)
    };
    auto fetch = [&](Walk& w) {
        if (w.i < items) { w.s0 = __ldg(p.cu + w.b); w.s1 = __ldg(p.cu + w.b + 1); }
    };
    auto walk_init = [&]() {
        Walk w{static_cast<int>(blockIdx.x), static_cast<int>(blockIdx.x) / per_b, static_cast<int>(blockIdx.x) % per_b, 0, 0};
        fetch(w);
        return w;
    };
    // returns the current item in (b, qb, h, start, len) and advances w to the next one (issuing its loads)
    auto take = [&](Walk& w, int& b, int& qb, int& h, int& start, int& len) {
        b = w.b;
        qb = w.r / p.heads;
        h = w.r - qb * p.heads;
        start = w.s0;
        len = w.s1 - w.s0;
        w.i += gridDim.x;
        w.b += step_b;
        w.r += step_r;
        if (w.r >= per_b) { w.r -= per_b; ++w.b; }
        fetch(w);
        return qb * BLK < len;  // items whose query block lies past the sentence end are empty: skipped by every role
    };

    // The producer and MMA warps walk their loops as whole warps (uniform control flow) and elect one lane to issue:
    // under `if (lane == 0)` every single-thread instruction gets wrapped in a divergence-safe ELECT loop and the issue
    // rate of that one thread becomes the bottleneck.
    // d = 64, two CTAs per SM: they start at 128 registers per thread; the four single-thread-role warps hand theirs to the
    // softmax warps (each setmaxnreg at the head of its warpgroup's branch: ptxas budgets the code that follows it)
    constexpr bool REGSPLIT = (D == 64 && K3_CTAS_D64 == 2) || (D == 32 && K3_CTAS_D32 == 3);
    constexpr int REGS_CONTROL = D == 64 ? 40 : 32, REGS_SOFTMAX = D == 64 ? 216 : 128;  // 2 x 128 / 2 x 80 per thread pair at launch
    if (warp < SM_WARP0) {
    if constexpr (REGSPLIT) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_CONTROL));
    if (warp == PRODUCER_WARP) {
        uint32_t q_it = 0, kv_it = 0;
        Walk w = walk_init();
        while (w.i < items) {
            int b, qb, h, start, len;
            if (!take(w, b, qb, h, start, len)) continue;
            const uint32_t qs = q_it & 1;
            mbar_wait(bar_q_empty + 8 * qs, ((q_it >> 1) & 1) ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(bar_q_full + 8 * qs, TILE);
                tma_load_2d(smem_q + qs * TILE, &tmap_qkv, h * D, start + qb * BLK, bar_q_full + 8 * qs);
            }
            __syncwarp();
            ++q_it;
            const int nkb = (len + BLK - 1) / BLK;
            for (int kb = 0; kb < nkb; ++kb, ++kv_it) {
                const uint32_t s = kv_it & 1;
                mbar_wait(bar_kv_empty + 8 * s, ((kv_it >> 1) & 1) ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(bar_kv_full + 8 * s, 2 * TILE);
                    tma_load_2d(smem_kv + s * 2 * TILE, &tmap_qkv, p.H + h * D, start + kb * BLK, bar_kv_full + 8 * s);
                    tma_load_2d(smem_kv + s * 2 * TILE + TILE, &tmap_qkv, 2 * p.H + h * D, start + kb * BLK,
                                bar_kv_full + 8 * s);
                }
                __syncwarp();
            }
        }
    } else if (warp == MMA_WARP) {
        // The units of all of this CTA's items form one stream. S_k = Q.K_k^T is issued one unit ahead of O_k = P_k.V_k
        // ACROSS item boundaries, so the softmax warps never wait for the tensor pipe at the end of an item: two
        // cursors walk the same item sequence, `sc` for the S MMAs and `pc` (one unit behind) for the P.V MMAs.
        constexpr uint32_t idesc_s = umma_idesc_op(BLK, UNIT);
        constexpr uint32_t idesc_o = umma_idesc_op(BLK, D) | (1u << 16);  // B (= V) is MN-major
        const uint64_t dq0 = umma_desc_k<D>(smem_q), dk0 = umma_desc_k<D>(smem_kv);
        const uint64_t dv0 = umma_desc_mn<D>(smem_kv + TILE);
        struct Cursor {
            Walk w;
            bool valid;        // points at a unit
            int nunits, u;     // units of the current item, current unit (keys [u*64, u*64+64): block u/2, half u&1)
            uint32_t q_it, kv_it;  // ring positions of the current item's Q tile and first K/V block; q_it is also the
                                   // item's number in this CTA's stream (O buffer q_it & 1)
        };
        auto advance = [&](Cursor& c) {
            if (c.valid && c.u + 1 < c.nunits) { ++c.u; return true; }
            if (c.valid) { c.kv_it += (c.nunits + 1) >> 1; ++c.q_it; c.valid = false; }
            while (c.w.i < items) {
                int b, qb, h, start, len;
                if (!take(c.w, b, qb, h, start, len)) continue;
                c.nunits = (len + UNIT - 1) / UNIT;
                c.u = 0;
                c.valid = true;
                return true;
            }
            return false;
        };
        uint32_t us = 0;  // units whose S MMA has been issued (S buffer us & 1)
        uint32_t uo = 0;  // units whose P.V MMA has been issued (O buffer uo & 1)
        auto issue_s = [&](const Cursor& c) {
            const uint32_t qs = c.q_it & 1, kvi = c.kv_it + (c.u >> 1), s = kvi & 1;
            if (c.u == 0) mbar_wait(bar_q_full + 8 * qs, (c.q_it >> 1) & 1);       // first touch of this item's Q
            if ((c.u & 1) == 0) mbar_wait(bar_kv_full + 8 * s, (kvi >> 1) & 1);    // first touch of this K/V block
            mbar_wait(bar_s_empty + 8 * (us % NB), ((us / NB) & 1) ^ 1);           // softmax drained this S buffer
            tc_fence_after_sync();
            if (elect_one()) {
                const uint64_t dq = dq0 + ((qs * TILE) >> 4);
                const uint64_t dk = dk0 + ((s * 2 * TILE + (c.u & 1) * (UNIT * D * 2)) >> 4);
#pragma unroll
                for (int k = 0; k < D / 16; ++k)
                    umma_bf16(tmem_base + (us % NB) * UNIT, dq + 2 * k, dk + 2 * k, idesc_s, k > 0 ? 1u : 0u);
                umma_commit(bar_s_full + 8 * (us % NB));
                if (c.u == c.nunits - 1) umma_commit(bar_q_empty + 8 * qs);  // Q is only read by the S MMAs
            }
            __syncwarp();
            ++us;
        };
        auto issue_pv = [&](const Cursor& c) {
            const uint32_t kvi = c.kv_it + (c.u >> 1), s = kvi & 1, pb = uo % NB, ob = c.q_it % OB;
            mbar_wait(bar_p_full + 8 * pb, (uo / NB) & 1);         // P_u is in tensor memory (and O rescaled if it had to be)
            // the item's first unit overwrites its O buffer: the item two back has been finalised out of it
            if (c.u == 0) mbar_wait(bar_o_empty + 8 * ob, ((c.q_it / OB) & 1) ^ 1);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t tp = tmem_base + P_COL + pb * (UNIT / 2);
                const uint64_t dv = dv0 + ((s * 2 * TILE + (c.u & 1) * (UNIT * D * 2)) >> 4);
#pragma unroll
                for (int k = 0; k < UNIT / 16; ++k)  // P: 8 columns per 16 keys; V: 16 rows of D*2 B
                    umma_ts(tmem_base + O_COL + ob * D, tp + 8 * k, dv + ((k * 16 * D * 2) >> 4), idesc_o,
                            (c.u > 0 || k > 0) ? 1u : 0u);  // the units of an item accumulate in tensor memory
                umma_commit(bar_o_full + 8 * pb);  // "P.V of this unit has completed" (P buffer free, O readable)
                if ((c.u & 1) == 1 || c.u == c.nunits - 1) umma_commit(bar_kv_empty + 8 * s);  // block fully consumed
            }
            __syncwarp();
            ++uo;
        };
        Cursor sc{walk_init(), false, 0, 0, 0, 0}, pc{walk_init(), false, 0, 0, 0, 0};
        bool s_more = advance(sc);
        if (s_more) issue_s(sc);
        while (advance(pc)) {
            if (s_more) {
                s_more = advance(sc);
                if (s_more) issue_s(sc);
            }
            issue_pv(pc);
        }
    }
    } else {
        if constexpr (REGSPLIT) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_SOFTMAX));
        const int quad = warp & 3;
        const int row = quad * 32 + lane;  // query row inside the item = TMEM lane
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        uint32_t uc = 0;  // units processed so far by this CTA (S / P buffer uc & 1, phase (uc >> 1) & 1)
        uint32_t ic = 0;  // items started (O buffer ic & 1)

        // The units of an item accumulate in ONE tensor-memory accumulator (P.V with the accumulate flag), so there is no
        // running output in registers and nothing to fold in per unit. For that the exponentials of a unit use the
        // maximum the row had BEFORE it as long as the new maximum is not more than 2^8 above it (p <= 256 then: fine in
        // 16 bits, the row sum is fp32; softmax is shift-invariant, so this is exact arithmetic, not an approximation).
        // Only when some row of the warp jumps further the accumulator is rescaled in place - tcgen05.ld, multiply,
        // tcgen05.st - after the previous unit's P.V has completed.
        constexpr float LAZY = 8.0f;
        // The previous item, whose last P.V may still be running: finalised during the next item's first unit.
        struct Pending {
            bool valid;
            uint32_t ur, ob;  // stream index of its last unit, its O buffer
            float l;
            int start, len, qb, h;
        } pend{false, 0, 0, 0.f, 0, 0, 0, 0};
        auto finalise_pending = [&]() {
            mbar_wait(bar_o_full + 8 * (pend.ur % NB), (pend.ur / NB) & 1);  // the item's last P.V has completed
            tc_fence_after_sync();
            // O / l. Each lane owns one row of D 16-bit values (64 or 128 contiguous bytes of ctx): whole 32-byte sectors
            // straight from registers.
            const float inv = 1.0f / pend.l;
            const bool live = pend.qb * BLK + row < pend.len;
            __nv_bfloat16* dst = p.ctx + static_cast<size_t>(pend.start + pend.qb * BLK + row) * p.H + pend.h * D;
#pragma unroll
            for (int c = 0; c < D / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(t_lane + O_COL + pend.ob * D + c * 32, v);
                tmem_ld_wait();
                if (live) {
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        uint32_t w[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            w[j] = pack_op2(__uint_as_float(v[q * 16 + 2 * j]) * inv, __uint_as_float(v[q * 16 + 2 * j + 1]) * inv);
                        stg_v8(dst + c * 32 + q * 16, w);
                    }
                }
            }
            tc_fence_before_sync();
            mbar_arrive(bar_o_empty + 8 * pend.ob);
            pend.valid = false;
        };

        Walk w = walk_init();
        while (w.i < items) {
            int b, qb, h, start, len;
            if (!take(w, b, qb, h, start, len)) continue;
            float m = -INFINITY, l = 0.f;
            const int nunits = (len + UNIT - 1) / UNIT;
            const uint32_t ob = ic % OB;
            for (int u = 0; u < nunits; ++u, ++uc) {
                const uint32_t sb = uc % NB;
                const int nvalid = min(UNIT, len - u * UNIT);  // keys of this unit inside the sentence
                mbar_wait(bar_s_full + 8 * sb, (uc / NB) & 1);
                tc_fence_after_sync();
                uint32_t v[UNIT];
                tmem_ld_32x32(t_lane + sb * UNIT, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                tmem_ld_32x32(t_lane + sb * UNIT + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive(bar_s_empty + 8 * sb);  // S is in registers: the MMA warp may overwrite the buffer
                // Keys past the sentence end become -inf once, up front (exp2 then gives exactly 0), so that everything
                // below is straight-line code the compiler can interleave freely: 64 independent exp chains.
                if (nvalid < UNIT) {
#pragma unroll
                    for (int j = 0; j < UNIT; ++j)
                        if (j >= nvalid) v[j] = 0xff800000u;  // -inf
                }
                const float m_new = fmaxf(m, row_max(v) * p.scale_log2);
                if (u == 0) {
                    m = m_new;  // every unit has at least one key of the sentence: m is finite from here on
                } else if (__any_sync(0xffffffffu, m_new > m + LAZY)) {
                    // rare: rescale this warp's rows of the accumulator (rows that did not jump: factor 1)
                    const bool jump = m_new > m + LAZY;
                    const float corr = jump ? ex2(m - m_new) : 1.0f;
                    if (jump) m = m_new;
                    l *= corr;
                    mbar_wait(bar_o_full + 8 * ((uc - 1) % NB), ((uc - 1) / NB) & 1);  // the previous unit's P.V has completed
                    tc_fence_after_sync();
#pragma unroll
                    for (int c = 0; c < D / 32; ++c) {
                        uint32_t o[32];
                        tmem_ld_32x32(t_lane + O_COL + ob * D + c * 32, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * corr);
                        tmem_st_32x32(t_lane + O_COL + ob * D + c * 32, o);
                    }
                    tmem_st_wait();
                    tc_fence_before_sync();
                }
                // p = exp2(s*scale - m) (packed-pair FMA), rounded to 16 bits, two keys per word; row sum in four packed chains
                const uint64_t sc2 = pk2(p.scale_log2, p.scale_log2), nm2 = pk2(-m, -m);
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
                // the P buffer was last read by the P.V of the unit NB back
                mbar_wait(bar_o_full + 8 * sb, ((uc / NB) & 1) ^ 1);
                tc_fence_after_sync();
                // (single P buffer: the previous item's rows go out here, while no later P.V can complete - its barrier
                // would otherwise run a phase ahead of the wait in finalise_pending)
                if (NB == 1 && pend.valid) finalise_pending();
                tmem_st_32x32(t_lane + P_COL + sb * (UNIT / 2), pw);
                float lb;
                {
                    float a0, a1, a2, a3;
                    upk2(add2(lb2[0], lb2[1]), a0, a1);
                    upk2(add2(lb2[2], lb2[3]), a2, a3);
                    lb = (a0 + a1) + (a2 + a3);
                }
                l += lb;
                tmem_st_wait();
                tc_fence_before_sync();
                mbar_arrive(bar_p_full + 8 * sb);
                // while P_u . V_u runs: the previous item's rows go out
                if (pend.valid) finalise_pending();
            }
            pend.valid = true; pend.ur = uc - 1; pend.ob = ob; pend.l = l;
            pend.start = start; pend.len = len; pend.qb = qb; pend.h = h;
            ++ic;
        }
        if (pend.valid) finalise_pending();
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == ALLOC_WARP) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
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

template <int D>
cudaError_t launch(const __nv_bfloat16* qkv, const int* cu, int rows_cap, int batch, int max_seqlen, int heads,
                   __nv_bfloat16* ctx, int num_sms, cudaStream_t s) {
    auto fn = encode_fn();
    if (!fn) return cudaErrorInvalidValue;
    const int H = heads * D;
    alignas(64) CUtensorMap map;
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(3 * H), static_cast<cuuint64_t>(rows_cap)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(3 * H) * 2};
    cuuint32_t box[2] = {D, BLK};
    cuuint32_t estr[2] = {1, 1};
    if (fn(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(qkv), dims, strides, box, estr,
           CU_TENSOR_MAP_INTERLEAVE_NONE, D == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    auto kern = attention_tc_kernel<D>;
    constexpr int smem = AttnSmem<D>::BYTES;
    static PerDeviceOnce attr_set;  // function attributes belong to a (function, device) pair
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr_set.mark();
    }
    AttnDev p;
    p.cu = cu; p.batch = batch; p.heads = heads; p.H = H; p.q_blocks = (max_seqlen + BLK - 1) / BLK;
    p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    p.ctx = ctx;
    const long long items = static_cast<long long>(batch) * heads * p.q_blocks;
    if (items <= 0) return cudaSuccess;
    const int ctas = num_sms * (D == 32 ? K3_CTAS_D32 : K3_CTAS_D64);
    const int grid = static_cast<int>(items < ctas ? items : ctas);
    return launch_pdl(kern, dim3(grid), dim3(ATT_THREADS), smem, s, map, p);
}

}  // namespace

cudaError_t launch_attention_tc(const __nv_bfloat16* qkv, const int* cu, int rows_cap, int batch, int max_seqlen,
                                int heads, int head_dim, __nv_bfloat16* ctx, int num_sms, cudaStream_t s) {
    if (batch <= 0 || max_seqlen <= 0) return cudaSuccess;
    if (static_cast<long long>(batch) * heads * ((max_seqlen + BLK - 1) / BLK) > 0x7fffffffLL) return cudaErrorInvalidValue;
    if (head_dim == 32) return launch<32>(qkv, cu, rows_cap, batch, max_seqlen, heads, ctx, num_sms, s);
    if (head_dim == 64) return launch<64>(qkv, cu, rows_cap, batch, max_seqlen, heads, ctx, num_sms, s);
    return cudaErrorInvalidValue;
}

}  // operand-format namespace
}  // namespace b200
