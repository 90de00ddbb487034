// K3: varlen ("flash"-style) self-attention for the packed token layout.
//
// Replaces candle bert.rs BertSelfAttention::forward after the Q/K/V projections: reshape to
// [B,h,S,d], scores = q.k^T / sqrt(d), + (1-mask)*f32::MIN, softmax, .v, head merge -- which the
// reference materialises as [B,h,S,S] f32 plus two transposed copies (called from
// embedding_generator.rs:198). Here a CTA owns 64 query rows of one (sentence, head), streams the
// sentence's keys/values in 64-row tiles through shared memory and keeps the running max / sum /
// output in registers (online softmax, fp32), so neither the scores nor the mask ever exist in
// HBM. Padding is never touched: a sentence only has its `len` real keys, which is exactly what
// the additive f32::MIN mask achieves in the reference (masked keys get weight exp(-huge) = 0).
//
// Tensor-core path: mma.sync.m16n8k16 bf16 (HMMA) with ldmatrix operands. Attention is ~5 % of the
// path's FLOPs at seq 128, d = 32 under-fills any MMA K dimension, and the
// kernel is bound by exp2 and shared-memory traffic rather than the tensor pipe.
#include "kernels.h"

#include "common.cuh"

namespace b200 {

namespace {

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int BQ = 64;   // query rows per CTA (4 warps x 16)
constexpr int BKV = 64;  // keys per tile

template <int D>
__global__ void __launch_bounds__(128)
attention_bf16_kernel(const __nv_bfloat16* __restrict__ qkv, const int* __restrict__ cu, int H, int q_blocks,
                      float scale_log2, __nv_bfloat16* __restrict__ ctx) {
    constexpr int LD = D + 8;  // padded row: conflict-free ldmatrix
    __shared__ __align__(16) __nv_bfloat16 sQ[BQ * LD];
    __shared__ __align__(16) __nv_bfloat16 sK[BKV * LD];
    __shared__ __align__(16) __nv_bfloat16 sV[BKV * LD];

    const int b = blockIdx.x / q_blocks, head = blockIdx.y;
    const int start = cu[b], len = cu[b + 1] - start;
    const int q0 = (blockIdx.x - b * q_blocks) * BQ;
    if (q0 >= len) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t ld_qkv = static_cast<size_t>(3) * H;
    const __nv_bfloat16* base = qkv + static_cast<size_t>(start) * ld_qkv + head * D;

    constexpr int CPR = D / 8;  // 16-byte chunks per row
    // Q tile (rows past the sentence end are zero)
    for (int i = threadIdx.x; i < BQ * CPR; i += 128) {
        const int r = i / CPR, c = i % CPR;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (q0 + r < len) v = *reinterpret_cast<const uint4*>(base + static_cast<size_t>(q0 + r) * ld_qkv + c * 8);
        *reinterpret_cast<uint4*>(&sQ[r * LD + c * 8]) = v;
    }
    __syncthreads();
    uint32_t qf[D / 16][4];
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks)
        ldsm_x4(qf[ks], smem_u32(&sQ[(warp * 16 + (lane & 15)) * LD + ks * 16 + (lane >> 4) * 8]));

    float o[D / 8][4];
#pragma unroll
    for (int n = 0; n < D / 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;  // rows g and g+8 of this warp's 16

    for (int j0 = 0; j0 < len; j0 += BKV) {
        __syncthreads();  // previous tile fully consumed
        for (int i = threadIdx.x; i < BKV * CPR; i += 128) {
            const int r = i / CPR, c = i % CPR;
            uint4 k = make_uint4(0, 0, 0, 0), v = make_uint4(0, 0, 0, 0);
            if (j0 + r < len) {
                const __nv_bfloat16* p = base + static_cast<size_t>(j0 + r) * ld_qkv + c * 8;
                k = *reinterpret_cast<const uint4*>(p + H);
                v = *reinterpret_cast<const uint4*>(p + 2 * H);
            }
            *reinterpret_cast<uint4*>(&sK[r * LD + c * 8]) = k;
            *reinterpret_cast<uint4*>(&sV[r * LD + c * 8]) = v;
        }
        __syncthreads();

        // S = Q K^T (16 x 64 per warp)
        float s[BKV / 8][4];
#pragma unroll
        for (int nt = 0; nt < BKV / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
            for (int kp = 0; kp < D / 32; ++kp) {  // two 16-wide k steps per ldmatrix.x4
                uint32_t kf[4];
                ldsm_x4(kf, smem_u32(&sK[(nt * 8 + (lane & 7)) * LD + kp * 32 + (lane >> 3) * 8]));
                mma_bf16_16816(s[nt], qf[2 * kp], kf[0], kf[1]);
                mma_bf16_16816(s[nt], qf[2 * kp + 1], kf[2], kf[3]);
            }
        }
        // scale, mask the keys past the sentence end, online softmax
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < BKV / 8; ++nt) {
            const int key = j0 + nt * 8 + (lane & 3) * 2;
            const bool v0 = key < len, v1 = key + 1 < len;
            s[nt][0] = v0 ? s[nt][0] * scale_log2 : -INFINITY;
            s[nt][1] = v1 ? s[nt][1] * scale_log2 : -INFINITY;
            s[nt][2] = v0 ? s[nt][2] * scale_log2 : -INFINITY;
            s[nt][3] = v1 ? s[nt][3] * scale_log2 : -INFINITY;
            mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float c0 = exp2f(m0 - mx0), c1 = exp2f(m1 - mx1);
        m0 = mx0; m1 = mx1;
        l0 *= c0; l1 *= c1;
#pragma unroll
        for (int n = 0; n < D / 8; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }

        uint32_t pf[BKV / 16][4];  // P as A fragments (bf16)
#pragma unroll
        for (int nt = 0; nt < BKV / 8; ++nt) {
            const float p0 = exp2f(s[nt][0] - m0), p1 = exp2f(s[nt][1] - m0);
            const float p2 = exp2f(s[nt][2] - m1), p3 = exp2f(s[nt][3] - m1);
            l0 += p0 + p1;
            l1 += p2 + p3;
            pf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16x2(p0, p1);
            pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(p2, p3);
        }
        // O += P V
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {  // two 8-wide d tiles per ldmatrix.x4.trans
                uint32_t vf[4];
                ldsm_x4_trans(vf, smem_u32(&sV[(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LD + np * 16 +
                                               (lane >> 4) * 8]));
                mma_bf16_16816(o[2 * np], pf[kk], vf[0], vf[1]);
                mma_bf16_16816(o[2 * np + 1], pf[kk], vf[2], vf[3]);
            }
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = 1.0f / l0, inv1 = 1.0f / l1;
    const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
    __nv_bfloat16* out = ctx + static_cast<size_t>(start) * H + head * D + (lane & 3) * 2;
#pragma unroll
    for (int n = 0; n < D / 8; ++n) {
        if (r0 < len)
            *reinterpret_cast<uint32_t*>(out + static_cast<size_t>(r0) * H + n * 8) =
                pack_bf16x2(o[n][0] * inv0, o[n][1] * inv0);
        if (r1 < len)
            *reinterpret_cast<uint32_t*>(out + static_cast<size_t>(r1) * H + n * 8) =
                pack_bf16x2(o[n][2] * inv1, o[n][3] * inv1);
    }
}

}  // namespace

cudaError_t launch_attention_bf16(const __nv_bfloat16* qkv, const int* cu, int batch, int max_seqlen, int heads,
                                  int head_dim, __nv_bfloat16* ctx, cudaStream_t s) {
    if (batch <= 0 || max_seqlen <= 0) return cudaSuccess;
    const int H = heads * head_dim;
    const float scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(head_dim));
    const int q_blocks = (max_seqlen + BQ - 1) / BQ;
    const dim3 grid(static_cast<unsigned>(batch) * q_blocks, heads, 1);
    if (head_dim == 32)
        attention_bf16_kernel<32><<<grid, 128, 0, s>>>(qkv, cu, H, q_blocks, scale_log2, ctx);
    else if (head_dim == 64)
        attention_bf16_kernel<64><<<grid, 128, 0, s>>>(qkv, cu, H, q_blocks, scale_log2, ctx);
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

}  // namespace b200
