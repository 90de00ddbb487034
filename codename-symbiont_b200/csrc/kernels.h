// Host-side launchers of the non-GEMM kernels (elementwise.cu, attention_tc.cu, qkv_attention.cu, ffn_fused.cu,
// fp32_ref.cu). The 16-bit kernels exist once per operand format (fmt.h): b200::op_bf16::* and b200::op_f16::*.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "fmt.h"

namespace b200 {

struct EmbedArgs {
    const uint32_t* ids;       // padded [batch, seq_len] or packed [total_tokens]
    const uint32_t* type_ids;
    const int* cu;             // device cu_seqlens[batch + 1]
    int batch, seq_len, total_tokens;
    int padded;                // 1: ids are [batch, seq_len]; 0: packed
    const float *word, *pos, *type;  // fp32 tables (16-bit tables alone cost 6e-3 max-abs, precision_study.py)
    const float *gamma, *beta;
    float eps;
    int H, vocab, type_vocab, max_pos;
    void* out;                 // packed [tokens, H] in T
    void* out_lo;              // 16-bit mode: low half of the (hi, lo) pair (one e5m2 byte per element), or nullptr
    int* status;               // |= 2 on id / type id / position out of range
};

// K2f (ffn_fused.cu): x' = LayerNorm(gelu(a.W1^T + b1).W2^T + b2 + a) in one kernel; hidden 128 / 256 / 384, erf-GELU.
struct FfnFusedArgs {
    const void* a;       // [m, hidden] 16-bit: GEMM operand and high half of the residual
    const void* a_lo;    // [m, hidden] bytes: low half of the residual (e5m2), or nullptr
    const void* w1;      // [inter, hidden] 16-bit (intermediate.dense.weight)
    const void* w2;      // [hidden, inter] 16-bit (output.dense.weight)
    const float *b1, *b2, *gamma, *beta;
    float eps;
    int m, hidden, inter;
    const int* d_rows;   // optional device-side row count
    void* out;           // [m, hidden] 16-bit (hi)
    void* out_lo;        // [m, hidden] bytes (lo) or nullptr
};

// Everything below exists once per operand format; T = __nv_bfloat16 stands for "16-bit operand of this namespace".
#define B200_FORMAT_KERNEL_DECLS                                                                                        \
    /* attention mask [batch, seq_len] -> lens[batch], cu[batch+1]; status |= 1 if not a 0/1 prefix. */                 \
    cudaError_t launch_mask_lengths(const uint32_t* mask, int batch, int seq_len, int* lens, int* cu, int* status,      \
                                    cudaStream_t s);                                                                    \
    template <typename T>                                                                                               \
    cudaError_t launch_embed_layernorm(const EmbedArgs& a, cudaStream_t s);                                             \
    template <typename TY, typename T>                                                                                  \
    cudaError_t launch_add_layernorm(const TY* y, const T* residual, const uint8_t* residual_lo, int rows_cap,          \
                                     const int* d_rows, const float* gamma, const float* beta, float eps, int H,        \
                                     T* out, uint8_t* out_lo, cudaStream_t s);                                          \
    template <typename T>                                                                                               \
    cudaError_t launch_mean_pool(const T* hidden, const uint8_t* hidden_lo, const int* cu, int batch, int H, int normalize, \
                                 float* out, cudaStream_t s);                                                           \
    /* K3 on tcgen05 (attention_tc.cu): varlen flash attention over packed qkv [tokens, 3H]; head_dim 32 or 64; qkv    \
       has rows_cap rows (TMA bound), persistent grid. */                                                               \
    cudaError_t launch_attention_tc(const __nv_bfloat16* qkv, const int* cu, int rows_cap, int batch, int max_seqlen,   \
                                    int heads, int head_dim, __nv_bfloat16* ctx, int num_sms, cudaStream_t s);          \
    /* K23 (qkv_attention.cu): Q/K/V projection + attention of whole sentences in one kernel: ctx from x (hi)           \
       directly, the packed qkv tensor never exists. hidden 128 / 256 / 384 with head_dim 32, sentences of at most 128  \
       tokens. Its work items are tiles of whole sentences with at most 128 tokens together, listed by                  \
       launch_pack_tiles (tile_first [batch + 1], n_tiles device scalar, seg_off [batch / 32 + 2] scratch; cap = 128). */ \
    bool qkv_attention_supported(int hidden, int heads, int max_seqlen);                                                \
    cudaError_t launch_pack_tiles(const int* cu, int batch, int cap, int* seg_off, int* tile_first, int* n_tiles,       \
                                  cudaStream_t s);                                                                      \
    cudaError_t launch_qkv_attention(const __nv_bfloat16* x, const __nv_bfloat16* wqkv, const float* bqkv,              \
                                     const int* cu, const int* tile_first, const int* n_tiles, int rows_cap, int batch, \
                                     int max_seqlen, int hidden, int heads, __nv_bfloat16* ctx, int num_sms,            \
                                     cudaStream_t s);                                                                   \
    bool ffn_fused_supported(int hidden, int inter, int act);                                                           \
    cudaError_t ffn_fused_launch(const FfnFusedArgs& a, cudaStream_t stream, int num_sms);

B200_NS_BF16 { B200_FORMAT_KERNEL_DECLS }
B200_NS_F16 { B200_FORMAT_KERNEL_DECLS }

// fp32 validation kernels (correctness only).
cudaError_t launch_attention_f32(const float* qkv, const int* cu, int batch, int max_seqlen, int heads,
                                 int head_dim, float* ctx, cudaStream_t s);
// y[M,N] = act(x[M,K] . w[N,K]^T + bias); act: -1 none, else B200_EMBED_ACT_*.
cudaError_t launch_gemm_f32(const float* x, const float* w, const float* bias, int m_cap, const int* d_rows, int n,
                            int k, int act, float* y, cudaStream_t s);
// fp32 -> bf16 / fp16 (RNE) weight conversion
cudaError_t launch_f32_to_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t s);
cudaError_t launch_f32_to_f16(const float* src, __nv_bfloat16* dst /* fp16 bits */, size_t n, cudaStream_t s);

}  // namespace b200
