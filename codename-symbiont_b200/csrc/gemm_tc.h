// Host interface of the tcgen05 GEMM (gemm_tc.cu).
#pragma once
#include <cuda_runtime.h>

#include "fmt.h"

namespace b200 {

enum GemmEpilogue {
    EPI_BIAS = 0,       // y = x.W^T + b
    EPI_GELU_ERF = 1,   // y = gelu_erf(x.W^T + b)      (candle HiddenAct::Gelu)
    EPI_GELU_TANH = 2,  // tanh-form GELU               (HiddenAct::GeluApproximate)
    EPI_RELU = 3,       // relu                         (HiddenAct::Relu)
    EPI_RES_LN = 4,     // y = LayerNorm(x.W^T + b + residual); block_n * cluster size == n
};

struct GemmTcArgs {
    const void* a;         // [m, k] 16-bit row-major (activations)
    const void* b;         // [n, k] 16-bit row-major (HF Linear weight)
    int m, n, k;
    const int* d_rows;     // optional device int: rows actually present (<= m); tiles beyond are skipped
    const float* bias;     // [n]
    const void* residual;  // [m, n] 16-bit (EPI_RES_LN), high half of the residual stream
    const void* residual_lo;  // [m, n] bytes: low half (residual = hi + lo, lo an e5m2 byte per element), or nullptr
    const float* gamma;    // [n]     (EPI_RES_LN)
    const float* beta;     // [n]     (EPI_RES_LN)
    float eps;
    void* out;             // [m, n] 16-bit, or fp32 when out_f32
    void* out_lo;          // EPI_RES_LN: optional [m, n] bytes, low half of the output
    int out_f32;           // only with EPI_BIAS
    int epi;               // GemmEpilogue
    int pair;              // plain bf16 epilogues: 1 = clusters of 2 CTAs on adjacent row blocks, one cta_group::2 MMA per pair
    int block_n;           // 128, 192, 256 (EPI_RES_LN: n / cluster size, from gemm_tc_pick_block_n)
};

// One set per operand format (fmt.h).
#define B200_GEMM_TC_DECLS                                                                               \
    /* EPI_RES_LN: CTAs per cluster that share one row block (1, 2 or 4); 0 if this N cannot be fused. */ \
    int gemm_tc_pick_cluster(int n, int epi);                                                            \
    /* Largest supported tile width for this N and epilogue; 0 if none. */                               \
    int gemm_tc_pick_block_n(int n, int epi);                                                            \
    cudaError_t gemm_tc_launch(const GemmTcArgs& args, cudaStream_t stream, int num_sms);
B200_NS_BF16 { B200_GEMM_TC_DECLS }
B200_NS_F16 { B200_GEMM_TC_DECLS }

}  // namespace b200
