// Operand format of the 16-bit tensor-core path. The kernel sources (gemm_tc.cu, ffn_fused.cu, qkv_attention.cu,
// attention_tc.cu, elementwise.cu) are compiled twice: as they are (bf16 operands) and with -DB200_OPERAND_F16 (IEEE
// fp16 operands: tcgen05.mma kind::f16 runs both at the same rate; fp16 has 11 significand bits instead of 8, which is
// what brings the raw output within 2e-3 of the f32 reference - scripts/precision_study_f16.py). Each build puts its
// external functions into its own namespace, b200::op_bf16 or b200::op_f16, inline in the translation unit that
// defines it, and the engine picks per handle. Storage is 16 bits either way; `__nv_bfloat16*` in a signature of this
// library means "16-bit operand in the format of the enclosing namespace".
#pragma once

#if defined(B200_OPERAND_F16)
#define B200_NS_BF16 namespace op_bf16
#define B200_NS_F16 inline namespace op_f16
#define B200_NS_THIS inline namespace op_f16
#else
#define B200_NS_BF16 inline namespace op_bf16
#define B200_NS_F16 namespace op_f16
#define B200_NS_THIS inline namespace op_bf16
#endif
