// Pieces shared by the tcgen05 kernels (gemm_tc.cu, ffn_fused.cu): epilogue math, swizzled staging rows for
// TMA stores, tensor-map encoding.
#pragma once

#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"
#include "gemm_tc.h"

namespace b200 {
B200_NS_THIS {
namespace {

constexpr int CHUNK = 32;                   // epilogue chunk: 32 rows x 32 columns per warp
constexpr int CHUNK_BF16 = 32 * CHUNK * 2;  // 2 KB, 64-byte rows (SWIZZLE_64B)
constexpr int CHUNK_F32 = 32 * CHUNK * 4;   // 4 KB, 128-byte rows (SWIZZLE_128B)
constexpr int SMEM_LIMIT = 232448;          // 227 KB per CTA on sm_100

// gelu_erf(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf(x / sqrt 2) ~= tanh(x (c1 + c3 x^2 + c5 x^4 + c7 x^6)) on
// |x| <= 5 (clamped; beyond that both sides are 1 to 1e-6). Coefficients from scripts/fit_gelu.py: max abs error
// 1.2e-5 (2.5e-5 |x|), 40x tighter than the usual tanh-GELU and far below the bf16 rounding of the output.
// One MUFU.TANH + 8 FMA-pipe ops instead of ~25 for erff().
__device__ __forceinline__ float gelu_erf_fast(float x) {
    const float xc = fminf(fmaxf(x, -5.0f), 5.0f);
    const float x2 = xc * xc;
    float p = fmaf(x2, -8.645902750674527e-06f, -2.546073342664881e-04f);
    p = fmaf(x2, p, 3.672949335123125e-02f);
    p = fmaf(x2, p, 7.976898743684135e-01f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(xc * p));
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}

// ---- packed fp32 pairs (FFMA2 / FMUL2 / FADD2): same FMA-pipe time as two scalar ops, half the issue slots ----
__device__ __forceinline__ uint64_t pk2(float a, float b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// gelu_erf_fast on a pair; same arithmetic as the scalar form for |x| <= 5 (x^2 is clamped instead of x: beyond 5 the
// tanh argument only grows, so the result is 0.5 x (1 +- 1) either way).
__device__ __forceinline__ uint64_t gelu_erf_fast2(uint64_t x) {
    float a, b;
    upk2(mul2(x, x), a, b);
    const uint64_t x2 = pk2(fminf(a, 25.0f), fminf(b, 25.0f));
    uint64_t p = fma2(x2, pk2(-8.645902750674527e-06f, -8.645902750674527e-06f), pk2(-2.546073342664881e-04f, -2.546073342664881e-04f));
    p = fma2(x2, p, pk2(3.672949335123125e-02f, 3.672949335123125e-02f));
    p = fma2(x2, p, pk2(7.976898743684135e-01f, 7.976898743684135e-01f));
    upk2(mul2(x, p), a, b);
    float ta, tb;
    asm("tanh.approx.f32 %0, %1;" : "=f"(ta) : "f"(a));
    asm("tanh.approx.f32 %0, %1;" : "=f"(tb) : "f"(b));
    const uint64_t hx = mul2(x, pk2(0.5f, 0.5f));
    return fma2(hx, pk2(ta, tb), hx);
}
// (acc pair + bias pair) through the epilogue's activation -> two bf16 in one word
template <int EPI>
__device__ __forceinline__ uint32_t epi_pair_bf16(uint32_t acc0, uint32_t acc1, float b0, float b1);

template <int EPI>
__device__ __forceinline__ float epi_act(float x) {
    if constexpr (EPI == EPI_GELU_ERF) return gelu_erf_fast(x);
    else if constexpr (EPI == EPI_GELU_TANH) return activate(x, 1);
    else if constexpr (EPI == EPI_RELU) return fmaxf(x, 0.f);
    else return x;
}

template <int EPI>
__device__ __forceinline__ uint32_t epi_pair_bf16(uint32_t acc0, uint32_t acc1, float b0, float b1) {
    if constexpr (EPI == EPI_GELU_ERF || EPI == EPI_BIAS) {
        uint64_t x = add2(pk2(__uint_as_float(acc0), __uint_as_float(acc1)), pk2(b0, b1));
        if constexpr (EPI == EPI_GELU_ERF) x = gelu_erf_fast2(x);
        float a, b;
        upk2(x, a, b);
        return pack_op2(a, b);
    } else {
        return pack_op2(epi_act<EPI>(__uint_as_float(acc0) + b0), epi_act<EPI>(__uint_as_float(acc1) + b1));
    }
}

// Staging buffers are laid out as the TMA store box expects.
// fp32: 128-byte rows, SWIZZLE_128B (piece q of row r lands at piece q ^ (r & 7)).
__device__ __forceinline__ void stage_row_f32(uint32_t buf, int r, const float (&f)[32]) {
    const uint32_t row = buf + r * 128;
    const int sw = r & 7;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const uint32_t a = row + ((q ^ sw) << 4);
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(f[q * 4 + 0]), "f"(f[q * 4 + 1]),
                     "f"(f[q * 4 + 2]), "f"(f[q * 4 + 3])
                     : "memory");
    }
}
__device__ __forceinline__ void sts_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// 32 bytes (one full sector) per lane straight to global memory
__device__ __forceinline__ void stg_v8(void* ptr, const uint32_t* w) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// ---- residual + LayerNorm epilogue arithmetic on one 32-column chunk of this thread's row (packed fp32 pairs) ----
// Pass 1: x = acc + bias + residual_hi (+ residual_lo), accumulated into (sum, sum of squares) pairs; x replaces acc
// in v. bias_s: shared address of the chunk's 32 bias floats; hi_addr: shared addresses of the four 16-byte pieces (8
// 16-bit values each) of this thread's row of the residual chunk, lo_addr: of the two 16-byte pieces (16 e5m2 bytes each)
// of its low halves, whatever layout they sit in.
__device__ __forceinline__ void ln_accumulate_chunk(uint32_t (&v)[32], uint32_t bias_s, const uint32_t (&hi_addr)[4],
                                                    const uint32_t (&lo_addr)[2], bool has_lo, uint64_t& sum2,
                                                    uint64_t& sq2) {
    uint32_t lw[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // four low halves per word
    if (has_lo) {
        const uint4 l0 = lds_v4(lo_addr[0]), l1 = lds_v4(lo_addr[1]);
        lw[0] = l0.x; lw[1] = l0.y; lw[2] = l0.z; lw[3] = l0.w;
        lw[4] = l1.x; lw[5] = l1.y; lw[6] = l1.z; lw[7] = l1.w;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint4 r = lds_v4(hi_addr[q]);
        const uint4 b0 = lds_v4(bias_s + q * 32), b1 = lds_v4(bias_s + q * 32 + 16);
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
        const uint32_t bw[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint64_t x = add2(pk2(__uint_as_float(v[q * 8 + 2 * j]), __uint_as_float(v[q * 8 + 2 * j + 1])),
                              pk2(__uint_as_float(bw[2 * j]), __uint_as_float(bw[2 * j + 1])));
            { const float2 f = unpack_op2(rw[j]); x = add2(x, pk2(f.x, f.y)); }
            if (has_lo) { const float2 f = unpack_lo8x2(lw[2 * q + (j >> 1)] >> ((j & 1) * 16)); x = add2(x, pk2(f.x, f.y)); }
            sum2 = add2(sum2, x);
            sq2 = fma2(x, x, sq2);
            float a, b;
            upk2(x, a, b);
            v[q * 8 + 2 * j] = __float_as_uint(a);
            v[q * 8 + 2 * j + 1] = __float_as_uint(b);
        }
    }
}
// Pass 2: y = (x - mean) * rstd * gamma + beta -> hi = 16-bit(y), two values per word, and lo = e5m2(y - hi), four per
// word. gamma_s / beta_s: shared addresses of the chunk's 32 floats.
__device__ __forceinline__ void ln_normalize_chunk(const uint32_t (&x)[32], float mean, float rstd, uint32_t gamma_s,
                                                   uint32_t beta_s, uint32_t (&hi)[16], uint32_t (&lo)[8]) {
    const uint64_t r2 = pk2(rstd, rstd), nm2 = pk2(-mean * rstd, -mean * rstd), m1 = pk2(-1.0f, -1.0f);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const uint4 g = lds_v4(gamma_s + q * 16), be = lds_v4(beta_s + q * 16);
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, bw[4] = {be.x, be.y, be.z, be.w};
        float d[4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint64_t t = fma2(pk2(__uint_as_float(x[q * 4 + 2 * j]), __uint_as_float(x[q * 4 + 2 * j + 1])), r2, nm2);
            const uint64_t y = fma2(t, pk2(__uint_as_float(gw[2 * j]), __uint_as_float(gw[2 * j + 1])),
                                    pk2(__uint_as_float(bw[2 * j]), __uint_as_float(bw[2 * j + 1])));
            float a, b;
            upk2(y, a, b);
            const uint32_t h = pack_op2(a, b);
            hi[q * 2 + j] = h;
            const float2 hf = unpack_op2(h);
            upk2(fma2(pk2(hf.x, hf.y), m1, y), d[2 * j], d[2 * j + 1]);
        }
        lo[q] = pack_lo8x4(d[0], d[1], d[2], d[3]);
    }
}
// one staged 32 x 32 bf16 chunk (SWIZZLE_64B rows of 64 bytes): this lane's 16 words
__device__ __forceinline__ void stage_words_bf16(uint32_t buf, int r, const uint32_t (&w)[16]) {
    const uint32_t row = buf + r * 64;
    const int sw = (r >> 1) & 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) sts_v4(row + ((q ^ sw) << 4), w[q * 4], w[q * 4 + 1], w[q * 4 + 2], w[q * 4 + 3]);
}

// one staged 32 x 32 chunk of low halves (SWIZZLE_32B rows of 32 bytes; the buffer is 256-byte aligned): this lane's 8 words
__device__ __forceinline__ void stage_words_lo8(uint32_t buf, int r, const uint32_t (&w)[8]) {
    const uint32_t row = buf + r * 32;
    const int sw = (r >> 2) & 1;
#pragma unroll
    for (int q = 0; q < 2; ++q) sts_v4(row + ((q ^ sw) << 4), w[q * 4], w[q * 4 + 1], w[q * 4 + 2], w[q * 4 + 3]);
}
// shared addresses of this lane's two 16-byte pieces of such a chunk
__device__ __forceinline__ void lo8_piece_addrs(uint32_t buf, int r, uint32_t (&addr)[2]) {
    const uint32_t row = buf + r * 32;
    const int sw = (r >> 2) & 1;
    addr[0] = row + (sw << 4);
    addr[1] = row + ((1 ^ sw) << 4);
}

// ---- attention operand descriptors (attention_tc.cu, qkv_attention.cu) ----
// K-major operand, rows of 64 bytes (D = 32): SWIZZLE_64B, 8-row groups 512 B apart.
__device__ __forceinline__ uint64_t umma_desc_k_sw64(uint32_t a) {
    return static_cast<uint64_t>((a & 0x3FFFFu) >> 4) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
// MN-major operand (V as stored: [key, d], d contiguous). The swizzle atom is one row of D*2 bytes (MN) by 8 keys
// (K); SBO = distance between 8-key groups; LBO (between MN atoms) is unused because N = D fits one atom.
template <int D>
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t a) {
    constexpr uint64_t sbo = (8 * D * 2) >> 4;
    constexpr uint64_t layout = D == 32 ? 4 : 2;  // SWIZZLE_64B : SWIZZLE_128B
    return static_cast<uint64_t>((a & 0x3FFFFu) >> 4) | (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}
template <int D>
__device__ __forceinline__ uint64_t umma_desc_k(uint32_t a) {
    return D == 32 ? umma_desc_k_sw64(a) : umma_desc_sw128(a);
}

// max of three in one instruction (FMNMX3)
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
// row maximum of n fp32 bit patterns (n % 8 == 0): four chains of three-input maxima
template <int N>
__device__ __forceinline__ float row_max(const uint32_t (&v)[N]) {
    static_assert(N % 8 == 0, "eight values per round");
    float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
    for (int j = 0; j < N; j += 8) {
        m0 = fmax3(m0, __uint_as_float(v[j]), __uint_as_float(v[j + 1]));
        m1 = fmax3(m1, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        m2 = fmax3(m2, __uint_as_float(v[j + 4]), __uint_as_float(v[j + 5]));
        m3 = fmax3(m3, __uint_as_float(v[j + 6]), __uint_as_float(v[j + 7]));
    }
    return fmaxf(fmax3(m0, m1, m2), m3);
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

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

// Row-major [rows, cols] matrix of es-byte elements (1: bytes, 2: 16-bit operands, 4: fp32); box = box_cols x box_rows.
bool make_tmap_es(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_cols, uint32_t box_rows,
                  int es, CUtensorMapSwizzle swizzle) {
    auto fn = get_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * static_cast<uint64_t>(es)};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = es == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                           : (es == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8);
    CUresult r = fn(map, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}
bool make_tmap(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_cols, uint32_t box_rows,
               bool f32, CUtensorMapSwizzle swizzle) {
    return make_tmap_es(map, ptr, rows, cols, box_cols, box_rows, f32 ? 4 : 2, swizzle);
}
// the low halves of a [rows, cols] activation: one byte per element, boxes of 32 x 32 bytes, SWIZZLE_32B
bool make_tmap_lo8(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_cols, uint32_t box_rows) {
    return make_tmap_es(map, ptr, rows, cols, box_cols, box_rows, 1, CU_TENSOR_MAP_SWIZZLE_32B);
}

}  // namespace
}  // operand-format namespace
}  // namespace b200
