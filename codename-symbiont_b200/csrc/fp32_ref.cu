// fp32 validation mode (B200_EMBED_FLAG_FP32): the same path with fp32 weights, activations and
// plain SIMT arithmetic -- no tensor cores, no bf16 anywhere. It exists to separate "kernel bug" from
// "bf16 rounding" when checking against the oracle (gate: cosine >= 0.999999), not to be fast.
// Restates candle's f32 ops: Linear (matmul + bias), gelu_erf, softmax attention (bert.rs).
#include "kernels.h"

#include "common.cuh"

namespace b200 {

namespace {

// y[M,N] = act(x[M,K] . w[N,K]^T + bias). 64x64 tile, 16x16 threads, 4x4 outputs per thread, K step 16.
__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, int m_cap,
                const int* __restrict__ d_rows, int n, int k, int act, float* __restrict__ y) {
    __shared__ float sx[16][64 + 4];
    __shared__ float sw[16][64 + 4];
    int rows = m_cap;
    if (d_rows != nullptr) rows = min(rows, *d_rows);
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    if (m0 >= rows) return;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < k; k0 += 16) {
        for (int i = threadIdx.x; i < 64 * 16; i += 256) {
            const int r = i >> 4, c = i & 15;
            sx[c][r] = (m0 + r < rows && k0 + c < k) ? x[static_cast<size_t>(m0 + r) * k + k0 + c] : 0.f;
            sw[c][r] = (n0 + r < n && k0 + c < k) ? w[static_cast<size_t>(n0 + r) * k + k0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = sx[kk][ty * 4 + i]; b[i] = sw[kk][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = m0 + ty * 4 + i;
        if (r >= rows) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = n0 + tx * 4 + j;
            if (c < n) y[static_cast<size_t>(r) * n + c] = activate(acc[i][j] + bias[c], act);
        }
    }
}

// One CTA per (sentence, head); each warp owns query rows warp, warp+4, ...; scores of one row live in
// shared memory; exact max-subtracted softmax in fp32 over the sentence's real keys only.
__global__ void __launch_bounds__(128)
attention_f32_kernel(const float* __restrict__ qkv, const int* __restrict__ cu, int H, int D, int max_seqlen,
                     float* __restrict__ ctx) {
    extern __shared__ float sm[];  // [4][max_seqlen] scores + [4][D] query
    const int b = blockIdx.x, head = blockIdx.y;
    const int start = cu[b], len = cu[b + 1] - start;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* sc = sm + warp * max_seqlen;
    float* sq = sm + 4 * max_seqlen + warp * D;
    const size_t ld = static_cast<size_t>(3) * H;
    const float* base = qkv + static_cast<size_t>(start) * ld + head * D;
    const float inv_sqrt_d = 1.0f / sqrtf(static_cast<float>(D));
    for (int i = warp; i < len; i += 4) {
        for (int c = lane; c < D; c += 32) sq[c] = base[static_cast<size_t>(i) * ld + c];
        __syncwarp();
        float mx = -INFINITY;
        for (int j = lane; j < len; j += 32) {
            const float* kj = base + static_cast<size_t>(j) * ld + H;
            float s = 0.f;
            for (int c = 0; c < D; ++c) s = fmaf(sq[c], kj[c], s);
            s *= inv_sqrt_d;
            sc[j] = s;
            mx = fmaxf(mx, s);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int j = lane; j < len; j += 32) {
            const float p = expf(sc[j] - mx);
            sc[j] = p;
            sum += p;
        }
        sum = warp_sum(sum);
        __syncwarp();
        for (int c = lane; c < D; c += 32) {
            float o = 0.f;
            for (int j = 0; j < len; ++j) o = fmaf(sc[j] / sum, base[static_cast<size_t>(j) * ld + 2 * H + c], o);
            ctx[static_cast<size_t>(start + i) * H + head * D + c] = o;
        }
        __syncwarp();
    }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = __float2bfloat16_rn(src[i]);
}

}  // namespace

cudaError_t launch_gemm_f32(const float* x, const float* w, const float* bias, int m_cap, const int* d_rows, int n,
                            int k, int act, float* y, cudaStream_t s) {
    if (m_cap <= 0) return cudaSuccess;
    const dim3 grid((n + 63) / 64, (m_cap + 63) / 64);
    gemm_f32_kernel<<<grid, 256, 0, s>>>(x, w, bias, m_cap, d_rows, n, k, act, y);
    return cudaGetLastError();
}

cudaError_t launch_attention_f32(const float* qkv, const int* cu, int batch, int max_seqlen, int heads,
                                 int head_dim, float* ctx, cudaStream_t s) {
    if (batch <= 0 || max_seqlen <= 0) return cudaSuccess;
    const size_t smem = (4 * static_cast<size_t>(max_seqlen) + 4 * head_dim) * sizeof(float);
    attention_f32_kernel<<<dim3(batch, heads), 128, smem, s>>>(qkv, cu, heads * head_dim, head_dim, max_seqlen,
                                                              ctx);
    return cudaGetLastError();
}

__global__ void f32_to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, size_t n) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = __float2half_rn(src[i]);
}
cudaError_t launch_f32_to_f16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    f32_to_f16_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(src, reinterpret_cast<__half*>(dst), n);
    return cudaGetLastError();
}

cudaError_t launch_f32_to_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    f32_to_bf16_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, dst, n);
    return cudaGetLastError();
}

}  // namespace b200
