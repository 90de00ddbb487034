// Memory-bound kernels of the embedding path (all HBM-bound; one warp per token row, 8/16-byte
// vector accesses, fp32 statistics, warp-shuffle reductions):
//   lengths / cu_seqlens from the tokenizer's attention mask   (embedding_generator.rs:185, :190)
//   K1 embedding gather + LayerNorm       (candle bert.rs BertEmbeddings::forward)
//   K4 residual add + LayerNorm           (BertSelfOutput::forward / BertOutput::forward)
//   K5 masked mean-pool (+ L2 normalise)  (embedding_generator.rs:201-207; L2 is the north_star addition)
#include "kernels.h"

#include "common.cuh"

namespace b200 {
B200_NS_THIS {

namespace {

constexpr int MAX_CHUNKS = 8;  // hidden <= 1024: 32 lanes x 4 elements x 8 chunks

// 16-bit activations that feed a residual add are kept as a (hi, lo) pair: hi = the 16-bit value the next GEMM
// multiplies, lo = y - hi as an 8-bit e5m2 float (common.cuh); hi + lo is what the residual add and the pooling read.
// Rounding the residual stream to 16 bits alone would cost 8e-3 (bf16) / 2.9e-3 (fp16) max-abs on short sentences
// (scripts/precision_study.py, precision_study_f16.py).
template <typename T>
__device__ __forceinline__ void store_split(T* hi, uint8_t* lo, int col, float4 y);
template <>
__device__ __forceinline__ void store_split<float>(float* hi, uint8_t*, int col, float4 y) {
    Vec4<float>::store(hi + col, y);
}
template <>
__device__ __forceinline__ void store_split<__nv_bfloat16>(__nv_bfloat16* hi, uint8_t* lo, int col, float4 y) {
    const uint2 h = make_uint2(pack_op2(y.x, y.y), pack_op2(y.z, y.w));
    *reinterpret_cast<uint2*>(hi + col) = h;
    if (lo != nullptr) {
        const float2 a = unpack_op2(h.x), b = unpack_op2(h.y);
        *reinterpret_cast<uint32_t*>(lo + col) = pack_lo8x4(y.x - a.x, y.y - a.y, y.z - b.x, y.w - b.y);
    }
}
template <typename T>
__device__ __forceinline__ float4 load_split(const T* hi, const uint8_t* lo, size_t off) {
    float4 a = Vec4<T>::load(hi + off);
    if (lo != nullptr) {
        const float4 b = unpack_lo8x4(*reinterpret_cast<const uint32_t*>(lo + off));
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    return a;
}

// Exact two-pass LayerNorm of one row held across a warp (chunk i of lane l = columns 128 i + 4 l ..+3).
template <typename T, int NC>
__device__ __forceinline__ void warp_layernorm_store(float4 (&v)[NC], int nchunk, int H, int lane,
                                                     const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, float eps, T* out_row,
                                                     uint8_t* out_lo_row) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NC; ++i)
        if (i < nchunk) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    const float mean = warp_sum(s) / static_cast<float>(H);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NC; ++i)
        if (i < nchunk) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    const float var = warp_sum(q) / static_cast<float>(H);
    const float rstd = 1.0f / sqrtf(var + eps);
#pragma unroll
    for (int i = 0; i < NC; ++i)
        if (i < nchunk) {
            const int col = i * 128 + lane * 4;
            const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + col));
            const float4 b = __ldg(reinterpret_cast<const float4*>(beta + col));
            float4 y;
            y.x = (v[i].x - mean) * rstd * g.x + b.x;
            y.y = (v[i].y - mean) * rstd * g.y + b.y;
            y.z = (v[i].z - mean) * rstd * g.z + b.z;
            y.w = (v[i].w - mean) * rstd * g.w + b.w;
            store_split<T>(out_row, out_lo_row, col, y);
        }
}

// ---- attention mask -> lengths ------------------------------------------------------------------
// One warp per sentence. The mask must be a right-padded 0/1 prefix (PaddingDirection::Right,
// embedding_generator.rs:85); anything else sets status bit 0.
__global__ void mask_lengths_kernel(const uint32_t* __restrict__ mask, int batch, int seq_len,
                                    int* __restrict__ lens, int* __restrict__ status) {
    pdl_trigger();
    pdl_wait();
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= batch) return;
    const uint32_t* row = mask + static_cast<size_t>(b) * seq_len;
    int cnt = 0;
    bool bad = false;
    for (int j = lane; j < seq_len; j += 32) {
        const uint32_t m = row[j];
        if (m > 1u) bad = true;
        if (j > 0 && m > row[j - 1]) bad = true;
        cnt += (m != 0u);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
        lens[b] = cnt;
        if (bad) atomicOr(status, 1);
    }
}

// Single-block exclusive scan: cu[0] = 0, cu[b+1] = cu[b] + lens[b].
__global__ void scan_lengths_kernel(const int* __restrict__ lens, int batch, int* __restrict__ cu) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) { carry = 0; cu[0] = 0; }
    __syncthreads();
    for (int base = 0; base < batch; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = i < batch ? lens[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = lane < (blockDim.x >> 5) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;  // inclusive
        }
        __syncthreads();
        const int prefix = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + x;  // inclusive
        if (i < batch) cu[i + 1] = prefix;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = prefix;
        __syncthreads();
    }
}

// Tiles for K23: consecutive whole sentences with at most `cap` tokens together (a longer sentence gets a tile of its
// own; K23 is only used when none is). tile_first[t] = first sentence of tile t, tile_first[n_tiles] = batch.
// Greedy packing is sequential, so it runs per SEGMENT of 128 sentences (a tile never crosses a segment boundary: at
// most one tile per segment is cut short) by one thread each; counts are scanned in shared memory by the single
// block, which then writes the compacted list. seg_off: [ceil(batch / 128) + 1] scratch (callers size it for segments of 32).
constexpr int PACK_SEG = 128;  // 4484 tiles for 16 384 corpus-length sentences with 32, 4351 with 128, 4310 without segments
__device__ __forceinline__ int pack_segment(const int* cu, int b0, int b1, int cap, int* __restrict__ out) {
    int n = 0, first = b0;
    int base = cu[b0];
    for (int b = b0; b < b1; ++b) {
        const int end = cu[b + 1];
        if (end - base > cap && b > first) {  // sentence b does not fit: it starts the next tile
            if (out) out[n] = first;
            ++n;
            first = b;
            base = cu[b];
        }
    }
    if (out) out[n] = first;
    return n + 1;
}
__global__ void pack_tiles_kernel(const int* cu, int batch, int cap, int stage_max, int* __restrict__ seg_off,
                                  int* __restrict__ tile_first, int* __restrict__ n_tiles) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    // the sequential walks below read cu twice, one dependent load after the other: for the batch sizes the callers send
    // (the engine <= 8192 sentences per launch, the corpus driver <= 16384) the block first copies it into shared memory
    // with coalesced loads; `stage_max` = the capacity of the dynamic shared array (0: read global memory)
    extern __shared__ int s_cu[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nseg = (batch + PACK_SEG - 1) / PACK_SEG;
    if (batch <= stage_max) {
        for (int i = threadIdx.x; i <= batch; i += blockDim.x) s_cu[i] = cu[i];
        cu = s_cu;
    }
    if (threadIdx.x == 0) { carry = 0; seg_off[0] = 0; }
    __syncthreads();
    for (int base = 0; base < nseg; base += blockDim.x) {  // count tiles per segment, exclusive scan into seg_off
        const int sg = base + threadIdx.x;
        const int v = sg < nseg ? pack_segment(cu, sg * PACK_SEG, min(batch, (sg + 1) * PACK_SEG), cap, nullptr) : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = lane < (blockDim.x >> 5) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const int prefix = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + x;  // inclusive
        if (sg < nseg) seg_off[sg + 1] = prefix;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = prefix;
        __syncthreads();
    }
    for (int sg = threadIdx.x; sg < nseg; sg += blockDim.x)  // same walk again, now writing the compacted list
        pack_segment(cu, sg * PACK_SEG, min(batch, (sg + 1) * PACK_SEG), cap, tile_first + seg_off[sg]);
    if (threadIdx.x == 0) {
        const int n = carry;
        *n_tiles = n;
        tile_first[n] = batch;
    }
}

// ---- K1 -------------------------------------------------------------------------------------------
// One warp per TOK consecutive token slots. PADDED: slots are [batch, seq_len] positions, pad positions are skipped,
// the row is written packed at cu[b] + j. Packed: slots are tokens, sentence found by binary search in cu.
// The work per token is a chain of dependent loads (bounds -> id -> three table rows) followed by a short reduction:
// all TOK tokens' loads are issued before any of them is consumed, which is what keeps enough bytes in flight.
// five blocks per SM (51 registers; the 54 the compiler takes on its own allow four): 0.374 -> 0.337 ms at c2, six and
// eight blocks spill a little and give 0.345
#ifndef EMBED_MIN_BLOCKS
#define EMBED_MIN_BLOCKS 5
#endif
template <typename T, bool PADDED, int TOK, int NCH>  // NCH: 128-column chunks a token needs (H <= 128 NCH)
__global__ void __launch_bounds__(256, EMBED_MIN_BLOCKS)
embed_layernorm_kernel(const uint32_t* __restrict__ ids, const uint32_t* __restrict__ type_ids,
                       const int* __restrict__ cu, int batch, int seq_len, int total_tokens,
                       const float* __restrict__ word, const float* __restrict__ pos,
                       const float* __restrict__ type, const float* __restrict__ gamma,
                       const float* __restrict__ beta, float eps, int H, int vocab,
                       int type_vocab, int max_pos, T* __restrict__ out, uint8_t* __restrict__ out_lo,
                       int* __restrict__ status) {
    pdl_trigger();
    pdl_wait();
    const long long w0 = (static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * TOK;
    const int lane = threadIdx.x & 31;
    const long long slots = PADDED ? static_cast<long long>(batch) * seq_len : total_tokens;
    long long src[TOK], dst[TOK];
    int j[TOK];
    bool live[TOK];
#pragma unroll
    for (int t = 0; t < TOK; ++t) {
        const long long w = w0 + t;
        live[t] = w < slots;
        src[t] = w; dst[t] = w; j[t] = 0;
        if (!live[t]) continue;
        if (PADDED) {
            const int b = static_cast<int>(w / seq_len);
            j[t] = static_cast<int>(w - static_cast<long long>(b) * seq_len);
            const int start = __ldg(cu + b);
            live[t] = j[t] < __ldg(cu + b + 1) - start;
            dst[t] = start + j[t];
        } else {
            int lo = 0, hi = batch;  // largest b with cu[b] <= w
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (__ldg(cu + mid) <= w) lo = mid; else hi = mid;
            }
            j[t] = static_cast<int>(w - __ldg(cu + lo));
        }
    }
    uint32_t id[TOK], tt[TOK];
#pragma unroll
    for (int t = 0; t < TOK; ++t) {
        id[t] = live[t] ? __ldg(ids + src[t]) : 0u;
        tt[t] = live[t] ? __ldg(type_ids + src[t]) : 0u;
    }
    bool bad = false;
#pragma unroll
    for (int t = 0; t < TOK; ++t) {
        if (id[t] >= static_cast<uint32_t>(vocab)) { bad = true; id[t] = 0u; }
        if (tt[t] >= static_cast<uint32_t>(type_vocab)) { bad = true; tt[t] = 0u; }
        if (j[t] >= max_pos) { bad = true; j[t] = 0; }
    }
    if (bad && lane == 0) atomicOr(status, 2);
    constexpr int nchunk = NCH;  // H == 128 * NCH: no predicated chunks
    float4 v[TOK][NCH];
#pragma unroll
    for (int t = 0; t < TOK; ++t) {
        const float* wr = word + static_cast<size_t>(id[t]) * H;
        const float* tr = type + static_cast<size_t>(tt[t]) * H;
        const float* pr = pos + static_cast<size_t>(j[t]) * H;
#pragma unroll
        for (int i = 0; i < NCH; ++i)
            if (i < nchunk && live[t]) {
                const int col = i * 128 + lane * 4;
                const float4 a = Vec4<float>::load(wr + col), ty = Vec4<float>::load(tr + col), p = Vec4<float>::load(pr + col);
                v[t][i] = make_float4((a.x + ty.x) + p.x, (a.y + ty.y) + p.y, (a.z + ty.z) + p.z, (a.w + ty.w) + p.w);
            }
    }
#pragma unroll
    for (int t = 0; t < TOK; ++t)
        if (live[t])
            warp_layernorm_store<T, NCH>(v[t], nchunk, H, lane, gamma, beta, eps, out + static_cast<size_t>(dst[t]) * H,
                                    out_lo ? out_lo + static_cast<size_t>(dst[t]) * H : nullptr);
}

// ---- K4 -------------------------------------------------------------------------------------------
template <typename TY, typename T>
__global__ void add_layernorm_kernel(const TY* __restrict__ y, const T* __restrict__ residual,
                                     const uint8_t* __restrict__ residual_lo, int rows_cap,
                                     const int* __restrict__ d_rows, const float* __restrict__ gamma,
                                     const float* __restrict__ beta, float eps, int H, T* __restrict__ out,
                                     uint8_t* __restrict__ out_lo) {
    const long long r = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    int rows = rows_cap;
    if (d_rows != nullptr) rows = min(rows, *d_rows);
    if (r >= rows) return;
    const int nchunk = H >> 7;
    float4 v[MAX_CHUNKS];
#pragma unroll
    for (int i = 0; i < MAX_CHUNKS; ++i)
        if (i < nchunk) {
            const int col = i * 128 + lane * 4;
            const float4 a = Vec4<TY>::load(y + r * H + col), b = load_split<T>(residual, residual_lo, r * H + col);
            v[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
        }
    warp_layernorm_store<T, MAX_CHUNKS>(v, nchunk, H, lane, gamma, beta, eps, out + r * H, out_lo ? out_lo + r * H : nullptr);
}

// ---- K5 -------------------------------------------------------------------------------------------
// One block per sentence: thread (g, c) sums rows g, g+G, ... of 4 columns; groups are combined in
// shared memory; the divisor is len + 1e-9 in fp32 exactly as the reference; optional L2 norm.
template <typename T>
__global__ void mean_pool_kernel(const T* __restrict__ hidden, const uint8_t* __restrict__ hidden_lo,
                                 const int* __restrict__ cu, int H, int normalize, float* __restrict__ out) {
    extern __shared__ float4 part[];  // [G][H/4]
    __shared__ float red[32];
    const int b = blockIdx.x;
    const int cols4 = H >> 2;
    const int G = blockDim.x / cols4;
    const int g = threadIdx.x / cols4, c = threadIdx.x - g * cols4;
    pdl_trigger();
    pdl_wait();
    const int start = cu[b], len = cu[b + 1] - start;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g < G) {
        const size_t p = static_cast<size_t>(start) * H + c * 4;
        int j = g;
        for (; j + G < len; j += 2 * G) {  // two rows in flight
            const float4 a = load_split<T>(hidden, hidden_lo, p + static_cast<size_t>(j) * H);
            const float4 d = load_split<T>(hidden, hidden_lo, p + static_cast<size_t>(j + G) * H);
            acc.x += a.x + d.x; acc.y += a.y + d.y; acc.z += a.z + d.z; acc.w += a.w + d.w;
        }
        if (j < len) {
            const float4 a = load_split<T>(hidden, hidden_lo, p + static_cast<size_t>(j) * H);
            acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
        }
        part[g * cols4 + c] = acc;
    }
    __syncthreads();
    float ss = 0.f;
    float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
    if (threadIdx.x < cols4) {
        for (int k = 0; k < G; ++k) {
            const float4 a = part[k * cols4 + threadIdx.x];
            tot.x += a.x; tot.y += a.y; tot.z += a.z; tot.w += a.w;
        }
        const float den = static_cast<float>(len) + 1e-9f;
        tot.x /= den; tot.y /= den; tot.z /= den; tot.w /= den;
        ss = (tot.x * tot.x + tot.y * tot.y) + (tot.z * tot.z + tot.w * tot.w);
    }
    if (normalize) {
        ss = warp_sum(ss);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        float t = 0.f;
        for (int k = 0; k < (blockDim.x >> 5); ++k) t += red[k];
        const float inv = 1.0f / fmaxf(sqrtf(t), 1e-12f);
        tot.x *= inv; tot.y *= inv; tot.z *= inv; tot.w *= inv;
    }
    if (threadIdx.x < cols4) *reinterpret_cast<float4*>(out + static_cast<size_t>(b) * H + threadIdx.x * 4) = tot;
}

}  // namespace

cudaError_t launch_mask_lengths(const uint32_t* mask, int batch, int seq_len, int* lens, int* cu, int* status,
                                cudaStream_t s) {
    if (batch <= 0) return cudaSuccess;
    if (cudaError_t e = launch_pdl(mask_lengths_kernel, dim3((batch + 7) / 8), dim3(256), 0, s, mask, batch, seq_len, lens, status);
        e != cudaSuccess)
        return e;
    return launch_pdl(scan_lengths_kernel, dim3(1), dim3(1024), 0, s, static_cast<const int*>(lens), batch, cu);
}

cudaError_t launch_pack_tiles(const int* cu, int batch, int cap, int* seg_off, int* tile_first, int* n_tiles,
                              cudaStream_t s) {
    if (batch <= 0) return cudaSuccess;
    constexpr int STAGE_LIMIT = 32768;  // sentences whose cu fits the staging array (128 KB of dynamic shared memory)
    const int stage_max = batch <= STAGE_LIMIT ? batch : 0;
    const size_t smem = stage_max ? (static_cast<size_t>(stage_max) + 1) * sizeof(int) : 0;
    if (smem > 48 * 1024) {
        static PerDeviceOnce attr_set;  // function attributes belong to a (function, device) pair
        if (!attr_set.done()) {
            cudaError_t e = cudaFuncSetAttribute(pack_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (STAGE_LIMIT + 1) * static_cast<int>(sizeof(int)));
            if (e != cudaSuccess) return e;
            attr_set.mark();
        }
    }
    pack_tiles_kernel<<<1, 1024, smem, s>>>(cu, batch, cap, stage_max, seg_off, tile_first, n_tiles);
    return cudaGetLastError();
}

template <typename T>
cudaError_t launch_embed_layernorm(const EmbedArgs& a, cudaStream_t s) {
    const long long slots = a.padded ? static_cast<long long>(a.batch) * a.seq_len : a.total_tokens;
    if (slots <= 0) return cudaSuccess;
    // two tokens per warp (measured at H = 384: 0.51 ms with one, 0.41 with two, 0.50 with four per warp)
    const int nch = a.H / 128;
    if (a.H % 128 != 0 || nch < 1 || nch > MAX_CHUNKS) return cudaErrorInvalidValue;
    const int tok = nch <= 4 ? 2 : 1;
    const unsigned grid = static_cast<unsigned>((slots + 8LL * tok - 1) / (8LL * tok));
    cudaError_t launch_err = cudaSuccess;
#define B200_EMBED_LAUNCH(PAD, TOK, NCH)                                                                                   \
    launch_err = launch_pdl(embed_layernorm_kernel<T, PAD, TOK, NCH>, dim3(grid), dim3(256), 0, s,                       \
        a.ids, a.type_ids, a.cu, a.batch, a.seq_len, a.total_tokens, a.word, a.pos, a.type, a.gamma, a.beta, a.eps, a.H,  \
        a.vocab, a.type_vocab, a.max_pos, static_cast<T*>(a.out), static_cast<uint8_t*>(a.out_lo), a.status)
#define B200_EMBED_PICK(PAD)                                    \
    switch (nch) {                                              \
        case 1: B200_EMBED_LAUNCH(PAD, 2, 1); break;            \
        case 2: B200_EMBED_LAUNCH(PAD, 2, 2); break;            \
        case 3: B200_EMBED_LAUNCH(PAD, 2, 3); break;            \
        case 4: B200_EMBED_LAUNCH(PAD, 2, 4); break;            \
        case 5: B200_EMBED_LAUNCH(PAD, 1, 5); break;            \
        case 6: B200_EMBED_LAUNCH(PAD, 1, 6); break;            \
        case 7: B200_EMBED_LAUNCH(PAD, 1, 7); break;            \
        default: B200_EMBED_LAUNCH(PAD, 1, 8); break;           \
    }
    if (a.padded) { B200_EMBED_PICK(true); } else { B200_EMBED_PICK(false); }
#undef B200_EMBED_PICK
#undef B200_EMBED_LAUNCH
    return launch_err;
}
template cudaError_t launch_embed_layernorm<__nv_bfloat16>(const EmbedArgs&, cudaStream_t);
template cudaError_t launch_embed_layernorm<float>(const EmbedArgs&, cudaStream_t);

template <typename TY, typename T>
cudaError_t launch_add_layernorm(const TY* y, const T* residual, const uint8_t* residual_lo, int rows_cap,
                                 const int* d_rows, const float* gamma, const float* beta, float eps, int H, T* out,
                                 uint8_t* out_lo, cudaStream_t s) {
    if (rows_cap <= 0) return cudaSuccess;
    add_layernorm_kernel<TY, T><<<(rows_cap + 7) / 8, 256, 0, s>>>(y, residual, residual_lo, rows_cap, d_rows, gamma,
                                                                  beta, eps, H, out, out_lo);
    return cudaGetLastError();
}
template cudaError_t launch_add_layernorm<__nv_bfloat16, __nv_bfloat16>(const __nv_bfloat16*, const __nv_bfloat16*,
                                                                        const uint8_t*, int, const int*,
                                                                        const float*, const float*, float, int,
                                                                        __nv_bfloat16*, uint8_t*, cudaStream_t);
template cudaError_t launch_add_layernorm<float, __nv_bfloat16>(const float*, const __nv_bfloat16*,
                                                                const uint8_t*, int, const int*, const float*,
                                                                const float*, float, int, __nv_bfloat16*,
                                                                uint8_t*, cudaStream_t);
template cudaError_t launch_add_layernorm<float, float>(const float*, const float*, const uint8_t*, int, const int*,
                                                        const float*, const float*, float, int, float*, uint8_t*,
                                                        cudaStream_t);

template <typename T>
cudaError_t launch_mean_pool(const T* hidden, const uint8_t* hidden_lo, const int* cu, int batch, int H, int normalize,
                             float* out, cudaStream_t s) {
    if (batch <= 0) return cudaSuccess;
    const int cols4 = H / 4;
    int G = 512 / cols4;
    if (G < 1) G = 1;
    if (G > 8) G = 8;
    const int threads = ((G * cols4 + 31) / 32) * 32;
    return launch_pdl(mean_pool_kernel<T>, dim3(batch), dim3(threads), G * cols4 * sizeof(float4), s, hidden, hidden_lo, cu, H,
                      normalize, out);
}
template cudaError_t launch_mean_pool<__nv_bfloat16>(const __nv_bfloat16*, const uint8_t*, const int*, int, int,
                                                     int, float*, cudaStream_t);
template cudaError_t launch_mean_pool<float>(const float*, const uint8_t*, const int*, int, int, int, float*,
                                                     cudaStream_t);

}  // operand-format namespace
}  // namespace b200
