// f4: exact cosine top-k over the embeddings the path produces, resident in HBM, as the alternative to the
// Qdrant search hop of the consumer:
//   /root/reference/services/vector_memory_service/src/main.rs:34-37   collection: VectorParams{size, Distance::Cosine}
//   :140-202  upsert: one point per sentence, vectors = the embedding          -> b200_index_add / _add_device
//   :261-286  SearchPoints{vector, limit = top_k, score_threshold: None, ..}   -> b200_index_search
//   :319-403  results in descending score order, score = cosine similarity     -> ids + scores, payload lookup stays host-side
// Qdrant (un-vendored service; client crate qdrant-client in vector_memory_service/Cargo.toml) normalises every vector on
// insert and the query on search for Distance::Cosine and scores with a dot product (lib/segment/src/spaces/simple.rs,
// cosine_preprocess: unchanged if the length is < f32::EPSILON or within 1e-6 of 1); below its indexing threshold it scans
// every vector, above it it answers from an HNSW graph (approximate). This is the exact scan at any size.
//
// Work per search: one pass over the [n, dim] f32 matrix (HBM-bound: 4 * dim bytes per vector, shared by up to 4 queries),
// a score per vector and query, then a top-k selection that touches 4 bytes per vector. Kernels:
//   normalize_rows   one warp per vector, in place (insert and query preparation)
//   scan             one warp per block of 32 consecutive vectors, query fragments in registers, fp32 FMA, coalesced loads
//                    (512 B per warp instruction, streaming), one coalesced 128 B score store per block
//   rescore          exact f32 scores of a short candidate list (second stage when the optional bf16 copy of the rows serves
//                    the scan: same lane layout and operation order as scan, so the bits agree)
//   topk_filter /    blocks of 4096 candidates: the k best of each kept (k <= 256: threshold from the 256 per-thread bests, then
//   topk_level       a sort of the few survivors; above: a bitonic sort of the block), ties -> lower id; applied level by
//                    level until one block holds the answer
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>

#include "../../include/b200_embed.h"

namespace {

thread_local std::string g_index_err;
int fail(int code, const std::string& msg) {
    g_index_err = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    const int code = (e == cudaErrorMemoryAllocation) ? B200_EMBED_ERR_OOM
                     : (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? B200_EMBED_ERR_NO_DEVICE
                                                                                   : B200_EMBED_ERR_CUDA;
    return fail(code, std::string(what) + ": " + cudaGetErrorString(e));
}
#define IX_TRY(expr)                                          \
    do {                                                      \
        cudaError_t e__ = (expr);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, #expr); \
    } while (0)

constexpr int QT_MAX = 4;      // queries that share one pass over the matrix
constexpr int SEL_CHUNK = 4096;  // candidates per selection block
constexpr int SEL_THREADS = 256;
constexpr int K_MAX = 1024;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// cosine_preprocess, one warp per row, in place
__global__ void normalize_rows_kernel(float* __restrict__ v, int64_t n, int dim) {
    const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    float4* p = reinterpret_cast<float4*>(v + row * dim);
    const int n4 = dim >> 2;
    float s = 0.f;
    for (int j = lane; j < n4; j += 32) {
        const float4 x = p[j];
        s = fmaf(x.x, x.x, s); s = fmaf(x.y, x.y, s); s = fmaf(x.z, x.z, s); s = fmaf(x.w, x.w, s);
    }
    s = warp_sum(s);
    // Qdrant's cosine_preprocess tests the SQUARED length (is_length_zero_or_normalized): below f32::EPSILON, or within
    // 1e-6 of 1, leaves the vector as it is
    if (s < 1.1920929e-07f || fabsf(s - 1.0f) <= 1e-6f) return;
    const float len = sqrtf(s);
    for (int j = lane; j < n4; j += 32) {
        float4 x = p[j];
        x.x /= len; x.y /= len; x.z /= len; x.w /= len;
        p[j] = x;
    }
}

// scores[t][r] = <vecs[r], q[t]> for r < n. EPL = dim / 32 elements per lane (as EPL/4 float4: lane, lane+32, ...).
// Four consecutive elements of a row as floats: a 16-byte load of the f32 matrix, an 8-byte load of its bf16 copy.
__device__ __forceinline__ float4 load4(const float* row, int idx4) { return __ldcs(reinterpret_cast<const float4*>(row) + idx4); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16* row, int idx4) {
    const uint2 u = __ldcs(reinterpret_cast<const uint2*>(row) + idx4);
    return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16),
                       __uint_as_float(u.y & 0xffff0000u));
}

template <int EPL, int QT, typename ET>
__global__ void __launch_bounds__(256) scan_kernel(const ET* __restrict__ vecs, int64_t n, const float* __restrict__ queries,
                                                   float* __restrict__ scores, int64_t score_stride) {
    constexpr int V4 = EPL / 4, DIM = EPL * 32;
    // rows per step: all their loads are issued before any is consumed. Two rows of f32 or four of bf16 keep the same
    // number of bytes in flight per warp (four only while the row fragments fit in registers).
    constexpr int R = (sizeof(ET) == 2 && EPL <= 16 && QT <= 2) ? 4 : 2;
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    float4 q[QT][V4];
#pragma unroll
    for (int t = 0; t < QT; ++t)
#pragma unroll
        for (int j = 0; j < V4; ++j) q[t][j] = reinterpret_cast<const float4*>(queries + t * DIM)[j * 32 + lane];
    const int64_t nblocks = (n + 31) >> 5;
    for (int64_t b = warp_global; b < nblocks; b += warps) {
        const int64_t r0 = b << 5;
        const int rows = static_cast<int>(n - r0 < 32 ? n - r0 : 32);
        float mine[QT];
#pragma unroll
        for (int t = 0; t < QT; ++t) mine[t] = 0.f;
        for (int i = 0; i < rows; i += R) {
            float4 x[R][V4];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const ET* p = vecs + (r0 + (i + r < rows ? i + r : i)) * DIM;  // a row past the end re-reads row i (unused)
#pragma unroll
                for (int j = 0; j < V4; ++j) x[r][j] = load4(p, j * 32 + lane);
            }
#pragma unroll
            for (int t = 0; t < QT; ++t) {
                float a[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    a[r] = 0.f;
#pragma unroll
                    for (int j = 0; j < V4; ++j) {
                        a[r] = fmaf(x[r][j].x, q[t][j].x, a[r]); a[r] = fmaf(x[r][j].y, q[t][j].y, a[r]);
                        a[r] = fmaf(x[r][j].z, q[t][j].z, a[r]); a[r] = fmaf(x[r][j].w, q[t][j].w, a[r]);
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    a[r] = warp_sum(a[r]);
                    if (i + r < rows && lane == i + r) mine[t] = a[r];
                }
            }
        }
        if (lane < rows) {
#pragma unroll
            for (int t = 0; t < QT; ++t) scores[t * score_stride + r0 + lane] = mine[t];
        }
    }
}

// Exact scores of a short candidate list: one warp per (candidate, query), the same lane layout and the same order of
// fused multiply-adds as scan_kernel on the f32 matrix, so a rescored candidate gets bit for bit the score a full f32
// scan gives it.
template <int EPL>
__global__ void __launch_bounds__(256) rescore_kernel(const float* __restrict__ vecs, const float* __restrict__ queries,
                                                      const int64_t* __restrict__ ids, float* __restrict__ scores, int kp,
                                                      int stride) {
    constexpr int V4 = EPL / 4, DIM = EPL * 32;
    const int lane = threadIdx.x & 31;
    const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), t = blockIdx.y;
    if (c >= kp) return;
    const int64_t id = ids[t * stride + c];
    if (id < 0) {
        if (lane == 0) scores[t * stride + c] = -INFINITY;
        return;
    }
    const float* row = vecs + id * DIM;
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < V4; ++j) {
        const float4 x = load4(row, j * 32 + lane);
        const float4 q = reinterpret_cast<const float4*>(queries + t * DIM)[j * 32 + lane];
        a = fmaf(x.x, q.x, a); a = fmaf(x.y, q.y, a); a = fmaf(x.z, q.z, a); a = fmaf(x.w, q.w, a);
    }
    a = warp_sum(a);
    if (lane == 0) scores[t * stride + c] = a;
}

__global__ void to_bf16_rows_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t count) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < count) dst[i] = __float2bfloat16_rn(src[i]);
}

// (score, id) ordering of the results: higher score first, lower id first among equal scores; NaN never wins.
__device__ __forceinline__ bool better(float s, int64_t id, float bs, int64_t bid) {
    return s > bs || (s == bs && id < bid);
}

// One selection level. Block (blk, qi) takes candidates [blk * SEL_CHUNK, +SEL_CHUNK) of query qi (ids implicit = position
// when ids_in == nullptr), sorts them best-first in shared memory (bitonic network: 78 compare-exchange stages whatever
// k is; k rounds of block-wide argmax measured 1.5 us per round) and writes the first k to out[qi][blk * k ...]. Missing
// ones get (-inf, -1).
__global__ void __launch_bounds__(SEL_THREADS) topk_level_kernel(const float* __restrict__ sc_in, const int64_t* __restrict__ ids_in,
                                                                 int64_t count, int64_t in_stride, int k,
                                                                 float* __restrict__ sc_out, int64_t* __restrict__ ids_out,
                                                                 int64_t out_stride) {
    extern __shared__ __align__(16) unsigned char sel_smem[];  // SEL_CHUNK ids (8 B) then SEL_CHUNK scores (4 B)
    int64_t* s_id = reinterpret_cast<int64_t*>(sel_smem);
    float* s_sc = reinterpret_cast<float*>(sel_smem + SEL_CHUNK * sizeof(int64_t));
    const int tid = threadIdx.x;
    const int64_t base = static_cast<int64_t>(blockIdx.x) * SEL_CHUNK;
    const float* sc = sc_in + static_cast<int64_t>(blockIdx.y) * in_stride;
    const int64_t* ids = ids_in ? ids_in + static_cast<int64_t>(blockIdx.y) * in_stride : nullptr;
    const float NEG = -INFINITY;
    for (int i = tid; i < SEL_CHUNK; i += SEL_THREADS) {
        const int64_t g = base + i;
        float s = NEG;
        int64_t id = INT64_MAX;  // padding sorts behind every real candidate, also behind one that scores -inf
        if (g < count) {
            s = sc[g];
            id = ids ? ids[g] : g;
            if (!(s == s) || id < 0) { s = NEG; id = INT64_MAX; }  // NaN scores and padding never win
        }
        s_sc[i] = s;
        s_id[i] = id;
    }
    __syncthreads();
    for (int size = 2; size <= SEL_CHUNK; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < SEL_CHUNK / 2; t += SEL_THREADS) {
                const int i = 2 * t - (t & (stride - 1));  // the lower index of the t-th pair at this stride
                const int j = i + stride;
                const bool best_first = (i & size) == 0;
                const float si = s_sc[i], sj = s_sc[j];
                const int64_t ii = s_id[i], ij = s_id[j];
                const bool swap = best_first ? better(sj, ij, si, ii) : better(si, ii, sj, ij);
                if (swap) { s_sc[i] = sj; s_sc[j] = si; s_id[i] = ij; s_id[j] = ii; }
            }
            __syncthreads();
        }
    }
    float* o_sc = sc_out + static_cast<int64_t>(blockIdx.y) * out_stride + static_cast<int64_t>(blockIdx.x) * k;
    int64_t* o_id = ids_out + static_cast<int64_t>(blockIdx.y) * out_stride + static_cast<int64_t>(blockIdx.x) * k;
    for (int i = tid; i < k; i += SEL_THREADS) {
        const bool real = s_id[i] != INT64_MAX;
        o_sc[i] = real ? s_sc[i] : NEG;
        o_id[i] = real ? s_id[i] : -1;
    }
}

// The same level for k <= 256 without sorting the whole chunk. Each of the 256 threads owns 16 of the 4096 candidates and
// finds its best; the k-th best of those 256 "local bests" (one 256-element bitonic sort) is a lower bound T for the
// chunk's k-th best, because k candidates at least as good as T exist. Only candidates at least as good as T can be in
// the answer - at most 16 k, typically little more than k - so they are compacted (shared-memory counter; the order of
// arrival does not matter) and sorted on their own; the first k are written. More than 1024 survivors (many equal
// scores) fall back to the full sort of the chunk. 36 us -> ~10 us per level for k = 10.
constexpr int FILT_MAX = 1024;
__device__ __forceinline__ void bitonic_sort_pairs(float* sc, int64_t* id, int n, int tid, int nthreads) {  // best first
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < n / 2; t += nthreads) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool best_first = (i & size) == 0;
                const float si = sc[i], sj = sc[j];
                const int64_t ii = id[i], ij = id[j];
                const bool swap = best_first ? better(sj, ij, si, ii) : better(si, ii, sj, ij);
                if (swap) { sc[i] = sj; sc[j] = si; id[i] = ij; id[j] = ii; }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(SEL_THREADS) topk_filter_kernel(const float* __restrict__ sc_in, const int64_t* __restrict__ ids_in,
                                                                  int64_t count, int64_t in_stride, int k,
                                                                  float* __restrict__ sc_out, int64_t* __restrict__ ids_out,
                                                                  int64_t out_stride) {
    extern __shared__ __align__(16) unsigned char sel_smem[];  // chunk ids | chunk scores | survivor ids | survivor scores
    int64_t* s_id = reinterpret_cast<int64_t*>(sel_smem);
    float* s_sc = reinterpret_cast<float*>(sel_smem + SEL_CHUNK * sizeof(int64_t));
    int64_t* t_id = reinterpret_cast<int64_t*>(sel_smem + SEL_CHUNK * (sizeof(int64_t) + sizeof(float)));
    float* t_sc = reinterpret_cast<float*>(sel_smem + SEL_CHUNK * (sizeof(int64_t) + sizeof(float)) + FILT_MAX * sizeof(int64_t));
    __shared__ int n_surv;
    const int tid = threadIdx.x;
    const int64_t base = static_cast<int64_t>(blockIdx.x) * SEL_CHUNK;
    const float* sc = sc_in + static_cast<int64_t>(blockIdx.y) * in_stride;
    const int64_t* ids = ids_in ? ids_in + static_cast<int64_t>(blockIdx.y) * in_stride : nullptr;
    const float NEG = -INFINITY;
    float my_s = NEG;
    int64_t my_id = INT64_MAX;
    for (int i = tid; i < SEL_CHUNK; i += SEL_THREADS) {
        const int64_t g = base + i;
        float v = NEG;
        int64_t id = INT64_MAX;  // padding sorts behind every real candidate
        if (g < count) {
            v = sc[g];
            id = ids ? ids[g] : g;
            if (!(v == v) || id < 0) { v = NEG; id = INT64_MAX; }
        }
        s_sc[i] = v;
        s_id[i] = id;
        if (better(v, id, my_s, my_id)) { my_s = v; my_id = id; }
    }
    if (tid == 0) n_surv = 0;
    t_sc[tid] = my_s;
    t_id[tid] = my_id;
    __syncthreads();
    bitonic_sort_pairs(t_sc, t_id, SEL_THREADS, tid, SEL_THREADS);
    const float thr_s = t_sc[k - 1];
    const int64_t thr_id = t_id[k - 1];
    __syncthreads();  // everyone has read the threshold before the survivor list overwrites the array
    for (int i = tid; i < FILT_MAX; i += SEL_THREADS) { t_sc[i] = NEG; t_id[i] = INT64_MAX; }
    __syncthreads();
    for (int i = tid; i < SEL_CHUNK; i += SEL_THREADS) {
        const float v = s_sc[i];
        const int64_t id = s_id[i];
        if (id != INT64_MAX && !better(thr_s, thr_id, v, id)) {  // at least as good as the threshold
            const int at = atomicAdd(&n_surv, 1);
            if (at < FILT_MAX) { t_sc[at] = v; t_id[at] = id; }
        }
    }
    __syncthreads();
    const int ns = n_surv;
    float* o_sc = sc_out + static_cast<int64_t>(blockIdx.y) * out_stride + static_cast<int64_t>(blockIdx.x) * k;
    int64_t* o_id = ids_out + static_cast<int64_t>(blockIdx.y) * out_stride + static_cast<int64_t>(blockIdx.x) * k;
    const float* r_sc;
    const int64_t* r_id;
    if (ns <= FILT_MAX) {
        int n2 = 32;
        while (n2 < ns) n2 <<= 1;
        bitonic_sort_pairs(t_sc, t_id, n2, tid, SEL_THREADS);
        r_sc = t_sc;
        r_id = t_id;
    } else {
        bitonic_sort_pairs(s_sc, s_id, SEL_CHUNK, tid, SEL_THREADS);
        r_sc = s_sc;
        r_id = s_id;
    }
    for (int i = tid; i < k; i += SEL_THREADS) {
        const bool real = r_id[i] != INT64_MAX;
        o_sc[i] = real ? r_sc[i] : NEG;
        o_id[i] = real ? r_id[i] : -1;
    }
}

constexpr int K_FILTER_MAX = SEL_THREADS;
constexpr size_t FILT_SMEM = SEL_CHUNK * (sizeof(int64_t) + sizeof(float)) + FILT_MAX * (sizeof(int64_t) + sizeof(float));

constexpr size_t SEL_SMEM = SEL_CHUNK * (sizeof(int64_t) + sizeof(float));

template <int EPL, typename ET>
cudaError_t launch_scan_qt(int qt, const ET* vecs, int64_t n, const float* q, float* scores, int64_t stride, int grid,
                           cudaStream_t s) {
    if (qt == 1) scan_kernel<EPL, 1, ET><<<grid, 256, 0, s>>>(vecs, n, q, scores, stride);
    else if (qt == 2) scan_kernel<EPL, 2, ET><<<grid, 256, 0, s>>>(vecs, n, q, scores, stride);
    else if (qt == 3) scan_kernel<EPL, 3, ET><<<grid, 256, 0, s>>>(vecs, n, q, scores, stride);
    else scan_kernel<EPL, 4, ET><<<grid, 256, 0, s>>>(vecs, n, q, scores, stride);
    return cudaGetLastError();
}

template <typename ET>
cudaError_t launch_scan(int dim, int qt, const ET* vecs, int64_t n, const float* q, float* scores, int64_t stride,
                        int num_sms, cudaStream_t s) {
    // 8 warps per CTA, enough CTAs to fill every SM a few times over but never more warps than blocks of 32 rows
    const int64_t blocks32 = (n + 31) / 32;
    int64_t grid = (blocks32 + 7) / 8;
    const int64_t cap = static_cast<int64_t>(num_sms) * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    const int g = static_cast<int>(grid);
    switch (dim) {
        case 128: return launch_scan_qt<4, ET>(qt, vecs, n, q, scores, stride, g, s);
        case 256: return launch_scan_qt<8, ET>(qt, vecs, n, q, scores, stride, g, s);
        case 384: return launch_scan_qt<12, ET>(qt, vecs, n, q, scores, stride, g, s);
        case 512: return launch_scan_qt<16, ET>(qt, vecs, n, q, scores, stride, g, s);
        case 768: return launch_scan_qt<24, ET>(qt, vecs, n, q, scores, stride, g, s);
        case 1024: return launch_scan_qt<32, ET>(qt, vecs, n, q, scores, stride, g, s);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_rescore(int dim, const float* vecs, const float* q, const int64_t* ids, float* scores, int kp, int qt,
                           cudaStream_t s) {
    const dim3 grid((kp + 7) / 8, qt);
    switch (dim) {
        case 128: rescore_kernel<4><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        case 256: rescore_kernel<8><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        case 384: rescore_kernel<12><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        case 512: rescore_kernel<16><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        case 768: rescore_kernel<24><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        case 1024: rescore_kernel<32><<<grid, 256, 0, s>>>(vecs, q, ids, scores, kp, K_MAX); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace

struct b200_index {
    int device = 0, dim = 0, num_sms = 148;
    int64_t cap = 0, n = 0;
    float* vecs = nullptr;          // [cap, dim] f32, rows normalised on insert
    __nv_bfloat16* vecs16 = nullptr;  // optional bf16 copy of the rows for the first stage of a two-stage search
    float* d_q = nullptr;           // [QT_MAX, dim] normalised queries of the current pass
    float* d_scores = nullptr;      // [QT_MAX, cap]
    float* d_csc[2] = {nullptr, nullptr};    // candidate scores of the selection levels (ping-pong)
    int64_t* d_cid[2] = {nullptr, nullptr};  // candidate ids
    int64_t cand_cap = 0;           // per query
    float* d_out_sc = nullptr;      // [QT_MAX, K_MAX]
    int64_t* d_out_id = nullptr;
    float *h_q = nullptr, *h_out_sc = nullptr;  // pinned staging: queries in, results out
    int64_t* h_out_id = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_scan_ms = 0.f, last_select_ms = 0.f;
    int64_t exact_fallbacks = 0;  // two-stage searches that had to be repeated on the f32 matrix
    std::mutex mu;
};

namespace {

struct DevGuard {
    int prev = -1;
    explicit DevGuard(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); else prev = -1; }
    ~DevGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int64_t level_blocks(int64_t count) { return (count + SEL_CHUNK - 1) / SEL_CHUNK; }

// Normalises `q` queries already in ix->d_q, scans, selects; results land in ix->d_out_sc / d_out_id ([qt][k]).
int search_pass(b200_index* ix, int qt, int k, bool prescan) {
    cudaStream_t s = ix->stream;
    const int64_t n = ix->n;
    normalize_rows_kernel<<<(qt * 32 + 255) / 256, 256, 0, s>>>(ix->d_q, qt, ix->dim);
    IX_TRY(cudaGetLastError());
    IX_TRY(cudaEventRecord(ix->ev0, s));
    if (prescan) IX_TRY(launch_scan(ix->dim, qt, ix->vecs16, n, ix->d_q, ix->d_scores, ix->cap, ix->num_sms, s));
    else IX_TRY(launch_scan(ix->dim, qt, ix->vecs, n, ix->d_q, ix->d_scores, ix->cap, ix->num_sms, s));
    IX_TRY(cudaEventRecord(ix->ev1, s));
    // selection levels: scores -> candidates -> ... -> one block
    const float* sc_in = ix->d_scores;
    const int64_t* id_in = nullptr;
    int64_t count = n, in_stride = ix->cap;
    int pp = 0;
    for (;;) {
        const int64_t nb = level_blocks(count);
        const bool last = nb == 1;
        float* sc_out = last ? ix->d_out_sc : ix->d_csc[pp];
        int64_t* id_out = last ? ix->d_out_id : ix->d_cid[pp];
        const int64_t out_stride = last ? K_MAX : ix->cand_cap;
        const dim3 grid(static_cast<unsigned>(nb), static_cast<unsigned>(qt));
        if (k <= K_FILTER_MAX)
            topk_filter_kernel<<<grid, SEL_THREADS, FILT_SMEM, s>>>(sc_in, id_in, count, in_stride, k, sc_out, id_out, out_stride);
        else
            topk_level_kernel<<<grid, SEL_THREADS, SEL_SMEM, s>>>(sc_in, id_in, count, in_stride, k, sc_out, id_out, out_stride);
        IX_TRY(cudaGetLastError());
        if (last) break;
        sc_in = sc_out; id_in = id_out; count = nb * k; in_stride = out_stride;
        pp ^= 1;
    }
    return B200_EMBED_OK;
}

int ensure_capacity(b200_index* ix, int64_t want) {
    if (want <= ix->cap) return B200_EMBED_OK;
    int64_t cap = ix->cap > 0 ? ix->cap : 4096;
    while (cap < want) cap *= 2;
    // Every new buffer is allocated (and the rows copied) before any old one is touched: a failed allocation leaves the
    // index exactly as it was, still answering searches.
    const int64_t cand = level_blocks(cap) * K_MAX;  // the first level keeps at most K_MAX of every SEL_CHUNK candidates
    float *nv = nullptr, *nsc = nullptr, *ncsc[2] = {nullptr, nullptr};
    __nv_bfloat16* n16 = nullptr;
    int64_t* ncid[2] = {nullptr, nullptr};
    auto undo = [&](cudaError_t e, const char* what) {
        for (void* p : {static_cast<void*>(nv), static_cast<void*>(nsc), static_cast<void*>(n16), static_cast<void*>(ncsc[0]),
                        static_cast<void*>(ncsc[1]), static_cast<void*>(ncid[0]), static_cast<void*>(ncid[1])})
            if (p) cudaFree(p);
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? B200_EMBED_ERR_OOM : B200_EMBED_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    };
    cudaError_t e;
    if ((e = cudaMalloc(&nv, static_cast<size_t>(cap) * ix->dim * sizeof(float))) != cudaSuccess) return undo(e, "index rows");
    if (ix->vecs16 && (e = cudaMalloc(&n16, static_cast<size_t>(cap) * ix->dim * sizeof(__nv_bfloat16))) != cudaSuccess) return undo(e, "bf16 rows");
    if ((e = cudaMalloc(&nsc, static_cast<size_t>(cap) * QT_MAX * sizeof(float))) != cudaSuccess) return undo(e, "scores");
    for (int i = 0; i < 2; ++i) {
        if ((e = cudaMalloc(&ncsc[i], static_cast<size_t>(cand) * QT_MAX * sizeof(float))) != cudaSuccess) return undo(e, "candidate scores");
        if ((e = cudaMalloc(&ncid[i], static_cast<size_t>(cand) * QT_MAX * sizeof(int64_t))) != cudaSuccess) return undo(e, "candidate ids");
    }
    if (ix->n > 0) {
        if ((e = cudaMemcpyAsync(nv, ix->vecs, static_cast<size_t>(ix->n) * ix->dim * sizeof(float), cudaMemcpyDeviceToDevice, ix->stream)) != cudaSuccess)
            return undo(e, "copy rows");
        if (n16 && (e = cudaMemcpyAsync(n16, ix->vecs16, static_cast<size_t>(ix->n) * ix->dim * sizeof(__nv_bfloat16), cudaMemcpyDeviceToDevice,
                                        ix->stream)) != cudaSuccess)
            return undo(e, "copy bf16 rows");
    }
    if ((e = cudaStreamSynchronize(ix->stream)) != cudaSuccess) return undo(e, "copy rows");
    for (void* p : {static_cast<void*>(ix->vecs), static_cast<void*>(ix->vecs16), static_cast<void*>(ix->d_scores),
                    static_cast<void*>(ix->d_csc[0]), static_cast<void*>(ix->d_csc[1]), static_cast<void*>(ix->d_cid[0]),
                    static_cast<void*>(ix->d_cid[1])})
        if (p) cudaFree(p);
    ix->vecs = nv;
    ix->vecs16 = n16;
    ix->d_scores = nsc;
    for (int i = 0; i < 2; ++i) { ix->d_csc[i] = ncsc[i]; ix->d_cid[i] = ncid[i]; }
    ix->cand_cap = cand;
    ix->cap = cap;
    return B200_EMBED_OK;
}

}  // namespace

extern "C" {

const char* b200_index_last_error(void) { return g_index_err.c_str(); }

int32_t b200_index_create(int32_t dim, int64_t capacity_hint, int32_t device, b200_index** out) {
    if (!out) return fail(B200_EMBED_ERR_INVALID_ARG, "null out pointer");
    *out = nullptr;
    if (!(dim == 128 || dim == 256 || dim == 384 || dim == 512 || dim == 768 || dim == 1024))
        return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "index dimension must be 128, 256, 384, 512, 768 or 1024");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(B200_EMBED_ERR_NO_DEVICE, "no CUDA device: the index has no CPU fallback");
    if (device < 0 || device >= count) return fail(B200_EMBED_ERR_INVALID_ARG, "bad device ordinal");
    DevGuard guard(device);
    b200_index* ix = new b200_index();
    ix->device = device;
    ix->dim = dim;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ix->num_sms = prop.multiProcessorCount;
    auto bail = [&](int rc) { b200_index_destroy(ix); return rc; };
    if (cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(fail(B200_EMBED_ERR_CUDA, "stream"));
    if (cudaEventCreate(&ix->ev0) != cudaSuccess || cudaEventCreate(&ix->ev1) != cudaSuccess) return bail(fail(B200_EMBED_ERR_CUDA, "event"));
    if (cudaMalloc(&ix->d_q, sizeof(float) * QT_MAX * dim) != cudaSuccess ||
        cudaMalloc(&ix->d_out_sc, sizeof(float) * QT_MAX * K_MAX) != cudaSuccess ||
        cudaMalloc(&ix->d_out_id, sizeof(int64_t) * QT_MAX * K_MAX) != cudaSuccess)
        return bail(fail(B200_EMBED_ERR_OOM, "workspace"));
    if (cudaHostAlloc(&ix->h_q, sizeof(float) * QT_MAX * dim, cudaHostAllocDefault) != cudaSuccess ||
        cudaHostAlloc(&ix->h_out_sc, sizeof(float) * QT_MAX * K_MAX, cudaHostAllocDefault) != cudaSuccess ||
        cudaHostAlloc(&ix->h_out_id, sizeof(int64_t) * QT_MAX * K_MAX, cudaHostAllocDefault) != cudaSuccess)
        return bail(fail(B200_EMBED_ERR_OOM, "pinned staging"));
    if (cudaFuncSetAttribute(topk_level_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SEL_SMEM)) != cudaSuccess ||
        cudaFuncSetAttribute(topk_filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(FILT_SMEM)) != cudaSuccess)
        return bail(fail(B200_EMBED_ERR_CUDA, "shared memory attribute"));
    const int rc = ensure_capacity(ix, capacity_hint > 0 ? capacity_hint : 4096);
    if (rc) return bail(rc);
    *out = ix;
    return B200_EMBED_OK;
}

void b200_index_destroy(b200_index* ix) {
    if (!ix) return;
    DevGuard guard(ix->device);
    if (ix->stream) cudaStreamSynchronize(ix->stream);
    for (void* p : {static_cast<void*>(ix->vecs), static_cast<void*>(ix->d_q), static_cast<void*>(ix->d_scores),
                    static_cast<void*>(ix->d_csc[0]), static_cast<void*>(ix->d_csc[1]), static_cast<void*>(ix->d_cid[0]),
                    static_cast<void*>(ix->d_cid[1]), static_cast<void*>(ix->d_out_sc), static_cast<void*>(ix->d_out_id),
                    static_cast<void*>(ix->vecs16)})
        if (p) cudaFree(p);
    for (void* p : {static_cast<void*>(ix->h_q), static_cast<void*>(ix->h_out_sc), static_cast<void*>(ix->h_out_id)})
        if (p) cudaFreeHost(p);
    if (ix->ev0) cudaEventDestroy(ix->ev0);
    if (ix->ev1) cudaEventDestroy(ix->ev1);
    if (ix->stream) cudaStreamDestroy(ix->stream);
    delete ix;
}

int64_t b200_index_size(const b200_index* ix) { return ix ? ix->n : -1; }
int32_t b200_index_dim(const b200_index* ix) { return ix ? ix->dim : -1; }

static int32_t index_add(b200_index* ix, const float* vectors, int64_t n, bool on_device) {
    if (!ix || n < 0 || (n > 0 && !vectors)) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if (n == 0) return B200_EMBED_OK;
    std::lock_guard<std::mutex> lk(ix->mu);
    DevGuard guard(ix->device);
    const int rc = ensure_capacity(ix, ix->n + n);
    if (rc) return rc;
    float* dst = ix->vecs + ix->n * ix->dim;
    IX_TRY(cudaMemcpyAsync(dst, vectors, static_cast<size_t>(n) * ix->dim * sizeof(float),
                           on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ix->stream));
    const int64_t blocks = (n * 32 + 255) / 256;
    normalize_rows_kernel<<<static_cast<unsigned>(blocks), 256, 0, ix->stream>>>(dst, n, ix->dim);
    IX_TRY(cudaGetLastError());
    if (ix->vecs16) {
        const int64_t count = n * ix->dim;
        to_bf16_rows_kernel<<<static_cast<unsigned>((count + 255) / 256), 256, 0, ix->stream>>>(dst, ix->vecs16 + ix->n * ix->dim, count);
        IX_TRY(cudaGetLastError());
    }
    IX_TRY(cudaStreamSynchronize(ix->stream));
    ix->n += n;
    return B200_EMBED_OK;
}

int32_t b200_index_add(b200_index* ix, const float* vectors, int64_t n) { return index_add(ix, vectors, n, false); }
int32_t b200_index_add_device(b200_index* ix, const float* device_vectors, int64_t n) { return index_add(ix, device_vectors, n, true); }

int32_t b200_index_search(b200_index* ix, const float* queries, int32_t q, int32_t top_k, int64_t* ids_out, float* scores_out) {
    if (!ix || q < 0 || (q > 0 && (!queries || !ids_out || !scores_out))) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if (top_k < 1 || top_k > K_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "top_k must be between 1 and 1024");
    std::lock_guard<std::mutex> lk(ix->mu);
    DevGuard guard(ix->device);
    ix->last_scan_ms = ix->last_select_ms = 0.f;
    if (ix->n == 0) {
        for (int64_t i = 0; i < static_cast<int64_t>(q) * top_k; ++i) { ids_out[i] = -1; scores_out[i] = -INFINITY; }
        return B200_EMBED_OK;
    }
    // Two-stage search when the index carries a bf16 copy (b200_index_enable_prescan): the scan reads the
    // copy (half the bytes). Rounding a normalised row to bf16 moves its score by at most eps = 2^-8 (|x| |q| = 1, plus
    // slack for the f32 sums), so nothing whose approximate score is below (k-th best approximate score - 2 eps) can be
    // in the exact answer. The best kp >= 4 k approximate candidates are taken (kp = max(32, 4 k)); when the kp-th lies below that bound they
    // contain every possible member of the answer, they are rescored exactly from the f32 rows and sorted. Otherwise (a
    // crowd of near-equal scores) the pass is repeated on the f32 matrix. The result is the exact one either way.
    // It pays for a pass of ONE query with top_k <= 64 (measured: 0.57 -> 0.41 ms per call at 2 M x 384, 1.87 -> 1.04 ms
    // at 4 M x 768); passes of several queries share one f32 scan, which is cheaper per query than a bf16 scan each, and a
    // longer candidate list costs more to select than the scan saves.
    constexpr float EPS2 = 2.f * (1.f / 256.f + 1e-5f);
    for (int q0 = 0; q0 < q; q0 += QT_MAX) {
        const int qt = q - q0 < QT_MAX ? q - q0 : QT_MAX;
        const bool two_stage = ix->vecs16 != nullptr && top_k <= 64 && qt == 1;
        const int kp = two_stage ? std::max(32, 4 * top_k) : top_k;
        std::memcpy(ix->h_q, queries + static_cast<size_t>(q0) * ix->dim, sizeof(float) * qt * ix->dim);
        IX_TRY(cudaMemcpyAsync(ix->d_q, ix->h_q, sizeof(float) * qt * ix->dim, cudaMemcpyHostToDevice, ix->stream));
        bool exact_pass = !two_stage;
        if (two_stage) {
            int rc = search_pass(ix, qt, kp, true);
            if (rc) return rc;
            IX_TRY(cudaMemcpyAsync(ix->h_out_sc, ix->d_out_sc, sizeof(float) * qt * K_MAX, cudaMemcpyDeviceToHost, ix->stream));
            IX_TRY(cudaMemcpyAsync(ix->h_out_id, ix->d_out_id, sizeof(int64_t) * qt * K_MAX, cudaMemcpyDeviceToHost, ix->stream));
            IX_TRY(cudaStreamSynchronize(ix->stream));
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, ix->ev0, ix->ev1) == cudaSuccess) ix->last_scan_ms += ms;
            for (int t = 0; t < qt && !exact_pass; ++t) {
                const float* sc = ix->h_out_sc + static_cast<size_t>(t) * K_MAX;
                const int64_t* id = ix->h_out_id + static_cast<size_t>(t) * K_MAX;
                if (id[kp - 1] < 0) continue;  // fewer than kp vectors in all: every one of them is a candidate
                if (!(sc[kp - 1] < sc[top_k - 1] - EPS2)) exact_pass = true;  // the candidate list may be too short
            }
            if (!exact_pass) {
                IX_TRY(launch_rescore(ix->dim, ix->vecs, ix->d_q, ix->d_out_id, ix->d_out_sc, kp, qt, ix->stream));
                IX_TRY(cudaMemcpyAsync(ix->h_out_sc, ix->d_out_sc, sizeof(float) * qt * K_MAX, cudaMemcpyDeviceToHost, ix->stream));
                IX_TRY(cudaStreamSynchronize(ix->stream));
                for (int t = 0; t < qt; ++t) {
                    const float* sc = ix->h_out_sc + static_cast<size_t>(t) * K_MAX;
                    const int64_t* id = ix->h_out_id + static_cast<size_t>(t) * K_MAX;
                    int order[K_MAX];
                    int m = 0;
                    for (int c = 0; c < kp; ++c)
                        if (id[c] >= 0) order[m++] = c;
                    std::sort(order, order + m, [&](int x, int y) { return sc[x] > sc[y] || (sc[x] == sc[y] && id[x] < id[y]); });
                    float* so = scores_out + static_cast<size_t>(q0 + t) * top_k;
                    int64_t* io = ids_out + static_cast<size_t>(q0 + t) * top_k;
                    for (int c = 0; c < top_k; ++c) {
                        so[c] = c < m ? sc[order[c]] : -INFINITY;
                        io[c] = c < m ? id[order[c]] : -1;
                    }
                }
                continue;
            }
            ++ix->exact_fallbacks;
        }
        const int rc = search_pass(ix, qt, top_k, false);
        if (rc) return rc;
        IX_TRY(cudaMemcpyAsync(ix->h_out_sc, ix->d_out_sc, sizeof(float) * qt * K_MAX, cudaMemcpyDeviceToHost, ix->stream));
        IX_TRY(cudaMemcpyAsync(ix->h_out_id, ix->d_out_id, sizeof(int64_t) * qt * K_MAX, cudaMemcpyDeviceToHost, ix->stream));
        IX_TRY(cudaStreamSynchronize(ix->stream));
        for (int t = 0; t < qt; ++t) {
            std::memcpy(scores_out + static_cast<size_t>(q0 + t) * top_k, ix->h_out_sc + static_cast<size_t>(t) * K_MAX, sizeof(float) * top_k);
            std::memcpy(ids_out + static_cast<size_t>(q0 + t) * top_k, ix->h_out_id + static_cast<size_t>(t) * K_MAX, sizeof(int64_t) * top_k);
        }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ix->ev0, ix->ev1) == cudaSuccess) ix->last_scan_ms += ms;
    }
    return B200_EMBED_OK;
}

int32_t b200_index_enable_prescan(b200_index* ix) {
    if (!ix) return fail(B200_EMBED_ERR_INVALID_ARG, "null index");
    std::lock_guard<std::mutex> lk(ix->mu);
    if (ix->vecs16) return B200_EMBED_OK;
    DevGuard guard(ix->device);
    IX_TRY(cudaMalloc(&ix->vecs16, static_cast<size_t>(ix->cap) * ix->dim * sizeof(__nv_bfloat16)));
    const int64_t count = ix->n * ix->dim;
    if (count > 0) {
        to_bf16_rows_kernel<<<static_cast<unsigned>((count + 255) / 256), 256, 0, ix->stream>>>(ix->vecs, ix->vecs16, count);
        IX_TRY(cudaGetLastError());
        IX_TRY(cudaStreamSynchronize(ix->stream));
    }
    return B200_EMBED_OK;
}

int64_t b200_index_exact_fallbacks(const b200_index* ix) { return ix ? ix->exact_fallbacks : -1; }

float b200_index_last_scan_ms(const b200_index* ix) { return ix ? ix->last_scan_ms : -1.f; }

// ---- persistence: the rows as they are in HBM (already normalised), so that a restarted service does not have to embed
// its corpus again (the Qdrant collection the reference writes to is persistent). File: "B200IDX1", dim (i32), n (i64),
// then n * dim f32, little endian.
int32_t b200_index_save(b200_index* ix, const char* path) {
    if (!ix || !path) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(ix->mu);
    DevGuard guard(ix->device);
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(B200_EMBED_ERR_IO, std::string("cannot create ") + path);
    const int32_t dim = ix->dim;
    const int64_t n = ix->n;
    bool ok = std::fwrite("B200IDX1", 1, 8, f) == 8 && std::fwrite(&dim, 4, 1, f) == 1 && std::fwrite(&n, 8, 1, f) == 1;
    const int64_t rows_per_chunk = (int64_t(32) << 20) / (dim * 4);
    std::unique_ptr<float[]> buf(new float[static_cast<size_t>(rows_per_chunk) * dim]);
    int rc = B200_EMBED_OK;
    for (int64_t r = 0; ok && r < n; r += rows_per_chunk) {
        const int64_t m = n - r < rows_per_chunk ? n - r : rows_per_chunk;
        const cudaError_t e = cudaMemcpy(buf.get(), ix->vecs + r * dim, static_cast<size_t>(m) * dim * 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { rc = cuda_fail(e, "cudaMemcpy (index save)"); break; }
        ok = std::fwrite(buf.get(), 4, static_cast<size_t>(m) * dim, f) == static_cast<size_t>(m) * dim;
    }
    if (std::fclose(f) != 0) ok = false;
    if (rc) return rc;
    return ok ? B200_EMBED_OK : fail(B200_EMBED_ERR_IO, std::string("short write to ") + path);
}

int32_t b200_index_load(const char* path, int32_t device, b200_index** out) {
    if (!path || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    *out = nullptr;
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(B200_EMBED_ERR_IO, std::string("cannot open ") + path);
    char magic[8];
    int32_t dim = 0;
    int64_t n = 0;
    if (std::fread(magic, 1, 8, f) != 8 || std::memcmp(magic, "B200IDX1", 8) != 0 || std::fread(&dim, 4, 1, f) != 1 ||
        std::fread(&n, 8, 1, f) != 1 || n < 0) {
        std::fclose(f);
        return fail(B200_EMBED_ERR_IO, std::string(path) + " is not an index file");
    }
    b200_index* ix = nullptr;
    int rc = b200_index_create(dim, n, device, &ix);
    if (rc) { std::fclose(f); return rc; }
    {
        DevGuard guard(device);
        const int64_t rows_per_chunk = (int64_t(32) << 20) / (dim * 4);
        std::unique_ptr<float[]> buf(new float[static_cast<size_t>(rows_per_chunk) * dim]);
        for (int64_t r = 0; r < n; r += rows_per_chunk) {
            const int64_t m = n - r < rows_per_chunk ? n - r : rows_per_chunk;
            if (std::fread(buf.get(), 4, static_cast<size_t>(m) * dim, f) != static_cast<size_t>(m) * dim) {
                rc = fail(B200_EMBED_ERR_IO, std::string(path) + " is truncated");
                break;
            }
            // rows are stored normalised: copied as they are, bit for bit
            const cudaError_t e = cudaMemcpy(ix->vecs + r * dim, buf.get(), static_cast<size_t>(m) * dim * 4, cudaMemcpyHostToDevice);
            if (e != cudaSuccess) { rc = cuda_fail(e, "cudaMemcpy (index load)"); break; }
        }
    }
    std::fclose(f);
    if (rc) { b200_index_destroy(ix); return rc; }
    ix->n = n;
    *out = ix;
    return B200_EMBED_OK;
}

}  // extern "C"
