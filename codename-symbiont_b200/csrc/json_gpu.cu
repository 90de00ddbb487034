// The JSON items of the results formatted ON THE GPU (after the path): the same bytes as
// b200_json_sentence_items_ranges (json_out.cpp) - `{"sentence_text":"..","embedding":[f, f, ..]}` per sentence, what
// serde_json::to_vec writes for shared_models::SentenceEmbedding (/root/reference/libs/shared_models/src/lib.rs:39-43,
// published at /root/reference/services/preprocessing_service/src/main.rs:138-142) - from embeddings that are still in device
// memory. A 384-d sentence is ~4.5 KB of decimal text: on the host the shortest-digit formatting of 384 floats per
// sentence (16 cores: ~1 M sentences/s) caps an 8-GPU ingest at a twentieth of what the GPUs embed; here it is integer
// work spread over the GPU and only the finished text crosses PCIe.
//   fmt_kernel     one thread per float: Schubfach shortest digits (the routine of json_out.cpp, bit for bit: one 64 x 32
//                  multiply per interval end against the shared power-of-ten table) in ryu's layout -> a 16-byte slot + length
//   rowlen_kernel  one warp per sentence: length of its escaped text, of its array text, and so of the whole item
//   scan_kernel    exclusive prefix sum of the item lengths -> byte offset of every item
//   write_kernel   one block per sentence: the head with the sentence text escaped by one warp, then the floats of the row
//                  placed by a block-wide prefix sum, then `]}`
// HBM-bound integer / byte work: 4 B read and ~11.7 B written per float.
#include <cuda_runtime.h>
#include <stdint.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200_embed.h"
#include "pow10_table.h"

namespace {

thread_local std::string g_jg_err;
int fail(int code, const std::string& msg) {
    g_jg_err = msg;
    return code;
}

__constant__ uint64_t d_pow10[B200_POW10_MAX - B200_POW10_MIN + 1] = {B200_POW10_64_VALUES};

__device__ __forceinline__ int floor_log10_pow2(int e) { return (e * 1262611) >> 22; }
__device__ __forceinline__ int floor_log10_three_quarters_pow2(int e) { return (e * 1262611 - 524031) >> 22; }
__device__ __forceinline__ int floor_log2_pow10(int e) { return (e * 1741647) >> 19; }
__device__ __forceinline__ uint32_t round_to_odd(uint64_t g, uint32_t cp) {
    const uint32_t y1 = static_cast<uint32_t>(__umul64hi(g, static_cast<uint64_t>(cp)));  // the product is below 2^96
    const uint32_t y0 = static_cast<uint32_t>((g * static_cast<uint64_t>(cp)) >> 32);
    return y1 | (y0 > 1 ? 1u : 0u);
}

// decimal significand (<= 9 digits) and exponent of a finite v > 0: v = sig * 10^exp (json_out.cpp schubfach_f32)
__device__ uint32_t schubfach_f32(uint32_t bits, int* exp10) {
    const uint32_t frac = bits & 0x7FFFFFu, E = (bits >> 23) & 0xFFu;
    uint32_t c;
    int q;
    if (E != 0) {
        c = frac | 0x800000u;
        q = static_cast<int>(E) - 150;
        if (0 <= -q && -q < 24 && (c & ((1u << -q) - 1)) == 0) {
            *exp10 = 0;
            return c >> -q;
        }
    } else {
        c = frac;
        q = -149;
    }
    const bool even = (c & 1) == 0;
    const bool lower_closer = frac == 0 && E > 1;
    const uint32_t cbl = 4 * c - 2 + (lower_closer ? 1 : 0), cb = 4 * c, cbr = 4 * c + 2;
    const int k = lower_closer ? floor_log10_three_quarters_pow2(q) : floor_log10_pow2(q);
    const int h = q + floor_log2_pow10(-k) + 1;
    const uint64_t g = d_pow10[-k - B200_POW10_MIN];
    const uint32_t vbl = round_to_odd(g, cbl << h), vb = round_to_odd(g, cb << h), vbr = round_to_odd(g, cbr << h);
    const uint32_t lower = vbl + (even ? 0 : 1), upper = vbr - (even ? 0 : 1);
    const uint32_t s = vb / 4;
    if (s >= 10) {
        const uint32_t sp = s / 10;
        const bool up_inside = lower <= 40 * sp, wp_inside = 40 * sp + 40 <= upper;
        if (up_inside != wp_inside) {
            *exp10 = k + 1;
            return sp + (wp_inside ? 1 : 0);
        }
    }
    const bool u_inside = lower <= 4 * s, w_inside = 4 * s + 4 <= upper;
    *exp10 = k;
    if (u_inside != w_inside) return s + (w_inside ? 1 : 0);
    const uint32_t mid = 4 * s + 2;
    const bool round_up = vb > mid || (vb == mid && (s & 1));
    return s + (round_up ? 1 : 0);
}

// ryu "pretty" layout (json_out.cpp format_f32): characters into out[0..16), returns their number (<= 16)
__device__ int format_f32(float v, char* out) {
    uint32_t bits = __float_as_uint(v);
    if (((bits >> 23) & 0xFFu) == 0xFFu) { out[0] = 'n'; out[1] = 'u'; out[2] = 'l'; out[3] = 'l'; return 4; }
    int i = 0;
    if (bits >> 31) { out[i++] = '-'; bits &= 0x7FFFFFFFu; }
    if (bits == 0) { out[i] = '0'; out[i + 1] = '.'; out[i + 2] = '0'; return i + 3; }
    int e10;
    uint32_t sig = schubfach_f32(bits, &e10);
    char d[10];
    int n = sig >= 100000000u ? 9 : sig >= 10000000u ? 8 : sig >= 1000000u ? 7 : sig >= 100000u ? 6 : sig >= 10000u ? 5
            : sig >= 1000u ? 4 : sig >= 100u ? 3 : sig >= 10u ? 2 : 1;
    for (int j = n - 1; j >= 0; --j) { d[j] = static_cast<char>('0' + sig % 10u); sig /= 10u; }
    int kk = e10 + n;  // value = 0.d1..dn x 10^kk
    while (n > 1 && d[n - 1] == '0') --n;
    if (-6 < kk && kk <= 0) {  // 0.001234
        out[i++] = '0'; out[i++] = '.';
        for (int j = 0; j < -kk; ++j) out[i++] = '0';
        for (int j = 0; j < n; ++j) out[i++] = d[j];
        return i;
    }
    const int k = kk - n;
    if (0 <= k && kk <= 13) {  // 12340000000.0
        for (int j = 0; j < n; ++j) out[i++] = d[j];
        for (int j = n; j < kk; ++j) out[i++] = '0';
        out[i++] = '.'; out[i++] = '0';
        return i;
    }
    if (0 < kk && kk <= 13) {  // 12.34
        for (int j = 0; j < kk; ++j) out[i++] = d[j];
        out[i++] = '.';
        for (int j = kk; j < n; ++j) out[i++] = d[j];
        return i;
    }
    out[i++] = d[0];  // scientific: d[.ddd]e[-]X
    if (n > 1) {
        out[i++] = '.';
        for (int j = 1; j < n; ++j) out[i++] = d[j];
    }
    out[i++] = 'e';
    int e = kk - 1;
    if (e < 0) { out[i++] = '-'; e = -e; }
    if (e >= 10) out[i++] = static_cast<char>('0' + e / 10);
    out[i++] = static_cast<char>('0' + e % 10);
    return i;
}

__global__ void fmt_kernel(const float* __restrict__ v, int64_t count, uint4* __restrict__ slots, uint8_t* __restrict__ lens) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= count) return;
    union { char c[16]; uint4 q; } u;
    u.q = make_uint4(0, 0, 0, 0);
    const int n = format_f32(__ldg(v + i), u.c);
    slots[i] = u.q;
    lens[i] = static_cast<uint8_t>(n);
}

// serde_json's escapes: \" \\ \b \f \n \r \t take 2 characters, the other control characters 6 (\u00XX), all else 1
__device__ __forceinline__ int esc_size(unsigned char c) {
    if (c == '"' || c == '\\') return 2;
    if (c >= 0x20) return 1;
    return (c == '\b' || c == '\f' || c == '\n' || c == '\r' || c == '\t') ? 2 : 6;
}
constexpr int HEAD_PRE = 18, HEAD_POST = 15;  // {"sentence_text":"   and   ","embedding":[

// item i = [,] head_i + floats + "]}": length of the escaped sentence text, of the array text (values and the commas between
// them) and so of the whole item
__global__ void rowlen_kernel(const uint8_t* __restrict__ lens, const char* __restrict__ text, const int64_t* __restrict__ begin,
                              const int64_t* __restrict__ end, int64_t text_base, int n, int dim, int first,
                              int32_t* __restrict__ head_len, int64_t* __restrict__ item_len) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= n) return;
    const uint8_t* p = lens + static_cast<int64_t>(row) * dim;
    int s = 0;
    for (int j = lane; j < dim; j += 32) s += p[j];
    const char* t = text + (begin[row] - text_base);
    const int tl = static_cast<int>(end[row] - begin[row]);
    int e = 0;
    for (int j = lane; j < tl; j += 32) e += esc_size(static_cast<unsigned char>(t[j]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        e += __shfl_xor_sync(0xffffffffu, e, o);
    }
    if (lane == 0) {
        const int hl = ((row > 0 || !first) ? 1 : 0) + HEAD_PRE + e + HEAD_POST;
        head_len[row] = hl;
        item_len[row] = static_cast<int64_t>(hl) + s + (dim > 0 ? dim - 1 : 0) + 2;
    }
}

// exclusive prefix sum of item_len[0, n) into item_off[0, n]; one block
__global__ void scan_kernel(const int64_t* __restrict__ item_len, int n, int64_t* __restrict__ item_off) {
    __shared__ int64_t part[1024];
    const int t = threadIdx.x, per = (n + blockDim.x - 1) / blockDim.x;
    const int lo = min(t * per, n), hi = min(lo + per, n);
    int64_t s = 0;
    for (int i = lo; i < hi; ++i) s += item_len[i];
    part[t] = s;
    __syncthreads();
    if (t == 0) {
        int64_t run = 0;
        for (int i = 0; i < static_cast<int>(blockDim.x); ++i) { const int64_t x = part[i]; part[i] = run; run += x; }
        item_off[n] = run;
    }
    __syncthreads();
    int64_t run = part[t];
    for (int i = lo; i < hi; ++i) { item_off[i] = run; run += item_len[i]; }
}

constexpr int WR_THREADS = 128;
__global__ void __launch_bounds__(WR_THREADS)
write_kernel(const uint4* __restrict__ slots, const uint8_t* __restrict__ lens, const char* __restrict__ text,
             const int64_t* __restrict__ begin, const int64_t* __restrict__ end, int64_t text_base, int first,
             const int32_t* __restrict__ head_len, const int64_t* __restrict__ item_off, int dim, char* __restrict__ out) {
    const int row = blockIdx.x, t = threadIdx.x;
    char* o = out + item_off[row];
    const int hl = head_len[row];
    if (t < 32) {  // warp 0: [,]{"sentence_text":"<escaped text>","embedding":[
        const char pre[HEAD_PRE + 1] = "{\"sentence_text\":\"";
        const char post[HEAD_POST + 1] = "\",\"embedding\":[";
        int pos = 0;
        if (row > 0 || !first) { if (t == 0) o[0] = ','; pos = 1; }
        if (t < HEAD_PRE) o[pos + t] = pre[t];
        pos += HEAD_PRE;
        const char* src = text + (begin[row] - text_base);
        const int tl = static_cast<int>(end[row] - begin[row]);
        for (int c0 = 0; c0 < tl; c0 += 32) {
            const int j = c0 + t;
            const unsigned char c = j < tl ? static_cast<unsigned char>(src[j]) : 0;
            const int sz = j < tl ? esc_size(c) : 0;
            int incl = sz;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, incl, d);
                if (t >= d) incl += y;
            }
            char* w = o + pos + incl - sz;
            if (sz == 1) w[0] = static_cast<char>(c);
            else if (sz == 2) {
                w[0] = '\\';
                w[1] = c == '"' ? '"' : c == '\\' ? '\\' : c == '\b' ? 'b' : c == '\f' ? 'f' : c == '\n' ? 'n' : c == '\r' ? 'r' : 't';
            } else if (sz == 6) {
                const char* hex = "0123456789abcdef";
                w[0] = '\\'; w[1] = 'u'; w[2] = '0'; w[3] = '0'; w[4] = hex[c >> 4]; w[5] = hex[c & 15];
            }
            pos += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (t < HEAD_POST) o[pos + t] = post[t];
    }
    o += hl;
    __shared__ int warp_sum[WR_THREADS / 32];
    __shared__ int base;
    if (t == 0) base = 0;
    __syncthreads();
    for (int j0 = 0; j0 < dim; j0 += WR_THREADS) {
        const int j = j0 + t;
        union { char c[16]; uint4 q; } u;
        int len = 0;
        if (j < dim) {
            const int64_t idx = static_cast<int64_t>(row) * dim + j;
            u.q = slots[idx];
            len = lens[idx] + (j + 1 < dim ? 1 : 0);  // the value and the comma after it
        }
        // block-wide exclusive scan of len
        int incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, d);
            if ((t & 31) >= d) incl += y;
        }
        if ((t & 31) == 31) warp_sum[t >> 5] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < (t >> 5); ++w) wbase += warp_sum[w];
        const int pos = base + wbase + incl - len;
        if (j < dim) {
            const int n = lens[static_cast<int64_t>(row) * dim + j];
            for (int k = 0; k < n; ++k) o[pos + k] = u.c[k];
            if (j + 1 < dim) o[pos + n] = ',';
        }
        __syncthreads();
        if (t == WR_THREADS - 1) base = pos + len;
        __syncthreads();
    }
    if (t == 0) { o[base] = ']'; o[base + 1] = '}'; }
}

}  // namespace

struct b200_jsongpu {
    int device = 0;
    cudaStream_t stream = nullptr;
    // device buffers, grown on demand
    uint4* slots = nullptr;
    uint8_t* lens = nullptr;
    size_t float_cap = 0;
    char* d_heads = nullptr;  // the span of the caller's text that holds the sentences
    size_t heads_cap = 0;
    int64_t *d_head_off = nullptr /* begin | end */, *d_item_len = nullptr, *d_item_off = nullptr;
    int32_t* d_head_len = nullptr;
    size_t row_cap = 0;
    char* d_out = nullptr;
    size_t out_cap = 0;
    float* d_emb = nullptr;  // staging of host embeddings (b200_jsongpu_sentence_items_host)
    size_t emb_cap = 0;
    int64_t* h_total = nullptr;  // pinned
    std::recursive_mutex mu;  // one call at a time per handle
};

namespace {

#define JG_TRY(expr)                                                                                    \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return fail(e__ == cudaErrorMemoryAllocation ? B200_EMBED_ERR_OOM : B200_EMBED_ERR_CUDA,    \
                        std::string(#expr) + ": " + cudaGetErrorString(e__));                           \
    } while (0)

struct DevGuard {
    int prev = -1;
    explicit DevGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        cudaSetDevice(dev);
    }
    ~DevGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

template <typename T>
int grow_dev(T** p, size_t* cap, size_t want) {
    if (want <= *cap) return B200_EMBED_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const size_t n = want + want / 4;
    JG_TRY(cudaMalloc(p, n * sizeof(T)));
    *cap = n;
    return B200_EMBED_OK;
}
}  // namespace

extern "C" {

const char* b200_jsongpu_last_error(void) { return g_jg_err.c_str(); }

int32_t b200_jsongpu_create(int32_t device, b200_jsongpu** out) {
    if (!out) return fail(B200_EMBED_ERR_INVALID_ARG, "null out pointer");
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(B200_EMBED_ERR_NO_DEVICE, "no CUDA device: the GPU formatter has no CPU fallback (use b200_json_sentence_items_ranges)");
    if (device < 0 || device >= count) return fail(B200_EMBED_ERR_INVALID_ARG, "bad device ordinal");
    DevGuard guard(device);
    b200_jsongpu* g = new (std::nothrow) b200_jsongpu();
    if (!g) return fail(B200_EMBED_ERR_OOM, "out of memory");
    g->device = device;
    if (cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaHostAlloc(&g->h_total, sizeof(int64_t), cudaHostAllocDefault) != cudaSuccess) {
        b200_jsongpu_destroy(g);
        return fail(B200_EMBED_ERR_CUDA, "stream / pinned word");
    }
    *out = g;
    return B200_EMBED_OK;
}

void b200_jsongpu_destroy(b200_jsongpu* g) {
    if (!g) return;
    DevGuard guard(g->device);
    if (g->stream) cudaStreamSynchronize(g->stream);
    for (void* p : {static_cast<void*>(g->slots), static_cast<void*>(g->lens), static_cast<void*>(g->d_heads),
                    static_cast<void*>(g->d_head_off), static_cast<void*>(g->d_item_len), static_cast<void*>(g->d_item_off),
                    static_cast<void*>(g->d_head_len), static_cast<void*>(g->d_out), static_cast<void*>(g->d_emb)})
        if (p) cudaFree(p);
    if (g->h_total) cudaFreeHost(g->h_total);
    if (g->stream) cudaStreamDestroy(g->stream);
    delete g;
}

int64_t b200_jsongpu_sentence_items(b200_jsongpu* g, const char* text, const int64_t* begin, const int64_t* end,
                                    const float* device_embeddings, int32_t n, int32_t dim, int32_t first, char* out,
                                    int64_t out_cap) {
    if (!g || n < 0 || dim <= 0 || (n > 0 && (!text || !begin || !end || !device_embeddings)))
        return fail(B200_EMBED_ERR_INVALID_ARG, "jsongpu_sentence_items: bad argument");
    if (n == 0) return 0;
    std::lock_guard<std::recursive_mutex> lk(g->mu);
    DevGuard guard(g->device);
    static const bool trace = getenv("B200_JSONGPU_TRACE") != nullptr;  // stage times on stderr
    const auto t_begin = std::chrono::steady_clock::now();
    auto ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    try {
        // the sentence texts go to the device as they are: one span of `text` covering every range, plus the ranges
        int64_t lo = begin[0], hi = end[0];
        for (int32_t i = 0; i < n; ++i) {
            if (begin[i] < 0 || end[i] < begin[i]) return fail(B200_EMBED_ERR_INVALID_ARG, "jsongpu_sentence_items: bad sentence range");
            lo = begin[i] < lo ? begin[i] : lo;
            hi = end[i] > hi ? end[i] : hi;
        }
        const size_t span = static_cast<size_t>(hi - lo);
        int rc = B200_EMBED_OK;
        const int64_t count = static_cast<int64_t>(n) * dim;
        const double t_heads = ms();
        // device buffers
        if (static_cast<size_t>(count) > g->float_cap) {  // slots and lengths grow together
            if (g->slots) cudaFree(g->slots);
            if (g->lens) cudaFree(g->lens);
            g->slots = nullptr; g->lens = nullptr; g->float_cap = 0;
            const size_t want = static_cast<size_t>(count) + static_cast<size_t>(count) / 4;
            JG_TRY(cudaMalloc(&g->slots, want * sizeof(uint4)));
            JG_TRY(cudaMalloc(&g->lens, want));
            g->float_cap = want;
        }
        if ((rc = grow_dev(&g->d_heads, &g->heads_cap, span + 1))) return rc;
        if (static_cast<size_t>(n) + 1 > g->row_cap) {
            for (void* p : {static_cast<void*>(g->d_head_off), static_cast<void*>(g->d_item_len), static_cast<void*>(g->d_item_off),
                            static_cast<void*>(g->d_head_len)})
                if (p) cudaFree(p);
            g->d_head_off = g->d_item_len = g->d_item_off = nullptr;
            g->d_head_len = nullptr;
            g->row_cap = 0;
            const size_t want = static_cast<size_t>(n) + static_cast<size_t>(n) / 4 + 2;
            JG_TRY(cudaMalloc(&g->d_head_off, want * 16));  // begin | end
            JG_TRY(cudaMalloc(&g->d_item_len, want * 8));
            JG_TRY(cudaMalloc(&g->d_item_off, want * 8));
            JG_TRY(cudaMalloc(&g->d_head_len, want * 4));
            g->row_cap = want;
        }
        cudaStream_t s = g->stream;
        int64_t* d_begin = g->d_head_off;
        int64_t* d_end = g->d_head_off + g->row_cap;
        if (span) JG_TRY(cudaMemcpyAsync(g->d_heads, text + lo, span, cudaMemcpyHostToDevice, s));
        JG_TRY(cudaMemcpyAsync(d_begin, begin, static_cast<size_t>(n) * 8, cudaMemcpyHostToDevice, s));
        JG_TRY(cudaMemcpyAsync(d_end, end, static_cast<size_t>(n) * 8, cudaMemcpyHostToDevice, s));
        fmt_kernel<<<static_cast<unsigned>((count + 255) / 256), 256, 0, s>>>(device_embeddings, count, g->slots, g->lens);
        rowlen_kernel<<<static_cast<unsigned>((static_cast<int64_t>(n) * 32 + 255) / 256), 256, 0, s>>>(
            g->lens, g->d_heads, d_begin, d_end, lo, n, dim, first, g->d_head_len, g->d_item_len);
        scan_kernel<<<1, 1024, 0, s>>>(g->d_item_len, n, g->d_item_off);
        JG_TRY(cudaGetLastError());
        JG_TRY(cudaMemcpyAsync(g->h_total, g->d_item_off + n, 8, cudaMemcpyDeviceToHost, s));
        JG_TRY(cudaStreamSynchronize(s));
        const int64_t total = *g->h_total;
        const double t_sized = ms();
        if (!out || out_cap < total) return total;  // size query, like the host entry
        if ((rc = grow_dev(&g->d_out, &g->out_cap, static_cast<size_t>(total)))) return rc;
        write_kernel<<<static_cast<unsigned>(n), WR_THREADS, 0, s>>>(g->slots, g->lens, g->d_heads, d_begin, d_end, lo, first,
                                                                     g->d_head_len, g->d_item_off, dim, g->d_out);
        JG_TRY(cudaGetLastError());
        JG_TRY(cudaGetLastError());
        if (trace) cudaStreamSynchronize(s);
        const double t_written = ms();
        JG_TRY(cudaMemcpyAsync(out, g->d_out, static_cast<size_t>(total), cudaMemcpyDeviceToHost, s));
        JG_TRY(cudaStreamSynchronize(s));
        if (trace)
            std::fprintf(stderr, "[jsongpu] set-up %.2f ms, format + lengths + scan %.2f, write %.2f, D2H of %.1f MB %.2f\n", t_heads,
                         t_sized - t_heads, t_written - t_sized, total / 1e6, ms() - t_written);
        return total;
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("jsongpu_sentence_items: ") + e.what());
    }
}

int64_t b200_jsongpu_sentence_items_host(b200_jsongpu* g, const char* text, const int64_t* begin, const int64_t* end,
                                         const float* host_embeddings, int32_t n, int32_t dim, int32_t first, char* out,
                                         int64_t out_cap) {
    if (!g || n < 0 || dim <= 0 || (n > 0 && !host_embeddings)) return fail(B200_EMBED_ERR_INVALID_ARG, "jsongpu_sentence_items_host: bad argument");
    if (n == 0) return 0;
    std::lock_guard<std::recursive_mutex> lk(g->mu);
    {
        DevGuard guard(g->device);
        const size_t count = static_cast<size_t>(n) * dim;
        const int rc = grow_dev(&g->d_emb, &g->emb_cap, count);
        if (rc) return rc;
        JG_TRY(cudaMemcpyAsync(g->d_emb, host_embeddings, count * sizeof(float), cudaMemcpyHostToDevice, g->stream));
    }
    return b200_jsongpu_sentence_items(g, text, begin, end, g->d_emb, n, dim, first, out, out_cap);
}

}  // extern "C"
