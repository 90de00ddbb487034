// libb200_embed.so: model state, workspaces, the forward schedule and the C ABI (include/b200_embed.h).
//
// This file is the B200-side counterpart of the reference's `EmbeddingGenerator`
// (/root/reference/services/preprocessing_service/src/embedding_generator.rs): `new` (:17-132) maps to
// b200_embed_create, the body of the chunk loop in `generate_sentence_embeddings` (:189-214) maps to
// b200_embed_forward*. The schedule per call (packed tokens, T = sum of real lengths):
//   mask -> lengths -> cu_seqlens | K1 embed+LN | per layer: K2 QKV GEMM, K3 attention,
//   K2 attn-out GEMM (+residual+LN), K2 FFN1 GEMM (+GELU), K2 FFN2 GEMM (+residual+LN) | K5 pool.
// Every grid is sized from the host-known upper bound batch*seq_len and every kernel reads the
// real token count from device memory (cu_seqlens[batch]), so a call never synchronises the host
// between its kernels. There is no CPU fallback anywhere in this file.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200_embed.h"
#include "gemm_tc.h"
#include "kernels.h"
#include "weights_io.h"

using bf16 = __nv_bfloat16;  // 16-bit operand storage: bf16 or fp16 bits, whichever format the handle computes in

namespace {

// The 16-bit kernels of one operand format (fmt.h): every handle points at one of the two sets.
struct Ops {
    cudaError_t (*gemm_tc_launch)(const b200::GemmTcArgs&, cudaStream_t, int);
    int (*pick_block_n)(int, int);
    cudaError_t (*mask_lengths)(const uint32_t*, int, int, int*, int*, int*, cudaStream_t);
    cudaError_t (*pack_tiles)(const int*, int, int, int*, int*, int*, cudaStream_t);
    cudaError_t (*embed_ln)(const b200::EmbedArgs&, cudaStream_t);
    cudaError_t (*add_ln_f32)(const float*, const bf16*, const uint8_t*, int, const int*, const float*, const float*, float, int,
                              bf16*, uint8_t*, cudaStream_t);
    cudaError_t (*add_ln_16)(const bf16*, const bf16*, const uint8_t*, int, const int*, const float*, const float*, float, int,
                             bf16*, uint8_t*, cudaStream_t);
    cudaError_t (*mean_pool)(const bf16*, const uint8_t*, const int*, int, int, int, float*, cudaStream_t);
    cudaError_t (*attention_tc)(const bf16*, const int*, int, int, int, int, int, bf16*, int, cudaStream_t);
    bool (*qkv_attention_supported)(int, int, int);
    cudaError_t (*qkv_attention)(const bf16*, const bf16*, const float*, const int*, const int*, const int*, int, int, int, int,
                                 int, bf16*, int, cudaStream_t);
    bool (*ffn_fused_supported)(int, int, int);
    cudaError_t (*ffn_fused_launch)(const b200::FfnFusedArgs&, cudaStream_t, int);
    cudaError_t (*convert_weights)(const float*, bf16*, size_t, cudaStream_t);
};
#define B200_OPS(NS, CONVERT)                                                                                          \
    Ops {                                                                                                              \
        &b200::NS::gemm_tc_launch, &b200::NS::gemm_tc_pick_block_n, &b200::NS::launch_mask_lengths,                    \
            &b200::NS::launch_pack_tiles, &b200::NS::launch_embed_layernorm<bf16>,                                     \
            &b200::NS::launch_add_layernorm<float, bf16>, &b200::NS::launch_add_layernorm<bf16, bf16>,                 \
            &b200::NS::launch_mean_pool<bf16>, &b200::NS::launch_attention_tc, &b200::NS::qkv_attention_supported,     \
            &b200::NS::launch_qkv_attention, &b200::NS::ffn_fused_supported, &b200::NS::ffn_fused_launch, CONVERT      \
    }
const Ops kOpsBf16 = B200_OPS(op_bf16, &b200::launch_f32_to_bf16);
const Ops kOpsF16 = B200_OPS(op_f16, &b200::launch_f32_to_f16);
#undef B200_OPS
namespace f32ops = b200::op_bf16;  // the float instantiations are the same code in both sets

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    const int code = (e == cudaErrorMemoryAllocation) ? B200_EMBED_ERR_OOM
                     : (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? B200_EMBED_ERR_NO_DEVICE
                                                                                   : B200_EMBED_ERR_CUDA;
    return fail(code, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU_TRY(expr)                                         \
    do {                                                     \
        cudaError_t e__ = (expr);                            \
        if (e__ != cudaSuccess) return cuda_fail(e__, #expr); \
    } while (0)

constexpr int MAX_CONTEXTS = 4;
constexpr size_t ALIGN_ELEMS = 128;  // sub-allocation granularity inside the weight arenas

struct LayerW {
    // fp32 master (device)
    float *qkv_w, *qkv_b, *ao_w, *ao_b, *ao_g, *ao_be, *i_w, *i_b, *o_w, *o_b, *o_g, *o_be;
    // bf16 copies of the matrices (device)
    bf16 *qkv_w16, *ao_w16, *i_w16, *o_w16;
};

struct Workspace {
    size_t cap = 0;  // rows (tokens)
    void *x = nullptr, *qkv = nullptr, *ctx = nullptr, *a = nullptr, *inter = nullptr;
    void *x_lo = nullptr, *a_lo = nullptr;  // 16-bit mode: low halves of the residual stream (x = hi + lo), one e5m2 byte per element
    float* y32 = nullptr;  // pre-LayerNorm fp32 rows when the LN is not fused into the GEMM
    void release() {
        for (void* p : {x, qkv, ctx, a, inter, x_lo, a_lo, static_cast<void*>(y32)})
            if (p) cudaFree(p);
        x = qkv = ctx = a = inter = x_lo = a_lo = nullptr;
        y32 = nullptr;
        cap = 0;
    }
};

// A small call is launch-bound (about twenty kernels of 5-15 us): its kernel sequence is captured once per
// (batch, seq_len, flags) and replayed as one CUDA graph. Every pointer in it belongs to the context, so a graph stays
// valid until one of the context's buffers is reallocated (g_alloc_gen moves).
struct GraphEntry {
    uint64_t key = 0;
    cudaGraphExec_t exec = nullptr;  // nullptr: the shape has been seen once (and run the plain way), not captured yet
    int launches = 0;
};
std::atomic<uint64_t> g_alloc_gen{1};
thread_local int64_t t_launches = 0;  // kernels this thread has enqueued (what a capture recorded = its difference)

struct ProfRecord {
    int kind;
    cudaEvent_t start, stop;
};

struct Context {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    Workspace ws16, ws32;
    uint32_t *d_ids = nullptr, *d_mask = nullptr, *d_tt = nullptr;
    size_t in_cap = 0;
    int *d_lens = nullptr, *d_cu = nullptr;
    int *d_tile_first = nullptr, *d_seg_off = nullptr, *d_n_tiles = nullptr;  // K23 work list (launch_pack_tiles)
    size_t b_cap = 0;
    float* d_out = nullptr;
    size_t out_cap = 0;
    int* d_status = nullptr;
    int* h_status = nullptr;  // pinned
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    std::vector<ProfRecord> prof;
    std::vector<GraphEntry> graphs;
    uint64_t graph_gen = 0;
    bool busy = false;
    bool created = false;
};

}  // namespace

struct b200_embed {
    b200_embed_config cfg{};
    int device = 0;
    int num_sms = 148;
    int H = 0, I = 0, L = 0, heads = 0, D = 0;
    size_t cap_tokens = 0;
    bool fuse_ln = false;
    int bn_qkv = 0, bn_ffn1 = 0, bn_h = 0, bn_ln = 0;
    int act_epi = b200::EPI_GELU_ERF;
    bool gemm_pair = true;   // plain bf16 GEMMs as CTA pairs (one cta_group::2 MMA per pair); B200_EMBED_GEMM_PAIR=0 for A/B
    bool ffn_fused = false;  // K2f replaces the FFN1 / FFN2 GEMM pair: on whenever it supports the shape (hidden 128 / 256 /
                             // 384, erf-GELU); B200_EMBED_FFN_FUSED=0 for A/B
    // The fused kernels parallelise over sentences (K23) / pairs of 128-row blocks (K2f): with too few of those for the
    // SMs the unfused kernels, which also split the output columns, finish a call sooner (measured on MiniLM, one
    // B200: K2f wins from 64 x 128 tokens, K23 from 256 sentences; a single query takes 0.39 ms unfused, 0.55 fused).
    // 0 = auto by call size, 1 = always when the shape is supported (B200_EMBED_FFN_FUSED / _QKV_ATTN = 1), used by tests.
    bool force_fused_ffn = false, force_qkv_attn = false;
    bool qkv_attn = true;    // K23 replaces the QKV GEMM + attention pair whenever it supports the call (hidden 128 / 256 / 384,
                             // head_dim 32, seq_len <= 128); B200_EMBED_QKV_ATTN=0 for A/B
    std::atomic<bool> use_graphs{true};  // small host-buffer calls replay a captured CUDA graph; B200_EMBED_GRAPHS=0 for
                                         // A/B; cleared for good if a capture ever fails (any calling thread may do that)
    const Ops* ops = &kOpsF16;  // kernels of the handle's operand format (cfg.operand_format)
    bool f16 = true;
    // weights
    float* arena32 = nullptr;
    bf16* arena16 = nullptr;
    float *word = nullptr, *pos = nullptr, *type = nullptr, *emb_g = nullptr, *emb_b = nullptr;
    std::vector<LayerW> layers;
    // contexts
    std::mutex mu;
    std::condition_variable cv;
    Context ctxs[MAX_CONTEXTS];
    // tickets of b200_embed_submit: ticket id -> context index
    int64_t next_ticket = 1;
    int64_t ticket_of[MAX_CONTEXTS] = {0, 0, 0, 0};
    // accounting
    std::atomic<int64_t> launches{0};
    std::atomic<int> profile{0};
    std::mutex stats_mu;
    b200_embed_kind_stats stats[B200_EMBED_NUM_KINDS] = {};
};

namespace {

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int init_context(b200_embed* h, Context& c) {
    if (c.created) return B200_EMBED_OK;
    CU_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    CU_TRY(cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming));
    CU_TRY(cudaMalloc(&c.d_status, sizeof(int)));
    CU_TRY(cudaMemset(c.d_status, 0, sizeof(int)));
    CU_TRY(cudaHostAlloc(&c.h_status, sizeof(int), cudaHostAllocDefault));
    *c.h_status = 0;
    c.created = true;
    (void)h;
    return B200_EMBED_OK;
}

void destroy_context(Context& c) {
    if (!c.created) return;
    cudaStreamSynchronize(c.stream);
    for (GraphEntry& g : c.graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    c.ws16.release();
    c.ws32.release();
    for (void* p : {static_cast<void*>(c.d_ids), static_cast<void*>(c.d_mask), static_cast<void*>(c.d_tt),
                    static_cast<void*>(c.d_lens), static_cast<void*>(c.d_cu), static_cast<void*>(c.d_out),
                    static_cast<void*>(c.d_status), static_cast<void*>(c.d_tile_first), static_cast<void*>(c.d_seg_off),
                    static_cast<void*>(c.d_n_tiles)})
        if (p) cudaFree(p);
    if (c.h_status) cudaFreeHost(c.h_status);
    for (cudaEvent_t e : c.ev_pool) cudaEventDestroy(e);
    cudaEventDestroy(c.done);
    cudaStreamDestroy(c.stream);
    c = Context();
}

// Blocks until a context is free (blocking = true) or returns nullptr.
Context* lease(b200_embed* h, bool blocking, int* index = nullptr) {
    std::unique_lock<std::mutex> lk(h->mu);
    for (;;) {
        for (int i = 0; i < MAX_CONTEXTS; ++i)
            if (!h->ctxs[i].busy) {
                h->ctxs[i].busy = true;
                if (index) *index = i;
                return &h->ctxs[i];
            }
        if (!blocking) return nullptr;
        h->cv.wait(lk);
    }
}
void release(b200_embed* h, Context* c) {
    {
        std::lock_guard<std::mutex> lk(h->mu);
        c->busy = false;
    }
    h->cv.notify_one();
}

int ensure_workspace(b200_embed* h, Workspace& w, size_t rows, size_t elt, cudaStream_t s) {
    rows = round_up(rows, 128);
    if (rows <= w.cap) return B200_EMBED_OK;
    g_alloc_gen.fetch_add(1);
    w.release();
    const size_t H = h->H, I = h->I;
    CU_TRY(cudaMalloc(&w.x, rows * H * elt));
    CU_TRY(cudaMalloc(&w.qkv, rows * 3 * H * elt));
    CU_TRY(cudaMalloc(&w.ctx, rows * H * elt));
    CU_TRY(cudaMalloc(&w.a, rows * H * elt));
    CU_TRY(cudaMalloc(&w.inter, rows * I * elt));
    if (elt == 2) {
        CU_TRY(cudaMalloc(&w.x_lo, rows * H));
        CU_TRY(cudaMalloc(&w.a_lo, rows * H));
    }
    if (elt == 4 || !h->fuse_ln) CU_TRY(cudaMalloc(&w.y32, rows * H * sizeof(float)));
    // Rows past the real token count are read (never used) by 128-row tiles; keep them finite. The memsets
    // go on the stream that will use the buffers: the contexts' streams do not synchronise with stream 0.
    CU_TRY(cudaMemsetAsync(w.x, 0, rows * H * elt, s));
    CU_TRY(cudaMemsetAsync(w.qkv, 0, rows * 3 * H * elt, s));
    CU_TRY(cudaMemsetAsync(w.ctx, 0, rows * H * elt, s));
    CU_TRY(cudaMemsetAsync(w.a, 0, rows * H * elt, s));
    CU_TRY(cudaMemsetAsync(w.inter, 0, rows * I * elt, s));
    if (w.x_lo) CU_TRY(cudaMemsetAsync(w.x_lo, 0, rows * H, s));
    if (w.a_lo) CU_TRY(cudaMemsetAsync(w.a_lo, 0, rows * H, s));
    w.cap = rows;
    return B200_EMBED_OK;
}

template <typename T>
int ensure_buf(T** p, size_t* cap, size_t n) {
    if (n <= *cap) return B200_EMBED_OK;
    g_alloc_gen.fetch_add(1);
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const size_t want = n + n / 4;
    CU_TRY(cudaMalloc(p, want * sizeof(T)));
    *cap = want;
    return B200_EMBED_OK;
}

int ensure_input_bufs(Context& c, size_t n) {
    if (n <= c.in_cap) return B200_EMBED_OK;
    g_alloc_gen.fetch_add(1);
    for (uint32_t** p : {&c.d_ids, &c.d_mask, &c.d_tt}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    c.in_cap = 0;
    const size_t want = n + n / 4;
    CU_TRY(cudaMalloc(&c.d_ids, want * 4));
    CU_TRY(cudaMalloc(&c.d_mask, want * 4));
    CU_TRY(cudaMalloc(&c.d_tt, want * 4));
    c.in_cap = want;
    return B200_EMBED_OK;
}

int ensure_batch_bufs(Context& c, size_t nb) {
    if (nb + 2 <= c.b_cap) return B200_EMBED_OK;
    g_alloc_gen.fetch_add(1);
    for (int** p : {&c.d_lens, &c.d_cu, &c.d_tile_first, &c.d_seg_off, &c.d_n_tiles}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    c.b_cap = 0;
    const size_t want = nb + nb / 4 + 2;
    CU_TRY(cudaMalloc(&c.d_lens, want * 4));
    CU_TRY(cudaMalloc(&c.d_cu, want * 4));
    CU_TRY(cudaMalloc(&c.d_tile_first, want * 4));
    CU_TRY(cudaMalloc(&c.d_seg_off, (want / 32 + 2) * 4));
    CU_TRY(cudaMalloc(&c.d_n_tiles, 4));
    c.b_cap = want;
    return B200_EMBED_OK;
}

// ---- profiling ------------------------------------------------------------------------------
cudaEvent_t next_event(Context& c) {
    if (c.ev_used == c.ev_pool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
        c.ev_pool.push_back(e);
    }
    return c.ev_pool[c.ev_used++];
}

template <typename F>
cudaError_t timed(b200_embed* h, Context& c, cudaStream_t s, int kind, int nlaunch, F&& f) {
    h->launches.fetch_add(nlaunch, std::memory_order_relaxed);
    t_launches += nlaunch;
    if (!h->profile.load(std::memory_order_relaxed)) return f();
    cudaEvent_t e0 = next_event(c), e1 = next_event(c);
    if (!e0 || !e1) return f();
    cudaEventRecord(e0, s);
    const cudaError_t r = f();
    cudaEventRecord(e1, s);
    c.prof.push_back({kind, e0, e1});
    return r;
}

void collect_profile(b200_embed* h, Context& c) {
    if (c.prof.empty()) { c.ev_used = 0; return; }
    std::lock_guard<std::mutex> lk(h->stats_mu);
    for (const ProfRecord& r : c.prof) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.start, r.stop) == cudaSuccess) {
            h->stats[r.kind].total_ms += ms;
            h->stats[r.kind].launches += 1;
        }
    }
    c.prof.clear();
    c.ev_used = 0;
}

// ---- the schedule ---------------------------------------------------------------------------
struct Job {
    const uint32_t *d_ids, *d_mask, *d_tt;  // device; padded [batch, seq] or packed
    const int* d_cu;                        // packed input: device cu_seqlens; padded: nullptr
    int batch, seq_len, tokens_upper;
    uint32_t flags;
    float* d_out;  // device [batch, H]
};

int gemm16(b200_embed* h, Context& c, cudaStream_t s, int kind, const void* x, const bf16* w, const float* bias,
           int rows_cap, const int* d_rows, int n, int k, int epi, int block_n, const void* residual,
           const void* residual_lo, const float* gamma, const float* beta, void* out, void* out_lo, int out_f32) {
    b200::GemmTcArgs g{};
    g.a = x; g.b = w; g.m = rows_cap; g.n = n; g.k = k; g.d_rows = d_rows; g.bias = bias; g.residual = residual;
    g.residual_lo = residual_lo; g.out_lo = out_lo;
    g.gamma = gamma; g.beta = beta; g.eps = h->cfg.layer_norm_eps; g.out = out; g.out_f32 = out_f32; g.epi = epi;
    g.block_n = block_n;
    g.pair = (h->gemm_pair && epi != b200::EPI_RES_LN && !out_f32) ? 1 : 0;
    CU_TRY(timed(h, c, s, kind, 1, [&] { return h->ops->gemm_tc_launch(g, s, h->num_sms); }));
    return B200_EMBED_OK;
}

int run_chunk(b200_embed* h, Context& c, cudaStream_t s, const Job& j) {
    const bool fp32 = (j.flags & B200_EMBED_FLAG_FP32) != 0;
    const int H = h->H, I = h->I, rows = j.tokens_upper;
    const float eps = h->cfg.layer_norm_eps;
    Workspace& w = fp32 ? c.ws32 : c.ws16;
    int rc = ensure_workspace(h, w, rows, fp32 ? 4 : 2, s);
    if (rc) return rc;
    if ((rc = ensure_batch_bufs(c, static_cast<size_t>(j.batch)))) return rc;
    const int* cu = j.d_cu;
    if (!cu) {
        CU_TRY(timed(h, c, s, B200_EMBED_KIND_LENGTHS, 2, [&] {
            return h->ops->mask_lengths(j.d_mask, j.batch, j.seq_len, c.d_lens, c.d_cu, c.d_status, s);
        }));
        cu = c.d_cu;
    }
    const int* d_rows = cu + j.batch;
    // K23 (fused Q/K/V projection + attention) works on tiles of whole sentences: list them once per call
    const bool use_k23 = !fp32 && h->qkv_attn && h->ops->qkv_attention_supported(H, h->heads, j.seq_len) &&
                         (h->force_qkv_attn || j.batch >= 256);
    if (use_k23)
        CU_TRY(timed(h, c, s, B200_EMBED_KIND_LENGTHS, 1, [&] {
            return h->ops->pack_tiles(cu, j.batch, 128, c.d_seg_off, c.d_tile_first, c.d_n_tiles, s);
        }));

    b200::EmbedArgs e{};
    e.ids = j.d_ids; e.type_ids = j.d_tt; e.cu = cu; e.batch = j.batch; e.seq_len = j.seq_len;
    e.total_tokens = j.tokens_upper; e.padded = j.d_cu ? 0 : 1; e.gamma = h->emb_g; e.beta = h->emb_b; e.eps = eps;
    e.H = H; e.vocab = h->cfg.vocab_size; e.type_vocab = h->cfg.type_vocab_size;
    e.max_pos = h->cfg.max_position_embeddings; e.out = w.x; e.out_lo = fp32 ? nullptr : w.x_lo; e.status = c.d_status;
    e.word = h->word; e.pos = h->pos; e.type = h->type;
    CU_TRY(timed(h, c, s, B200_EMBED_KIND_EMBED_LN, 1, [&] {
        return fp32 ? f32ops::launch_embed_layernorm<float>(e, s) : h->ops->embed_ln(e, s);
    }));

    for (int l = 0; l < h->L; ++l) {
        const LayerW& lw = h->layers[l];
        if (fp32) {
            float *x = static_cast<float*>(w.x), *qkv = static_cast<float*>(w.qkv), *ctx = static_cast<float*>(w.ctx),
                  *a = static_cast<float*>(w.a), *inter = static_cast<float*>(w.inter);
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_GEMM_QKV, 1, [&] {
                return b200::launch_gemm_f32(x, lw.qkv_w, lw.qkv_b, rows, d_rows, 3 * H, H, -1, qkv, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_ATTENTION, 1, [&] {
                return b200::launch_attention_f32(qkv, cu, j.batch, j.seq_len, h->heads, h->D, ctx, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_GEMM_AO, 1, [&] {
                return b200::launch_gemm_f32(ctx, lw.ao_w, lw.ao_b, rows, d_rows, H, H, -1, w.y32, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_ADD_LN, 1, [&] {
                return f32ops::launch_add_layernorm<float, float>(w.y32, x, nullptr, rows, d_rows, lw.ao_g, lw.ao_be, eps, H, a, nullptr, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_GEMM_FFN1, 1, [&] {
                return b200::launch_gemm_f32(a, lw.i_w, lw.i_b, rows, d_rows, I, H, h->cfg.hidden_act, inter, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_GEMM_FFN2, 1, [&] {
                return b200::launch_gemm_f32(inter, lw.o_w, lw.o_b, rows, d_rows, H, I, -1, w.y32, s);
            }));
            CU_TRY(timed(h, c, s, B200_EMBED_KIND_ADD_LN, 1, [&] {
                return f32ops::launch_add_layernorm<float, float>(w.y32, a, nullptr, rows, d_rows, lw.o_g, lw.o_be, eps, H, x, nullptr, s);
            }));
        } else {
            bf16 *x = static_cast<bf16*>(w.x), *qkv = static_cast<bf16*>(w.qkv), *ctx = static_cast<bf16*>(w.ctx),
                 *a = static_cast<bf16*>(w.a), *inter = static_cast<bf16*>(w.inter);
            uint8_t *x_lo = static_cast<uint8_t*>(w.x_lo), *a_lo = static_cast<uint8_t*>(w.a_lo);
            if (use_k23) {
                CU_TRY(timed(h, c, s, B200_EMBED_KIND_QKV_ATTN, 1, [&] {
                    return h->ops->qkv_attention(x, lw.qkv_w16, lw.qkv_b, cu, c.d_tile_first, c.d_n_tiles, rows, j.batch,
                                                      j.seq_len, H, h->heads, ctx, h->num_sms, s);
                }));
            } else {
                rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_QKV, x, lw.qkv_w16, lw.qkv_b, rows, d_rows, 3 * H, H,
                            b200::EPI_BIAS, h->bn_qkv, nullptr, nullptr, nullptr, nullptr, qkv, nullptr, 0);
                if (rc) return rc;
                CU_TRY(timed(h, c, s, B200_EMBED_KIND_ATTENTION, 1, [&] {
                    return h->ops->attention_tc(qkv, cu, rows, j.batch, j.seq_len, h->heads, h->D, ctx, h->num_sms, s);
                }));
            }
            if (h->fuse_ln) {
                rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_AO, ctx, lw.ao_w16, lw.ao_b, rows, d_rows, H, H,
                            b200::EPI_RES_LN, h->bn_ln, x, x_lo, lw.ao_g, lw.ao_be, a, a_lo, 0);
                if (rc) return rc;
            } else {
                rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_AO, ctx, lw.ao_w16, lw.ao_b, rows, d_rows, H, H,
                            b200::EPI_BIAS, h->bn_h, nullptr, nullptr, nullptr, nullptr, w.y32, nullptr, 1);
                if (rc) return rc;
                CU_TRY(timed(h, c, s, B200_EMBED_KIND_ADD_LN, 1, [&] {
                    return h->ops->add_ln_f32(w.y32, x, x_lo, rows, d_rows, lw.ao_g, lw.ao_be, eps, H, a, a_lo, s);
                }));
            }
            if (h->ffn_fused && (h->force_fused_ffn || rows >= 64 * 128)) {
                b200::FfnFusedArgs f{};
                f.a = a; f.a_lo = a_lo; f.w1 = lw.i_w16; f.w2 = lw.o_w16; f.b1 = lw.i_b; f.b2 = lw.o_b; f.gamma = lw.o_g;
                f.beta = lw.o_be; f.eps = eps; f.m = rows; f.hidden = H; f.inter = I; f.d_rows = d_rows; f.out = x;
                f.out_lo = x_lo;
                CU_TRY(timed(h, c, s, B200_EMBED_KIND_FFN_FUSED, 1, [&] { return h->ops->ffn_fused_launch(f, s, h->num_sms); }));
                continue;
            }
            rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_FFN1, a, lw.i_w16, lw.i_b, rows, d_rows, I, H, h->act_epi,
                        h->bn_ffn1, nullptr, nullptr, nullptr, nullptr, inter, nullptr, 0);
            if (rc) return rc;
            if (h->fuse_ln) {
                rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_FFN2, inter, lw.o_w16, lw.o_b, rows, d_rows, H, I,
                            b200::EPI_RES_LN, h->bn_ln, a, a_lo, lw.o_g, lw.o_be, x, x_lo, 0);
                if (rc) return rc;
            } else {
                rc = gemm16(h, c, s, B200_EMBED_KIND_GEMM_FFN2, inter, lw.o_w16, lw.o_b, rows, d_rows, H, I,
                            b200::EPI_BIAS, h->bn_h, nullptr, nullptr, nullptr, nullptr, w.y32, nullptr, 1);
                if (rc) return rc;
                CU_TRY(timed(h, c, s, B200_EMBED_KIND_ADD_LN, 1, [&] {
                    return h->ops->add_ln_f32(w.y32, a, a_lo, rows, d_rows, lw.o_g, lw.o_be, eps, H, x, x_lo, s);
                }));
            }
        }
    }
    const int normalize = (j.flags & B200_EMBED_FLAG_NORMALIZE) ? 1 : 0;
    CU_TRY(timed(h, c, s, B200_EMBED_KIND_POOL, 1, [&] {
        return fp32 ? f32ops::launch_mean_pool<float>(static_cast<const float*>(w.x), nullptr, cu, j.batch, H, normalize, j.d_out, s)
                    : h->ops->mean_pool(static_cast<const bf16*>(w.x), static_cast<const uint8_t*>(w.x_lo), cu, j.batch, H,
                                        normalize, j.d_out, s);
    }));
    return B200_EMBED_OK;
}

void drop_graphs(Context& c) {
    for (GraphEntry& g : c.graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    c.graphs.clear();
}

// run_chunk, replayed from a CUDA graph when the call is small enough to be launch-bound and has been seen before.
int run_chunk_cached(b200_embed* h, Context& c, cudaStream_t s, const Job& j) {
    constexpr int GRAPH_MAX_TOKENS = 16384;
    constexpr size_t GRAPH_MAX_ENTRIES = 64;
    if (!h->use_graphs || h->profile.load(std::memory_order_relaxed) || j.d_cu != nullptr || j.tokens_upper > GRAPH_MAX_TOKENS)
        return run_chunk(h, c, s, j);
    if (c.graph_gen != g_alloc_gen.load()) {
        drop_graphs(c);
        c.graph_gen = g_alloc_gen.load();
    }
    const uint64_t key = static_cast<uint64_t>(j.batch) | (static_cast<uint64_t>(j.seq_len) << 24) | (static_cast<uint64_t>(j.flags) << 48);
    GraphEntry* e = nullptr;
    for (GraphEntry& g : c.graphs)
        if (g.key == key) { e = &g; break; }
    if (!e) {  // first sight: run the plain way (this also makes every allocation and one-time attribute call)
        if (c.graphs.size() >= GRAPH_MAX_ENTRIES) drop_graphs(c);
        const int rc = run_chunk(h, c, s, j);
        if (!rc && c.graph_gen == g_alloc_gen.load()) c.graphs.push_back(GraphEntry{key, nullptr, 0});
        else c.graph_gen = 0;  // something was reallocated: start over next time
        return rc;
    }
    if (!e->exec) {
        const int64_t before = t_launches;
        if (cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed) != cudaSuccess) { cudaGetLastError(); return run_chunk(h, c, s, j); }
        const int rc = run_chunk(h, c, s, j);
        cudaGraph_t graph = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(s, &graph);
        const int captured = static_cast<int>(t_launches - before);
        h->launches.fetch_sub(captured);  // nothing has run yet
        cudaGraphExec_t exec = nullptr;
        const bool ok = !rc && ce == cudaSuccess && graph && c.graph_gen == g_alloc_gen.load() &&
                        cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
        if (!ok) {  // fall back to plain launches for good
            cudaGetLastError();
            if (exec) cudaGraphExecDestroy(exec);
            h->use_graphs = false;
            return rc ? rc : run_chunk(h, c, s, j);
        }
        e->exec = exec;
        e->launches = captured;
    }
    h->launches.fetch_add(e->launches, std::memory_order_relaxed);
    CU_TRY(cudaGraphLaunch(e->exec, s));
    return B200_EMBED_OK;
}

int check_status(Context& c) {
    const int st = *c.h_status;
    if (st & 1) return fail(B200_EMBED_ERR_INVALID_ARG, "attention_mask is not a right-padded 0/1 prefix");
    if (st & 2) return fail(B200_EMBED_ERR_INVALID_ARG, "input id, token type id or position out of range");
    return B200_EMBED_OK;
}

int validate_common(const b200_embed* h, const void* a, const void* b, const void* c, int batch, int seq_len,
                    const float* out) {
    if (!h) return fail(B200_EMBED_ERR_INVALID_ARG, "null handle");
    if (batch < 0 || seq_len < 0) return fail(B200_EMBED_ERR_INVALID_ARG, "negative batch or seq_len");
    if (batch == 0) return B200_EMBED_OK;
    if (!a || !b || !c || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null buffer");
    if (seq_len == 0) return fail(B200_EMBED_ERR_INVALID_ARG, "seq_len is 0");
    if (seq_len > h->cfg.max_position_embeddings)
        return fail(B200_EMBED_ERR_INVALID_ARG, "seq_len " + std::to_string(seq_len) + " exceeds max_position_embeddings " +
                                                    std::to_string(h->cfg.max_position_embeddings));
    return B200_EMBED_OK;
}

// Enqueues a padded host-buffer request on context c (H2D, kernels, D2H, status copy); no sync.
int enqueue_padded_host(b200_embed* h, Context& c, const uint32_t* ids, const uint32_t* mask, const uint32_t* tt,
                        int batch, int seq_len, uint32_t flags, float* out) {
    cudaStream_t s = c.stream;
    CU_TRY(cudaStreamWaitEvent(s, c.done, 0));
    CU_TRY(cudaMemsetAsync(c.d_status, 0, sizeof(int), s));
    size_t per = h->cap_tokens / static_cast<size_t>(seq_len);
    if (per < 1) per = 1;
    for (int b0 = 0; b0 < batch; b0 += static_cast<int>(per)) {
        const int nb = static_cast<int>(std::min<size_t>(per, static_cast<size_t>(batch - b0)));
        const size_t n = static_cast<size_t>(nb) * seq_len, off = static_cast<size_t>(b0) * seq_len;
        int rc = ensure_input_bufs(c, n);
        if (!rc) rc = ensure_buf(&c.d_out, &c.out_cap, static_cast<size_t>(nb) * h->H);
        if (!rc) rc = ensure_batch_bufs(c, static_cast<size_t>(nb));
        if (rc) return rc;
        CU_TRY(cudaMemcpyAsync(c.d_ids, ids + off, n * 4, cudaMemcpyHostToDevice, s));
        CU_TRY(cudaMemcpyAsync(c.d_mask, mask + off, n * 4, cudaMemcpyHostToDevice, s));
        CU_TRY(cudaMemcpyAsync(c.d_tt, tt + off, n * 4, cudaMemcpyHostToDevice, s));
        Job j{c.d_ids, c.d_mask, c.d_tt, nullptr, nb, seq_len, static_cast<int>(n), flags, c.d_out};
        rc = run_chunk_cached(h, c, s, j);
        if (rc) return rc;
        CU_TRY(cudaMemcpyAsync(out + static_cast<size_t>(b0) * h->H, c.d_out, static_cast<size_t>(nb) * h->H * 4,
                               cudaMemcpyDeviceToHost, s));
    }
    CU_TRY(cudaMemcpyAsync(c.h_status, c.d_status, sizeof(int), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(c.done, s));
    return B200_EMBED_OK;
}

int finish(b200_embed* h, Context& c) {
    const cudaError_t e = cudaStreamSynchronize(c.stream);
    collect_profile(h, c);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize");
    return check_status(c);
}

int upload(float* dst, const std::vector<float>& src) {
    CU_TRY(cudaMemcpy(dst, src.data(), src.size() * 4, cudaMemcpyHostToDevice));
    return B200_EMBED_OK;
}

std::string layer_name(int l, const char* suffix) { return "encoder.layer." + std::to_string(l) + "." + suffix; }

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char* b200_embed_last_error(void) { return g_err.c_str(); }

int32_t b200_embed_create(const b200_embed_config* cfg, const char* path, uint64_t seed, int32_t device,
                          b200_embed** out) {
    if (!cfg || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null config or out pointer");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(B200_EMBED_ERR_NO_DEVICE, std::string("no CUDA device (there is no CPU fallback): ") +
                                                  (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0 || device >= ndev) return fail(B200_EMBED_ERR_NO_DEVICE, "device ordinal out of range");
    const int H = cfg->hidden_size, I = cfg->intermediate_size, L = cfg->num_hidden_layers,
              heads = cfg->num_attention_heads;
    if (H <= 0 || heads <= 0 || L <= 0 || I <= 0 || cfg->vocab_size <= 0 || cfg->max_position_embeddings <= 0 ||
        cfg->type_vocab_size <= 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "non-positive model dimension");
    if (H % 128 != 0 || H > 1024) return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "hidden_size must be a multiple of 128, <= 1024");
    if (H % heads != 0 || (H / heads != 32 && H / heads != 64))
        return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "head dim must be 32 or 64");
    if (I % 128 != 0) return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "intermediate_size must be a multiple of 128");
    if (cfg->hidden_act < 0 || cfg->hidden_act > 2) return fail(B200_EMBED_ERR_INVALID_ARG, "unknown hidden_act");
    if (cfg->operand_format < 0 || cfg->operand_format > 2) return fail(B200_EMBED_ERR_INVALID_ARG, "unknown operand_format");

    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(B200_EMBED_ERR_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major) +
                                                  std::to_string(prop.minor) + "; this library is built for sm_100a only");
    DeviceGuard guard(device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");

    std::unique_ptr<b200_embed> h(new b200_embed());
    h->cfg = *cfg;
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->H = H; h->I = I; h->L = L; h->heads = heads; h->D = H / heads;
    h->cap_tokens = cfg->max_tokens_per_launch > 0 ? static_cast<size_t>(cfg->max_tokens_per_launch) : (1u << 20);
    if (h->cap_tokens < static_cast<size_t>(cfg->max_position_embeddings)) h->cap_tokens = cfg->max_position_embeddings;
    h->f16 = cfg->operand_format != B200_EMBED_OPERANDS_BF16;
    h->ops = h->f16 ? &kOpsF16 : &kOpsBf16;
    h->bn_ln = h->ops->pick_block_n(H, b200::EPI_RES_LN);
    h->fuse_ln = h->bn_ln != 0;
    if (const char* v = getenv("B200_EMBED_FUSE_LN")) h->fuse_ln = h->fuse_ln && std::string(v) != "0";
    h->act_epi = b200::EPI_GELU_ERF + cfg->hidden_act;
    h->ffn_fused = h->fuse_ln && h->ops->ffn_fused_supported(H, I, cfg->hidden_act);
    if (const char* v = getenv("B200_EMBED_FFN_FUSED")) {
        h->ffn_fused = h->ffn_fused && std::string(v) != "0";
        h->force_fused_ffn = std::string(v) == "1";
    }
    if (const char* v = getenv("B200_EMBED_QKV_ATTN")) {
        h->qkv_attn = std::string(v) != "0";
        h->force_qkv_attn = std::string(v) == "1";
    }
    if (const char* v = getenv("B200_EMBED_GEMM_PAIR")) h->gemm_pair = std::string(v) != "0";
    if (const char* v = getenv("B200_EMBED_GRAPHS")) h->use_graphs = std::string(v) != "0";
    h->bn_qkv = h->ops->pick_block_n(3 * H, b200::EPI_BIAS);
    h->bn_ffn1 = h->ops->pick_block_n(I, h->act_epi);
    h->bn_h = h->ops->pick_block_n(H, b200::EPI_GELU_ERF);  // 256 or 128 (fp32-out instantiations)
    if (!h->bn_qkv || !h->bn_ffn1 || !h->bn_h) return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "no GEMM tiling for this shape");

    // ---- weight arenas ----
    const size_t V = cfg->vocab_size, P = cfg->max_position_embeddings, T = cfg->type_vocab_size;
    auto al = [](size_t n) { return round_up(n, ALIGN_ELEMS); };
    const size_t per_layer32 = al(3 * (size_t)H * H) + al(3 * (size_t)H) + al((size_t)H * H) + 3 * al(H) + al((size_t)I * H) + al(I) +
                               al((size_t)H * I) + 3 * al(H);
    const size_t total32 = al(V * H) + al(P * H) + al(T * H) + 2 * al(H) + L * per_layer32;
    const size_t per_layer16 = al(3 * (size_t)H * H) + al((size_t)H * H) + al((size_t)I * H) + al((size_t)H * I);
    const size_t total16 = L * per_layer16;
    CU_TRY(cudaMalloc(&h->arena32, total32 * sizeof(float)));
    e = cudaMalloc(&h->arena16, total16 * sizeof(bf16));
    if (e != cudaSuccess) { cudaFree(h->arena32); return cuda_fail(e, "cudaMalloc(bf16 weights)"); }
    struct Cleanup {
        b200_embed* h; bool armed = true;
        ~Cleanup() { if (armed) { cudaFree(h->arena32); cudaFree(h->arena16); } }
    } cleanup{h.get()};

    float* p32 = h->arena32;
    bf16* p16 = h->arena16;
    auto take32 = [&](size_t n) { float* r = p32; p32 += al(n); return r; };
    auto take16 = [&](size_t n) { bf16* r = p16; p16 += al(n); return r; };

    b200::SafetensorsFile file;
    std::string err;
    const bool from_file = path != nullptr && path[0] != '\0';
    if (from_file && !file.open(path, &err)) return fail(B200_EMBED_ERR_IO, err);
    std::vector<float> host;
    // loads tensor `name` (rows x cols) into dst (device)
    auto load = [&](const std::string& name, std::vector<int64_t> shape, float* dst) -> int {
        size_t n = 1;
        for (int64_t d : shape) n *= static_cast<size_t>(d);
        host.resize(n);
        if (from_file) {
            if (!file.read_f32(name, shape, host.data(), &err)) return fail(B200_EMBED_ERR_IO, err);
        } else {
            b200::random_init_tensor(name, seed, host.data(), n);
        }
        return upload(dst, host);
    };
    int rc;
    h->word = take32(V * H); h->pos = take32(P * H); h->type = take32(T * H); h->emb_g = take32(H); h->emb_b = take32(H);
    if ((rc = load("embeddings.word_embeddings.weight", {(int64_t)V, H}, h->word))) return rc;
    if ((rc = load("embeddings.position_embeddings.weight", {(int64_t)P, H}, h->pos))) return rc;
    if ((rc = load("embeddings.token_type_embeddings.weight", {(int64_t)T, H}, h->type))) return rc;
    if ((rc = load("embeddings.LayerNorm.weight", {H}, h->emb_g))) return rc;
    if ((rc = load("embeddings.LayerNorm.bias", {H}, h->emb_b))) return rc;
    h->layers.resize(L);
    for (int l = 0; l < L; ++l) {
        LayerW& w = h->layers[l];
        w.qkv_w = take32(3 * (size_t)H * H); w.qkv_b = take32(3 * (size_t)H);
        w.ao_w = take32((size_t)H * H); w.ao_b = take32(H); w.ao_g = take32(H); w.ao_be = take32(H);
        w.i_w = take32((size_t)I * H); w.i_b = take32(I);
        w.o_w = take32((size_t)H * I); w.o_b = take32(H); w.o_g = take32(H); w.o_be = take32(H);
        w.qkv_w16 = take16(3 * (size_t)H * H); w.ao_w16 = take16((size_t)H * H);
        w.i_w16 = take16((size_t)I * H); w.o_w16 = take16((size_t)H * I);
        const char* qkv_names[3] = {"query", "key", "value"};
        for (int t = 0; t < 3; ++t) {  // Q, K, V stacked into one [3H, H] matrix / [3H] bias
            const std::string base = std::string("attention.self.") + qkv_names[t];
            if ((rc = load(layer_name(l, (base + ".weight").c_str()), {H, H}, w.qkv_w + (size_t)t * H * H))) return rc;
            if ((rc = load(layer_name(l, (base + ".bias").c_str()), {H}, w.qkv_b + (size_t)t * H))) return rc;
        }
        if ((rc = load(layer_name(l, "attention.output.dense.weight"), {H, H}, w.ao_w))) return rc;
        if ((rc = load(layer_name(l, "attention.output.dense.bias"), {H}, w.ao_b))) return rc;
        if ((rc = load(layer_name(l, "attention.output.LayerNorm.weight"), {H}, w.ao_g))) return rc;
        if ((rc = load(layer_name(l, "attention.output.LayerNorm.bias"), {H}, w.ao_be))) return rc;
        if ((rc = load(layer_name(l, "intermediate.dense.weight"), {I, H}, w.i_w))) return rc;
        if ((rc = load(layer_name(l, "intermediate.dense.bias"), {I}, w.i_b))) return rc;
        if ((rc = load(layer_name(l, "output.dense.weight"), {H, I}, w.o_w))) return rc;
        if ((rc = load(layer_name(l, "output.dense.bias"), {H}, w.o_b))) return rc;
        if ((rc = load(layer_name(l, "output.LayerNorm.weight"), {H}, w.o_g))) return rc;
        if ((rc = load(layer_name(l, "output.LayerNorm.bias"), {H}, w.o_be))) return rc;
    }
    // 16-bit copies in the handle's operand format (round to nearest even of the fp32 master)
    for (int l = 0; l < L; ++l) {
        LayerW& w = h->layers[l];
        CU_TRY(h->ops->convert_weights(w.qkv_w, w.qkv_w16, 3 * (size_t)H * H, nullptr));
        CU_TRY(h->ops->convert_weights(w.ao_w, w.ao_w16, (size_t)H * H, nullptr));
        CU_TRY(h->ops->convert_weights(w.i_w, w.i_w16, (size_t)I * H, nullptr));
        CU_TRY(h->ops->convert_weights(w.o_w, w.o_w16, (size_t)H * I, nullptr));
    }
    CU_TRY(cudaDeviceSynchronize());
    for (int i = 0; i < MAX_CONTEXTS; ++i)
        if ((rc = init_context(h.get(), h->ctxs[i]))) {
            for (int k = 0; k < MAX_CONTEXTS; ++k) destroy_context(h->ctxs[k]);
            return rc;
        }
    cleanup.armed = false;
    *out = h.release();
    return B200_EMBED_OK;
}

void b200_embed_destroy(b200_embed* h) {
    if (!h) return;
    DeviceGuard guard(h->device);
    for (int i = 0; i < MAX_CONTEXTS; ++i) destroy_context(h->ctxs[i]);
    cudaFree(h->arena32);
    cudaFree(h->arena16);
    delete h;
}

int32_t b200_embed_hidden_size(const b200_embed* h) { return h ? h->H : 0; }
int32_t b200_embed_max_positions(const b200_embed* h) { return h ? h->cfg.max_position_embeddings : 0; }
int64_t b200_embed_kernel_launches(const b200_embed* h) { return h ? h->launches.load() : 0; }

int32_t b200_embed_forward(b200_embed* h, const uint32_t* ids, const uint32_t* mask, const uint32_t* tt,
                           int32_t batch, int32_t seq_len, uint32_t flags, float* out) {
    int rc = validate_common(h, ids, mask, tt, batch, seq_len, out);
    if (rc || batch == 0) return rc;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    Context* c = lease(h, true);
    rc = enqueue_padded_host(h, *c, ids, mask, tt, batch, seq_len, flags, out);
    const int rc2 = finish(h, *c);
    release(h, c);
    return rc ? rc : rc2;
}

int32_t b200_embed_submit(b200_embed* h, const uint32_t* ids, const uint32_t* mask, const uint32_t* tt,
                          int32_t batch, int32_t seq_len, uint32_t flags, float* out, int64_t* ticket) {
    if (!ticket) return fail(B200_EMBED_ERR_INVALID_ARG, "null ticket pointer");
    *ticket = 0;
    int rc = validate_common(h, ids, mask, tt, batch, seq_len, out);
    if (rc) return rc;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    int idx = -1;
    Context* c = lease(h, false, &idx);
    if (!c) return fail(B200_EMBED_ERR_INVALID_ARG, "too many requests in flight (max 4): call b200_embed_wait first");
    if (batch > 0) {
        rc = enqueue_padded_host(h, *c, ids, mask, tt, batch, seq_len, flags, out);
        if (rc) {
            cudaStreamSynchronize(c->stream);
            collect_profile(h, *c);
            release(h, c);
            return rc;
        }
    } else {
        *c->h_status = 0;
    }
    {
        std::lock_guard<std::mutex> lk(h->mu);
        *ticket = h->next_ticket++;
        h->ticket_of[idx] = *ticket;
    }
    return B200_EMBED_OK;
}

int32_t b200_embed_wait(b200_embed* h, int64_t ticket) {
    if (!h || ticket <= 0) return fail(B200_EMBED_ERR_INVALID_ARG, "bad handle or ticket");
    int idx = -1;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        for (int i = 0; i < MAX_CONTEXTS; ++i)
            if (h->ticket_of[i] == ticket) { idx = i; h->ticket_of[i] = 0; }
    }
    if (idx < 0) return fail(B200_EMBED_ERR_INVALID_ARG, "unknown ticket");
    DeviceGuard guard(h->device);
    Context* c = &h->ctxs[idx];
    const int rc = finish(h, *c);
    release(h, c);
    return rc;
}

// Validates a packed request on the host; B200_EMBED_OK with batch == 0 means "nothing to do".
static int validate_varlen(const b200_embed* h, const uint32_t* ids, const uint32_t* tt, const int32_t* cu, int32_t batch,
                           int32_t total, int32_t max_seqlen, const float* out) {
    if (!h) return fail(B200_EMBED_ERR_INVALID_ARG, "null handle");
    if (batch < 0 || total < 0) return fail(B200_EMBED_ERR_INVALID_ARG, "negative batch or total_tokens");
    if (batch == 0) return B200_EMBED_OK;
    if (!ids || !tt || !cu || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null buffer");
    if (cu[0] != 0 || cu[batch] != total) return fail(B200_EMBED_ERR_INVALID_ARG, "cu_seqlens must start at 0 and end at total_tokens");
    if (max_seqlen > h->cfg.max_position_embeddings) return fail(B200_EMBED_ERR_INVALID_ARG, "max_seqlen exceeds max_position_embeddings");
    for (int b = 0; b < batch; ++b) {
        const int len = cu[b + 1] - cu[b];
        if (len < 0 || len > max_seqlen) return fail(B200_EMBED_ERR_INVALID_ARG, "cu_seqlens not monotone or a sentence longer than max_seqlen");
    }
    return B200_EMBED_OK;
}

// Enqueues a packed host-buffer request on context c (H2D, kernels, D2H, status copy); no sync.
static int enqueue_packed_host(b200_embed* h, Context* c, const uint32_t* ids, const uint32_t* tt, const int32_t* cu,
                               int32_t batch, int32_t max_seqlen, uint32_t flags, float* out) {
    cudaStream_t s = c->stream;
    CU_TRY(cudaStreamWaitEvent(s, c->done, 0));
    CU_TRY(cudaMemsetAsync(c->d_status, 0, sizeof(int), s));
    std::vector<int32_t> local;
    int b0 = 0;
    while (b0 < batch) {
        int b1 = b0 + 1;  // at least one sentence per chunk
        while (b1 < batch && static_cast<size_t>(cu[b1 + 1] - cu[b0]) <= h->cap_tokens) ++b1;
        const int nb = b1 - b0, ntok = cu[b1] - cu[b0];
        local.resize(nb + 1);
        for (int i = 0; i <= nb; ++i) local[i] = cu[b0 + i] - cu[b0];
        int r = ensure_input_bufs(*c, (size_t)ntok + 1);
        if (!r) r = ensure_buf(&c->d_out, &c->out_cap, (size_t)nb * h->H);
        if (!r) r = ensure_batch_bufs(*c, (size_t)nb);
        if (r) return r;
        if (ntok > 0) {
            CU_TRY(cudaMemcpyAsync(c->d_ids, ids + cu[b0], (size_t)ntok * 4, cudaMemcpyHostToDevice, s));
            CU_TRY(cudaMemcpyAsync(c->d_tt, tt + cu[b0], (size_t)ntok * 4, cudaMemcpyHostToDevice, s));
        }
        // `local` is pageable: this copy is staged before the call returns, so reuse is safe
        CU_TRY(cudaMemcpyAsync(c->d_cu, local.data(), (size_t)(nb + 1) * 4, cudaMemcpyHostToDevice, s));
        Job j{c->d_ids, nullptr, c->d_tt, c->d_cu, nb, max_seqlen, ntok, flags, c->d_out};
        if ((r = run_chunk(h, *c, s, j))) return r;
        CU_TRY(cudaMemcpyAsync(out + (size_t)b0 * h->H, c->d_out, (size_t)nb * h->H * 4, cudaMemcpyDeviceToHost, s));
        b0 = b1;
    }
    CU_TRY(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(int), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaEventRecord(c->done, s));
    return B200_EMBED_OK;
}

int32_t b200_embed_forward_varlen(b200_embed* h, const uint32_t* ids, const uint32_t* tt, const int32_t* cu,
                                  int32_t batch, int32_t total, int32_t max_seqlen, uint32_t flags, float* out) {
    int rc = validate_varlen(h, ids, tt, cu, batch, total, max_seqlen, out);
    if (rc || batch == 0) return rc;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    Context* c = lease(h, true);
    rc = enqueue_packed_host(h, c, ids, tt, cu, batch, max_seqlen, flags, out);
    const int rc2 = finish(h, *c);
    release(h, c);
    return rc ? rc : rc2;
}

int32_t b200_embed_submit_varlen(b200_embed* h, const uint32_t* ids, const uint32_t* tt, const int32_t* cu, int32_t batch,
                                 int32_t total, int32_t max_seqlen, uint32_t flags, float* out, int64_t* ticket) {
    if (!ticket) return fail(B200_EMBED_ERR_INVALID_ARG, "null ticket pointer");
    *ticket = 0;
    int rc = validate_varlen(h, ids, tt, cu, batch, total, max_seqlen, out);
    if (rc) return rc;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    int idx = -1;
    Context* c = lease(h, false, &idx);
    if (!c) return fail(B200_EMBED_ERR_INVALID_ARG, "too many requests in flight (max 4): call b200_embed_wait first");
    if (batch > 0) {
        rc = enqueue_packed_host(h, c, ids, tt, cu, batch, max_seqlen, flags, out);
        if (rc) {
            cudaStreamSynchronize(c->stream);
            collect_profile(h, *c);
            release(h, c);
            return rc;
        }
    } else {
        *c->h_status = 0;
    }
    {
        std::lock_guard<std::mutex> lk(h->mu);
        *ticket = h->next_ticket++;
        h->ticket_of[idx] = *ticket;
    }
    return B200_EMBED_OK;
}

int32_t b200_embed_forward_device(b200_embed* h, const uint32_t* d_ids, const uint32_t* d_mask, const uint32_t* d_tt,
                                  int32_t batch, int32_t seq_len, uint32_t flags, float* d_out, void* stream) {
    int rc = validate_common(h, d_ids, d_mask, d_tt, batch, seq_len, d_out);
    if (rc || batch == 0) return rc;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    Context* c = lease(h, true);
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c->stream;
    auto body = [&]() -> int {
        CU_TRY(cudaStreamWaitEvent(s, c->done, 0));
        if (!stream) CU_TRY(cudaMemsetAsync(c->d_status, 0, sizeof(int), s));
        size_t per = h->cap_tokens / static_cast<size_t>(seq_len);
        if (per < 1) per = 1;
        for (int b0 = 0; b0 < batch; b0 += static_cast<int>(per)) {
            const int nb = static_cast<int>(std::min<size_t>(per, static_cast<size_t>(batch - b0)));
            const size_t off = static_cast<size_t>(b0) * seq_len;
            int r = ensure_batch_bufs(*c, (size_t)nb);
            if (r) return r;
            Job j{d_ids + off, d_mask + off, d_tt + off, nullptr, nb, seq_len, static_cast<int>(nb * (size_t)seq_len),
                  flags, d_out + static_cast<size_t>(b0) * h->H};
            if ((r = run_chunk(h, *c, s, j))) return r;
        }
        if (!stream) CU_TRY(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(int), cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaEventRecord(c->done, s));
        return B200_EMBED_OK;
    };
    rc = body();
    int rc2 = B200_EMBED_OK;
    if (!stream) rc2 = finish(h, *c);
    else if (h->profile.load()) { cudaStreamSynchronize(s); collect_profile(h, *c); }
    release(h, c);
    return rc ? rc : rc2;
}

int32_t b200_embed_host_alloc(size_t bytes, void** out) {
    if (!out) return fail(B200_EMBED_ERR_INVALID_ARG, "null out pointer");
    *out = nullptr;
    if (bytes == 0) return B200_EMBED_OK;
    const cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) return cuda_fail(e, "cudaHostAlloc");
    return B200_EMBED_OK;
}
void b200_embed_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int32_t b200_embed_export_safetensors(const b200_embed* h, const char* path) {
    if (!h || !path) return fail(B200_EMBED_ERR_INVALID_ARG, "null handle or path");
    DeviceGuard guard(h->device);
    const int H = h->H, I = h->I;
    std::vector<std::vector<float>> keep;
    std::vector<b200::NamedTensor> list;
    auto add = [&](const std::string& name, std::vector<int64_t> shape, const float* dev) -> int {
        size_t n = 1;
        for (int64_t d : shape) n *= static_cast<size_t>(d);
        keep.emplace_back(n);
        CU_TRY(cudaMemcpy(keep.back().data(), dev, n * 4, cudaMemcpyDeviceToHost));
        list.push_back({name, shape, nullptr});
        return B200_EMBED_OK;
    };
    int rc;
    if ((rc = add("embeddings.word_embeddings.weight", {h->cfg.vocab_size, H}, h->word))) return rc;
    if ((rc = add("embeddings.position_embeddings.weight", {h->cfg.max_position_embeddings, H}, h->pos))) return rc;
    if ((rc = add("embeddings.token_type_embeddings.weight", {h->cfg.type_vocab_size, H}, h->type))) return rc;
    if ((rc = add("embeddings.LayerNorm.weight", {H}, h->emb_g))) return rc;
    if ((rc = add("embeddings.LayerNorm.bias", {H}, h->emb_b))) return rc;
    for (int l = 0; l < h->L; ++l) {
        const LayerW& w = h->layers[l];
        const char* qkv_names[3] = {"query", "key", "value"};
        for (int t = 0; t < 3; ++t) {
            const std::string base = std::string("attention.self.") + qkv_names[t];
            if ((rc = add(layer_name(l, (base + ".weight").c_str()), {H, H}, w.qkv_w + (size_t)t * H * H))) return rc;
            if ((rc = add(layer_name(l, (base + ".bias").c_str()), {H}, w.qkv_b + (size_t)t * H))) return rc;
        }
        if ((rc = add(layer_name(l, "attention.output.dense.weight"), {H, H}, w.ao_w))) return rc;
        if ((rc = add(layer_name(l, "attention.output.dense.bias"), {H}, w.ao_b))) return rc;
        if ((rc = add(layer_name(l, "attention.output.LayerNorm.weight"), {H}, w.ao_g))) return rc;
        if ((rc = add(layer_name(l, "attention.output.LayerNorm.bias"), {H}, w.ao_be))) return rc;
        if ((rc = add(layer_name(l, "intermediate.dense.weight"), {I, H}, w.i_w))) return rc;
        if ((rc = add(layer_name(l, "intermediate.dense.bias"), {I}, w.i_b))) return rc;
        if ((rc = add(layer_name(l, "output.dense.weight"), {H, I}, w.o_w))) return rc;
        if ((rc = add(layer_name(l, "output.dense.bias"), {H}, w.o_b))) return rc;
        if ((rc = add(layer_name(l, "output.LayerNorm.weight"), {H}, w.o_g))) return rc;
        if ((rc = add(layer_name(l, "output.LayerNorm.bias"), {H}, w.o_be))) return rc;
    }
    for (size_t i = 0; i < list.size(); ++i) list[i].data = keep[i].data();
    std::string err;
    if (!b200::write_safetensors(path, list, &err)) return fail(B200_EMBED_ERR_IO, err);
    return B200_EMBED_OK;
}

int32_t b200_embed_profile_enable(b200_embed* h, int32_t on) {
    if (!h) return fail(B200_EMBED_ERR_INVALID_ARG, "null handle");
    h->profile.store(on ? 1 : 0);
    return B200_EMBED_OK;
}

int32_t b200_embed_profile_read(b200_embed* h, b200_embed_kind_stats* out, int32_t reset) {
    if (!h || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null handle or out");
    std::lock_guard<std::mutex> lk(h->stats_mu);
    memcpy(out, h->stats, sizeof(h->stats));
    if (reset) memset(h->stats, 0, sizeof(h->stats));
    return B200_EMBED_OK;
}

// ---- single kernels on device pointers ------------------------------------------------------------
int32_t b200_op_embed_layernorm(b200_embed* h, const uint32_t* ids, const uint32_t* tt, const int32_t* cu,
                                int32_t batch, int32_t total, int32_t dtype, void* out, void* out_lo, void* stream) {
    if (!h || !ids || !tt || !cu || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null argument");
    DeviceGuard guard(h->device);
    Context& c = h->ctxs[0];
    b200::EmbedArgs e{};
    e.ids = ids; e.type_ids = tt; e.cu = cu; e.batch = batch; e.seq_len = 0; e.total_tokens = total; e.padded = 0;
    e.gamma = h->emb_g; e.beta = h->emb_b; e.eps = h->cfg.layer_norm_eps; e.H = h->H; e.vocab = h->cfg.vocab_size;
    e.type_vocab = h->cfg.type_vocab_size; e.max_pos = h->cfg.max_position_embeddings; e.out = out;
    e.out_lo = dtype ? nullptr : out_lo;
    e.status = c.d_status;
    e.word = h->word; e.pos = h->pos; e.type = h->type;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->launches.fetch_add(1);
    CU_TRY(dtype ? f32ops::launch_embed_layernorm<float>(e, s) : h->ops->embed_ln(e, s));
    return B200_EMBED_OK;
}

int32_t b200_op_linear(b200_embed* h, int32_t layer, int32_t which, const void* x, const void* residual,
                       const void* residual_lo, int32_t rows, int32_t dtype, void* y, void* y_lo, void* stream) {
    if (!h || !x || !y || layer < 0 || layer >= h->L || which < 0 || which > 3)
        return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if ((which == 1 || which == 3) && !residual) return fail(B200_EMBED_ERR_INVALID_ARG, "residual required");
    DeviceGuard guard(h->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const LayerW& w = h->layers[layer];
    const int H = h->H, I = h->I;
    const float eps = h->cfg.layer_norm_eps;
    const float* w32[4] = {w.qkv_w, w.ao_w, w.i_w, w.o_w};
    const bf16* w16[4] = {w.qkv_w16, w.ao_w16, w.i_w16, w.o_w16};
    const float* bias[4] = {w.qkv_b, w.ao_b, w.i_b, w.o_b};
    const int n[4] = {3 * H, H, I, H}, k[4] = {H, H, H, I};
    const float* gam = which == 1 ? w.ao_g : w.o_g;
    const float* bet = which == 1 ? w.ao_be : w.o_be;
    const bool ln = which == 1 || which == 3;
    if (dtype) {
        float* tmp = static_cast<float*>(y);
        if (ln) CU_TRY(cudaMalloc(&tmp, (size_t)rows * H * 4));
        h->launches.fetch_add(ln ? 2 : 1);
        cudaError_t e = b200::launch_gemm_f32(static_cast<const float*>(x), w32[which], bias[which], rows, nullptr,
                                              n[which], k[which], which == 2 ? h->cfg.hidden_act : -1, tmp, s);
        if (e == cudaSuccess && ln)
            e = f32ops::launch_add_layernorm<float, float>(tmp, static_cast<const float*>(residual), nullptr, rows, nullptr,
                                                         gam, bet, eps, H, static_cast<float*>(y), nullptr, s);
        if (ln) { cudaStreamSynchronize(s); cudaFree(tmp); }
        CU_TRY(e);
        return B200_EMBED_OK;
    }
    Context& c = h->ctxs[0];
    if (!ln) {
        const int epi = which == 2 ? h->act_epi : b200::EPI_BIAS;
        return gemm16(h, c, s, which == 0 ? B200_EMBED_KIND_GEMM_QKV : B200_EMBED_KIND_GEMM_FFN1, x, w16[which],
                      bias[which], rows, nullptr, n[which], k[which], epi, which == 0 ? h->bn_qkv : h->bn_ffn1, nullptr,
                      nullptr, nullptr, nullptr, y, nullptr, 0);
    }
    if (h->fuse_ln)
        return gemm16(h, c, s, which == 1 ? B200_EMBED_KIND_GEMM_AO : B200_EMBED_KIND_GEMM_FFN2, x, w16[which],
                      bias[which], rows, nullptr, H, k[which], b200::EPI_RES_LN, h->bn_ln, residual, residual_lo, gam, bet,
                      y, y_lo, 0);
    float* tmp = nullptr;
    CU_TRY(cudaMalloc(&tmp, (size_t)rows * H * 4));
    int rc = gemm16(h, c, s, which == 1 ? B200_EMBED_KIND_GEMM_AO : B200_EMBED_KIND_GEMM_FFN2, x, w16[which],
                    bias[which], rows, nullptr, H, k[which], b200::EPI_BIAS, h->bn_h, nullptr, nullptr, nullptr, nullptr,
                    tmp, nullptr, 1);
    if (!rc) {
        h->launches.fetch_add(1);
        const cudaError_t e = h->ops->add_ln_f32(
            tmp, static_cast<const bf16*>(residual), static_cast<const uint8_t*>(residual_lo), rows, nullptr, gam, bet, eps, H,
            static_cast<bf16*>(y), static_cast<uint8_t*>(y_lo), s);
        if (e != cudaSuccess) rc = cuda_fail(e, "add_layernorm");
    }
    cudaStreamSynchronize(s);
    cudaFree(tmp);
    return rc;
}

int32_t b200_op_ffn(b200_embed* h, int32_t layer, const void* x, const void* x_lo, int32_t rows, void* y, void* y_lo,
                    void* stream) {
    if (!h || !x || !y || layer < 0 || layer >= h->L || rows < 0) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if (!h->ops->ffn_fused_supported(h->H, h->I, h->cfg.hidden_act))
        return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "fused FFN needs hidden size 128 / 256 / 384 and erf-GELU");
    DeviceGuard guard(h->device);
    const LayerW& w = h->layers[layer];
    b200::FfnFusedArgs f{};
    f.a = x; f.a_lo = x_lo; f.w1 = w.i_w16; f.w2 = w.o_w16; f.b1 = w.i_b; f.b2 = w.o_b; f.gamma = w.o_g; f.beta = w.o_be;
    f.eps = h->cfg.layer_norm_eps; f.m = rows; f.hidden = h->H; f.inter = h->I; f.out = y; f.out_lo = y_lo;
    h->launches.fetch_add(1);
    CU_TRY(h->ops->ffn_fused_launch(f, static_cast<cudaStream_t>(stream), h->num_sms));
    return B200_EMBED_OK;
}

int32_t b200_op_attention(b200_embed* h, const void* qkv, const int32_t* cu, int32_t batch, int32_t rows,
                          int32_t max_seqlen, int32_t dtype, void* ctx, void* stream) {
    if (!h || !qkv || !cu || !ctx) return fail(B200_EMBED_ERR_INVALID_ARG, "null argument");
    DeviceGuard guard(h->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->launches.fetch_add(1);
    if (dtype != 0 && dtype != 1) return fail(B200_EMBED_ERR_INVALID_ARG, "b200_op_attention: dtype must be 0 or 1");
    CU_TRY(dtype == 1 ? b200::launch_attention_f32(static_cast<const float*>(qkv), cu, batch, max_seqlen, h->heads, h->D,
                                                   static_cast<float*>(ctx), s)
                      : h->ops->attention_tc(static_cast<const bf16*>(qkv), cu, rows, batch, max_seqlen, h->heads, h->D,
                                             static_cast<bf16*>(ctx), h->num_sms, s));
    return B200_EMBED_OK;
}

int32_t b200_op_qkv_attention(b200_embed* h, int32_t layer, const void* x, const int32_t* cu, int32_t batch, int32_t rows,
                              int32_t max_seqlen, void* ctx, void* stream) {
    if (!h || !x || !cu || !ctx || layer < 0 || layer >= h->L) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if (!h->ops->qkv_attention_supported(h->H, h->heads, max_seqlen))
        return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "fused QKV + attention: hidden 128/256/384, head_dim 32, seq <= 128");
    DeviceGuard guard(h->device);
    const LayerW& w = h->layers[layer];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int* tmp = nullptr;  // tile list of this call: tile_first [batch + 1] | n_tiles | seg_off [batch / 32 + 2]
    const size_t n_ints = static_cast<size_t>(batch) + 2 + static_cast<size_t>(batch) / 32 + 2;
    CU_TRY(cudaMalloc(&tmp, n_ints * 4));
    h->launches.fetch_add(2);
    cudaError_t e = h->ops->pack_tiles(cu, batch, 128, tmp + batch + 2, tmp, tmp + batch + 1, s);
    if (e == cudaSuccess)
        e = h->ops->qkv_attention(static_cast<const bf16*>(x), w.qkv_w16, w.qkv_b, cu, tmp, tmp + batch + 1, rows, batch,
                                       max_seqlen, h->H, h->heads, static_cast<bf16*>(ctx), h->num_sms, s);
    cudaStreamSynchronize(s);
    cudaFree(tmp);
    CU_TRY(e);
    return B200_EMBED_OK;
}

int32_t b200_op_add_layernorm(b200_embed* h, int32_t layer, int32_t which, const void* y, int32_t y_is_f32,
                              const void* residual, const void* residual_lo, int32_t rows, int32_t dtype, void* out,
                              void* out_lo, void* stream) {
    if (!h || !y || !residual || !out || layer < 0 || layer >= h->L || (which != 1 && which != 3))
        return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    DeviceGuard guard(h->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const LayerW& w = h->layers[layer];
    const float* gam = which == 1 ? w.ao_g : w.o_g;
    const float* bet = which == 1 ? w.ao_be : w.o_be;
    const float eps = h->cfg.layer_norm_eps;
    h->launches.fetch_add(1);
    if (dtype)
        CU_TRY((f32ops::launch_add_layernorm<float, float>(static_cast<const float*>(y), static_cast<const float*>(residual),
                                                         nullptr, rows, nullptr, gam, bet, eps, h->H,
                                                         static_cast<float*>(out), nullptr, s)));
    else if (y_is_f32)
        CU_TRY((h->ops->add_ln_f32(static_cast<const float*>(y), static_cast<const bf16*>(residual),
                                                        static_cast<const uint8_t*>(residual_lo), rows, nullptr, gam, bet, eps,
                                                        h->H, static_cast<bf16*>(out), static_cast<uint8_t*>(out_lo), s)));
    else
        CU_TRY((h->ops->add_ln_16(static_cast<const bf16*>(y), static_cast<const bf16*>(residual),
                                                       static_cast<const uint8_t*>(residual_lo), rows, nullptr, gam, bet, eps,
                                                       h->H, static_cast<bf16*>(out), static_cast<uint8_t*>(out_lo), s)));
    return B200_EMBED_OK;
}

int32_t b200_op_mean_pool(b200_embed* h, const void* hidden, const void* hidden_lo, const int32_t* cu, int32_t batch,
                          int32_t dtype, int32_t normalize, float* out, void* stream) {
    if (!h || !hidden || !cu || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null argument");
    DeviceGuard guard(h->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->launches.fetch_add(1);
    CU_TRY(dtype ? f32ops::launch_mean_pool<float>(static_cast<const float*>(hidden), nullptr, cu, batch, h->H, normalize, out, s)
                 : h->ops->mean_pool(static_cast<const bf16*>(hidden), static_cast<const uint8_t*>(hidden_lo), cu, batch, h->H,
                                     normalize, out, s));
    return B200_EMBED_OK;
}

static int op_gemm16(const Ops& ops, const void* a, const void* b, const float* bias, int32_t m, int32_t n, int32_t k,
                     int32_t act, int32_t block_n, void* c, void* stream) {
    if (!a || !b || !bias || !c || m < 0 || n <= 0 || k <= 0) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    if (act < -1 || act > 2) return fail(B200_EMBED_ERR_INVALID_ARG, "unknown activation");
    int dev = 0, sms = 148;
    CU_TRY(cudaGetDevice(&dev));
    CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int epi = act < 0 ? b200::EPI_BIAS : b200::EPI_GELU_ERF + act;
    if (block_n == 0) block_n = ops.pick_block_n(n, epi);
    if (block_n == 0 || n % block_n) return fail(B200_EMBED_ERR_UNSUPPORTED_SHAPE, "n has no supported tile width");
    b200::GemmTcArgs g{};
    g.a = a; g.b = b; g.m = m; g.n = n; g.k = k; g.bias = bias; g.out = c; g.epi = epi; g.block_n = block_n;
    const char* pv = getenv("B200_EMBED_GEMM_PAIR");
    g.pair = (pv && std::string(pv) == "0") ? 0 : 1;
    CU_TRY(ops.gemm_tc_launch(g, static_cast<cudaStream_t>(stream), sms));
    return B200_EMBED_OK;
}
int32_t b200_op_gemm_bf16(const void* a, const void* b, const float* bias, int32_t m, int32_t n, int32_t k,
                          int32_t act, int32_t block_n, void* c, void* stream) {
    return op_gemm16(kOpsBf16, a, b, bias, m, n, k, act, block_n, c, stream);
}
int32_t b200_op_gemm_f16(const void* a, const void* b, const float* bias, int32_t m, int32_t n, int32_t k,
                         int32_t act, int32_t block_n, void* c, void* stream) {
    return op_gemm16(kOpsF16, a, b, bias, m, n, k, act, block_n, c, stream);
}
int32_t b200_embed_operand_format(const b200_embed* h) {
    return !h ? 0 : (h->f16 ? B200_EMBED_OPERANDS_FP16 : B200_EMBED_OPERANDS_BF16);
}

}  // extern "C"

// =============================================================================================
// b200_op_linear_noalloc: the unfused Linear + residual + LayerNorm step without any allocation
// =============================================================================================
// The pre-LayerNorm fp32 rows that b200_op_linear keeps in a cudaMalloc'ed temporary live here instead, in
// device memory that is part of the loaded library image. Every GPU that loads the library carries these
// 4 MiB until the process exits, used or not.
static __device__ __align__(256) float g_noalloc_scratch[B200_OP_NOALLOC_MAX_ELEMS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {};

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel can use: ask the
// runtime for the device address once per device (with that device current) and hand that to the kernels.
int noalloc_scratch(int device, float** out) {
    if (device < 0 || device >= kMaxDevices) return fail(B200_EMBED_ERR_INVALID_ARG, "device ordinal too large");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        CU_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[device] = static_cast<float*>(p);
    }
    *out = g_noalloc_ptr[device];
    return B200_EMBED_OK;
}
}  // namespace

extern "C" int32_t b200_op_linear_noalloc(b200_embed* h, int32_t layer, int32_t which, const void* x,
                                          const void* residual, const void* residual_lo, int32_t rows, int32_t dtype,
                                          void* y, void* y_lo, void* stream) {
    if (!h || !x || !residual || !y || layer < 0 || layer >= h->L || (which != 1 && which != 3) || rows < 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "b200_op_linear_noalloc: bad argument (which must be 1 or 3)");
    if (dtype != 0 && dtype != 1) return fail(B200_EMBED_ERR_INVALID_ARG, "b200_op_linear_noalloc: dtype must be 0 or 1");
    const int H = h->H, I = h->I;
    if (static_cast<long long>(rows) * H > static_cast<long long>(B200_OP_NOALLOC_MAX_ELEMS))
        return fail(B200_EMBED_ERR_INVALID_ARG,
                    "b200_op_linear_noalloc: rows*hidden = " + std::to_string(static_cast<long long>(rows) * H) +
                        " exceeds the built-in scratch of " + std::to_string(B200_OP_NOALLOC_MAX_ELEMS) +
                        " floats (4 MiB); use b200_op_linear or split the rows");
    if (rows == 0) return B200_EMBED_OK;
    DeviceGuard guard(h->device);
    if (!guard.ok) return fail(B200_EMBED_ERR_CUDA, "cudaSetDevice failed");
    float* scratch = nullptr;
    int rc = noalloc_scratch(h->device, &scratch);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const LayerW& w = h->layers[layer];
    const float eps = h->cfg.layer_norm_eps;
    const int k = which == 1 ? H : I;
    const float* gam = which == 1 ? w.ao_g : w.o_g;
    const float* bet = which == 1 ? w.ao_be : w.o_be;
    const float* bias = which == 1 ? w.ao_b : w.o_b;
    h->launches.fetch_add(2);
    if (dtype == 1) {
        CU_TRY(b200::launch_gemm_f32(static_cast<const float*>(x), which == 1 ? w.ao_w : w.o_w, bias, rows, nullptr, H, k,
                                     -1, scratch, s));
        CU_TRY((f32ops::launch_add_layernorm<float, float>(scratch, static_cast<const float*>(residual), nullptr, rows,
                                                         nullptr, gam, bet, eps, H, static_cast<float*>(y), nullptr, s)));
        return B200_EMBED_OK;
    }
    b200::GemmTcArgs g{};
    g.a = x; g.b = which == 1 ? w.ao_w16 : w.o_w16; g.m = rows; g.n = H; g.k = k; g.bias = bias; g.eps = eps;
    g.out = scratch; g.out_f32 = 1; g.epi = b200::EPI_BIAS; g.block_n = h->bn_h;
    CU_TRY(h->ops->gemm_tc_launch(g, s, h->num_sms));
    CU_TRY((h->ops->add_ln_f32(scratch, static_cast<const bf16*>(residual),
                                                    static_cast<const uint8_t*>(residual_lo), rows, nullptr, gam, bet, eps, H,
                                                    static_cast<bf16*>(y), static_cast<uint8_t*>(y_lo), s)));
    return B200_EMBED_OK;
}
