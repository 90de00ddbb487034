/*
 * b200_embed.h -- C ABI of libb200_embed.so: a B200-native (sm_100a) replacement for the one
 * data-parallel hot path of makkenzo/codename-symbiont, the sentence-embedding forward pass in
 * `preprocessing_service` (tokenised batch -> BERT encoder -> masked mean-pool [-> L2 norm]).
 *
 * Reference interface replaced (paths relative to the reference repository root):
 *   services/preprocessing_service/src/embedding_generator.rs
 *     :17   EmbeddingGenerator::new(model_id, revision, force_cpu)      -> b200_embed_create
 *     :134  generate_sentence_embeddings(&self, &[String])              -> b200_embed_forward*
 *     :189-191 Tensor::from_vec(ids / mask / type_ids, (b, P))           (the three u32 inputs)
 *     :198  BertModel::forward(ids, type_ids, Some(mask))               (the encoder)
 *     :201-207 masked mean-pool                                         (fused epilogue)
 *     :214  to_vec2::<f32>()                                            (float* out, [batch][hidden])
 * Everything above that call (NATS handling, tokenizers, sentence split) and below it
 * (libs/shared_models payloads, the Qdrant upsert) can stay unchanged; see INTEGRATION.md for the
 * Rust binding a maintainer would add. The host stages either side of the call have optional
 * replacements at the end of this header: b200_ingest_* (sentence split + WordPiece tokeniser,
 * main.rs:28-63, embedding_generator.rs:73-99, :160-187) and b200_json_* (the serde_json payloads).
 *
 * Conventions: every entry point returns B200_EMBED_OK (0) or a negative B200_EMBED_ERR_* code and
 * never throws or aborts across the boundary; the message of the last failure on the calling
 * thread is b200_embed_last_error(). Pointers are plain host pointers unless the name says
 * `device`. The caller owns all input and output buffers for the duration of the call; the
 * library owns weights and workspaces. There is NO CPU fallback: without a CUDA device every
 * compute entry returns B200_EMBED_ERR_NO_DEVICE.
 */
#ifndef B200_EMBED_H_
#define B200_EMBED_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_EMBED_OK 0
#define B200_EMBED_ERR_INVALID_ARG (-1)       /* bad pointer/size, id >= vocab, mask not a 0/1 prefix */
#define B200_EMBED_ERR_CUDA (-2)              /* a CUDA call or kernel failed */
#define B200_EMBED_ERR_OOM (-3)               /* device or pinned allocation failed */
#define B200_EMBED_ERR_UNSUPPORTED_SHAPE (-4) /* model shape outside what the kernels are built for */
#define B200_EMBED_ERR_NO_DEVICE (-5)         /* no CUDA device: there is no CPU fallback */
#define B200_EMBED_ERR_IO (-6)                /* safetensors file unreadable / tensor missing */

/* flags for the forward entry points */
#define B200_EMBED_FLAG_NORMALIZE 1u /* append v / max(||v||_2, 1e-12). The reference does NOT normalise
                                        (embedding_generator.rs:201-207), so the drop-in default is 0. */
#define B200_EMBED_FLAG_FP32 2u      /* fp32 validation mode: fp32 weights, activations and SIMT GEMMs
                                        (slow; parity gate cosine >= 0.999999). Default is the 16-bit tensor-core path. */

/* hidden_act values: candle HiddenAct::{Gelu, GeluApproximate, Relu} */
#define B200_EMBED_ACT_GELU_ERF 0
#define B200_EMBED_ACT_GELU_TANH 1
#define B200_EMBED_ACT_RELU 2

typedef struct b200_embed b200_embed;

/* The config.json fields candle's bert::Config reads (embedding_generator.rs:70-71). */
typedef struct b200_embed_config {
    int32_t vocab_size;
    int32_t hidden_size;        /* multiple of 128, <= 1024 */
    int32_t num_hidden_layers;
    int32_t num_attention_heads; /* head dim must be 32 or 64 */
    int32_t intermediate_size;  /* multiple of 64 */
    int32_t max_position_embeddings;
    int32_t type_vocab_size;
    int32_t hidden_act;         /* B200_EMBED_ACT_* */
    float layer_norm_eps;
    int32_t max_tokens_per_launch; /* workspace capacity in tokens; 0 = default (1 Mi). Larger calls are
                                      split into sentence chunks internally, order preserved. */
    int32_t operand_format;     /* B200_EMBED_OPERANDS_*: what the tensor-core GEMMs multiply. 0 = default = fp16 */
} b200_embed_config;

/* Operand format of the 16-bit path (weights, activations fed to tcgen05.mma kind::f16, the (hi, lo) residual pairs).
 * Both run at the same tensor-core rate and move the same bytes; accumulation, LayerNorm statistics, softmax and pooling
 * are fp32 either way. fp16 (11 significand bits) keeps the un-normalised output within max-abs 2e-3 of the f32
 * reference (measured 7e-4 on MiniLM-L6), which bf16 (8 bits) cannot: rounding the weight matrices to bf16 alone moves it
 * by 2.6-3.1e-3. fp16's range (6e-5 .. 65504 at full precision) covers BERT weights and post-LayerNorm activations; a
 * checkpoint with activations beyond it should pick bf16. */
#define B200_EMBED_OPERANDS_DEFAULT 0
#define B200_EMBED_OPERANDS_BF16 1
#define B200_EMBED_OPERANDS_FP16 2

/* -- lifetime: replaces EmbeddingGenerator::new (embedding_generator.rs:17-132) ------------------
 * safetensors_path: a HuggingFace BERT `model.safetensors` (F32/F16/BF16; names as in
 * BertModel::load, optional `bert.` prefix, legacy gamma/beta LayerNorm names) or NULL for the
 * seeded random init used by benchmarks (spec in codename-symbiont_b200/weights.py).
 * Converts weights to the operand format (cfg->operand_format) once, packs Q/K/V into one [3H,H] matrix, keeps the f32 master for
 * the validation mode. The handle is thread-safe: concurrent forward calls each lease a
 * stream + workspace (the reference shares one Arc<EmbeddingGenerator>, main.rs:314, :373). */
int32_t b200_embed_create(const b200_embed_config* cfg, const char* safetensors_path, uint64_t seed,
                          int32_t device, b200_embed** out);
void b200_embed_destroy(b200_embed* h);

int32_t b200_embed_hidden_size(const b200_embed* h);
/* B200_EMBED_OPERANDS_BF16 or _FP16: the format this handle computes in (and that the b200_op_* device-pointer entry
 * points below expect their 16-bit tensors in). */
int32_t b200_embed_operand_format(const b200_embed* h);
int32_t b200_embed_max_positions(const b200_embed* h);
/* thread-local, never NULL */
const char* b200_embed_last_error(void);

/* -- the hot path: replaces embedding_generator.rs:189-214 ---------------------------------------
 * input_ids / attention_mask / token_type_ids: [batch, seq_len] row-major u32 exactly as
 * `tokenizers` produced them (embedding_generator.rs:179-187); attention_mask must be a
 * right-padded 0/1 prefix per row (PaddingDirection::Right, :85). seq_len <=
 * max_position_embeddings. Padded positions are never computed (they cannot reach the output,
 * SURVEY.md section 3.4). out: [batch, hidden] f32 row-major. batch == 0 is OK and writes nothing
 * (empty in => empty out, :135-137). Blocks until `out` is complete. */
int32_t b200_embed_forward(b200_embed* h, const uint32_t* input_ids, const uint32_t* attention_mask,
                           const uint32_t* token_type_ids, int32_t batch, int32_t seq_len, uint32_t flags,
                           float* out);

/* Packed variant for callers that already dropped the padding: ids/type ids of all sentences back
 * to back, cu_seqlens[batch+1] exclusive prefix sums (cu_seqlens[batch] == total_tokens). */
int32_t b200_embed_forward_varlen(b200_embed* h, const uint32_t* ids_packed, const uint32_t* type_ids_packed,
                                  const int32_t* cu_seqlens, int32_t batch, int32_t total_tokens,
                                  int32_t max_seqlen, uint32_t flags, float* out);

/* Asynchronous pair for the ingest pipeline (main.rs:360-393 spawns one task per message): submit
 * returns a ticket at once; the buffers must stay valid and (for overlap) pinned until
 * b200_embed_wait(ticket) returns. Same semantics as b200_embed_forward otherwise. */
int32_t b200_embed_submit(b200_embed* h, const uint32_t* input_ids, const uint32_t* attention_mask,
                          const uint32_t* token_type_ids, int32_t batch, int32_t seq_len, uint32_t flags,
                          float* out, int64_t* ticket);
int32_t b200_embed_wait(b200_embed* h, int64_t ticket);
/* The same for packed input (the form b200_ingest_encode_packed produces): b200_embed_forward_varlen split into submit and
 * b200_embed_wait. cu_seqlens is read before the call returns; ids_packed, type_ids_packed and out must stay valid (and,
 * for overlap, pinned) until the wait. The corpus driver below keeps two of these in flight per GPU. */
int32_t b200_embed_submit_varlen(b200_embed* h, const uint32_t* ids_packed, const uint32_t* type_ids_packed,
                                 const int32_t* cu_seqlens, int32_t batch, int32_t total_tokens, int32_t max_seqlen,
                                 uint32_t flags, float* out, int64_t* ticket);

/* Same computation with all four buffers already resident in device memory (replaces the
 * candle-on-CUDA tensors of :189-198 when the caller keeps data on the GPU). Enqueues on `stream`
 * (a cudaStream_t, NULL = the handle's own stream). With stream == NULL the call synchronises and reports input
 * errors (ids / type ids out of range, a mask that is not a right-padded prefix) like b200_embed_forward. With a
 * caller stream it returns after enqueueing and input errors are NOT reported: out-of-range ids are clamped to
 * row 0 on the device, so validate on the host (or call once with stream == NULL) when the inputs are untrusted. */
int32_t b200_embed_forward_device(b200_embed* h, const uint32_t* device_ids, const uint32_t* device_mask,
                                  const uint32_t* device_type_ids, int32_t batch, int32_t seq_len,
                                  uint32_t flags, float* device_out, void* stream);

/* Pinned host memory for the submit/wait pipeline (cudaHostAlloc / cudaFreeHost). */
int32_t b200_embed_host_alloc(size_t bytes, void** out);
void b200_embed_host_free(void* p);

/* -- introspection ---------------------------------------------------------------------------------
 * Writes the f32 master weights as safetensors (HF names), so a checker can load exactly the
 * weights the handle computes with (replaces nothing; VarBuilder is read-only, :111). */
int32_t b200_embed_export_safetensors(const b200_embed* h, const char* path);
/* Number of kernels this handle has launched so far. */
int64_t b200_embed_kernel_launches(const b200_embed* h);

/* Per-kernel device timing. While enabled, each forward records CUDA events around every kernel
 * on its launching stream and adds them to per-kind totals (kinds below). Off by default. */
#define B200_EMBED_KIND_LENGTHS 0   /* mask -> lengths -> cu_seqlens */
#define B200_EMBED_KIND_EMBED_LN 1  /* K1 gather + LayerNorm            (bert.rs BertEmbeddings::forward) */
#define B200_EMBED_KIND_GEMM_QKV 2  /* K2 fused Q/K/V projection        (BertSelfAttention::forward) */
#define B200_EMBED_KIND_ATTENTION 3 /* K3 varlen flash attention */
#define B200_EMBED_KIND_GEMM_AO 4   /* K2 attention.output.dense (+res+LN fused when it fits) (BertSelfOutput) */
#define B200_EMBED_KIND_GEMM_FFN1 5 /* K2 intermediate.dense + GELU     (BertIntermediate) */
#define B200_EMBED_KIND_GEMM_FFN2 6 /* K2 output.dense (+res+LN fused)  (BertOutput) */
#define B200_EMBED_KIND_ADD_LN 7    /* K4 residual + LayerNorm when not fused */
#define B200_EMBED_KIND_POOL 8      /* K5 masked mean-pool (+L2)        (embedding_generator.rs:201-207) */
#define B200_EMBED_KIND_FFN_FUSED 9 /* K2f intermediate.dense + GELU + output.dense + residual + LayerNorm in one kernel
                                       (BertIntermediate + BertOutput); replaces kinds 5 and 6 when hidden <= 384 */
#define B200_EMBED_KIND_QKV_ATTN 10 /* K23 Q/K/V projection + attention of whole sentences in one kernel (BertSelfAttention);
                                       replaces kinds 2 and 3 when hidden <= 384, head_dim 32, seq_len <= 128 */
#define B200_EMBED_NUM_KINDS 11
typedef struct b200_embed_kind_stats {
    double total_ms;  /* sum of event-timed durations */
    int64_t launches;
} b200_embed_kind_stats;
int32_t b200_embed_profile_enable(b200_embed* h, int32_t on);
int32_t b200_embed_profile_read(b200_embed* h, b200_embed_kind_stats* out /* [B200_EMBED_NUM_KINDS] */,
                                int32_t reset);

/* -- single kernels on device pointers (parity tests and roofline runs) ---------------------------
 * All pointers are device pointers; `stream` is a cudaStream_t (NULL = default stream). dtype:
 * 0 = 16-bit activations in the handle's operand format / tensor-core GEMM, 1 = fp32 validation kernels. Row counts are
 * tokens. */

/* Residual-stream tensors in the 16-bit mode are (hi, lo) pairs of [rows, H] arrays: hi = round16(v), in the operand
 * format, is what the next GEMM multiplies; lo = v - hi as ONE BYTE per element (an e5m2 float: fp16's exponent range,
 * 2 mantissa bits, round to nearest even, saturating), so hi + lo is what residual adds and the pooling read at 3 bytes
 * per element. Every `*_lo` argument points to rows x H bytes, may be NULL (plain 16-bit) and is ignored when dtype = 1.
 * ("bf16" in the comments below means "the handle's 16-bit operand format".) */

/* K1. out[t,:] = LN(word[ids[t]] + type[tt[t]] + pos[position of t]) for the packed tokens of
 * `batch` sentences; cu_seqlens int32[batch+1] on device. (bert.rs BertEmbeddings::forward) */
int32_t b200_op_embed_layernorm(b200_embed* h, const uint32_t* ids_packed, const uint32_t* type_ids_packed,
                                const int32_t* cu_seqlens, int32_t batch, int32_t total_tokens, int32_t dtype,
                                void* out, void* out_lo, void* stream);
/* K2. y = epilogue(x . W^T + b) with the weights of encoder layer `layer`:
 * which = 0 QKV (N=3H), 1 attention.output.dense (+ residual + LayerNorm), 2 intermediate.dense (+ act),
 * 3 output.dense (+ residual + LayerNorm). residual is read only for which = 1, 3. */
int32_t b200_op_linear(b200_embed* h, int32_t layer, int32_t which, const void* x, const void* residual,
                       const void* residual_lo, int32_t rows, int32_t dtype, void* y, void* y_lo, void* stream);
/* K2f. y = LayerNorm(gelu(x . Wi^T + bi) . Wo^T + bo + x) with the feed-forward weights of `layer`, in one kernel
 * (bf16 tensor cores; hidden size 128 / 256 / 384 with erf-GELU, else B200_EMBED_ERR_UNSUPPORTED_SHAPE).
 * x / x_lo and y / y_lo are (hi, lo) pairs as above. (bert.rs BertIntermediate::forward + BertOutput::forward) */
int32_t b200_op_ffn(b200_embed* h, int32_t layer, const void* x, const void* x_lo, int32_t rows, void* y, void* y_lo,
                    void* stream);

/* The same Linear + residual + LayerNorm step as b200_op_linear(which = 1 or 3) for callers that must not
 * allocate. It has no counterpart in the reference. b200_op_linear parks the fp32 pre-LayerNorm rows in a
 * cudaMalloc'ed temporary when the LayerNorm is not fused into the GEMM (hidden sizes other than 128, 256, 384,
 * 512, 768 and 1024 - e.g. 640 - and the fp32 mode); cudaMalloc and cudaFree are off limits for a caller that has already handed every byte
 * of the GPU to its own pool, and they synchronise the whole device, which a caller with other streams in
 * flight in a latency-bound loop cannot afford. Here those rows go to a fixed 4 MiB array that is part of the
 * library image instead (resident on every GPU that loads the library until the process exits), and the call
 * runs the GEMM and the add + LayerNorm kernel back to back on `stream`; it never calls cudaMalloc, cudaFree
 * or any allocating library routine.
 *   - rows * hidden_size must not exceed B200_OP_NOALLOC_MAX_ELEMS (else B200_EMBED_ERR_INVALID_ARG).
 *   - The first call on a device looks up the array's device address (cudaGetSymbolAddress); make that first
 *     call outside any stream capture.
 *   - All calls share the one array and it keeps whatever the last call left there. Nothing in it has to start
 *     at zero (every row that is read is written first), but two calls that overlap on the GPU -- different
 *     streams, different handles, different threads -- race on it: serialise them.
 * Arguments and results are those of b200_op_linear; which = 0 / 2 use no temporary and are not offered. */
#define B200_OP_NOALLOC_MAX_ELEMS (1 << 20)
int32_t b200_op_linear_noalloc(b200_embed* h, int32_t layer, int32_t which, const void* x, const void* residual,
                               const void* residual_lo, int32_t rows, int32_t dtype, void* y, void* y_lo,
                               void* stream);

/* K3. ctx = softmax(q k^T / sqrt(d)) v per sentence and head over packed qkv [rows, 3H].
 * dtype 0: tcgen05 kernel (the one the forward uses), 1: fp32 validation. */
int32_t b200_op_attention(b200_embed* h, const void* qkv, const int32_t* cu_seqlens, int32_t batch, int32_t rows,
                          int32_t max_seqlen, int32_t dtype, void* ctx, void* stream);
/* K23. ctx = attention(x . Wqkv^T + bqkv) of layer `layer` straight from the bf16 activations x [rows, H]: the fused
 * form of b200_op_linear(which = 0) + b200_op_attention(dtype = 0), used by the forward whenever it applies (hidden
 * 128 / 256 / 384, head_dim 32, max_seqlen <= 128; otherwise B200_EMBED_ERR_UNSUPPORTED_SHAPE). */
int32_t b200_op_qkv_attention(b200_embed* h, int32_t layer, const void* x, const int32_t* cu_seqlens, int32_t batch,
                              int32_t rows, int32_t max_seqlen, void* ctx, void* stream);
/* K4. out = LN(y + residual) * gamma + beta with the LayerNorm of (layer, which = 1 or 3).
 * y is fp32 when dtype = 0 came from an unfused GEMM (y_is_f32 != 0). */
int32_t b200_op_add_layernorm(b200_embed* h, int32_t layer, int32_t which, const void* y, int32_t y_is_f32,
                              const void* residual, const void* residual_lo, int32_t rows, int32_t dtype, void* out,
                              void* out_lo, void* stream);
/* K5. out[b,:] = sum_j hidden[cu[b]+j,:] / (len_b + 1e-9) [then L2 normalise]. out f32 [batch,H]. */
int32_t b200_op_mean_pool(b200_embed* h, const void* hidden, const void* hidden_lo, const int32_t* cu_seqlens,
                          int32_t batch, int32_t dtype, int32_t normalize, float* out, void* stream);
/* Generic bf16 tensor-core GEMM used by K2, exposed for shape sweeps: C[M,N] = A[M,K] . B[N,K]^T + bias,
 * bf16 in, fp32 accumulate, bf16 out. N % 128 == 0, K % 8 == 0. act: -1 none, else B200_EMBED_ACT_*. */
int32_t b200_op_gemm_bf16(const void* a, const void* b, const float* bias, int32_t m, int32_t n, int32_t k,
                          int32_t act, int32_t block_n, void* c, void* stream);
/* The same kernel with IEEE fp16 operands and output. */
int32_t b200_op_gemm_f16(const void* a, const void* b, const float* bias, int32_t m, int32_t n, int32_t k,
                         int32_t act, int32_t block_n, void* c, void* stream);

/* -- wire format of the results (host only, no GPU involved) --------------------------------------------
 * The JSON text serde_json::to_vec produces for the messages the service publishes after the embedding call
 * (preprocessing_service/src/main.rs:138-142, :238-245): compact, struct fields in declaration order, f32 in the
 * shortest decimal that round-trips (ryu "pretty" layout), non-finite floats as null, UTF-8 strings with JSON
 * escapes. Replaces serde's derive(Serialize) for shared_models::TextWithEmbeddingsMessage
 * (libs/shared_models/src/lib.rs:45-52, SentenceEmbedding :39-43) and QueryEmbeddingResult (:66-72); the schema is
 * unchanged. Each call returns the number of bytes of the document; it is written to `out` (no terminator) when
 * out != NULL and out_cap is large enough - call with out = NULL to size the buffer. Negative on bad arguments. */
int64_t b200_json_text_with_embeddings(const char* original_id, const char* source_url, const char* const* sentences,
                                       const float* embeddings /* [n * dim] row-major */, int32_t n, int32_t dim,
                                       const char* model_name, uint64_t timestamp_ms,
                                       int32_t threads /* 0 = all host cores */, char* out, int64_t out_cap);
/* Same document, the sentences given as byte ranges [begin[i], end[i]) of one UTF-8 text - the form
 * b200_ingest_split_sentences produces - so that no per-sentence string is built between the two calls. */
int64_t b200_json_text_with_embeddings_ranges(const char* original_id, const char* source_url, const char* text,
                                              const int64_t* begin, const int64_t* end, const float* embeddings,
                                              int32_t n, int32_t dim, const char* model_name, uint64_t timestamp_ms,
                                              int32_t threads, char* out, int64_t out_cap);
int64_t b200_json_query_embedding_result(const char* request_id, const float* embedding_or_null, int32_t dim,
                                         const char* model_name_or_null, const char* error_message_or_null, char* out,
                                         int64_t out_cap);
/* One f32 as serde_json writes it; out24 must hold 24 bytes. Returns the number of characters. */
int32_t b200_json_format_f32(float value, char* out24);
/* The same document in pieces, for a caller that formats a long message chunk by chunk (or publishes one message per
 * chunk under a payload limit): the envelope `{"original_id":..,"source_url":..,"embeddings_data":[` and
 * `],"model_name":..,"timestamp_ms":..}` (lengths always reported; written when the capacities suffice), and the
 * comma-separated items of n sentences given as byte ranges (first != 0: no comma in front of the first item). The items
 * call returns their size and writes them when out_cap suffices: pass n * (17 * dim + 80) + 6 * (text bytes) to be sure. */
int64_t b200_json_envelope(const char* original_id, const char* source_url, const char* model_name, uint64_t timestamp_ms,
                           char* head, int64_t head_cap, int64_t* head_len, char* tail, int64_t tail_cap, int64_t* tail_len);
int64_t b200_json_sentence_items_ranges(const char* text, const int64_t* begin, const int64_t* end, const float* embeddings,
                                        int32_t n, int32_t dim, int32_t first, int32_t threads, char* out, int64_t out_cap);

/* The items of n sentences formatted ON THE GPU, from embeddings still in device memory (e.g. the output of
 * b200_embed_forward_device): byte for byte what b200_json_sentence_items_ranges writes, so only the finished text (about
 * 4.5 KB per 384-d sentence) crosses PCIe and no host core formats a float. The sentence texts are escaped on the host
 * (a few hundred bytes per sentence). `out` is a host buffer (pinned memory makes the copy faster); the call returns the
 * size of the items and writes them when out_cap suffices - call with out = NULL to size it. One handle per GPU; calls on
 * one handle are serialised. CUDA only: without a device use the host entry above. Errors: b200_jsongpu_last_error(). */
typedef struct b200_jsongpu b200_jsongpu;
const char* b200_jsongpu_last_error(void);
int32_t b200_jsongpu_create(int32_t device, b200_jsongpu** out);
void b200_jsongpu_destroy(b200_jsongpu* g);
int64_t b200_jsongpu_sentence_items(b200_jsongpu* g, const char* text, const int64_t* begin, const int64_t* end,
                                    const float* device_embeddings, int32_t n, int32_t dim, int32_t first, char* out,
                                    int64_t out_cap);
/* The same for embeddings that are already back on the host (they are uploaded first, 4 bytes per value against the ~12 of
 * their text): for callers of the host-buffer entry points that want the formatting off their cores. */
int64_t b200_jsongpu_sentence_items_host(b200_jsongpu* g, const char* text, const int64_t* begin, const int64_t* end,
                                         const float* host_embeddings, int32_t n, int32_t dim, int32_t first, char* out,
                                         int64_t out_cap);

/* -- ingest front-end: raw text -> sentences -> token arrays (host only, no GPU involved) ------------------
 * What runs on the host before the embedding call. Errors of these entries are reported through
 * b200_ingest_last_error() (calling thread), with the same B200_EMBED_ERR_* codes. */
typedef struct b200_ingest b200_ingest;
const char* b200_ingest_last_error(void);

/* Loads a HuggingFace tokenizer.json. Two pipelines are understood:
 *   - BERT WordPiece (BertNormalizer, BertPreTokenizer, WordPiece, BertProcessing or TemplateProcessing): the BASELINE models;
 *   - Unigram / SentencePiece as XLM-R ships it (normalizer Precompiled charsmap [+ Replace " {2,}" -> " "], pre-tokenizer
 *     [WhitespaceSplit +] Metaspace, model Unigram with unk_id, TemplateProcessing "<s> $A </s>"): the tokenizer of the model
 *     the service deploys, paraphrase-multilingual-mpnet-base-v2 (preprocessing_service/src/main.rs:305).
 * Replaces Tokenizer::from_file plus the padding / truncation set-up of preprocessing_service/src/embedding_generator.rs:73-99:
 * padding is Fixed(seq_len) on the right with the id of "[PAD]" (0 when the vocabulary has none - the XLM-R case, exactly as
 * the reference falls back), type id 0; truncation is LongestFirst to seq_len (seq_len is passed to b200_ingest_encode).
 * B200_EMBED_ERR_IO when the file cannot be read, B200_EMBED_ERR_UNSUPPORTED_SHAPE for any other tokenizer (BPE models,
 * byte fallback, other normalizers / pre-tokenizers, added tokens with lstrip / rstrip / single_word / normalized). */
int32_t b200_ingest_create(const char* tokenizer_json_path, b200_ingest** out);
void b200_ingest_destroy(b200_ingest* t);
int32_t b200_ingest_vocab_size(const b200_ingest* t); /* largest id + 1, added tokens included */
int32_t b200_ingest_pad_id(const b200_ingest* t);

/* The sentence splitter of process_raw_text_message (preprocessing_service/src/main.rs:28-63): whitespace runs
 * collapsed to one space (split_whitespace + join), a cut after every '.', '?' or '!', every piece trimmed, a
 * non-empty remainder kept. Writes the cleaned text to `cleaned` (cleaned_cap >= len is always enough, no
 * terminator) and sentence i as the byte range [begin[i], end[i]) of it. Returns the number of sentences; when it
 * exceeds max_sentences only the first max_sentences ranges are written. 0 means the cleaned text is empty (the
 * reference answers "Cleaned text is empty for id: .."). Negative on bad arguments. */
int64_t b200_ingest_split_sentences(const char* raw_text, int64_t len, char* cleaned, int64_t cleaned_cap,
                                    int64_t* cleaned_len, int64_t* begin, int64_t* end, int64_t max_sentences);
/* The same with the number of worker threads given (0 = every host core, what the call above uses): a driver that splits
 * several texts side by side passes 1 so that the calls do not each start a thread per core. */
int64_t b200_ingest_split_sentences_threads(const char* raw_text, int64_t len, char* cleaned, int64_t cleaned_cap,
                                            int64_t* cleaned_len, int64_t* begin, int64_t* end, int64_t max_sentences,
                                            int32_t threads);

/* tokenizer.encode_batch(inputs, true) (embedding_generator.rs:160-187): sentence i is text[begin[i], end[i]) in
 * UTF-8; row i of ids / mask / type_ids ([n, seq_len] u32, row-major) receives [CLS] tokens [SEP] then padding,
 * exactly what b200_embed_forward takes. lengths_or_null[i] = number of real tokens of row i. `threads` = 0 uses
 * every host core. */
int32_t b200_ingest_encode(const b200_ingest* t, const char* text, const int64_t* begin, const int64_t* end, int32_t n,
                           int32_t seq_len, int32_t threads, uint32_t* ids, uint32_t* mask, uint32_t* type_ids,
                           int32_t* lengths_or_null);
/* The same tokens without the padding, in the form b200_embed_forward_varlen takes: the real tokens of all sentences
 * back to back in ids_packed / type_ids_packed and cu_seqlens[n + 1] (exclusive prefix sums). max_len is the truncation
 * length (max_position_embeddings). Returns the total number of tokens; when it exceeds capacity_tokens only
 * cu_seqlens is written - call again with room for that many. For sentences of ~20 tokens this is 1/6 of the bytes the
 * padded arrays take at seq_len 128, on the host and over PCIe. Negative on bad arguments. */
int64_t b200_ingest_encode_packed(const b200_ingest* t, const char* text, const int64_t* begin, const int64_t* end,
                                  int32_t n, int32_t max_len, int32_t threads, uint32_t* ids_packed,
                                  uint32_t* type_ids_packed, int64_t capacity_tokens, int32_t* cu_seqlens);

/* -- resident cosine index: the search step after the path (f4; CUDA, no CPU fallback) ---------------------------
 * Exact cosine top-k over embeddings kept in HBM, as an alternative to the Qdrant hop of vector_memory_service
 * (vector_memory_service/src/main.rs:34-37 collection with Distance::Cosine; :140-202 upsert of one vector per
 * sentence; :261-286 SearchPoints{vector, limit = top_k, no filter, no score threshold}; :319-403 results in
 * descending score order). Like Qdrant for Distance::Cosine, vectors are normalised on insert and the query on
 * search, and the score is their dot product (f32). Row numbers (insertion order, from 0) are the ids: the service
 * keeps its id -> payload map as it keeps Qdrant's point ids today. Errors: b200_index_last_error(). */
typedef struct b200_index b200_index;
const char* b200_index_last_error(void);
/* dim: 128, 256, 384, 512, 768 or 1024. capacity_hint: vectors to reserve room for (the index grows by doubling). */
int32_t b200_index_create(int32_t dim, int64_t capacity_hint, int32_t device, b200_index** out);
void b200_index_destroy(b200_index* ix);
int64_t b200_index_size(const b200_index* ix);
int32_t b200_index_dim(const b200_index* ix);
/* Appends n vectors ([n, dim] f32 row-major, host or device memory); they get ids size() .. size() + n - 1. */
int32_t b200_index_add(b200_index* ix, const float* vectors, int64_t n);
int32_t b200_index_add_device(b200_index* ix, const float* device_vectors, int64_t n);
/* q queries ([q, dim] f32, host) -> ids_out / scores_out [q, top_k] (host), best first; equal scores are ordered by
 * id; with fewer than top_k vectors in the index the rest is (-1, -inf). 1 <= top_k <= 1024. Blocks until done. */
int32_t b200_index_search(b200_index* ix, const float* queries, int32_t q, int32_t top_k, int64_t* ids_out,
                          float* scores_out);
/* Device time of the scan kernel(s) of the last search on this index, in ms (CUDA events on the index's stream). */
float b200_index_last_scan_ms(const b200_index* ix);
/* Optional: keep a bf16 copy of the rows (+50 % memory) and answer single-query searches with top_k <= 64 in two stages
 * (the case of the reference: one query per request; other calls keep the one-stage f32 scan): the scan reads the
 * copy (half the bytes), the best max(32, 4 top_k) candidates are rescored from the f32 rows. Rounding a unit row to
 * bf16 moves a score by at most 2^-8, so a candidate list whose last approximate score lies more than 2^-7 below the
 * top_k-th provably holds the whole answer; otherwise the search is repeated on the f32 matrix
 * (b200_index_exact_fallbacks counts those). Ids and scores are the same bits as without the copy. */
int32_t b200_index_enable_prescan(b200_index* ix);
int64_t b200_index_exact_fallbacks(const b200_index* ix);
/* The index as a file and back (the Qdrant collection of the reference is persistent; this is how a restarted service
 * avoids embedding its corpus again): the normalised rows, bit for bit, so a loaded index answers exactly like the saved
 * one. B200_EMBED_ERR_IO when the file cannot be written / read or is not an index file. */
int32_t b200_index_save(b200_index* ix, const char* path);
int32_t b200_index_load(const char* path, int32_t device, b200_index** out);

/* -- raw text in, JSON payload out (f1 -> path -> f3 in one call) ---------------------------------------------------
 * What process_raw_text_message and the publish of its result do for one RawTextMessage
 * (preprocessing_service/src/main.rs:28-142): clean and split the text, tokenise, embed, write the
 * TextWithEmbeddingsMessage JSON - with the three stages overlapped over chunks of `chunk_sentences` sentences (0 = 8192):
 * the calling thread tokenises chunk c + 1 while a second thread has the GPU embed chunk c and a third formats chunk c - 1
 * into `out`. Needs a WordPiece b200_ingest handle. out_cap must cover an upper bound that depends on the number of
 * sentences; when it does not (or out is NULL) the call returns B200_EMBED_ERR_INVALID_ARG with the bound in *out_len
 * and the sentence count in *n_sentences - allocate and call again. An empty cleaned text is the reference's
 * "Cleaned text is empty" error. Message of the last failure: b200_pipeline_last_error(). */
const char* b200_pipeline_last_error(void);
int32_t b200_pipeline_text_to_json(b200_embed* h, const b200_ingest* t, const char* raw_text, int64_t len,
                                   const char* original_id, const char* source_url, const char* model_name,
                                   uint64_t timestamp_ms, int32_t chunk_sentences, uint32_t flags, char* out,
                                   int64_t out_cap, int64_t* out_len, int64_t* n_sentences);

/* -- corpus ingest over all GPUs of one box (BASELINE config 5) --------------------------------------------------------------
 * The multi-GPU form of the service's ingest loop, which spawns one task per incoming message onto one shared generator
 * (preprocessing_service/src/main.rs:355-393) and tokenises inside the embedding call (embedding_generator.rs:146-164): one
 * process, one model replica + one host worker thread per GPU, a pool of tokeniser threads, no data-path collective (each
 * GPU reads host memory and its own HBM only). A SHARD is the sentences of one message or a window of a corpus, given as
 * byte ranges of one UTF-8 text (what b200_ingest_split_sentences produces). b200_corpus_submit returns a ticket at once;
 * the shard is tokenised in chunks by the pool, its sentences are ordered by token count and cut into micro-batches
 * (<= max_batch_sentences sentences, <= max_batch_tokens tokens) that the GPU workers take from one queue on demand, two
 * requests in flight per GPU (b200_embed_submit_varlen) over pinned staging buffers; finished rows are scattered back by
 * sentence index, so out[i] belongs to sentence i whatever GPU computed it. text / begin / end / out must stay valid until
 * b200_corpus_wait(ticket) returns; any number of shards may be in flight, from any threads. Needs a b200_ingest
 * tokenizer (borrowed, must outlive the corpus handle). Errors: b200_corpus_last_error(). */
#define B200_CORPUS_MAX_GPUS 16
typedef struct b200_corpus b200_corpus;
typedef struct b200_corpus_stats {
    int64_t sentences, tokens, micro_batches;
    int32_t tokenizer_threads, gpus;
    double tokenize_busy_s; /* summed over the tokeniser threads */
    double plan_s;          /* length sort + micro-batch cut */
    double gather_s, gpu_wait_s, scatter_s; /* summed over the GPU worker threads: staging copies, blocked in wait, result copies */
    double per_gpu_active_s[B200_CORPUS_MAX_GPUS]; /* first submit .. last retire since the last reset */
    int64_t per_gpu_sentences[B200_CORPUS_MAX_GPUS], per_gpu_tokens[B200_CORPUS_MAX_GPUS], per_gpu_batches[B200_CORPUS_MAX_GPUS];
} b200_corpus_stats;
const char* b200_corpus_last_error(void);
/* devices: the CUDA ordinals to use (1 .. 16). tokenizer_threads 0 = host cores minus one per GPU. max_batch_sentences 0 =
 * 16384, max_batch_tokens 0 = 512 Ki (pinned staging per GPU: 2 x (8 B per token + (4 * hidden + 4) B per sentence)). */
int32_t b200_corpus_create(const b200_embed_config* cfg, const char* safetensors_path, uint64_t seed, const int32_t* devices,
                           int32_t n_devices, const b200_ingest* tokenizer, int32_t tokenizer_threads,
                           int32_t max_batch_sentences, int32_t max_batch_tokens, b200_corpus** out);
void b200_corpus_destroy(b200_corpus* c);
int32_t b200_corpus_submit(b200_corpus* c, const char* text, const int64_t* begin, const int64_t* end, int64_t n, uint32_t flags,
                           float* out /* [n, hidden] f32, host */, int64_t* ticket);
int32_t b200_corpus_wait(b200_corpus* c, int64_t ticket);
int32_t b200_corpus_stats_read(b200_corpus* c, b200_corpus_stats* out, int32_t reset);
/* The planner on its own (host arithmetic, no GPU): order[n] = the sentences by ascending token count (stable), cut into
 * micro-batches of at most max_batch_sentences sentences and max_batch_tokens tokens; micro-batch k is
 * order[batch_begin[k] .. batch_begin[k + 1]). Returns the number of micro-batches (batch_begin is written when
 * batch_begin_cap >= that + 1); negative on bad arguments. */
int64_t b200_corpus_plan(const int32_t* lengths, int64_t n, int32_t max_len, int32_t max_batch_sentences,
                         int64_t max_batch_tokens, int32_t* order, int64_t* batch_begin, int64_t batch_begin_cap);
/* The plan b200_corpus_submit itself makes, same arguments and results: sentences longer than 128 tokens as above; those
 * of at most 128 tokens (the shape the fused QKV + attention kernel takes) cut into micro-batches in INPUT order and,
 * inside each micro-batch, ordered tile by tile - best-fit-decreasing packing into 128-token tiles, the tiles one after the
 * other - so that the kernel's tiles of whole consecutive sentences come out full. */
int64_t b200_corpus_plan_packed(const int32_t* lengths, int64_t n, int32_t max_len, int32_t max_batch_sentences,
                                int64_t max_batch_tokens, int32_t* order, int64_t* batch_begin, int64_t batch_begin_cap);

#ifdef __cplusplus
}
#endif
#endif /* B200_EMBED_H_ */
