/*
 * oracle/bert_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, OpenMP) of the one hot path of makkenzo/codename-symbiont:
 * the sentence-embedding forward of `preprocessing_service`.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library. The CUDA product (codename-symbiont_b200/csrc) never
 * links, loads or calls it.
 *
 * PARITY STATUS: "parity unpinned" by the reference itself. The reference ships no
 * golden vector, known-answer test or fixture for this path
 * (libs/shared_models/src/lib.rs:123-537 only round-trips JSON schemas), and the
 * arithmetic lives in un-vendored crates that cannot be built here (no cargo/rustc):
 *   candle-transformers = "0.9.1"  (models/bert.rs: BertModel)     Cargo.toml:22
 *   candle-nn           = "0.9.1"  (Linear, Embedding, LayerNorm, ops::softmax)  :21
 *   candle-core         = "0.9.1"  (matmul via `gemm`, gelu_erf)   :20
 * (services/preprocessing_service/Cargo.toml). The restatement follows candle's
 * published BERT op order (which itself ports HuggingFace BERT) and is pinned in THIS
 * repo against HuggingFace `transformers` BertModel run in the build container
 * (oracle/make_golden.py -> tests/golden/ fixtures).
 *
 * Reference call sites restated here (paths relative to /root/reference):
 *   services/preprocessing_service/src/embedding_generator.rs
 *     :146-148  chunks of 8 sentences                      -> oracle_generate()
 *     :83-91    fixed right padding to max_position_embeddings (done by the caller)
 *     :189-191  ids / mask / type_ids as (b, P) u32 tensors
 *     :198      BertModel::forward(ids, type_ids, Some(mask)) -> oracle_bert_forward()
 *     :201-207  masked mean-pool with +1e-9                -> oracle_mean_pool()
 *   candle-transformers 0.9.1 models/bert.rs (not vendored):
 *     BertEmbeddings::forward        -> oracle_embeddings()
 *     get_extended_attention_mask    -> (1-m)*f32::MIN additive bias, oracle_attention()
 *     BertSelfAttention::forward     -> oracle_linear() x3 + oracle_attention()
 *     BertSelfOutput / BertOutput    -> oracle_linear() + oracle_add_layer_norm()
 *     BertIntermediate (gelu_erf)    -> oracle_linear() + oracle_gelu_erf()
 *
 * Build: see oracle/Makefile. `real` is float (liboracle_f32.so, the candle dtype,
 * DTYPE=F32) or double (-DORACLE_F64, liboracle_f64.so, used to judge which f32
 * implementation is closer to the exact result).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef ORACLE_F64
typedef double real;
#define R_SQRT sqrt
#define R_EXP exp
#define R_ERF erf
#define R_TANH tanh
#else
typedef float real;
#define R_SQRT sqrtf
#define R_EXP expf
#define R_ERF erff
#define R_TANH tanhf
#endif

enum { ORACLE_ACT_GELU_ERF = 0, ORACLE_ACT_GELU_TANH = 1, ORACLE_ACT_RELU = 2 };

typedef struct {
    int32_t vocab_size;
    int32_t hidden_size;
    int32_t num_layers;
    int32_t num_heads;
    int32_t intermediate_size;
    int32_t max_position_embeddings;
    int32_t type_vocab_size;
    int32_t hidden_act; /* ORACLE_ACT_* ; candle HiddenAct::{Gelu,GeluApproximate,Relu} */
    float layer_norm_eps;
} oracle_config;

/* One encoder layer; every matrix is [out, in] row-major f32 (HF / candle Linear layout). */
typedef struct {
    const float *q_w, *q_b, *k_w, *k_b, *v_w, *v_b;
    const float *ao_w, *ao_b, *ao_ln_g, *ao_ln_b; /* attention.output.dense / LayerNorm */
    const float *i_w, *i_b;                       /* intermediate.dense */
    const float *o_w, *o_b, *o_ln_g, *o_ln_b;     /* output.dense / LayerNorm */
} oracle_layer;

typedef struct {
    const float *word_emb, *pos_emb, *type_emb, *emb_ln_g, *emb_ln_b;
    const oracle_layer *layers;
} oracle_weights;

int oracle_real_bytes(void) { return (int)sizeof(real); }

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---- LayerNorm over the last dim; biased variance; (x-mu)/sqrt(var+eps)*g+b ----
 * candle-nn LayerNorm (ops::layer_norm): mean and variance in f32 over hidden_size. */
static void layer_norm_rows(real *x, int64_t rows, int H, const float *g, const float *b, float eps) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; ++r) {
        real *p = x + r * H;
        real mu = 0;
        for (int c = 0; c < H; ++c) mu += p[c];
        mu /= (real)H;
        real var = 0;
        for (int c = 0; c < H; ++c) { real d = p[c] - mu; var += d * d; }
        var /= (real)H;
        real rstd = (real)1 / R_SQRT(var + (real)eps);
        for (int c = 0; c < H; ++c) p[c] = (p[c] - mu) * rstd * (real)g[c] + (real)b[c];
    }
}

/* BertEmbeddings::forward: word[ids] + type[tt], broadcast_add pos[0..S-1], LayerNorm.
 * Position ids are 0..S-1 for every row (candle; NOT the RoBERTa padding_idx offset). */
void oracle_embeddings(const oracle_config *cfg, const oracle_weights *w, const uint32_t *ids,
                       const uint32_t *type_ids, int B, int S, real *out) {
    const int H = cfg->hidden_size;
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < (int64_t)B * S; ++t) {
        const int j = (int)(t % S);
        const float *we = w->word_emb + (int64_t)ids[t] * H;
        const float *te = w->type_emb + (int64_t)type_ids[t] * H;
        const float *pe = w->pos_emb + (int64_t)j * H;
        real *o = out + t * H;
        for (int c = 0; c < H; ++c) o[c] = ((real)we[c] + (real)te[c]) + (real)pe[c];
    }
    layer_norm_rows(out, (int64_t)B * S, H, w->emb_ln_g, w->emb_ln_b, cfg->layer_norm_eps);
}

/* candle-nn Linear: y = x . W^T + b, W [N,K] row-major. wt is W transposed to [K,N] so the
 * inner loop runs along N (vectorisable axpy); accumulation over k is in increasing order. */
static void transpose_w(const float *w, int N, int K, real *wt) {
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k)
        for (int n = 0; n < N; ++n) wt[(int64_t)k * N + n] = (real)w[(int64_t)n * K + k];
}

static void linear_t(const real *x, int64_t M, int K, const real *wt, const float *bias, int N, real *y) {
#pragma omp parallel for schedule(static)
    for (int64_t m0 = 0; m0 < M; m0 += 4) {
        const int mr = (int)((M - m0) < 4 ? (M - m0) : 4);
        for (int i = 0; i < mr; ++i) {
            real *yr = y + (m0 + i) * N;
            for (int n = 0; n < N; ++n) yr[n] = 0;
        }
        for (int k = 0; k < K; ++k) {
            const real *wr = wt + (int64_t)k * N;
            for (int i = 0; i < mr; ++i) {
                const real a = x[(m0 + i) * K + k];
                real *yr = y + (m0 + i) * N;
                for (int n = 0; n < N; ++n) yr[n] += a * wr[n];
            }
        }
        for (int i = 0; i < mr; ++i) {
            real *yr = y + (m0 + i) * N;
            for (int n = 0; n < N; ++n) yr[n] += (real)bias[n];
        }
    }
}

void oracle_linear(const real *x, int64_t M, int K, const float *w, const float *bias, int N, real *y) {
    real *wt = (real *)malloc(sizeof(real) * (size_t)N * K);
    transpose_w(w, N, K, wt);
    linear_t(x, M, K, wt, bias, N, y);
    free(wt);
}

/* candle gelu_erf: 0.5*x*(1+erf(x/sqrt(2))); gelu (GeluApproximate): tanh form; relu. */
void oracle_activation(real *x, int64_t n, int act) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        const real v = x[i];
        if (act == ORACLE_ACT_GELU_ERF) {
            x[i] = (real)0.5 * v * ((real)1 + R_ERF(v * (real)0.70710678118654752440));
        } else if (act == ORACLE_ACT_GELU_TANH) {
            const real c = (real)0.79788456080286535588; /* sqrt(2/pi) */
            x[i] = (real)0.5 * v * ((real)1 + R_TANH(c * (v + (real)0.044715 * v * v * v)));
        } else {
            x[i] = v > 0 ? v : 0;
        }
    }
}

/* BertSelfAttention core after the three Linear layers.
 * q,k,v: [B,S,H] with head h at columns h*d..(h+1)*d. scores = q.k^T / sqrt(d);
 * scores += (1-mask[b,j]) * f32::MIN (get_extended_attention_mask); softmax over keys with
 * max subtraction (candle_nn::ops::softmax); ctx = P.v, heads merged back to [B,S,H]. */
void oracle_attention(const real *q, const real *k, const real *v, const uint32_t *mask, int B, int S,
                      int H, int heads, real *ctx) {
    const int d = H / heads;
    const real inv_sqrt_d = (real)(1.0 / sqrt((double)d));
    const real neg = (real)(-FLT_MAX); /* f32::MIN */
#pragma omp parallel
    {
        real *p = (real *)malloc(sizeof(real) * (size_t)S);
#pragma omp for collapse(2) schedule(static)
        for (int b = 0; b < B; ++b) {
            for (int h = 0; h < heads; ++h) {
                for (int i = 0; i < S; ++i) {
                    const real *qi = q + ((int64_t)b * S + i) * H + h * d;
                    real mx = -INFINITY;
                    for (int j = 0; j < S; ++j) {
                        const real *kj = k + ((int64_t)b * S + j) * H + h * d;
                        real s = 0;
                        for (int c = 0; c < d; ++c) s += qi[c] * kj[c];
                        s = s * inv_sqrt_d;
                        s = s + ((real)1 - (real)mask[(int64_t)b * S + j]) * neg;
                        p[j] = s;
                        if (s > mx) mx = s;
                    }
                    real sum = 0;
                    for (int j = 0; j < S; ++j) { p[j] = R_EXP(p[j] - mx); sum += p[j]; }
                    real *o = ctx + ((int64_t)b * S + i) * H + h * d;
                    for (int c = 0; c < d; ++c) o[c] = 0;
                    for (int j = 0; j < S; ++j) {
                        const real pj = p[j] / sum;
                        const real *vj = v + ((int64_t)b * S + j) * H + h * d;
                        for (int c = 0; c < d; ++c) o[c] += pj * vj[c];
                    }
                }
            }
        }
        free(p);
    }
}

/* y = LayerNorm(y + residual): BertSelfOutput::forward / BertOutput::forward. */
void oracle_add_layer_norm(real *y, const real *residual, int64_t rows, int H, const float *g, const float *b,
                           float eps) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < rows * H; ++i) y[i] += residual[i];
    layer_norm_rows(y, rows, H, g, b, eps);
}

void oracle_layer_norm(real *x, int64_t rows, int H, const float *g, const float *b, float eps) {
    layer_norm_rows(x, rows, H, g, b, eps);
}

/* BertModel::forward(ids, type_ids, Some(mask)) -> last hidden state [B,S,H].
 * stop_after_layer < 0: run all layers. hidden_out is `real`. */
int oracle_bert_forward(const oracle_config *cfg, const oracle_weights *w, const uint32_t *ids,
                        const uint32_t *type_ids, const uint32_t *mask, int B, int S, real *hidden_out) {
    const int H = cfg->hidden_size, I = cfg->intermediate_size, L = cfg->num_layers;
    const int64_t M = (int64_t)B * S;
    if (S > cfg->max_position_embeddings) return -1;
    for (int64_t t = 0; t < M; ++t) {
        if (ids[t] >= (uint32_t)cfg->vocab_size) return -2;
        if (type_ids[t] >= (uint32_t)cfg->type_vocab_size) return -3;
    }
    real *x = hidden_out;
    real *q = (real *)malloc(sizeof(real) * (size_t)M * H);
    real *k = (real *)malloc(sizeof(real) * (size_t)M * H);
    real *v = (real *)malloc(sizeof(real) * (size_t)M * H);
    real *ctx = (real *)malloc(sizeof(real) * (size_t)M * H);
    real *a = (real *)malloc(sizeof(real) * (size_t)M * H);
    real *inter = (real *)malloc(sizeof(real) * (size_t)M * I);
    if (!q || !k || !v || !ctx || !a || !inter) return -4;

    oracle_embeddings(cfg, w, ids, type_ids, B, S, x);
    for (int l = 0; l < L; ++l) {
        const oracle_layer *ly = &w->layers[l];
        oracle_linear(x, M, H, ly->q_w, ly->q_b, H, q);
        oracle_linear(x, M, H, ly->k_w, ly->k_b, H, k);
        oracle_linear(x, M, H, ly->v_w, ly->v_b, H, v);
        oracle_attention(q, k, v, mask, B, S, H, cfg->num_heads, ctx);
        oracle_linear(ctx, M, H, ly->ao_w, ly->ao_b, H, a);
        oracle_add_layer_norm(a, x, M, H, ly->ao_ln_g, ly->ao_ln_b, cfg->layer_norm_eps);
        oracle_linear(a, M, H, ly->i_w, ly->i_b, I, inter);
        oracle_activation(inter, M * I, cfg->hidden_act);
        oracle_linear(inter, M, I, ly->o_w, ly->o_b, H, x);
        oracle_add_layer_norm(x, a, M, H, ly->o_ln_g, ly->o_ln_b, cfg->layer_norm_eps);
    }
    free(q); free(k); free(v); free(ctx); free(a); free(inter);
    return 0;
}

/* embedding_generator.rs:201-207: out[b,:] = sum_j x[b,j,:]*m[b,j] / (sum_j m[b,j] + 1e-9).
 * The 1e-9 is added in f32 exactly as the reference does (Tensor::from_slice(&[1e-9f32])). */
void oracle_mean_pool(const real *hidden, const uint32_t *mask, int B, int S, int H, float *out) {
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; ++b) {
        real cnt = 0;
        for (int j = 0; j < S; ++j) cnt += (real)mask[(int64_t)b * S + j];
        const real den = cnt + (real)1e-9f;
        for (int c = 0; c < H; ++c) {
            real s = 0;
            for (int j = 0; j < S; ++j)
                s += hidden[((int64_t)b * S + j) * H + c] * (real)mask[(int64_t)b * S + j];
            out[(int64_t)b * H + c] = (float)(s / den);
        }
    }
}

/* north_star addition (absent from the reference): v / max(||v||_2, 1e-12). */
void oracle_l2_normalize(float *out, int B, int H) {
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; ++b) {
        double ss = 0;
        for (int c = 0; c < H; ++c) ss += (double)out[(int64_t)b * H + c] * out[(int64_t)b * H + c];
        double n = sqrt(ss);
        if (n < 1e-12) n = 1e-12;
        for (int c = 0; c < H; ++c) out[(int64_t)b * H + c] = (float)(out[(int64_t)b * H + c] / n);
    }
}

/* EmbeddingGenerator::generate_sentence_embeddings after tokenisation
 * (embedding_generator.rs:134-223): N sentences already padded to `P` tokens each, processed
 * in chunks of `chunk` (reference: 8), order preserved, out [N,H] f32.
 * normalize != 0 appends the north_star L2 step. Returns 0 or a negative error. */
int oracle_generate(const oracle_config *cfg, const oracle_weights *w, const uint32_t *ids,
                    const uint32_t *type_ids, const uint32_t *mask, int N, int P, int chunk, int normalize,
                    float *out) {
    const int H = cfg->hidden_size;
    if (N == 0) return 0; /* :135-137 */
    if (chunk <= 0) chunk = 8;
    real *hidden = (real *)malloc(sizeof(real) * (size_t)chunk * P * H);
    if (!hidden) return -4;
    for (int s0 = 0; s0 < N; s0 += chunk) {
        const int b = (N - s0) < chunk ? (N - s0) : chunk;
        const int64_t off = (int64_t)s0 * P;
        int rc = oracle_bert_forward(cfg, w, ids + off, type_ids + off, mask + off, b, P, hidden);
        if (rc) { free(hidden); return rc; }
        oracle_mean_pool(hidden, mask + off, b, P, H, out + (int64_t)s0 * H);
    }
    free(hidden);
    if (normalize) oracle_l2_normalize(out, N, H);
    return 0;
}
