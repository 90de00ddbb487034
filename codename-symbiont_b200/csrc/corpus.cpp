// Corpus ingest over all GPUs of one box (BASELINE config 5): one process, one model replica and one
// host worker thread per GPU, a pool of tokeniser threads, no data-path collective.
//
// The reference fans incoming messages out as one task per message onto one shared generator
// (/root/reference/services/preprocessing_service/src/main.rs:355-393) and tokenises inside the embedding call, eight
// sentences at a time (embedding_generator.rs:146-164). Here a SHARD (the sentences of one message, or a window of a
// corpus) is submitted and returns a ticket at once:
//   tokeniser pool   chunks of 2048 sentences -> packed token ids (b200_ingest_encode_packed, one thread per chunk)
//   planner          when a shard's last chunk is done: counting sort of its sentences by token count, cut into
//                    micro-batches under a sentence and a token budget (short sentences share 128-row tiles on the GPU,
//                    so batches of similar lengths pack best), appended to ONE queue
//   GPU workers      one thread per GPU pulls the next micro-batch from that queue (work is handed out on demand, so GPUs
//                    finish together whatever the length mix), gathers its tokens into a pinned slot, submits it
//                    (b200_embed_submit_varlen, two in flight per GPU: the copy-in / compute / copy-out of one batch
//                    overlaps the host work of the next) and scatters the finished rows back to the shard's output by
//                    sentence index
// Results are in input order whatever GPU computed them. Each GPU reads host memory and its own HBM only.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/b200_embed.h"

namespace {

thread_local std::string g_corpus_err;
int fail(int code, const std::string& msg) {
    g_corpus_err = msg;
    return code;
}
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

constexpr int CHUNK_SENTENCES = 2048;
constexpr int MAX_GPUS = B200_CORPUS_MAX_GPUS;

struct Chunk {
    std::vector<uint32_t> ids, types;
    std::vector<int32_t> cu;  // [sentences of the chunk + 1]
};

struct Shard {
    int64_t ticket = 0;
    const char* text = nullptr;
    const int64_t *begin = nullptr, *end = nullptr;
    int64_t n = 0;
    uint32_t flags = 0;
    float* out = nullptr;
    std::vector<Chunk> chunks;
    std::atomic<int> chunks_left{0};
    std::atomic<int> batches_left{0};
    int rc = B200_EMBED_OK;  // guarded by the corpus mutex
    std::string err;
    bool done = false;
};

struct MicroBatch {
    Shard* shard = nullptr;
    std::vector<int32_t> sentences;  // indices into the shard, ascending token count
    int64_t tokens = 0;
    int max_len = 0;
};

struct Slot {  // pinned staging of one micro-batch
    uint32_t *ids = nullptr, *types = nullptr;
    int32_t* cu = nullptr;
    float* out = nullptr;
    int64_t ticket = 0;
    std::unique_ptr<MicroBatch> mb;
    double t_submit = 0;
};

}  // namespace

struct b200_corpus {
    int dim = 0, max_len = 0, n_gpus = 0;
    int max_batch_sentences = 0;
    int64_t max_batch_tokens = 0;
    const b200_ingest* tok = nullptr;
    std::vector<b200_embed*> handles;
    std::vector<std::thread> tok_threads, gpu_threads;
    std::mutex mu;
    std::condition_variable cv_tok, cv_gpu, cv_done;
    std::deque<std::pair<Shard*, int>> tok_q;
    std::deque<std::unique_ptr<MicroBatch>> gpu_q;
    std::unordered_map<int64_t, std::unique_ptr<Shard>> shards;
    int64_t next_ticket = 1;
    bool stop = false;
    b200_corpus_stats stats{};  // guarded by mu
    double gpu_first[MAX_GPUS] = {}, gpu_last[MAX_GPUS] = {};  // first submit / last retire since the last reset (guarded by mu)
    std::vector<std::vector<Slot>> slots;  // [gpu][slot] pinned staging
};

namespace {

void shard_error(b200_corpus* c, Shard* sh, int code, const std::string& msg) {
    std::lock_guard<std::mutex> lk(c->mu);
    if (sh->rc == B200_EMBED_OK) { sh->rc = code; sh->err = msg; }
}

void shard_batch_done(b200_corpus* c, Shard* sh) {
    if (sh->batches_left.fetch_sub(1) == 1) {
        std::lock_guard<std::mutex> lk(c->mu);
        sh->done = true;
        c->cv_done.notify_all();
    }
}

// The plan of one shard: `order` = its sentences by ascending token count (counting sort, stable: equal lengths keep
// their input order), cut into runs of at most max_sentences sentences and max_tokens tokens; begins[k] .. begins[k + 1]
// is micro-batch k. Pure host arithmetic (also exported as b200_corpus_plan).
void plan_lengths(const int32_t* len, int64_t n, int32_t max_len, int32_t max_sentences, int64_t max_tokens,
                  std::vector<int32_t>& order, std::vector<int64_t>& begins) {
    std::vector<int64_t> count(static_cast<size_t>(max_len) + 2, 0);
    for (int64_t i = 0; i < n; ++i) ++count[static_cast<size_t>(std::min(std::max(len[i], 0), max_len)) + 1];
    for (size_t l = 1; l < count.size(); ++l) count[l] += count[l - 1];
    order.resize(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        order[static_cast<size_t>(count[static_cast<size_t>(std::min(std::max(len[i], 0), max_len))]++)] = static_cast<int32_t>(i);
    begins.clear();
    int64_t tokens = 0, sentences = 0;
    for (int64_t k = 0; k < n; ++k) {
        const int l = len[order[static_cast<size_t>(k)]];
        if (k == 0 || sentences >= max_sentences || tokens + l > max_tokens) {
            begins.push_back(k);
            tokens = 0;
            sentences = 0;
        }
        tokens += l;
        ++sentences;
    }
    begins.push_back(n);
}

// The engine's fused attention (K23) works on tiles of 128 token rows that hold WHOLE consecutive sentences, packed
// greedily in the order they are sent. Sorted by length that order fills the tiles to 83 % (corpus length law; 84 % in
// random order); sending a micro-batch tile by tile - best-fit-decreasing bin packing into 128-token bins, then the bins one
// after the other - fills them to 99 % (96 % after the packing kernel's 128-sentence segments). `sentences`: ascending by
// length on entry. Only the order inside the micro-batch changes; results are scattered back by sentence index as before.
void order_by_tiles(std::vector<int32_t>& sentences, const int32_t* len, int cap) {
    const size_t n = sentences.size();
    std::vector<std::vector<int32_t>> by_room(static_cast<size_t>(cap) + 1);  // open bins by remaining room
    std::vector<std::vector<int32_t>> bins;
    bins.reserve(n / 3 + 1);
    for (size_t k = n; k-- > 0;) {  // longest first
        const int32_t i = sentences[k];
        const int l = std::min(std::max(static_cast<int>(len[static_cast<size_t>(i)]), 1), cap);
        int room = l;
        while (room <= cap && by_room[static_cast<size_t>(room)].empty()) ++room;  // the fullest bin it still fits
        int32_t b;
        if (room <= cap) {
            b = by_room[static_cast<size_t>(room)].back();
            by_room[static_cast<size_t>(room)].pop_back();
        } else {
            b = static_cast<int32_t>(bins.size());
            bins.emplace_back();
            room = cap;
        }
        bins[static_cast<size_t>(b)].push_back(i);
        if (room - l > 0) by_room[static_cast<size_t>(room - l)].push_back(b);
    }
    size_t at = 0;
    for (const auto& bin : bins)
        for (int32_t i : bin) sentences[at++] = i;
}

// The plan the corpus driver uses (also exported as b200_corpus_plan_packed): plan_lengths, then the sentences of at most
// 128 tokens recut in input order and every micro-batch of such sentences put into tile order.
void plan_packed(const int32_t* len, int64_t n, int32_t max_len, int32_t max_sentences, int64_t max_tokens,
                 std::vector<int32_t>& order, std::vector<int64_t>& begins) {
    plan_lengths(len, n, max_len, max_sentences, max_tokens, order, begins);
    // Sentences of at most 128 tokens go through K23, whose tiles want long and short sentences side by side (order_by_tiles
    // below): they are cut into micro-batches in INPUT order - a length-sorted micro-batch holds nothing short enough to fill
    // the gaps its long sentences leave (89 % full tiles instead of 99 %). Longer sentences keep the length order.
    {
        int64_t n_short = 0;
        while (n_short < n && len[order[static_cast<size_t>(n_short)]] <= 128) ++n_short;
        if (n_short > 0) {
            std::sort(order.begin(), order.begin() + n_short);  // back to input order
            std::vector<int64_t> cut;
            int64_t tokens = 0, sentences = 0;
            for (int64_t k = 0; k < n_short; ++k) {
                const int l = len[order[static_cast<size_t>(k)]];
                if (k == 0 || sentences >= max_sentences || tokens + l > max_tokens) {
                    cut.push_back(k);
                    tokens = 0;
                    sentences = 0;
                }
                tokens += l;
                ++sentences;
            }
            // the long sentences' micro-batches: those of the length-ordered plan that start at or after n_short, plus a
            // cut at n_short itself
            if (n_short < n) cut.push_back(n_short);
            for (int64_t b : begins)
                if (b > n_short) cut.push_back(b);
            if (cut.back() != n) cut.push_back(n);
            begins.swap(cut);
        }
    }
    for (size_t b = 0; b + 1 < begins.size(); ++b) {
        int32_t longest = 0;
        for (int64_t k = begins[b]; k < begins[b + 1]; ++k) longest = std::max(longest, len[order[static_cast<size_t>(k)]]);
        if (longest <= 128) {  // the shape K23 takes
            std::vector<int32_t> part(order.begin() + begins[b], order.begin() + begins[b + 1]);
            std::sort(part.begin(), part.end(), [&](int32_t x, int32_t y) { return len[x] != len[y] ? len[x] < len[y] : x < y; });
            order_by_tiles(part, len, 128);
            std::copy(part.begin(), part.end(), order.begin() + begins[b]);
        }
    }
}

// All chunks of the shard are tokenised: order its sentences by token count and cut the micro-batches.
void plan_shard(b200_corpus* c, Shard* sh) {
    const int64_t n = sh->n;
    std::vector<int32_t> len(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        const Chunk& ch = sh->chunks[static_cast<size_t>(i / CHUNK_SENTENCES)];
        const int k = static_cast<int>(i % CHUNK_SENTENCES);
        len[static_cast<size_t>(i)] = ch.cu[static_cast<size_t>(k) + 1] - ch.cu[static_cast<size_t>(k)];
    }
    std::vector<int32_t> order;
    std::vector<int64_t> begins;
    plan_packed(len.data(), n, c->max_len, c->max_batch_sentences, c->max_batch_tokens, order, begins);
    std::vector<std::unique_ptr<MicroBatch>> made;
    for (size_t b = 0; b + 1 < begins.size(); ++b) {
        std::unique_ptr<MicroBatch> mb(new MicroBatch());
        mb->shard = sh;
        mb->sentences.assign(order.begin() + begins[b], order.begin() + begins[b + 1]);
        for (int32_t i : mb->sentences) {
            mb->tokens += len[static_cast<size_t>(i)];
            mb->max_len = std::max(mb->max_len, len[static_cast<size_t>(i)]);
        }
        made.push_back(std::move(mb));
    }
    sh->batches_left.store(static_cast<int>(made.size()));
    {
        std::lock_guard<std::mutex> lk(c->mu);
        c->stats.micro_batches += static_cast<int64_t>(made.size());
        for (auto& m : made) c->gpu_q.push_back(std::move(m));
    }
    c->cv_gpu.notify_all();
}

void tokenizer_loop(b200_corpus* c) {
    for (;;) {
        std::pair<Shard*, int> job;
        {
            std::unique_lock<std::mutex> lk(c->mu);
            c->cv_tok.wait(lk, [&] { return c->stop || !c->tok_q.empty(); });
            if (c->tok_q.empty()) return;  // stop
            job = c->tok_q.front();
            c->tok_q.pop_front();
        }
        Shard* sh = job.first;
        const double t0 = now_s();
        int64_t tokens = 0;
        try {
            const int64_t lo = static_cast<int64_t>(job.second) * CHUNK_SENTENCES;
            const int nb = static_cast<int>(std::min<int64_t>(CHUNK_SENTENCES, sh->n - lo));
            Chunk& ch = sh->chunks[static_cast<size_t>(job.second)];
            ch.cu.resize(static_cast<size_t>(nb) + 1);
            int64_t room = static_cast<int64_t>(nb) * 40 + 64;
            for (;;) {
                ch.ids.resize(static_cast<size_t>(room));
                ch.types.resize(static_cast<size_t>(room));
                const int64_t total = b200_ingest_encode_packed(c->tok, sh->text, sh->begin + lo, sh->end + lo, nb, c->max_len, 1,
                                                                ch.ids.data(), ch.types.data(), room, ch.cu.data());
                if (total < 0) {
                    shard_error(c, sh, static_cast<int>(total), b200_ingest_last_error());
                    for (auto& v : ch.cu) v = 0;  // an empty chunk keeps the shard's bookkeeping consistent
                    break;
                }
                tokens = total;
                if (total <= room) break;
                room = total;
            }
        } catch (const std::exception& e) {
            shard_error(c, sh, B200_EMBED_ERR_OOM, std::string("tokeniser thread: ") + e.what());
        }
        {
            std::lock_guard<std::mutex> lk(c->mu);
            c->stats.tokenize_busy_s += now_s() - t0;
            c->stats.tokens += tokens;
        }
        if (sh->chunks_left.fetch_sub(1) == 1) {
            const double t1 = now_s();
            try {
                plan_shard(c, sh);
            } catch (const std::exception& e) {
                shard_error(c, sh, B200_EMBED_ERR_OOM, std::string("planner: ") + e.what());
                std::lock_guard<std::mutex> lk(c->mu);
                sh->done = true;
                c->cv_done.notify_all();
            }
            std::lock_guard<std::mutex> lk(c->mu);
            c->stats.plan_s += now_s() - t1;
        }
    }
}

// Finishes the request in `s`: waits for the GPU, scatters the rows to the shard's output.
void retire_slot(b200_corpus* c, int g, Slot& s) {
    if (!s.mb) return;
    Shard* sh = s.mb->shard;
    const double t0 = now_s();
    const int rc = b200_embed_wait(c->handles[static_cast<size_t>(g)], s.ticket);
    const double t1 = now_s();
    if (rc != B200_EMBED_OK) {
        shard_error(c, sh, rc, b200_embed_last_error());
    } else {
        const size_t row = static_cast<size_t>(c->dim) * sizeof(float);
        const size_t nb = s.mb->sentences.size();
        for (size_t k = 0; k < nb; ++k)
            std::memcpy(sh->out + static_cast<size_t>(s.mb->sentences[k]) * c->dim, s.out + k * c->dim, row);
    }
    const double t2 = now_s();
    {
        std::lock_guard<std::mutex> lk(c->mu);
        c->stats.gpu_wait_s += t1 - t0;
        c->stats.scatter_s += t2 - t1;
        c->stats.per_gpu_sentences[g] += static_cast<int64_t>(s.mb->sentences.size());
        c->stats.per_gpu_tokens[g] += s.mb->tokens;
        c->stats.per_gpu_batches[g] += 1;
        c->stats.sentences += static_cast<int64_t>(s.mb->sentences.size());
        c->gpu_last[g] = t2;
    }
    shard_batch_done(c, sh);
    s.mb.reset();
}

void gpu_loop(b200_corpus* c, int g, Slot* slots, int nslots) {
    int k = 0;
    for (;;) {
        std::unique_ptr<MicroBatch> mb;
        {
            std::unique_lock<std::mutex> lk(c->mu);
            // while this GPU still has requests in flight do not sleep on an empty queue: retire them first
            bool inflight = false;
            for (int i = 0; i < nslots; ++i) inflight = inflight || slots[i].mb != nullptr;
            if (!inflight) c->cv_gpu.wait(lk, [&] { return c->stop || !c->gpu_q.empty(); });
            if (!c->gpu_q.empty()) {
                mb = std::move(c->gpu_q.front());
                c->gpu_q.pop_front();
            } else if (!inflight) {
                break;  // stop
            }
        }
        if (!mb) {  // nothing queued: drain what is in flight, oldest first
            for (int i = 0; i < nslots; ++i) retire_slot(c, g, slots[(k + i) % nslots]);
            continue;
        }
        Slot& s = slots[k % nslots];
        retire_slot(c, g, s);  // the slot's previous request (submitted nslots batches ago)
        Shard* sh = mb->shard;
        const double t0 = now_s();
        try {
            int64_t at = 0;
            const size_t nb = mb->sentences.size();
            s.cu[0] = 0;
            for (size_t i = 0; i < nb; ++i) {
                const int32_t idx = mb->sentences[i];
                const Chunk& ch = sh->chunks[static_cast<size_t>(idx / CHUNK_SENTENCES)];
                const int j = idx % CHUNK_SENTENCES;
                const int32_t a = ch.cu[static_cast<size_t>(j)], l = ch.cu[static_cast<size_t>(j) + 1] - a;
                std::memcpy(s.ids + at, ch.ids.data() + a, static_cast<size_t>(l) * 4);
                std::memcpy(s.types + at, ch.types.data() + a, static_cast<size_t>(l) * 4);
                at += l;
                s.cu[i + 1] = static_cast<int32_t>(at);
            }
            const double t1 = now_s();
            int rc = B200_EMBED_OK;
            {
                std::lock_guard<std::mutex> lk(c->mu);
                rc = sh->rc;  // a shard that already failed is not computed any further
                c->stats.gather_s += t1 - t0;
                if (c->gpu_first[g] == 0) c->gpu_first[g] = t0;
            }
            if (rc == B200_EMBED_OK)
                rc = b200_embed_submit_varlen(c->handles[static_cast<size_t>(g)], s.ids, s.types, s.cu, static_cast<int32_t>(nb),
                                              static_cast<int32_t>(at), mb->max_len, sh->flags, s.out, &s.ticket);
            if (rc != B200_EMBED_OK) {
                shard_error(c, sh, rc, b200_embed_last_error());
                shard_batch_done(c, sh);
            } else {
                s.mb = std::move(mb);
                s.t_submit = t1;
            }
        } catch (const std::exception& e) {
            shard_error(c, sh, B200_EMBED_ERR_OOM, std::string("GPU worker: ") + e.what());
            shard_batch_done(c, sh);
        }
        ++k;
    }
}

void free_slots(std::vector<Slot>& slots) {
    for (Slot& s : slots) {
        b200_embed_host_free(s.ids);
        b200_embed_host_free(s.types);
        b200_embed_host_free(s.cu);
        b200_embed_host_free(s.out);
    }
}

}  // namespace

extern "C" {

const char* b200_corpus_last_error(void) { return g_corpus_err.c_str(); }

int32_t b200_corpus_create(const b200_embed_config* cfg, const char* safetensors_path, uint64_t seed, const int32_t* devices,
                           int32_t n_devices, const b200_ingest* tokenizer, int32_t tokenizer_threads,
                           int32_t max_batch_sentences, int32_t max_batch_tokens, b200_corpus** out) {
    if (!cfg || !devices || !tokenizer || !out || n_devices < 1 || n_devices > MAX_GPUS)
        return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_create: bad argument (1 .. 16 devices)");
    *out = nullptr;
    try {
        std::unique_ptr<b200_corpus> c(new b200_corpus());
        c->tok = tokenizer;
        c->n_gpus = n_devices;
        c->max_batch_sentences = max_batch_sentences > 0 ? max_batch_sentences : 16384;
        c->max_batch_tokens = max_batch_tokens > 0 ? max_batch_tokens : 512 * 1024;
        c->max_len = cfg->max_position_embeddings;
        if (c->max_batch_tokens < c->max_len) c->max_batch_tokens = c->max_len;
        auto destroy_handles = [&] { for (b200_embed* h : c->handles) b200_embed_destroy(h); };
        for (int g = 0; g < n_devices; ++g) {
            b200_embed* h = nullptr;
            const int rc = b200_embed_create(cfg, safetensors_path, seed, devices[g], &h);
            if (rc != B200_EMBED_OK) {
                destroy_handles();
                return fail(rc, std::string("corpus_create: device ") + std::to_string(devices[g]) + ": " + b200_embed_last_error());
            }
            c->handles.push_back(h);
        }
        c->dim = b200_embed_hidden_size(c->handles[0]);
        c->slots.resize(static_cast<size_t>(n_devices));
        constexpr int NSLOTS = 2;
        for (int g = 0; g < n_devices; ++g) {
            for (int i = 0; i < NSLOTS; ++i) {
                Slot s;
                void *a = nullptr, *b = nullptr, *d = nullptr, *e = nullptr;
                const size_t tok_bytes = static_cast<size_t>(c->max_batch_tokens) * 4;
                if (b200_embed_host_alloc(tok_bytes, &a) || b200_embed_host_alloc(tok_bytes, &b) ||
                    b200_embed_host_alloc((static_cast<size_t>(c->max_batch_sentences) + 1) * 4, &d) ||
                    b200_embed_host_alloc(static_cast<size_t>(c->max_batch_sentences) * c->dim * 4, &e)) {
                    b200_embed_host_free(a); b200_embed_host_free(b); b200_embed_host_free(d); b200_embed_host_free(e);
                    for (auto& v : c->slots) free_slots(v);
                    destroy_handles();
                    return fail(B200_EMBED_ERR_OOM, std::string("corpus_create: pinned staging: ") + b200_embed_last_error());
                }
                s.ids = static_cast<uint32_t*>(a); s.types = static_cast<uint32_t*>(b);
                s.cu = static_cast<int32_t*>(d); s.out = static_cast<float*>(e);
                c->slots[static_cast<size_t>(g)].push_back(std::move(s));
            }
        }
        int nt = tokenizer_threads > 0 ? tokenizer_threads : static_cast<int>(std::thread::hardware_concurrency()) - n_devices;
        if (nt < 1) nt = 1;
        b200_corpus* raw = c.get();
        for (int i = 0; i < nt; ++i) c->tok_threads.emplace_back(tokenizer_loop, raw);
        for (int g = 0; g < n_devices; ++g)
            c->gpu_threads.emplace_back(gpu_loop, raw, g, c->slots[static_cast<size_t>(g)].data(), NSLOTS);
        c->stats.tokenizer_threads = nt;
        c->stats.gpus = n_devices;
        *out = c.release();
        return B200_EMBED_OK;
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("corpus_create: ") + e.what());
    }
}

void b200_corpus_destroy(b200_corpus* c) {
    if (!c) return;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        c->stop = true;
    }
    c->cv_tok.notify_all();
    c->cv_gpu.notify_all();
    for (auto& t : c->tok_threads) t.join();
    for (auto& t : c->gpu_threads) t.join();
    for (auto& v : c->slots) free_slots(v);
    for (b200_embed* h : c->handles) b200_embed_destroy(h);
    delete c;
}

int32_t b200_corpus_submit(b200_corpus* c, const char* text, const int64_t* begin, const int64_t* end, int64_t n, uint32_t flags,
                           float* out, int64_t* ticket) {
    if (!c || !ticket || n < 0 || (n > 0 && (!text || !begin || !end || !out)))
        return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_submit: bad argument");
    if (n > INT32_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_submit: more than 2^31 sentences in one shard");
    *ticket = 0;
    try {
        std::unique_ptr<Shard> sh(new Shard());
        sh->text = text; sh->begin = begin; sh->end = end; sh->n = n; sh->flags = flags; sh->out = out;
        const int nchunks = static_cast<int>((n + CHUNK_SENTENCES - 1) / CHUNK_SENTENCES);
        sh->chunks.resize(static_cast<size_t>(nchunks));
        sh->chunks_left.store(nchunks);
        Shard* raw = sh.get();
        {
            std::lock_guard<std::mutex> lk(c->mu);
            sh->ticket = c->next_ticket++;
            *ticket = sh->ticket;
            if (n == 0) sh->done = true;  // empty in, empty out
            for (int k = 0; k < nchunks; ++k) c->tok_q.emplace_back(raw, k);
            c->shards[sh->ticket] = std::move(sh);
        }
        c->cv_tok.notify_all();
        return B200_EMBED_OK;
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("corpus_submit: ") + e.what());
    }
}

int32_t b200_corpus_wait(b200_corpus* c, int64_t ticket) {
    if (!c || ticket <= 0) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_wait: bad handle or ticket");
    std::unique_ptr<Shard> sh;
    {
        std::unique_lock<std::mutex> lk(c->mu);
        auto it = c->shards.find(ticket);
        if (it == c->shards.end()) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_wait: unknown ticket");
        Shard* raw = it->second.get();
        c->cv_done.wait(lk, [&] { return raw->done; });
        sh = std::move(it->second);
        c->shards.erase(it);
    }
    if (sh->rc != B200_EMBED_OK) return fail(sh->rc, sh->err);
    return B200_EMBED_OK;
}

int64_t b200_corpus_plan(const int32_t* lengths, int64_t n, int32_t max_len, int32_t max_batch_sentences,
                         int64_t max_batch_tokens, int32_t* order, int64_t* batch_begin, int64_t batch_begin_cap) {
    if (n < 0 || (n > 0 && (!lengths || !order)) || max_len <= 0 || max_batch_sentences <= 0 || max_batch_tokens <= 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_plan: bad argument");
    if (n > INT32_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_plan: more than 2^31 sentences");
    try {
        std::vector<int32_t> ord;
        std::vector<int64_t> begins;
        plan_lengths(lengths, n, max_len, max_batch_sentences, max_batch_tokens, ord, begins);
        if (n > 0) std::memcpy(order, ord.data(), static_cast<size_t>(n) * sizeof(int32_t));
        const int64_t nb = static_cast<int64_t>(begins.size()) - 1;
        if (batch_begin && batch_begin_cap >= nb + 1) std::memcpy(batch_begin, begins.data(), static_cast<size_t>(nb + 1) * sizeof(int64_t));
        return nb;
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("corpus_plan: ") + e.what());
    }
}

int64_t b200_corpus_plan_packed(const int32_t* lengths, int64_t n, int32_t max_len, int32_t max_batch_sentences,
                                int64_t max_batch_tokens, int32_t* order, int64_t* batch_begin, int64_t batch_begin_cap) {
    if (n < 0 || (n > 0 && (!lengths || !order)) || max_len <= 0 || max_batch_sentences <= 0 || max_batch_tokens <= 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_plan_packed: bad argument");
    if (n > INT32_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_plan_packed: more than 2^31 sentences");
    try {
        std::vector<int32_t> ord;
        std::vector<int64_t> begins;
        plan_packed(lengths, n, max_len, max_batch_sentences, max_batch_tokens, ord, begins);
        if (n > 0) std::memcpy(order, ord.data(), static_cast<size_t>(n) * sizeof(int32_t));
        const int64_t nb = static_cast<int64_t>(begins.size()) - 1;
        if (batch_begin && batch_begin_cap >= nb + 1) std::memcpy(batch_begin, begins.data(), static_cast<size_t>(nb + 1) * sizeof(int64_t));
        return nb;
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("corpus_plan_packed: ") + e.what());
    }
}

int32_t b200_corpus_stats_read(b200_corpus* c, b200_corpus_stats* out, int32_t reset) {
    if (!c || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "corpus_stats_read: null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    *out = c->stats;
    for (int g = 0; g < c->n_gpus; ++g) out->per_gpu_active_s[g] = c->gpu_last[g] > c->gpu_first[g] ? c->gpu_last[g] - c->gpu_first[g] : 0.0;
    if (reset) {
        for (int g = 0; g < MAX_GPUS; ++g) c->gpu_first[g] = c->gpu_last[g] = 0;
        const int nt = c->stats.tokenizer_threads, ng = c->stats.gpus;
        c->stats = b200_corpus_stats{};
        c->stats.tokenizer_threads = nt;
        c->stats.gpus = ng;
    }
    return B200_EMBED_OK;
}

}  // extern "C"
