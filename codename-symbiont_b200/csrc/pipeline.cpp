// Raw text in, the JSON payload out: what process_raw_text_message + the publish of its result do for one RawTextMessage
// (/root/reference/services/preprocessing_service/src/main.rs:28-142), with the stages this library provides overlapped
// over chunks of sentences instead of run one after the other:
//     split (main.rs:28-63)  ->  per chunk:  tokenise (embedding_generator.rs:160-187)
//                                            -> embed (embedding_generator.rs:189-214, on the GPU)
//                                            -> JSON items (serde_json::to_vec, main.rs:138)
// The calling thread tokenises chunk c + 1 while a second thread has the GPU embed chunk c and a third formats the items
// of chunk c - 1 straight into the caller's buffer; every stage is internally multi-threaded where it is CPU work. Host
// only: everything below goes through the C ABI of the other files.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/b200_embed.h"

namespace {

thread_local std::string g_pipe_err;
int fail(int code, const std::string& msg) {
    g_pipe_err = msg;
    return code;
}

// Pinned staging buffers for the embeddings of a chunk, kept for the life of the process: pinning memory costs about as
// much per megabyte as copying it several times, so a service that converts message after message must not pay it per call.
struct PinnedPool {
    static constexpr size_t kMaxIdleBytes = size_t(1) << 30;  // idle buffers kept at most; the smallest go first beyond that
    std::mutex mu;
    std::vector<std::pair<void*, size_t>> free_list, in_use;
    void* acquire(size_t bytes) {
        {
            std::lock_guard<std::mutex> lk(mu);
            size_t best = free_list.size();
            for (size_t i = 0; i < free_list.size(); ++i)  // best fit: the smallest idle buffer that is large enough
                if (free_list[i].second >= bytes && (best == free_list.size() || free_list[i].second < free_list[best].second)) best = i;
            if (best < free_list.size()) {
                in_use.push_back(free_list[best]);
                void* p = free_list[best].first;
                free_list.erase(free_list.begin() + static_cast<long>(best));
                return p;
            }
        }
        void* p = nullptr;
        if (b200_embed_host_alloc(bytes, &p) != B200_EMBED_OK) return nullptr;
        std::lock_guard<std::mutex> lk(mu);
        in_use.emplace_back(p, bytes);
        return p;
    }
    void release(void* p) {
        std::vector<void*> drop;
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t i = 0; i < in_use.size(); ++i)
                if (in_use[i].first == p) {
                    free_list.push_back(in_use[i]);
                    in_use.erase(in_use.begin() + static_cast<long>(i));
                    break;
                }
            size_t idle = 0;
            for (auto& f : free_list) idle += f.second;
            while (idle > kMaxIdleBytes && !free_list.empty()) {  // a grown chunk size leaves small buffers behind: unpin them
                size_t k = 0;
                for (size_t i = 1; i < free_list.size(); ++i)
                    if (free_list[i].second < free_list[k].second) k = i;
                idle -= free_list[k].second;
                drop.push_back(free_list[k].first);
                free_list.erase(free_list.begin() + static_cast<long>(k));
            }
        }
        for (void* d : drop) b200_embed_host_free(d);
    }
};
PinnedPool g_pool;

struct Chunk {
    int lo = 0, hi = 0;                   // sentences [lo, hi)
    std::vector<uint32_t> ids, types;     // packed tokens
    std::vector<int32_t> cu;              // [hi - lo + 1]
    int64_t tokens = 0;
    int max_len = 0;
};

}  // namespace

extern "C" {

const char* b200_pipeline_last_error(void) { return g_pipe_err.c_str(); }

static int32_t pipeline_text_to_json_impl(b200_embed* h, const b200_ingest* t, const char* raw_text, int64_t len,
                                   const char* original_id, const char* source_url, const char* model_name,
                                   uint64_t timestamp_ms, int32_t chunk_sentences, uint32_t flags, char* out, int64_t out_cap,
                                   int64_t* out_len, int64_t* n_sentences) {
    if (!h || !t || !raw_text || len < 0 || !out_len || !n_sentences) return fail(B200_EMBED_ERR_INVALID_ARG, "bad argument");
    *out_len = 0;
    *n_sentences = 0;
    if (chunk_sentences <= 0) chunk_sentences = 8192;
    const bool trace = getenv("B200_PIPELINE_TRACE") != nullptr;  // stage timestamps (ms since entry) on stderr
    const auto t_begin = std::chrono::steady_clock::now();
    auto now_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    const int dim = b200_embed_hidden_size(h), P = b200_embed_max_positions(h);
    // 1. split
    std::unique_ptr<char[]> cleaned(new char[static_cast<size_t>(len) + 1]);
    int64_t clen = 0, cap = len / 16 + 16;
    std::vector<int64_t> begin, end;
    for (;;) {
        begin.resize(static_cast<size_t>(cap));
        end.resize(static_cast<size_t>(cap));
        const int64_t n = b200_ingest_split_sentences(raw_text, len, cleaned.get(), len, &clen, begin.data(), end.data(), cap);
        if (n < 0) return fail(static_cast<int>(n), b200_ingest_last_error());
        if (n <= cap) { cap = n; break; }
        cap = n;
    }
    const int64_t n = cap;
    *n_sentences = n;
    if (trace) std::fprintf(stderr, "[pipeline] split done %.1f ms, %lld sentences\n", now_ms(), static_cast<long long>(n));
    if (n == 0) return fail(B200_EMBED_ERR_INVALID_ARG, "Cleaned text is empty");  // main.rs:33-39
    if (n > INT32_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "too many sentences for one message");
    // capacity the caller must provide: every float takes at most 17 characters, every text byte at most 6
    int64_t hl = 0, tl = 0;
    b200_json_envelope(original_id, source_url, model_name, timestamp_ms, nullptr, 0, &hl, nullptr, 0, &tl);
    const int64_t bound = hl + tl + n * (static_cast<int64_t>(dim) * 17 + 80) + 6 * clen + 64;
    if (!out || out_cap < bound) {
        *out_len = bound;
        return fail(B200_EMBED_ERR_INVALID_ARG, "output buffer too small: " + std::to_string(bound) + " bytes needed (reported in out_len)");
    }
    // the embeddings of chunk c live in pinned staging buffer c % 3 until its JSON items are written
    constexpr int NBUF = 3;
    const size_t buf_bytes = static_cast<size_t>(std::min<int64_t>(n, chunk_sentences)) * dim * sizeof(float);
    float* ebuf[NBUF] = {nullptr, nullptr, nullptr};
    for (int i = 0; i < NBUF; ++i) {
        ebuf[i] = static_cast<float*>(g_pool.acquire(buf_bytes));
        if (!ebuf[i]) {
            for (int j = 0; j < i; ++j) g_pool.release(ebuf[j]);
            return fail(B200_EMBED_ERR_OOM, b200_embed_last_error());
        }
    }
    b200_json_envelope(original_id, source_url, model_name, timestamp_ms, out, hl, &hl, out + out_cap - tl, tl, &tl);  // tail parked at the end

    // 2. three stages over chunks, handed on through counters
    const int nchunks = static_cast<int>((n + chunk_sentences - 1) / chunk_sentences);
    std::vector<Chunk> chunks(static_cast<size_t>(nchunks));
    std::mutex mu;
    std::condition_variable cv;
    int tokenised = 0, embedded = 0, formatted = 0;
    std::atomic<int> rc{B200_EMBED_OK};  // written under mu, read by the stage loops without it
    std::string err;
    auto set_error = [&](int code, const char* msg) {
        std::lock_guard<std::mutex> lk(mu);
        if (rc.load() == B200_EMBED_OK) { err = msg; rc.store(code); }
        cv.notify_all();
    };
    std::thread gpu([&] {
      try {
        for (int c = 0; c < nchunks; ++c) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return (tokenised > c && formatted >= c - (NBUF - 1)) || rc != B200_EMBED_OK; });
                if (rc != B200_EMBED_OK) return;
            }
            Chunk& ch = chunks[static_cast<size_t>(c)];
            const int r = b200_embed_forward_varlen(h, ch.ids.data(), ch.types.data(), ch.cu.data(), ch.hi - ch.lo,
                                                    static_cast<int32_t>(ch.tokens), ch.max_len, flags, ebuf[c % NBUF]);
            if (r != B200_EMBED_OK) { set_error(r, b200_embed_last_error()); return; }
            if (trace) std::fprintf(stderr, "[pipeline] chunk %d embedded %.1f ms\n", c, now_ms());
            std::vector<uint32_t>().swap(ch.ids);
            std::vector<uint32_t>().swap(ch.types);
            {
                std::lock_guard<std::mutex> lk(mu);
                embedded = c + 1;
            }
            cv.notify_all();
        }
      } catch (const std::exception& e) { set_error(B200_EMBED_ERR_OOM, e.what()); }
    });
    int64_t at = hl;  // write position in `out` (only the JSON thread moves it)
    std::thread json;
    try {
    json = std::thread([&] {
      try {
        for (int c = 0; c < nchunks; ++c) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return embedded > c || rc != B200_EMBED_OK; });
                if (rc != B200_EMBED_OK) return;
            }
            const Chunk& ch = chunks[static_cast<size_t>(c)];
            const int64_t w = b200_json_sentence_items_ranges(cleaned.get(), begin.data() + ch.lo, end.data() + ch.lo,
                                                              ebuf[c % NBUF], ch.hi - ch.lo, dim, c == 0 ? 1 : 0,
                                                              0, out + at, out_cap - tl - at);
            if (w < 0 || at + w > out_cap - tl) { set_error(B200_EMBED_ERR_INVALID_ARG, "JSON items did not fit the output buffer"); return; }
            at += w;
            {
                std::lock_guard<std::mutex> lk(mu);
                formatted = c + 1;
            }
            cv.notify_all();
            if (trace) std::fprintf(stderr, "[pipeline] chunk %d formatted %.1f ms\n", c, now_ms());
        }
      } catch (const std::exception& e) { set_error(B200_EMBED_ERR_OOM, e.what()); }
    });
    } catch (const std::exception& e) {  // the second thread could not be started: stop the first, report
        set_error(B200_EMBED_ERR_OOM, e.what());
        gpu.join();
        for (int i = 0; i < NBUF; ++i) g_pool.release(ebuf[i]);
        return fail(rc.load(), err);
    }
    try {
    for (int c = 0; c < nchunks; ++c) {  // this thread tokenises, at most two chunks ahead of the GPU
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return embedded >= c - 1 || rc != B200_EMBED_OK; });
            if (rc != B200_EMBED_OK) break;
        }
        Chunk& ch = chunks[static_cast<size_t>(c)];
        ch.lo = static_cast<int>(static_cast<int64_t>(c) * chunk_sentences);
        ch.hi = static_cast<int>(std::min<int64_t>(n, static_cast<int64_t>(ch.lo) + chunk_sentences));
        const int nb = ch.hi - ch.lo;
        ch.cu.resize(static_cast<size_t>(nb) + 1);
        int64_t room = static_cast<int64_t>(nb) * 40 + 64;
        for (;;) {
            ch.ids.resize(static_cast<size_t>(room));
            ch.types.resize(static_cast<size_t>(room));
            const int64_t total = b200_ingest_encode_packed(t, cleaned.get(), begin.data() + ch.lo, end.data() + ch.lo, nb, P, 0,
                                                            ch.ids.data(), ch.types.data(), room, ch.cu.data());
            if (total < 0) { set_error(static_cast<int>(total), b200_ingest_last_error()); break; }
            ch.tokens = total;
            if (total <= room) break;
            room = total;
        }
        if (rc != B200_EMBED_OK) break;
        ch.max_len = 1;
        for (int i = 0; i < nb; ++i) ch.max_len = std::max(ch.max_len, ch.cu[static_cast<size_t>(i) + 1] - ch.cu[static_cast<size_t>(i)]);
        if (trace) std::fprintf(stderr, "[pipeline] chunk %d tokenised %.1f ms\n", c, now_ms());
        {
            std::lock_guard<std::mutex> lk(mu);
            tokenised = c + 1;
        }
        cv.notify_all();
    }
    } catch (const std::exception& e) { set_error(B200_EMBED_ERR_OOM, e.what()); }
    gpu.join();
    json.join();
    for (int i = 0; i < NBUF; ++i) g_pool.release(ebuf[i]);
    if (rc != B200_EMBED_OK) return fail(rc, err);
    std::memmove(out + at, out + out_cap - tl, static_cast<size_t>(tl));
    *out_len = at + tl;
    return B200_EMBED_OK;
}

int32_t b200_pipeline_text_to_json(b200_embed* h, const b200_ingest* t, const char* raw_text, int64_t len,
                                   const char* original_id, const char* source_url, const char* model_name,
                                   uint64_t timestamp_ms, int32_t chunk_sentences, uint32_t flags, char* out, int64_t out_cap,
                                   int64_t* out_len, int64_t* n_sentences) {
    try {  // nothing may leave through the C boundary (allocation failures of the buffers set up before the stages start)
        return pipeline_text_to_json_impl(h, t, raw_text, len, original_id, source_url, model_name, timestamp_ms, chunk_sentences,
                                          flags, out, out_cap, out_len, n_sentences);
    } catch (const std::exception& e) {
        return fail(B200_EMBED_ERR_OOM, std::string("pipeline: ") + e.what());
    }
}

}  // extern "C"
