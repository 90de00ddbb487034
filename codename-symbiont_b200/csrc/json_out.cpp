// Wire format of the embeddings: the JSON text serde_json::to_vec produces for
// shared_models::TextWithEmbeddingsMessage and QueryEmbeddingResult
// (/root/reference/libs/shared_models/src/lib.rs:39-52, :66-72; published at
// /root/reference/services/preprocessing_service/src/main.rs:138-142). A 384-d sentence is ~4 KB of decimal text, a
// 768-d one ~8 KB, so after the GPU path the formatter is the next cost on the host. Host-only code, no CUDA.
//
// serde_json writes compact JSON (no spaces), struct fields in declaration order, f32 through the `ryu` crate
// (shortest decimal that round-trips, "pretty" layout: 1.0, 0.001234, 1.2345e-7, 1e21), non-finite floats as null,
// strings with the escapes \" \\ \b \f \n \r \t and \u00XX for the other control characters, everything else as
// UTF-8. Neither crate is vendored in the reference tree: the layout rules are restated from ryu 1.x
// (src/pretty/mod.rs, format32) and checked in tests/ against a Python restatement and by exact f32 round trip.
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200_embed.h"

namespace {

// shortest round-trip digits of a finite non-zero float: value = 0.d1d2...dn x 10^kk  (n = ndig)
inline int shortest_digits(float v, char* digits, int* kk) {
    char buf[32];
    const auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    // d[.ddd]e[+-]XX : the exponent is the last 2 digits, the sign before them
    const char* end = r.ptr;
    const int e = (end[-2] - '0') * 10 + (end[-1] - '0');
    const bool neg = end[-3] == '-';
    const int mant_len = static_cast<int>(end - buf) - 4;  // without e+XX
    digits[0] = buf[0];
    int n = 1;
    if (mant_len > 1) {  // skip the '.'
        std::memcpy(digits + 1, buf + 2, static_cast<size_t>(mant_len - 2));
        n = mant_len - 1;
    }
    *kk = (neg ? -e : e) + 1;
    return n;
}

// ryu "pretty" layout for f32 (format32): returns characters written
inline int format_f32(float v, char* out) {
    if (!std::isfinite(v)) { std::memcpy(out, "null", 4); return 4; }
    int i = 0;
    if (std::signbit(v)) { out[i++] = '-'; v = -v; }
    if (v == 0.0f) { std::memcpy(out + i, "0.0", 3); return i + 3; }
    char d[16];
    int kk;
    const int n = shortest_digits(v, d, &kk);
    const int k = kk - n;  // value = digits x 10^k
    if (0 <= k && kk <= 13) {  // 1234e7 -> 12340000000.0
        std::memcpy(out + i, d, n);
        for (int j = n; j < kk; ++j) out[i + j] = '0';
        out[i + kk] = '.';
        out[i + kk + 1] = '0';
        return i + kk + 2;
    }
    if (0 < kk && kk <= 13) {  // 1234e-2 -> 12.34
        std::memcpy(out + i, d, kk);
        out[i + kk] = '.';
        std::memcpy(out + i + kk + 1, d + kk, n - kk);
        return i + n + 1;
    }
    if (-6 < kk && kk <= 0) {  // 1234e-6 -> 0.001234
        out[i++] = '0';
        out[i++] = '.';
        for (int j = 0; j < -kk; ++j) out[i++] = '0';
        std::memcpy(out + i, d, n);
        return i + n;
    }
    // scientific: d[.ddd]e[-]X
    out[i++] = d[0];
    if (n > 1) {
        out[i++] = '.';
        std::memcpy(out + i, d + 1, n - 1);
        i += n - 1;
    }
    out[i++] = 'e';
    int e = kk - 1;
    if (e < 0) { out[i++] = '-'; e = -e; }
    if (e >= 10) out[i++] = static_cast<char>('0' + e / 10);
    out[i++] = static_cast<char>('0' + e % 10);
    return i;
}

inline void append_string(std::string& o, const char* s) {
    static const char* hex = "0123456789abcdef";
    o.push_back('"');
    for (const unsigned char* p = reinterpret_cast<const unsigned char*>(s ? s : ""); *p; ++p) {
        const unsigned char c = *p;
        switch (c) {
            case '"': o += "\\\""; break;
            case '\\': o += "\\\\"; break;
            case '\b': o += "\\b"; break;
            case '\f': o += "\\f"; break;
            case '\n': o += "\\n"; break;
            case '\r': o += "\\r"; break;
            case '\t': o += "\\t"; break;
            default:
                if (c < 0x20) {
                    o += "\\u00";
                    o.push_back(hex[c >> 4]);
                    o.push_back(hex[c & 15]);
                } else {
                    o.push_back(static_cast<char>(c));
                }
        }
    }
    o.push_back('"');
}

// Every float takes at most 16 characters ("-1.17549435e-38"): grow once, then write through a raw pointer (appending
// float by float through std::string cost twice the time of the digit generation itself).
inline void append_vector(std::string& o, const float* v, int dim) {
    const size_t at = o.size();
    o.resize(at + static_cast<size_t>(dim) * 17 + 2);
    char* p = &o[at];
    *p++ = '[';
    for (int j = 0; j < dim; ++j) {
        if (j) *p++ = ',';
        p += format_f32(v[j], p);
    }
    *p++ = ']';
    o.resize(static_cast<size_t>(p - o.data()));
}

inline void append_u64(std::string& o, uint64_t v) {
    char buf[24];
    const auto r = std::to_chars(buf, buf + sizeof(buf), v);
    o.append(buf, r.ptr);
}

inline int64_t deliver(const std::string& s, char* out, int64_t cap) {
    const int64_t n = static_cast<int64_t>(s.size());
    if (out != nullptr && cap >= n) std::memcpy(out, s.data(), s.size());
    return n;
}

}  // namespace

extern "C" {

int32_t b200_json_format_f32(float value, char* out24) { return out24 ? format_f32(value, out24) : -1; }

int64_t b200_json_text_with_embeddings(const char* original_id, const char* source_url, const char* const* sentences,
                                       const float* embeddings, int32_t n, int32_t dim, const char* model_name,
                                       uint64_t timestamp_ms, int32_t threads, char* out, int64_t out_cap) {
    if (n < 0 || dim < 0 || (n > 0 && (!sentences || !embeddings))) return -1;
    // one {"sentence_text":..,"embedding":[..]} object per sentence; every thread formats a contiguous slice of the
    // sentences into its own buffer, the buffers are concatenated in order
    int t = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (t < 1) t = 1;
    if (t > n / 8) t = n / 8 > 0 ? n / 8 : 1;
    std::vector<std::string> parts(static_cast<size_t>(t));
    auto work = [&](int k) {
        const int lo = static_cast<int>(static_cast<int64_t>(n) * k / t), hi = static_cast<int>(static_cast<int64_t>(n) * (k + 1) / t);
        std::string& o = parts[static_cast<size_t>(k)];
        o.reserve(static_cast<size_t>(hi - lo) * (static_cast<size_t>(dim) * 12 + 96));
        for (int i = lo; i < hi; ++i) {
            if (i) o.push_back(',');
            o += "{\"sentence_text\":";
            append_string(o, sentences[i]);
            o += ",\"embedding\":";
            append_vector(o, embeddings + static_cast<size_t>(i) * dim, dim);
            o.push_back('}');
        }
    };
    if (t <= 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int k = 0; k < t; ++k) pool.emplace_back(work, k);
        for (auto& th : pool) th.join();
    }
    size_t total = 256;
    for (const auto& s : parts) total += s.size();
    std::string o;
    o.reserve(total + std::strlen(original_id ? original_id : "") * 6 + std::strlen(source_url ? source_url : "") * 6);
    o += "{\"original_id\":";
    append_string(o, original_id);
    o += ",\"source_url\":";
    append_string(o, source_url);
    o += ",\"embeddings_data\":[";
    for (const auto& s : parts) o += s;
    o += "],\"model_name\":";
    append_string(o, model_name);
    o += ",\"timestamp_ms\":";
    append_u64(o, timestamp_ms);
    o.push_back('}');
    return deliver(o, out, out_cap);
}

int64_t b200_json_query_embedding_result(const char* request_id, const float* embedding_or_null, int32_t dim,
                                         const char* model_name_or_null, const char* error_message_or_null, char* out,
                                         int64_t out_cap) {
    if (dim < 0) return -1;
    std::string o;
    o.reserve(static_cast<size_t>(dim) * 12 + 128);
    o += "{\"request_id\":";
    append_string(o, request_id);
    o += ",\"embedding\":";
    if (embedding_or_null) append_vector(o, embedding_or_null, dim); else o += "null";
    o += ",\"model_name\":";
    if (model_name_or_null) append_string(o, model_name_or_null); else o += "null";
    o += ",\"error_message\":";
    if (error_message_or_null) append_string(o, error_message_or_null); else o += "null";
    o.push_back('}');
    return deliver(o, out, out_cap);
}

}  // extern "C"
