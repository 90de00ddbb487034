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
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200_embed.h"
#include "pow10_table.h"

namespace {

// shortest round-trip digits of a finite non-zero float: value = 0.d1d2...dn x 10^kk  (n = ndig), through the C++
// library (its to_chars is shortest round-trip, Ryu inside). `digits` must hold 24 bytes: copies are fixed-size.
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
        std::memcpy(digits + 1, buf + 2, 16);
        n = mant_len - 1;
    }
    *kk = (neg ? -e : e) + 1;
    return n;
}

// The same digits without the library call: Schubfach (R. Giulietti, "The Schubfach way to render doubles", 2020) for
// binary32. v = c * 2^q; k = floor(log10(2^q)) picks the decade in which the rounding interval of v is at least one unit
// wide, the interval ends and v are scaled by 10^-k with one 64 x 32 bit multiplication each against a 64-bit upper
// approximation g of the power of ten (kPow10_64, exact values from integer arithmetic), rounded to odd so that
// comparisons stay exact; the answer is the shortest decimal in the interval, the closest to v among equals, ties to
// even. Checked against the library for every finite float (scripts/probes/f32_digits_check.cpp: all 2^31 - 2^23
// positive values). About 3x faster than std::to_chars + re-layout here.
constexpr int kPow10Min = -31, kPow10Max = 45;
const uint64_t kPow10_64[kPow10Max - kPow10Min + 1] = {B200_POW10_64_VALUES};
inline int floor_log10_pow2(int e) { return (e * 1262611) >> 22; }
inline int floor_log10_three_quarters_pow2(int e) { return (e * 1262611 - 524031) >> 22; }
inline int floor_log2_pow10(int e) { return (e * 1741647) >> 19; }
inline uint32_t round_to_odd(uint64_t g, uint32_t cp) {
    const unsigned __int128 p = static_cast<unsigned __int128>(g) * cp;
    const uint32_t y1 = static_cast<uint32_t>(p >> 64);
    const uint32_t y0 = static_cast<uint32_t>(static_cast<uint64_t>(p) >> 32);
    return y1 | (y0 > 1 ? 1u : 0u);
}

// -> decimal significand (<= 9 digits, may end in zeros) and exponent: v = sig * 10^exp. v finite, > 0.
inline uint32_t schubfach_f32(float v, int* exp10) {
    uint32_t bits;
    std::memcpy(&bits, &v, 4);
    const uint32_t frac = bits & 0x7FFFFFu, E = (bits >> 23) & 0xFFu;
    uint32_t c;
    int q;
    if (E != 0) {
        c = frac | 0x800000u;
        q = static_cast<int>(E) - 150;
        if (0 <= -q && -q < 24 && (c & ((1u << -q) - 1)) == 0) {  // a small integer
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
    const uint64_t g = kPow10_64[-k - kPow10Min];
    const uint32_t vbl = round_to_odd(g, cbl << h), vb = round_to_odd(g, cb << h), vbr = round_to_odd(g, cbr << h);
    const uint32_t lower = vbl + (even ? 0 : 1), upper = vbr - (even ? 0 : 1);
    const uint32_t s = vb / 4;
    if (s >= 10) {  // a shorter candidate one decade up?
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

// digits of a finite non-zero float (see shortest_digits): n <= 9, value = 0.d1..dn x 10^kk. `digits` must hold 24 bytes.
inline int shortest_digits_fast(float v, char* digits, int* kk) {
    static const char kPairs[] =
        "00010203040506070809101112131415161718192021222324252627282930313233343536373839404142434445464748495051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";
    int e10;
    const uint32_t sig = schubfach_f32(v, &e10);
    // nine digits, zero-padded on the left, two at a time
    char t[24];
    const uint32_t top = sig / 100000000u, rest = sig - top * 100000000u;
    const uint32_t hi = rest / 10000u, lo = rest - hi * 10000u;
    t[0] = static_cast<char>('0' + top);
    std::memcpy(t + 1, kPairs + 2 * (hi / 100u), 2);
    std::memcpy(t + 3, kPairs + 2 * (hi % 100u), 2);
    std::memcpy(t + 5, kPairs + 2 * (lo / 100u), 2);
    std::memcpy(t + 7, kPairs + 2 * (lo % 100u), 2);
    int n = sig >= 100000000u ? 9 : sig >= 10000000u ? 8 : sig >= 1000000u ? 7 : sig >= 100000u ? 6 : sig >= 10000u ? 5
            : sig >= 1000u ? 4 : sig >= 100u ? 3 : sig >= 10u ? 2 : 1;
    std::memcpy(digits, t + 9 - n, 9);
    e10 += n;  // value = 0.d1..dn x 10^(n + e10)
    while (n > 1 && digits[n - 1] == '0') --n;  // small integers and carried roundings end in zeros
    *kk = e10;
    return n;
}

// ryu "pretty" layout for f32 (format32): returns characters written (at most 16). Copies are fixed-size so that no
// memcpy call is made per float: `out` must have 32 writable bytes; what lies beyond the returned length is scratch.
inline int format_f32(float v, char* out) {
    if (!std::isfinite(v)) { std::memcpy(out, "null", 4); return 4; }
    int i = 0;
    if (std::signbit(v)) { out[i++] = '-'; v = -v; }
    if (v == 0.0f) { std::memcpy(out + i, "0.0", 3); return i + 3; }
    char d[32];
    int kk;
    const int n = shortest_digits_fast(v, d, &kk);  // n <= 9
    if (-6 < kk && kk <= 0) {  // 1234e-6 -> 0.001234 : where embedding components live, so tested first
        std::memcpy(out + i, "0.000000", 8);
        i += 2 - kk;
        std::memcpy(out + i, d, 16);
        return i + n;
    }
    const int k = kk - n;  // value = digits x 10^k
    if (0 <= k && kk <= 13) {  // 1234e7 -> 12340000000.0
        std::memcpy(out + i, d, 16);
        for (int j = n; j < kk; ++j) out[i + j] = '0';
        out[i + kk] = '.';
        out[i + kk + 1] = '0';
        return i + kk + 2;
    }
    if (0 < kk && kk <= 13) {  // 1234e-2 -> 12.34
        std::memcpy(out + i, d, 16);
        out[i + kk] = '.';
        std::memcpy(out + i + kk + 1, d + kk, 8);
        return i + n + 1;
    }
    // scientific: d[.ddd]e[-]X
    out[i++] = d[0];
    if (n > 1) {
        out[i++] = '.';
        std::memcpy(out + i, d + 1, 8);
        i += n - 1;
    }
    out[i++] = 'e';
    int e = kk - 1;
    if (e < 0) { out[i++] = '-'; e = -e; }
    if (e >= 10) out[i++] = static_cast<char>('0' + e / 10);
    out[i++] = static_cast<char>('0' + e % 10);
    return i;
}

// JSON string escapes of serde_json (ESCAPE table of serde_json/src/ser.rs): \" \\ \b \f \n \r \t, \u00XX for the other
// control characters, everything else as is. Writes at p, returns the new end; at most 6 * n + 2 characters.
inline char* put_string(char* p, const char* s, size_t n) {
    static const char* hex = "0123456789abcdef";
    *p++ = '"';
    for (size_t i = 0; i < n; ++i) {
        const unsigned char c = static_cast<unsigned char>(s[i]);
        if (c >= 0x20 && c != '"' && c != '\\') { *p++ = static_cast<char>(c); continue; }
        *p++ = '\\';
        switch (c) {
            case '"': *p++ = '"'; break;
            case '\\': *p++ = '\\'; break;
            case '\b': *p++ = 'b'; break;
            case '\f': *p++ = 'f'; break;
            case '\n': *p++ = 'n'; break;
            case '\r': *p++ = 'r'; break;
            case '\t': *p++ = 't'; break;
            default:
                *p++ = 'u'; *p++ = '0'; *p++ = '0';
                *p++ = hex[c >> 4];
                *p++ = hex[c & 15];
        }
    }
    *p++ = '"';
    return p;
}
inline char* put_cstr(char* p, const char* s) { return put_string(p, s ? s : "", s ? std::strlen(s) : 0); }
inline char* put_lit(char* p, const char* lit) {
    const size_t n = std::strlen(lit);
    std::memcpy(p, lit, n);
    return p + n;
}
// Every float takes at most 16 characters ("-1.17549435e-38")
inline char* put_vector(char* p, const float* v, int dim) {
    *p++ = '[';
    for (int j = 0; j < dim; ++j) {
        if (j) *p++ = ',';
        p += format_f32(v[j], p);
    }
    *p++ = ']';
    return p;
}
inline size_t vector_bound(int dim) { return static_cast<size_t>(dim) * 17 + 2 + 32; }  // + the scratch format_f32 may touch
inline size_t string_bound(size_t n) { return n * 6 + 2; }

struct Sentences {  // either NUL-terminated strings or byte ranges of one text
    const char* const* strs;
    const char* text;
    const int64_t* begin;
    const int64_t* end;
    const char* ptr(int i) const { return strs ? (strs[i] ? strs[i] : "") : text + begin[i]; }
    size_t len(int i) const { return strs ? (strs[i] ? std::strlen(strs[i]) : 0) : static_cast<size_t>(end[i] - begin[i]); }
};

// The document is produced in slices: every thread formats a contiguous run of sentences into its own (uninitialised)
// buffer sized by the upper bounds above, then copies it to its final offset in `out`; only the sizes are exchanged.
// The items `{"sentence_text":..,"embedding":[..]}` of sentences [0, n), comma-separated, formatted in slices: every thread
// formats a contiguous run of sentences into its own (uninitialised) buffer sized by the upper bounds above. `first`
// tells whether item 0 is the first of its array (no comma in front).
struct ItemParts {
    int t = 1;
    std::vector<std::unique_ptr<char[]>> parts;
    std::vector<size_t> len;
    size_t total = 0;
    template <typename F>
    void run(F&& fn) const {
        if (t <= 1) { fn(0); return; }
        std::vector<std::thread> pool;
        for (int k = 0; k < t; ++k) pool.emplace_back(fn, k);
        for (auto& th : pool) th.join();
    }
    void copy_to(char* out) const {  // the slices to their final offsets, in parallel
        std::vector<size_t> off(static_cast<size_t>(t) + 1, 0);
        for (int k = 0; k < t; ++k) off[static_cast<size_t>(k) + 1] = off[static_cast<size_t>(k)] + len[static_cast<size_t>(k)];
        run([&](int k) { std::memcpy(out + off[static_cast<size_t>(k)], parts[static_cast<size_t>(k)].get(), len[static_cast<size_t>(k)]); });
    }
};

void format_items(const Sentences& sn, const float* embeddings, int32_t n, int32_t dim, int32_t threads, bool first, ItemParts* ip) {
    int t = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (t < 1) t = 1;
    if (t > n / 8) t = n / 8 > 0 ? n / 8 : 1;
    ip->t = t;
    ip->parts.resize(static_cast<size_t>(t));
    ip->len.assign(static_cast<size_t>(t), 0);
    ip->run([&](int k) {
        const int lo = static_cast<int>(static_cast<int64_t>(n) * k / t), hi = static_cast<int>(static_cast<int64_t>(n) * (k + 1) / t);
        size_t bound = 0;
        for (int i = lo; i < hi; ++i) bound += string_bound(sn.len(i)) + vector_bound(dim) + 40;
        ip->parts[static_cast<size_t>(k)].reset(new char[bound + 1]);
        char* p = ip->parts[static_cast<size_t>(k)].get();
        char* const p0 = p;
        for (int i = lo; i < hi; ++i) {
            if (i || !first) *p++ = ',';
            p = put_lit(p, "{\"sentence_text\":");
            p = put_string(p, sn.ptr(i), sn.len(i));
            p = put_lit(p, ",\"embedding\":");
            p = put_vector(p, embeddings + static_cast<size_t>(i) * dim, dim);
            *p++ = '}';
        }
        ip->len[static_cast<size_t>(k)] = static_cast<size_t>(p - p0);
    });
    ip->total = 0;
    for (size_t l : ip->len) ip->total += l;
}

// `{"original_id":..,"source_url":..,"embeddings_data":[` and `],"model_name":..,"timestamp_ms":..}`
void envelope(const char* original_id, const char* source_url, const char* model_name, uint64_t timestamp_ms, std::string* head,
              std::string* tail) {
    std::vector<char> hb(string_bound(std::strlen(original_id ? original_id : "")) + string_bound(std::strlen(source_url ? source_url : "")) + 64);
    char* h = hb.data();
    h = put_lit(h, "{\"original_id\":");
    h = put_cstr(h, original_id);
    h = put_lit(h, ",\"source_url\":");
    h = put_cstr(h, source_url);
    h = put_lit(h, ",\"embeddings_data\":[");
    head->assign(hb.data(), static_cast<size_t>(h - hb.data()));
    std::vector<char> tb(string_bound(std::strlen(model_name ? model_name : "")) + 64);
    char* q = tb.data();
    q = put_lit(q, "],\"model_name\":");
    q = put_cstr(q, model_name);
    q = put_lit(q, ",\"timestamp_ms\":");
    q = std::to_chars(q, q + 24, timestamp_ms).ptr;
    *q++ = '}';
    tail->assign(tb.data(), static_cast<size_t>(q - tb.data()));
}

int64_t text_with_embeddings(const char* original_id, const char* source_url, const Sentences& sn, const float* embeddings,
                             int32_t n, int32_t dim, const char* model_name, uint64_t timestamp_ms, int32_t threads, char* out,
                             int64_t out_cap) {
    ItemParts ip;
    format_items(sn, embeddings, n, dim, threads, true, &ip);
    std::string head, tail;
    envelope(original_id, source_url, model_name, timestamp_ms, &head, &tail);
    const int64_t total = static_cast<int64_t>(head.size() + ip.total + tail.size());
    if (out != nullptr && out_cap >= total) {
        std::memcpy(out, head.data(), head.size());
        ip.copy_to(out + head.size());
        std::memcpy(out + head.size() + ip.total, tail.data(), tail.size());
    }
    return total;
}

}  // namespace

extern "C" {

int32_t b200_json_format_f32(float value, char* out24) {
    if (!out24) return -1;
    char tmp[48];
    const int n = format_f32(value, tmp);
    std::memcpy(out24, tmp, static_cast<size_t>(n));
    return n;
}

int64_t b200_json_text_with_embeddings(const char* original_id, const char* source_url, const char* const* sentences,
                                       const float* embeddings, int32_t n, int32_t dim, const char* model_name,
                                       uint64_t timestamp_ms, int32_t threads, char* out, int64_t out_cap) {
    if (n < 0 || dim < 0 || (n > 0 && (!sentences || !embeddings))) return -1;
    const Sentences sn{sentences, nullptr, nullptr, nullptr};
    return text_with_embeddings(original_id, source_url, sn, embeddings, n, dim, model_name, timestamp_ms, threads, out, out_cap);
}

int64_t b200_json_text_with_embeddings_ranges(const char* original_id, const char* source_url, const char* text,
                                              const int64_t* begin, const int64_t* end, const float* embeddings, int32_t n,
                                              int32_t dim, const char* model_name, uint64_t timestamp_ms, int32_t threads,
                                              char* out, int64_t out_cap) {
    if (n < 0 || dim < 0 || (n > 0 && (!text || !begin || !end || !embeddings))) return -1;
    for (int32_t i = 0; i < n; ++i)
        if (begin[i] < 0 || end[i] < begin[i]) return -1;
    const Sentences sn{nullptr, text, begin, end};
    return text_with_embeddings(original_id, source_url, sn, embeddings, n, dim, model_name, timestamp_ms, threads, out, out_cap);
}

int64_t b200_json_envelope(const char* original_id, const char* source_url, const char* model_name, uint64_t timestamp_ms,
                           char* head, int64_t head_cap, int64_t* head_len, char* tail, int64_t tail_cap, int64_t* tail_len) {
    if (!head_len || !tail_len) return -1;
    std::string hs, ts;
    envelope(original_id, source_url, model_name, timestamp_ms, &hs, &ts);
    *head_len = static_cast<int64_t>(hs.size());
    *tail_len = static_cast<int64_t>(ts.size());
    if (head && head_cap >= *head_len) std::memcpy(head, hs.data(), hs.size());
    if (tail && tail_cap >= *tail_len) std::memcpy(tail, ts.data(), ts.size());
    return *head_len + *tail_len;
}

int64_t b200_json_sentence_items_ranges(const char* text, const int64_t* begin, const int64_t* end, const float* embeddings,
                                        int32_t n, int32_t dim, int32_t first, int32_t threads, char* out, int64_t out_cap) {
    if (n < 0 || dim < 0 || (n > 0 && (!text || !begin || !end || !embeddings))) return -1;
    for (int32_t i = 0; i < n; ++i)
        if (begin[i] < 0 || end[i] < begin[i]) return -1;
    const Sentences sn{nullptr, text, begin, end};
    ItemParts ip;
    format_items(sn, embeddings, n, dim, threads, first != 0, &ip);
    if (out != nullptr && out_cap >= static_cast<int64_t>(ip.total)) ip.copy_to(out);
    return static_cast<int64_t>(ip.total);
}

int64_t b200_json_query_embedding_result(const char* request_id, const float* embedding_or_null, int32_t dim,
                                         const char* model_name_or_null, const char* error_message_or_null, char* out,
                                         int64_t out_cap) {
    if (dim < 0) return -1;
    auto len = [](const char* s) { return s ? std::strlen(s) : size_t(0); };
    std::vector<char> buf(string_bound(len(request_id)) + vector_bound(dim) + string_bound(len(model_name_or_null)) +
                          string_bound(len(error_message_or_null)) + 96);
    char* p = buf.data();
    p = put_lit(p, "{\"request_id\":");
    p = put_cstr(p, request_id);
    p = put_lit(p, ",\"embedding\":");
    p = embedding_or_null ? put_vector(p, embedding_or_null, dim) : put_lit(p, "null");
    p = put_lit(p, ",\"model_name\":");
    p = model_name_or_null ? put_cstr(p, model_name_or_null) : put_lit(p, "null");
    p = put_lit(p, ",\"error_message\":");
    p = error_message_or_null ? put_cstr(p, error_message_or_null) : put_lit(p, "null");
    *p++ = '}';
    const int64_t total = static_cast<int64_t>(p - buf.data());
    if (out != nullptr && out_cap >= total) std::memcpy(out, buf.data(), static_cast<size_t>(total));
    return total;
}

}  // extern "C"
