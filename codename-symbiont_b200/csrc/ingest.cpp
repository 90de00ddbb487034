// Host ingest front-end: raw text -> sentences -> [n, P] u32 ids / mask / type_ids, the three arrays
// the embedding path consumes. Host-only code, multi-threaded, no CUDA. It restates two things:
//
//  * the sentence splitter of the text processor
//    (/root/reference/services/preprocessing_service/src/main.rs:28-63): split_whitespace + join(" "), cut after every
//    '.', '?' or '!', trim, keep a non-empty remainder, fall back to the whole text;
//  * what `tokenizer.encode_batch(inputs, true)` does for a BERT WordPiece tokenizer.json as the embedding generator
//    configures it (/root/reference/services/preprocessing_service/src/embedding_generator.rs:73-99 padding Fixed(P),
//    pad id of "[PAD]" or 0, truncation LongestFirst to P; :160-187 the call and the three arrays). The algorithm
//    lives in the HuggingFace `tokenizers` crate (Cargo.toml: tokenizers = "0.21.1"), which is not vendored in the
//    reference tree: added-token extraction, BertNormalizer (clean_text, handle_chinese_chars, strip_accents,
//    lowercase), BertPreTokenizer (whitespace split, punctuation isolated), WordPiece greedy longest-match with
//    "##" continuation and [UNK] for an unmatched or over-long word, [CLS] .. [SEP] template, truncation, padding.
//    The per-code-point tables (bert_unicode_tables.h) are probed out of that library by
//    scripts/gen_unicode_tables.py; tests/test_ingest.py compares ids and masks bit-exactly with the library.
//
// Not covered (b200_ingest_create fails loudly): Unigram / BPE models (the production model's XLM-R tokenizer is
// Unigram), added tokens with lstrip / rstrip / single_word / normalized set, normalizer combinations other than
// lowercase+strip_accents together or neither.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/b200_embed.h"
#include "ingest_unigram.h"
#include "bert_unicode_tables.h"

namespace {

thread_local std::string g_ingest_err;
int fail(int code, const std::string& msg) {
    g_ingest_err = msg;
    return code;
}

// ---- UTF-8 ---------------------------------------------------------------------------------------------------------
// Decodes one scalar at p (p < end). Input comes from Rust `String`s on the reference side, so it is valid UTF-8; a
// malformed byte is passed through as U+FFFD of length 1 (which clean_text then removes).
inline uint32_t utf8_next(const unsigned char* p, const unsigned char* end, int* len) {
    const uint32_t b0 = p[0];
    if (b0 < 0x80) { *len = 1; return b0; }
    if (b0 >= 0xC2 && b0 < 0xE0 && p + 1 < end && (p[1] & 0xC0) == 0x80) {
        *len = 2;
        return ((b0 & 0x1F) << 6) | (p[1] & 0x3F);
    }
    if (b0 >= 0xE0 && b0 < 0xF0 && p + 2 < end && (p[1] & 0xC0) == 0x80 && (p[2] & 0xC0) == 0x80) {
        *len = 3;
        return ((b0 & 0x0F) << 12) | ((p[1] & 0x3F) << 6) | (p[2] & 0x3F);
    }
    if (b0 >= 0xF0 && b0 < 0xF5 && p + 3 < end && (p[1] & 0xC0) == 0x80 && (p[2] & 0xC0) == 0x80 && (p[3] & 0xC0) == 0x80) {
        *len = 4;
        return ((b0 & 0x07) << 18) | ((p[1] & 0x3F) << 12) | ((p[2] & 0x3F) << 6) | (p[3] & 0x3F);
    }
    *len = 1;
    return 0xFFFD;
}

inline void utf8_put(std::string& o, uint32_t c) {
    if (c < 0x80) {
        o.push_back(static_cast<char>(c));
    } else if (c < 0x800) {
        o.push_back(static_cast<char>(0xC0 | (c >> 6)));
        o.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    } else if (c < 0x10000) {
        o.push_back(static_cast<char>(0xE0 | (c >> 12)));
        o.push_back(static_cast<char>(0x80 | ((c >> 6) & 0x3F)));
        o.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    } else {
        o.push_back(static_cast<char>(0xF0 | (c >> 18)));
        o.push_back(static_cast<char>(0x80 | ((c >> 12) & 0x3F)));
        o.push_back(static_cast<char>(0x80 | ((c >> 6) & 0x3F)));
        o.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    }
}

// Rust char::is_whitespace (the White_Space property): what split_whitespace and trim use in main.rs:28-47
inline bool rust_is_whitespace(uint32_t c) {
    if (c < 0x80) return c == ' ' || (c >= 0x09 && c <= 0x0D);
    return c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 ||
           c == 0x202F || c == 0x205F || c == 0x3000;
}

using b200::bert_tables::Range;
inline bool in_ranges(const Range* r, int n, uint32_t c) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        if (c < r[mid].lo) hi = mid - 1;
        else if (c > r[mid].hi) lo = mid + 1;
        else return true;
    }
    return false;
}

// per-code-point class, flattened once for the BMP (one byte each), ranges above it
enum : uint8_t { C_KEEP = 0, C_REMOVED = 1, C_SPACE = 2, C_CJK = 3, C_PUNCT = 4, C_HAS_LOWER = 8 };
struct CharTable {
    std::vector<uint8_t> bmp;
    CharTable() : bmp(0x10000, 0) {
        using namespace b200::bert_tables;
        auto fill = [&](const Range* r, int n, uint8_t v, bool is_flag) {
            for (int i = 0; i < n; ++i)
                for (uint32_t c = r[i].lo; c <= r[i].hi && c < 0x10000; ++c)
                    bmp[c] = is_flag ? static_cast<uint8_t>(bmp[c] | v) : v;
        };
        fill(kRemoved, kRemovedCount, C_REMOVED, false);
        fill(kToSpace, kToSpaceCount, C_SPACE, false);
        fill(kCjk, kCjkCount, C_CJK, false);
        fill(kPunct, kPunctCount, C_PUNCT, true);
        for (int i = 0; i < kLowerCount; ++i)
            if (kLowerKey[i] < 0x10000) bmp[kLowerKey[i]] |= C_HAS_LOWER;
        for (uint32_t c = 0xAC00; c <= 0xD7A3; ++c) bmp[c] |= C_HAS_LOWER;  // Hangul syllables: arithmetic NFD
    }
    uint8_t get(uint32_t c) const {
        using namespace b200::bert_tables;
        if (c < 0x10000) return bmp[c];
        uint8_t v = C_KEEP;
        if (in_ranges(kRemoved, kRemovedCount, c)) v = C_REMOVED;
        else if (in_ranges(kToSpace, kToSpaceCount, c)) v = C_SPACE;
        else if (in_ranges(kCjk, kCjkCount, c)) v = C_CJK;
        if (in_ranges(kPunct, kPunctCount, c)) v |= C_PUNCT;
        if (std::binary_search(kLowerKey, kLowerKey + kLowerCount, c)) v |= C_HAS_LOWER;
        return v;
    }
};
const CharTable& char_table() {
    static const CharTable t;
    return t;
}

// lowercase(strip_Mn(NFD(c))) appended to o as UTF-8 (c is known to have a mapping)
inline void put_lowered(std::string& o, uint32_t c) {
    using namespace b200::bert_tables;
    if (c >= 0xAC00 && c <= 0xD7A3) {
        const uint32_t s = c - 0xAC00;
        utf8_put(o, 0x1100 + s / 588);
        utf8_put(o, 0x1161 + (s % 588) / 28);
        if (s % 28) utf8_put(o, 0x11A7 + s % 28);
        return;
    }
    const uint32_t* k = std::lower_bound(kLowerKey, kLowerKey + kLowerCount, c);
    const int i = static_cast<int>(k - kLowerKey);
    for (uint32_t j = kLowerOff[i]; j < kLowerOff[i + 1]; ++j) utf8_put(o, kLowerOut[j]);
}

// ---- JSON (tokenizer.json) -----------------------------------------------------------------------------------------
struct JVal {
    enum Type { NUL, BOOL, NUM, STR, ARR, OBJ } type = NUL;
    bool b = false;
    double n = 0;
    std::string s;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;
    const JVal* find(const char* key) const {
        if (type != OBJ) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool is_true(const char* key, bool dflt) const {
        const JVal* v = find(key);
        return (v && v->type == BOOL) ? v->b : dflt;
    }
};

struct JParser {
    const char* p;
    const char* end;
    std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
    bool fail_at(const char* what) {
        if (err.empty()) err = std::string("tokenizer.json: ") + what;
        return false;
    }
    static int hex4(const char* q) {
        int v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = q[i];
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else return -1;
        }
        return v;
    }
    bool string(std::string& out) {
        ws();
        if (p >= end || *p != '"') return fail_at("expected a string");
        ++p;
        out.clear();
        while (p < end && *p != '"') {
            if (*p != '\\') { out.push_back(*p++); continue; }
            if (++p >= end) return fail_at("bad escape");
            const char e = *p++;
            switch (e) {
                case '"': out.push_back('"'); break;
                case '\\': out.push_back('\\'); break;
                case '/': out.push_back('/'); break;
                case 'b': out.push_back('\b'); break;
                case 'f': out.push_back('\f'); break;
                case 'n': out.push_back('\n'); break;
                case 'r': out.push_back('\r'); break;
                case 't': out.push_back('\t'); break;
                case 'u': {
                    if (end - p < 4) return fail_at("bad \\u escape");
                    int c = hex4(p);
                    if (c < 0) return fail_at("bad \\u escape");
                    p += 4;
                    if (c >= 0xD800 && c <= 0xDBFF && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
                        const int lo = hex4(p + 2);
                        if (lo >= 0xDC00 && lo <= 0xDFFF) {
                            c = 0x10000 + ((c - 0xD800) << 10) + (lo - 0xDC00);
                            p += 6;
                        }
                    }
                    utf8_put(out, static_cast<uint32_t>(c));
                    break;
                }
                default: return fail_at("bad escape");
            }
        }
        if (p >= end) return fail_at("unterminated string");
        ++p;
        return true;
    }
    bool value(JVal& v, int depth = 0) {
        if (depth > 64) return fail_at("nesting too deep");
        ws();
        if (p >= end) return fail_at("unexpected end");
        const char c = *p;
        if (c == '"') { v.type = JVal::STR; return string(v.s); }
        if (c == '{') {
            ++p;
            v.type = JVal::OBJ;
            ws();
            if (p < end && *p == '}') { ++p; return true; }
            for (;;) {
                std::string key;
                if (!string(key)) return false;
                ws();
                if (p >= end || *p != ':') return fail_at("expected ':'");
                ++p;
                v.obj.emplace_back(std::move(key), JVal());
                if (!value(v.obj.back().second, depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail_at("expected ',' or '}'");
            }
        }
        if (c == '[') {
            ++p;
            v.type = JVal::ARR;
            ws();
            if (p < end && *p == ']') { ++p; return true; }
            for (;;) {
                v.arr.emplace_back();
                if (!value(v.arr.back(), depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail_at("expected ',' or ']'");
            }
        }
        if (end - p >= 4 && !std::memcmp(p, "true", 4)) { p += 4; v.type = JVal::BOOL; v.b = true; return true; }
        if (end - p >= 5 && !std::memcmp(p, "false", 5)) { p += 5; v.type = JVal::BOOL; v.b = false; return true; }
        if (end - p >= 4 && !std::memcmp(p, "null", 4)) { p += 4; v.type = JVal::NUL; return true; }
        char* e = nullptr;
        v.n = strtod(p, &e);
        if (e == p) return fail_at("unexpected character");
        p = e;
        v.type = JVal::NUM;
        return true;
    }
};

// ---- vocabulary: open-addressing table keyed by the token bytes, looked up without building a std::string -----------
struct Vocab {
    std::vector<int32_t> slot;  // index into keys, -1 empty
    std::vector<std::string> keys;
    std::vector<int32_t> ids;
    uint32_t mask = 0;
    size_t max_len = 0;
    static uint64_t hash(const char* s, size_t n) {
        uint64_t h = 0xCBF29CE484222325ull;
        for (size_t i = 0; i < n; ++i) { h ^= static_cast<unsigned char>(s[i]); h *= 0x100000001B3ull; }
        return h ^ (h >> 29);
    }
    void add(const std::string& k, int32_t id) { keys.push_back(k); ids.push_back(id); if (k.size() > max_len) max_len = k.size(); }
    void finish() {
        uint32_t cap = 16;
        while (cap < keys.size() * 2 + 2) cap <<= 1;
        mask = cap - 1;
        slot.assign(cap, -1);
        for (size_t i = 0; i < keys.size(); ++i) {
            uint32_t h = static_cast<uint32_t>(hash(keys[i].data(), keys[i].size())) & mask;
            while (slot[h] >= 0) {
                if (keys[static_cast<size_t>(slot[h])] == keys[i]) break;  // duplicate key: the later id wins, as in a map
                h = (h + 1) & mask;
            }
            slot[h] = static_cast<int32_t>(i);
        }
    }
    int32_t find(const char* s, size_t n) const {
        if (n > max_len || slot.empty()) return -1;
        uint32_t h = static_cast<uint32_t>(hash(s, n)) & mask;
        while (slot[h] >= 0) {
            const std::string& k = keys[static_cast<size_t>(slot[h])];
            if (k.size() == n && !std::memcmp(k.data(), s, n)) return ids[static_cast<size_t>(slot[h])];
            h = (h + 1) & mask;
        }
        return -1;
    }
};

// ---- the tokenizer -------------------------------------------------------------------------------------------------
struct Special { int32_t id; uint32_t type_id; };
struct Added { std::string content; int32_t id; };

}  // namespace

struct b200_ingest {
    int kind = 0;  // 0 WordPiece (BERT pipeline), 1 Unigram (SentencePiece / XLM-R pipeline, ingest_unigram.h)
    b200::unigram::Model uni;
    b200::unigram::Precompiled pre;
    bool has_pre = false, collapse_spaces = false, ws_split = false, meta_prepend = true;
    std::string meta = "\xE2\x96\x81";  // Metaspace replacement (U+2581)
    Vocab whole;  // every vocabulary entry: looked up for the first piece of a word
    Vocab cont;   // entries that start with the continuation prefix, stored without it
    std::vector<Added> added;  // matched on the raw text before normalisation, leftmost-longest
    bool added_first_byte[256] = {};
    bool clean_text = true, handle_chinese = true, lowercase = true;
    int32_t unk_id = -1, pad_id = 0, vocab_size = 0;
    int max_chars = 100;
    std::vector<Special> prefix, suffix;  // template around the sentence ([CLS] .. [SEP])
    uint32_t seq_type_id = 0;
    uint8_t ascii_action[128];  // 0 append, 1 skip, 2 boundary, 3 isolate
    char ascii_out[128];
};

namespace {

enum : uint8_t { A_APPEND = 0, A_SKIP = 1, A_BOUNDARY = 2, A_ISOLATE = 3 };

// The Unigram model, its normalizer and its pre-tokenizer out of tokenizer.json (the shapes an XLM-R / SentencePiece
// conversion produces; anything else is refused by name).
bool load_unigram(const JVal& root, const JVal& model, b200_ingest* t, std::string* err) {
    const JVal* vocab = model.find("vocab");
    if (!vocab || vocab->type != JVal::ARR || vocab->arr.empty()) { *err = "tokenizer.json: Unigram model without a vocab list"; return false; }
    if (model.is_true("byte_fallback", false)) { *err = "tokenizer.json: Unigram byte_fallback is not supported"; return false; }
    const JVal* unk = model.find("unk_id");
    if (!unk || unk->type != JVal::NUM) { *err = "tokenizer.json: Unigram model without unk_id"; return false; }
    std::vector<std::pair<std::string, int32_t>> keys;
    keys.reserve(vocab->arr.size());
    t->uni.score.resize(vocab->arr.size());
    double mn = 0;
    for (size_t i = 0; i < vocab->arr.size(); ++i) {
        const JVal& e = vocab->arr[i];
        if (e.type != JVal::ARR || e.arr.size() != 2 || e.arr[0].type != JVal::STR || e.arr[1].type != JVal::NUM) {
            *err = "tokenizer.json: Unigram vocab entry is not [piece, score]";
            return false;
        }
        keys.emplace_back(e.arr[0].s, static_cast<int32_t>(i));
        t->uni.score[i] = e.arr[1].n;
        if (i == 0 || e.arr[1].n < mn) mn = e.arr[1].n;
    }
    t->uni.min_score = mn;
    t->uni.unk_id = static_cast<int32_t>(unk->n);
    if (t->uni.unk_id < 0 || static_cast<size_t>(t->uni.unk_id) >= vocab->arr.size()) { *err = "tokenizer.json: unk_id out of range"; return false; }
    t->unk_id = t->uni.unk_id;
    t->uni.build(keys);

    auto one_normalizer = [&](const JVal& nz) -> bool {
        const JVal* ty = nz.find("type");
        if (!ty || ty->type != JVal::STR) { *err = "tokenizer.json: normalizer without a type"; return false; }
        if (ty->s == "Precompiled") {
            const JVal* cm = nz.find("precompiled_charsmap");
            std::string blob;
            if (!cm || cm->type != JVal::STR || !b200::unigram::base64_decode(cm->s, &blob) || !t->pre.load(blob)) {
                *err = "tokenizer.json: bad precompiled_charsmap";
                return false;
            }
            if (t->collapse_spaces) { *err = "tokenizer.json: Precompiled after Replace is not supported"; return false; }
            t->has_pre = true;
            return true;
        }
        if (ty->s == "Replace") {
            const JVal* pat = nz.find("pattern");
            const JVal* rx = pat ? pat->find("Regex") : nullptr;
            const JVal* content = nz.find("content");
            if (!rx || rx->type != JVal::STR || rx->s != " {2,}" || !content || content->type != JVal::STR || content->s != " ") {
                *err = "tokenizer.json: the only Replace normalizer supported is Regex \" {2,}\" -> \" \"";
                return false;
            }
            t->collapse_spaces = true;
            return true;
        }
        *err = "tokenizer.json: normalizer " + ty->s + " is not supported with a Unigram model";
        return false;
    };
    if (const JVal* nz = root.find("normalizer")) {
        if (nz->type == JVal::OBJ) {
            const JVal* ty = nz->find("type");
            if (ty && ty->type == JVal::STR && ty->s == "Sequence") {
                const JVal* list = nz->find("normalizers");
                if (!list || list->type != JVal::ARR) { *err = "tokenizer.json: bad normalizer Sequence"; return false; }
                for (const JVal& one : list->arr)
                    if (!one_normalizer(one)) return false;
            } else if (!one_normalizer(*nz)) return false;
        } else if (nz->type != JVal::NUL) { *err = "tokenizer.json: bad normalizer"; return false; }
    }
    bool have_meta = false;
    auto one_pretok = [&](const JVal& pt) -> bool {
        const JVal* ty = pt.find("type");
        if (!ty || ty->type != JVal::STR) { *err = "tokenizer.json: pre_tokenizer without a type"; return false; }
        if (ty->s == "WhitespaceSplit") {
            if (have_meta) { *err = "tokenizer.json: WhitespaceSplit after Metaspace is not supported"; return false; }
            t->ws_split = true;
            return true;
        }
        if (ty->s == "Metaspace") {
            if (const JVal* r = pt.find("replacement")) if (r->type == JVal::STR && !r->s.empty()) t->meta = r->s;
            if (const JVal* sp = pt.find("split")) if (sp->type == JVal::BOOL && !sp->b) { *err = "tokenizer.json: Metaspace split = false is not supported"; return false; }
            t->meta_prepend = pt.is_true("add_prefix_space", true);  // the older spelling
            if (const JVal* ps = pt.find("prepend_scheme")) {
                if (ps->type == JVal::STR && ps->s == "always") t->meta_prepend = true;
                else if (ps->type == JVal::STR && ps->s == "never") t->meta_prepend = false;
                else { *err = "tokenizer.json: Metaspace prepend_scheme must be always or never"; return false; }
            }
            have_meta = true;
            return true;
        }
        *err = "tokenizer.json: pre_tokenizer " + ty->s + " is not supported with a Unigram model";
        return false;
    };
    const JVal* pt = root.find("pre_tokenizer");
    if (!pt || pt->type != JVal::OBJ) { *err = "tokenizer.json: a Unigram model needs a Metaspace pre_tokenizer"; return false; }
    {
        const JVal* ty = pt->find("type");
        if (ty && ty->type == JVal::STR && ty->s == "Sequence") {
            const JVal* list = pt->find("pretokenizers");
            if (!list || list->type != JVal::ARR) { *err = "tokenizer.json: bad pre_tokenizer Sequence"; return false; }
            for (const JVal& one : list->arr)
                if (!one_pretok(one)) return false;
        } else if (!one_pretok(*pt)) return false;
    }
    if (!have_meta) { *err = "tokenizer.json: a Unigram model needs a Metaspace pre_tokenizer"; return false; }
    return true;
}

bool load_tokenizer(const char* path, b200_ingest* t, std::string* err) {
    std::string text;
    {
        FILE* f = std::fopen(path, "rb");
        if (!f) { *err = std::string("cannot open ") + path; return false; }
        char buf[1 << 16];
        size_t n;
        while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
        std::fclose(f);
    }
    JVal root;
    JParser jp{text.data(), text.data() + text.size(), {}};
    if (!jp.value(root)) { *err = jp.err; return false; }
    const JVal* model = root.find("model");
    if (!model || model->type != JVal::OBJ) { *err = "tokenizer.json: no model object"; return false; }
    const JVal* mtype = model->find("type");
    const JVal* vocab = model->find("vocab");
    const bool is_unigram = (mtype && mtype->type == JVal::STR && mtype->s == "Unigram") ||
                            (!mtype && vocab && vocab->type == JVal::ARR);
    int32_t max_id = -1;
    if (is_unigram) {
        if (!load_unigram(root, *model, t, err)) return false;
        t->kind = 1;
        max_id = static_cast<int32_t>(t->uni.score.size()) - 1;
    } else {
    // old files carry no "type"; a WordPiece model is the one whose vocab is an object and that has no merges
    if ((mtype && mtype->type == JVal::STR && mtype->s != "WordPiece") || !vocab || vocab->type != JVal::OBJ || model->find("merges")) {
        *err = "tokenizer.json: only WordPiece and Unigram models are supported (got " + (mtype && mtype->type == JVal::STR ? mtype->s : std::string("an untyped model")) + ")";
        return false;
    }
    std::string prefix = "##", unk = "[UNK]";
    if (const JVal* v = model->find("continuing_subword_prefix")) if (v->type == JVal::STR) prefix = v->s;
    if (const JVal* v = model->find("unk_token")) if (v->type == JVal::STR) unk = v->s;
    if (const JVal* v = model->find("max_input_chars_per_word")) if (v->type == JVal::NUM) t->max_chars = static_cast<int>(v->n);
    for (const auto& kv : vocab->obj) {
        if (kv.second.type != JVal::NUM) { *err = "tokenizer.json: vocab id is not a number"; return false; }
        const int32_t id = static_cast<int32_t>(kv.second.n);
        t->whole.add(kv.first, id);
        if (!prefix.empty() && kv.first.size() > prefix.size() && !kv.first.compare(0, prefix.size(), prefix))
            t->cont.add(kv.first.substr(prefix.size()), id);
        if (id > max_id) max_id = id;
    }
    if (prefix.empty()) t->cont = t->whole;
    t->whole.finish();
    t->cont.finish();
    t->unk_id = t->whole.find(unk.data(), unk.size());
    if (t->unk_id < 0) { *err = "tokenizer.json: unk_token " + unk + " is not in the vocabulary"; return false; }
    }  // WordPiece model

    if (const JVal* at = root.find("added_tokens")) {
        if (at->type == JVal::ARR)
            for (const JVal& a : at->arr) {
                const JVal* content = a.find("content");
                const JVal* id = a.find("id");
                if (!content || content->type != JVal::STR || !id || id->type != JVal::NUM || content->s.empty()) continue;
                if (a.is_true("single_word", false) || a.is_true("lstrip", false) || a.is_true("rstrip", false) || a.is_true("normalized", false)) {
                    *err = "tokenizer.json: added token " + content->s + " uses single_word / lstrip / rstrip / normalized, which is not supported";
                    return false;
                }
                t->added.push_back({content->s, static_cast<int32_t>(id->n)});
                t->added_first_byte[static_cast<unsigned char>(content->s[0])] = true;
                if (static_cast<int32_t>(id->n) > max_id) max_id = static_cast<int32_t>(id->n);
            }
    }
    t->vocab_size = max_id + 1;
    auto token_to_id = [&](const std::string& s) -> int32_t {  // added vocabulary first, then the model's
        for (const Added& a : t->added) if (a.content == s) return a.id;
        if (t->kind == 1) return t->uni.find(reinterpret_cast<const unsigned char*>(s.data()), s.size());
        return t->whole.find(s.data(), s.size());
    };
    const int32_t pad = token_to_id("[PAD]");  // embedding_generator.rs:76-82
    t->pad_id = pad >= 0 ? pad : 0;

    if (t->kind == 1) {
        // normalizer and pre-tokenizer were read by load_unigram
    } else if (const JVal* nz = root.find("normalizer")) {
        if (nz->type == JVal::OBJ) {
            const JVal* ty = nz->find("type");
            if (!ty || ty->type != JVal::STR || ty->s != "BertNormalizer") { *err = "tokenizer.json: only BertNormalizer is supported"; return false; }
            t->clean_text = nz->is_true("clean_text", true);
            t->handle_chinese = nz->is_true("handle_chinese_chars", true);
            t->lowercase = nz->is_true("lowercase", true);
            const JVal* sa = nz->find("strip_accents");
            const bool strip = (sa && sa->type == JVal::BOOL) ? sa->b : t->lowercase;  // null: follows lowercase
            if (strip != t->lowercase) { *err = "tokenizer.json: strip_accents and lowercase must be set together"; return false; }
        } else if (nz->type != JVal::NUL) { *err = "tokenizer.json: bad normalizer"; return false; }
        else { t->clean_text = t->handle_chinese = t->lowercase = false; }
    } else { t->clean_text = t->handle_chinese = t->lowercase = false; }
    if (t->kind == 1) {
    } else if (const JVal* pt = root.find("pre_tokenizer")) {
        const JVal* ty = pt->find("type");
        if (!ty || ty->type != JVal::STR || ty->s != "BertPreTokenizer") { *err = "tokenizer.json: only BertPreTokenizer is supported"; return false; }
    } else { *err = "tokenizer.json: no pre_tokenizer"; return false; }

    if (const JVal* pp = root.find("post_processor")) {
        const JVal* ty = pp->find("type");
        if (pp->type == JVal::NUL) {
        } else if (ty && ty->type == JVal::STR && ty->s == "BertProcessing") {
            const JVal* cls = pp->find("cls");
            const JVal* sep = pp->find("sep");
            if (!cls || !sep || cls->arr.size() != 2 || sep->arr.size() != 2) { *err = "tokenizer.json: bad BertProcessing"; return false; }
            t->prefix.push_back({static_cast<int32_t>(cls->arr[1].n), 0});
            t->suffix.push_back({static_cast<int32_t>(sep->arr[1].n), 0});
        } else if (ty && ty->type == JVal::STR && ty->s == "TemplateProcessing") {
            const JVal* single = pp->find("single");
            const JVal* specials = pp->find("special_tokens");
            if (!single || single->type != JVal::ARR) { *err = "tokenizer.json: bad TemplateProcessing"; return false; }
            bool seen_seq = false;
            for (const JVal& piece : single->arr) {
                if (const JVal* sq = piece.find("Sequence")) {
                    if (seen_seq) { *err = "tokenizer.json: template uses the sequence twice"; return false; }
                    seen_seq = true;
                    if (const JVal* ti = sq->find("type_id")) t->seq_type_id = static_cast<uint32_t>(ti->n);
                } else if (const JVal* sp = piece.find("SpecialToken")) {
                    const JVal* id = sp->find("id");
                    const JVal* ti = sp->find("type_id");
                    const JVal* entry = (id && specials) ? specials->find(id->s.c_str()) : nullptr;
                    const JVal* ids = entry ? entry->find("ids") : nullptr;
                    if (!ids || ids->type != JVal::ARR) { *err = "tokenizer.json: template special token without ids"; return false; }
                    for (const JVal& one : ids->arr)
                        (seen_seq ? t->suffix : t->prefix).push_back({static_cast<int32_t>(one.n), ti ? static_cast<uint32_t>(ti->n) : 0u});
                }
            }
            if (!seen_seq) { *err = "tokenizer.json: template without the sequence"; return false; }
        } else { *err = "tokenizer.json: unsupported post_processor"; return false; }
    }

    const CharTable& ct = char_table();
    for (int c = 0; c < 128; ++c) {
        const uint8_t k = ct.get(static_cast<uint32_t>(c));
        uint8_t act = A_APPEND;
        char out = static_cast<char>(c);
        if (t->clean_text && (k & 3) == C_REMOVED) act = A_SKIP;
        else if (c == ' ' || (t->clean_text ? (k & 3) == C_SPACE : rust_is_whitespace(static_cast<uint32_t>(c)))) act = A_BOUNDARY;
        else if (k & C_PUNCT) act = A_ISOLATE;
        else if (t->lowercase && c >= 'A' && c <= 'Z') out = static_cast<char>(c + 32);
        t->ascii_action[c] = act;
        t->ascii_out[c] = out;
    }
    return true;
}

struct Scratch {
    std::string norm, piece;   // Unigram: the normalised segment, the current pre-token with its Metaspace prefix
    std::vector<b200::unigram::Model::Node> lattice;
    std::string word;          // the current pre-token, normalised, UTF-8
    int word_chars = 0;
    std::vector<int32_t> toks; // the sentence's tokens before the template
    std::vector<int32_t> piece_toks;
};

// WordPiece on sc.word (models/wordpiece: greedy longest match, "##" continuation, the whole word -> [UNK] on a miss)
inline void wordpiece(const b200_ingest& t, Scratch& sc) {
    const std::string& w = sc.word;
    if (w.empty()) return;
    if (sc.word_chars > t.max_chars) { sc.toks.push_back(t.unk_id); return; }
    const size_t n = w.size();
    if (const int32_t id = t.whole.find(w.data(), n); id >= 0) { sc.toks.push_back(id); return; }  // whole word: the common case
    sc.piece_toks.clear();
    size_t start = 0;
    while (start < n) {
        const Vocab& v = start ? t.cont : t.whole;
        size_t end = std::min(n, start + v.max_len);
        int32_t id = -1;
        for (; end > start; --end) {
            if (end < n && (static_cast<unsigned char>(w[end]) & 0xC0) == 0x80) continue;  // not a character boundary
            id = v.find(w.data() + start, end - start);
            if (id >= 0) break;
        }
        if (id < 0) { sc.toks.push_back(t.unk_id); return; }
        sc.piece_toks.push_back(id);
        start = end;
    }
    sc.toks.insert(sc.toks.end(), sc.piece_toks.begin(), sc.piece_toks.end());
}

inline void flush_word(const b200_ingest& t, Scratch& sc) {
    if (!sc.word.empty()) {
        wordpiece(t, sc);
        sc.word.clear();
        sc.word_chars = 0;
    }
}

// one normalised character (after clean / CJK / lowercase) into the pre-tokeniser
inline void feed_char(const b200_ingest& t, Scratch& sc, uint32_t c, uint8_t klass) {
    if (c == ' ' || (!t.clean_text && rust_is_whitespace(c))) { flush_word(t, sc); return; }
    if (klass & C_PUNCT) {
        flush_word(t, sc);
        utf8_put(sc.word, c);
        sc.word_chars = 1;
        flush_word(t, sc);
        return;
    }
    utf8_put(sc.word, c);
    ++sc.word_chars;
}

// BertNormalizer + BertPreTokenizer + WordPiece over text[b, e), which holds no added token. Stops feeding once
// `budget` tokens exist: truncation keeps the first ones only.
// Unigram pipeline for one text segment between added tokens: Precompiled + Replace, WhitespaceSplit, Metaspace, Viterbi.
void encode_piece_unigram(const b200_ingest& t, Scratch& sc, const unsigned char* b, const unsigned char* e, size_t budget) {
    if (b >= e) return;
    sc.norm.clear();
    if (t.has_pre) t.pre.normalize(b, e, sc.norm);
    else sc.norm.assign(reinterpret_cast<const char*>(b), static_cast<size_t>(e - b));
    if (t.collapse_spaces) {  // Replace(Regex " {2,}", " ")
        size_t w = 0;
        for (size_t i = 0; i < sc.norm.size(); ++i) {
            if (sc.norm[i] == ' ' && w > 0 && sc.norm[w - 1] == ' ') continue;
            sc.norm[w++] = sc.norm[i];
        }
        sc.norm.resize(w);
    }
    const unsigned char* p = reinterpret_cast<const unsigned char*>(sc.norm.data());
    const unsigned char* end = p + sc.norm.size();
    const std::string& rep = t.meta;
    // one pre-token of the first stage (a whitespace-free run, or the whole segment) through Metaspace and the model
    auto metaspace = [&](const unsigned char* wb, const unsigned char* we) {
        sc.piece.clear();
        for (const unsigned char* q = wb; q < we; ++q) {
            if (*q == ' ') sc.piece += rep; else sc.piece.push_back(static_cast<char>(*q));
        }
        if (t.meta_prepend && sc.piece.compare(0, rep.size(), rep) != 0) sc.piece.insert(0, rep);
        // split, delimiter merged with what follows: a piece starts at every occurrence of the replacement
        const unsigned char* s0 = reinterpret_cast<const unsigned char*>(sc.piece.data());
        const size_t n = sc.piece.size();
        size_t start = 0;
        for (size_t i = 0; i <= n && sc.toks.size() < budget;) {
            const bool at_rep = i + rep.size() <= n && !std::memcmp(s0 + i, rep.data(), rep.size());
            if (i == n || (at_rep && i > start)) {
                if (i > start) t.uni.encode(s0 + start, i - start, sc.lattice, sc.toks);
                start = i;
                if (i == n) break;
            }
            i += at_rep ? rep.size() : 1;
        }
    };
    if (!t.ws_split) {
        metaspace(p, end);
        return;
    }
    while (p < end && sc.toks.size() < budget) {
        int len;
        // skip whitespace, then take the run up to the next whitespace character (char::is_whitespace)
        while (p < end) {
            const uint32_t c = utf8_next(p, end, &len);
            if (!rust_is_whitespace(c)) break;
            p += len;
        }
        const unsigned char* wb = p;
        while (p < end) {
            const uint32_t c = utf8_next(p, end, &len);
            if (rust_is_whitespace(c)) break;
            p += len;
        }
        if (p > wb) metaspace(wb, p);
    }
}

void encode_piece(const b200_ingest& t, Scratch& sc, const unsigned char* b, const unsigned char* e, size_t budget) {
    if (t.kind == 1) { encode_piece_unigram(t, sc, b, e, budget); return; }
    const CharTable& ct = char_table();
    std::string tmp;
    const unsigned char* p = b;
    while (p < e) {
        if (sc.toks.size() >= budget && sc.word.empty()) return;
        const unsigned char c0 = *p;
        if (c0 < 0x80) {
            ++p;
            switch (t.ascii_action[c0]) {
                case A_APPEND: sc.word.push_back(t.ascii_out[c0]); ++sc.word_chars; break;
                case A_SKIP: break;
                case A_BOUNDARY: flush_word(t, sc); break;
                default:
                    flush_word(t, sc);
                    sc.word.push_back(static_cast<char>(c0));
                    sc.word_chars = 1;
                    flush_word(t, sc);
            }
            continue;
        }
        int len;
        const uint32_t c = utf8_next(p, e, &len);
        p += len;
        const uint8_t k = ct.get(c);
        const uint8_t cls = k & 3;
        if (t.clean_text) {
            if (cls == C_REMOVED) continue;
            if (cls == C_SPACE) { flush_word(t, sc); continue; }
        }
        const bool cjk = t.handle_chinese && cls == C_CJK;
        if (cjk) flush_word(t, sc);
        if (t.lowercase && (k & C_HAS_LOWER)) {
            tmp.clear();
            put_lowered(tmp, c);
            const unsigned char* q = reinterpret_cast<const unsigned char*>(tmp.data());
            const unsigned char* qe = q + tmp.size();
            while (q < qe) {
                int l2;
                const uint32_t c2 = utf8_next(q, qe, &l2);
                q += l2;
                feed_char(t, sc, c2, ct.get(c2));
            }
        } else {
            feed_char(t, sc, c, k);
        }
        if (cjk) flush_word(t, sc);
    }
    flush_word(t, sc);
}

// One sentence -> row of ids / mask / type_ids (length P). Returns the number of real tokens.
// The sentence's tokens (without the template) into sc.toks; returns how many of them survive truncation to P.
size_t tokenize_sentence(const b200_ingest& t, Scratch& sc, const char* s, size_t n, int P) {
    const size_t n_special = t.prefix.size() + t.suffix.size();
    const size_t budget = static_cast<size_t>(P) - n_special;
    sc.toks.clear();
    sc.word.clear();
    sc.word_chars = 0;
    const unsigned char* b = reinterpret_cast<const unsigned char*>(s);
    const unsigned char* e = b + n;
    if (t.added.empty()) {
        encode_piece(t, sc, b, e, budget);
    } else {
        const unsigned char* seg = b;
        const unsigned char* p = b;
        while (p < e && sc.toks.size() < budget) {
            if (!t.added_first_byte[*p]) { ++p; continue; }
            const Added* best = nullptr;
            for (const Added& a : t.added)
                if (a.content.size() <= static_cast<size_t>(e - p) && (!best || a.content.size() > best->content.size()) &&
                    !std::memcmp(a.content.data(), p, a.content.size()))
                    best = &a;
            if (!best) { ++p; continue; }
            encode_piece(t, sc, seg, p, budget);
            sc.toks.push_back(best->id);
            p += best->content.size();
            seg = p;
        }
        if (sc.toks.size() < budget) encode_piece(t, sc, seg, e, budget);
    }
    return std::min(sc.toks.size(), budget);  // LongestFirst, direction right, one sequence
}

int encode_sentence(const b200_ingest& t, Scratch& sc, const char* s, size_t n, int P, uint32_t* ids, uint32_t* mask, uint32_t* types) {
    const size_t keep = tokenize_sentence(t, sc, s, n, P);
    int at = 0;
    for (const Special& sp : t.prefix) { ids[at] = static_cast<uint32_t>(sp.id); types[at] = sp.type_id; mask[at++] = 1; }
    for (size_t i = 0; i < keep; ++i) { ids[at] = static_cast<uint32_t>(sc.toks[i]); types[at] = t.seq_type_id; mask[at++] = 1; }
    for (const Special& sp : t.suffix) { ids[at] = static_cast<uint32_t>(sp.id); types[at] = sp.type_id; mask[at++] = 1; }
    const int real = at;
    for (; at < P; ++at) { ids[at] = static_cast<uint32_t>(t.pad_id); types[at] = 0; mask[at] = 0; }
    return real;
}

}  // namespace

extern "C" {

const char* b200_ingest_last_error(void) { return g_ingest_err.c_str(); }

int32_t b200_ingest_create(const char* tokenizer_json_path, b200_ingest** out) {
    if (!tokenizer_json_path || !out) return fail(B200_EMBED_ERR_INVALID_ARG, "null path or out pointer");
    *out = nullptr;
    std::unique_ptr<b200_ingest> t(new b200_ingest());
    std::string err;
    if (!load_tokenizer(tokenizer_json_path, t.get(), &err)) {
        const bool io = err.rfind("cannot open", 0) == 0;
        return fail(io ? B200_EMBED_ERR_IO : B200_EMBED_ERR_UNSUPPORTED_SHAPE, err);
    }
    *out = t.release();
    return B200_EMBED_OK;
}

void b200_ingest_destroy(b200_ingest* t) { delete t; }

int32_t b200_ingest_vocab_size(const b200_ingest* t) { return t ? t->vocab_size : -1; }
int32_t b200_ingest_pad_id(const b200_ingest* t) { return t ? t->pad_id : -1; }

// One chunk of the raw text [p, e): words joined by single spaces into `out` (never a leading or trailing space) and the
// offsets just after every '.', '?' or '!' in `cuts`. Chunks are cut at ASCII whitespace, so their outputs join with
// exactly one space and the cut offsets only need the chunk's base added.
static void clean_chunk(const unsigned char* p, const unsigned char* e, char* out, int64_t* out_len, std::vector<int64_t>* cuts) {
    int64_t o = 0;
    bool pending_space = false;
    while (p < e) {
        const unsigned char c0 = *p;
        if (c0 > ' ' && c0 < 0x80) {  // printable ASCII: the common case
            if (pending_space) { out[o++] = ' '; pending_space = false; }
            out[o++] = static_cast<char>(c0);
            ++p;
            if (c0 == '.' || c0 == '?' || c0 == '!') cuts->push_back(o);
            continue;
        }
        int l = 1;
        const uint32_t c = c0 < 0x80 ? static_cast<uint32_t>(c0) : utf8_next(p, e, &l);
        if (rust_is_whitespace(c)) {
            pending_space = o > 0;
        } else {
            if (pending_space) { out[o++] = ' '; pending_space = false; }
            std::memcpy(out + o, p, static_cast<size_t>(l));
            o += l;
        }
        p += l;
    }
    *out_len = o;
}

int64_t b200_ingest_split_sentences(const char* raw_text, int64_t len, char* cleaned, int64_t cleaned_cap,
                                    int64_t* cleaned_len, int64_t* begin, int64_t* end, int64_t max_sentences) {
    return b200_ingest_split_sentences_threads(raw_text, len, cleaned, cleaned_cap, cleaned_len, begin, end, max_sentences, 0);
}

int64_t b200_ingest_split_sentences_threads(const char* raw_text, int64_t len, char* cleaned, int64_t cleaned_cap,
                                            int64_t* cleaned_len, int64_t* begin, int64_t* end, int64_t max_sentences,
                                            int32_t threads) {
    if (!raw_text || len < 0 || !cleaned || cleaned_cap < len || !cleaned_len || max_sentences < 0 ||
        (max_sentences > 0 && (!begin || !end)))
        return fail(B200_EMBED_ERR_INVALID_ARG, "split_sentences: bad pointer or capacity (cleaned_cap must be >= len)");
    const unsigned char* base = reinterpret_cast<const unsigned char*>(raw_text);
    // split_whitespace().collect().join(" ") and the cuts after '.', '?', '!' in one pass; a long text is processed in
    // chunks (cut at ASCII whitespace) on several threads - the result does not depend on the chunking
    int nt = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (nt < 1) nt = 1;
    if (static_cast<int64_t>(nt) > len / (256 * 1024)) nt = static_cast<int>(len / (256 * 1024));
    if (nt < 1) nt = 1;
    std::vector<int64_t> bounds(1, 0);
    for (int k = 1; k < nt; ++k) {
        int64_t b = len * k / nt;
        if (b <= bounds.back()) b = bounds.back();
        while (b < len && !(base[b] == ' ' || (base[b] >= 0x09 && base[b] <= 0x0D))) ++b;
        if (b > bounds.back() && b < len) bounds.push_back(b);
    }
    bounds.push_back(len);
    const int nc = static_cast<int>(bounds.size()) - 1;
    std::vector<std::vector<int64_t>> cuts(static_cast<size_t>(nc));
    std::vector<int64_t> out_len(static_cast<size_t>(nc), 0);
    std::vector<std::unique_ptr<char[]>> bufs(static_cast<size_t>(nc));
    auto work = [&](int k) {
        const int64_t n = bounds[static_cast<size_t>(k) + 1] - bounds[static_cast<size_t>(k)];
        // the first chunk writes straight into the caller's buffer (its output starts at offset 0)
        char* out = cleaned;
        if (k > 0) {
            bufs[static_cast<size_t>(k)].reset(new char[static_cast<size_t>(n) + 1]);
            out = bufs[static_cast<size_t>(k)].get();
        }
        clean_chunk(base + bounds[static_cast<size_t>(k)], base + bounds[static_cast<size_t>(k) + 1], out,
                    &out_len[static_cast<size_t>(k)], &cuts[static_cast<size_t>(k)]);
    };
    if (nc == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int k = 0; k < nc; ++k) pool.emplace_back(work, k);
        for (auto& th : pool) th.join();
    }
    // offsets of the chunk outputs in the joined text: one space between two non-empty neighbours
    std::vector<int64_t> off(static_cast<size_t>(nc), 0);
    int64_t o = 0;
    for (int k = 0; k < nc; ++k) {
        if (out_len[static_cast<size_t>(k)] == 0) { off[static_cast<size_t>(k)] = o; continue; }
        if (o > 0) cleaned[o++] = ' ';
        off[static_cast<size_t>(k)] = o;
        o += out_len[static_cast<size_t>(k)];
    }
    if (nc > 1) {
        std::vector<std::thread> pool;
        for (int k = 1; k < nc; ++k)
            pool.emplace_back([&, k] { std::memcpy(cleaned + off[static_cast<size_t>(k)], bufs[static_cast<size_t>(k)].get(), static_cast<size_t>(out_len[static_cast<size_t>(k)])); });
        for (auto& th : pool) th.join();
    }
    *cleaned_len = o;
    if (o == 0) return 0;  // the caller reports "Cleaned text is empty" (main.rs:33-39)
    int64_t n = 0;
    auto push = [&](int64_t b, int64_t en) {  // [b, en) trimmed: only single spaces can be left at the ends
        while (b < en && cleaned[b] == ' ') ++b;
        while (en > b && cleaned[en - 1] == ' ') --en;
        if (b == en) return;
        if (n < max_sentences) { begin[n] = b; end[n] = en; }
        ++n;
    };
    int64_t start = 0;
    for (int k = 0; k < nc; ++k)
        for (const int64_t c : cuts[static_cast<size_t>(k)]) {
            const int64_t cut = off[static_cast<size_t>(k)] + c;
            push(start, cut);
            start = cut;
        }
    if (start < o) push(start, o);
    return n;
}

int32_t b200_ingest_encode(const b200_ingest* t, const char* text, const int64_t* begin, const int64_t* end, int32_t n,
                           int32_t seq_len, int32_t threads, uint32_t* ids, uint32_t* mask, uint32_t* type_ids,
                           int32_t* lengths_or_null) {
    if (!t) return fail(B200_EMBED_ERR_INVALID_ARG, "null tokenizer");
    if (n < 0 || (n > 0 && (!text || !begin || !end || !ids || !mask || !type_ids)))
        return fail(B200_EMBED_ERR_INVALID_ARG, "encode: null pointer");
    if (seq_len < static_cast<int32_t>(t->prefix.size() + t->suffix.size()) || seq_len <= 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "encode: seq_len is smaller than the special tokens of the template");
    for (int32_t i = 0; i < n; ++i)
        if (begin[i] < 0 || end[i] < begin[i]) return fail(B200_EMBED_ERR_INVALID_ARG, "encode: bad sentence range");
    int nt = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (nt < 1) nt = 1;
    if (nt > n / 16) nt = n / 16 > 0 ? n / 16 : 1;
    auto work = [&](int k) {
        Scratch sc;
        sc.word.reserve(256);
        sc.toks.reserve(static_cast<size_t>(seq_len) + 64);
        const int lo = static_cast<int>(static_cast<int64_t>(n) * k / nt), hi = static_cast<int>(static_cast<int64_t>(n) * (k + 1) / nt);
        for (int i = lo; i < hi; ++i) {
            const size_t row = static_cast<size_t>(i) * static_cast<size_t>(seq_len);
            const int real = encode_sentence(*t, sc, text + begin[i], static_cast<size_t>(end[i] - begin[i]), seq_len, ids + row,
                                             mask + row, type_ids + row);
            if (lengths_or_null) lengths_or_null[i] = real;
        }
    };
    if (nt <= 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int k = 0; k < nt; ++k) pool.emplace_back(work, k);
        for (auto& th : pool) th.join();
    }
    return B200_EMBED_OK;
}

int64_t b200_ingest_encode_packed(const b200_ingest* t, const char* text, const int64_t* begin, const int64_t* end, int32_t n,
                                  int32_t max_len, int32_t threads, uint32_t* ids_packed, uint32_t* type_ids_packed,
                                  int64_t capacity_tokens, int32_t* cu_seqlens) {
    if (!t) return fail(B200_EMBED_ERR_INVALID_ARG, "null tokenizer");
    if (n < 0 || !cu_seqlens || capacity_tokens < 0 || (n > 0 && (!text || !begin || !end)) ||
        (capacity_tokens > 0 && (!ids_packed || !type_ids_packed)))
        return fail(B200_EMBED_ERR_INVALID_ARG, "encode_packed: null pointer");
    if (max_len < static_cast<int32_t>(t->prefix.size() + t->suffix.size()) || max_len <= 0)
        return fail(B200_EMBED_ERR_INVALID_ARG, "encode_packed: max_len is smaller than the special tokens of the template");
    for (int32_t i = 0; i < n; ++i)
        if (begin[i] < 0 || end[i] < begin[i]) return fail(B200_EMBED_ERR_INVALID_ARG, "encode_packed: bad sentence range");
    int nt = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (nt < 1) nt = 1;
    if (nt > n / 16) nt = n / 16 > 0 ? n / 16 : 1;
    // every thread tokenises a contiguous run of sentences into its own buffers (lengths go straight into cu_seqlens),
    // then copies them to the offsets a prefix sum over the lengths gives
    struct Part { std::vector<uint32_t> ids, types; };
    std::vector<Part> parts(static_cast<size_t>(nt));
    auto span = [&](int k, int* lo, int* hi) {
        *lo = static_cast<int>(static_cast<int64_t>(n) * k / nt);
        *hi = static_cast<int>(static_cast<int64_t>(n) * (k + 1) / nt);
    };
    auto run = [&](auto&& fn) {
        if (nt <= 1) { fn(0); return; }
        std::vector<std::thread> pool;
        for (int k = 0; k < nt; ++k) pool.emplace_back(fn, k);
        for (auto& th : pool) th.join();
    };
    run([&](int k) {
        Scratch sc;
        sc.word.reserve(256);
        sc.toks.reserve(static_cast<size_t>(max_len) + 64);
        int lo, hi;
        span(k, &lo, &hi);
        Part& p = parts[static_cast<size_t>(k)];
        p.ids.reserve(static_cast<size_t>(hi - lo) * 32);
        p.types.reserve(static_cast<size_t>(hi - lo) * 32);
        for (int i = lo; i < hi; ++i) {
            const size_t keep = tokenize_sentence(*t, sc, text + begin[i], static_cast<size_t>(end[i] - begin[i]), max_len);
            for (const Special& sp : t->prefix) { p.ids.push_back(static_cast<uint32_t>(sp.id)); p.types.push_back(sp.type_id); }
            for (size_t j = 0; j < keep; ++j) { p.ids.push_back(static_cast<uint32_t>(sc.toks[j])); p.types.push_back(t->seq_type_id); }
            for (const Special& sp : t->suffix) { p.ids.push_back(static_cast<uint32_t>(sp.id)); p.types.push_back(sp.type_id); }
            cu_seqlens[i + 1] = static_cast<int32_t>(keep + t->prefix.size() + t->suffix.size());
        }
    });
    int64_t total = 0;
    cu_seqlens[0] = 0;
    for (int32_t i = 0; i < n; ++i) {
        total += cu_seqlens[i + 1];
        if (total > INT32_MAX) return fail(B200_EMBED_ERR_INVALID_ARG, "encode_packed: more than 2^31 tokens in one call");
        cu_seqlens[i + 1] = static_cast<int32_t>(total);
    }
    if (total > capacity_tokens) return total;  // nothing written: call again with room for `total` tokens
    run([&](int k) {
        int lo, hi;
        span(k, &lo, &hi);
        const Part& p = parts[static_cast<size_t>(k)];
        if (p.ids.empty()) return;
        std::memcpy(ids_packed + cu_seqlens[lo], p.ids.data(), p.ids.size() * sizeof(uint32_t));
        std::memcpy(type_ids_packed + cu_seqlens[lo], p.types.data(), p.types.size() * sizeof(uint32_t));
    });
    return total;
}

}  // extern "C"
