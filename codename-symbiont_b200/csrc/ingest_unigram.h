// Unigram (SentencePiece) front-end of b200_ingest: what `tokenizers` does for an XLM-R style tokenizer.json, the kind the
// service's deployed model ships (paraphrase-multilingual-mpnet-base-v2, /root/reference/services/preprocessing_service/src/
// main.rs:305; tokenizer set-up embedding_generator.rs:73-99, encode_batch :160-164):
//   normalizer     Sequence[ Precompiled(charsmap), Replace(Regex " {2,}" -> " ") ]   (either part optional)
//   pre_tokenizer  Sequence[ WhitespaceSplit, Metaspace(replacement, prepend always / add_prefix_space) ]  or Metaspace alone
//   model          Unigram(vocab [[piece, score]], unk_id), fused unknowns, no byte fallback
// The algorithms live in the un-vendored crates tokenizers 0.21 / spm_precompiled / unicode-segmentation; they are restated
// here from their published sources and checked bit for bit against the library's Python wheel (tests/test_ingest_unigram.py):
//   * Precompiled walks the text one extended grapheme cluster at a time (UAX #29; property tables: grapheme_tables.h). A
//     cluster shorter than 6 bytes is looked up whole in the charsmap's double-array trie and, on a hit, replaced by the
//     normalised string of the SHORTEST matching prefix; otherwise every character of it is looked up on its own.
//   * Unigram is the Viterbi walk of `encode_optimized`: best path per byte position over the vocabulary pieces found by a
//     common-prefix search, an unknown character costing min_score - 10, a later candidate winning only when strictly better;
//     consecutive unknown pieces are fused into one token.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "grapheme_tables.h"

namespace b200 {
namespace unigram {

// ---- grapheme clusters (UAX #29, extended) --------------------------------------------------------------------------------
enum : uint8_t { G_OTHER = 0, G_CR, G_LF, G_CONTROL, G_EXTEND, G_ZWJ, G_RI, G_PREPEND, G_SPACINGMARK, G_L, G_V, G_T, G_LV, G_LVT };

struct GraphemeTable {
    std::vector<uint8_t> bmp;  // class byte of every BMP code point
    GraphemeTable() : bmp(0x10000, 0) {
        for (int i = 0; i < kNumGraphemeRanges; ++i) {
            const GraphemeRange& r = kGraphemeRanges[i];
            for (uint32_t c = r.lo; c <= r.hi && c < 0x10000; ++c) bmp[c] = r.cls;
        }
    }
    uint8_t get(uint32_t c) const {
        if (c < 0x10000) return bmp[c];
        int lo = 0, hi = kNumGraphemeRanges - 1;
        while (lo <= hi) {
            const int mid = (lo + hi) >> 1;
            if (c < kGraphemeRanges[mid].lo) hi = mid - 1;
            else if (c > kGraphemeRanges[mid].hi) lo = mid + 1;
            else return kGraphemeRanges[mid].cls;
        }
        return 0;
    }
};
inline const GraphemeTable& grapheme_table() {
    static const GraphemeTable t;
    return t;
}

inline uint32_t u8_next(const unsigned char* p, const unsigned char* end, int* len) {
    const unsigned char c = *p;
    if (c < 0x80) { *len = 1; return c; }
    if ((c >> 5) == 6 && p + 1 < end) { *len = 2; return ((c & 0x1Fu) << 6) | (p[1] & 0x3Fu); }
    if ((c >> 4) == 14 && p + 2 < end) { *len = 3; return ((c & 0x0Fu) << 12) | ((p[1] & 0x3Fu) << 6) | (p[2] & 0x3Fu); }
    if ((c >> 3) == 30 && p + 3 < end) {
        *len = 4;
        return ((c & 0x07u) << 18) | ((p[1] & 0x3Fu) << 12) | ((p[2] & 0x3Fu) << 6) | (p[3] & 0x3Fu);
    }
    *len = 1;
    return 0xFFFD;
}

// End of the extended grapheme cluster that starts at p (p < end).
inline const unsigned char* next_grapheme(const unsigned char* p, const unsigned char* end) {
    const GraphemeTable& gt = grapheme_table();
    int len;
    uint32_t c = u8_next(p, end, &len);
    p += len;
    uint8_t k = gt.get(c);
    uint8_t prev = k & 15;
    // state carried along the cluster
    bool pict_seq = (k & 16) != 0;        // ExtPict Extend* so far (GB11)
    bool pict_zwj = false;                // ... followed by ZWJ
    int ri_count = prev == G_RI ? 1 : 0;  // GB12 / GB13
    bool incb_cons = ((k >> 5) & 3) == 1; // Consonant (Extend | Linker)* so far (GB9c)
    bool incb_linker = false;             // ... with at least one Linker
    while (p < end) {
        c = u8_next(p, end, &len);
        k = gt.get(c);
        const uint8_t cur = k & 15;
        const uint8_t icb = (k >> 5) & 3;
        bool join = false;
        if (prev == G_CR) join = cur == G_LF;                                                       // GB3
        else if (prev == G_CONTROL || prev == G_LF || cur == G_CONTROL || cur == G_CR || cur == G_LF) join = false;  // GB4, GB5
        else if (prev == G_L && (cur == G_L || cur == G_V || cur == G_LV || cur == G_LVT)) join = true;             // GB6
        else if ((prev == G_LV || prev == G_V) && (cur == G_V || cur == G_T)) join = true;          // GB7
        else if ((prev == G_LVT || prev == G_T) && cur == G_T) join = true;                         // GB8
        else if (cur == G_EXTEND || cur == G_ZWJ || cur == G_SPACINGMARK) join = true;              // GB9, GB9a
        else if (prev == G_PREPEND) join = true;                                                    // GB9b
        else if (incb_cons && incb_linker && icb == 1) join = true;                                 // GB9c
        else if (pict_zwj && (k & 16)) join = true;                                                 // GB11
        else if (prev == G_RI && cur == G_RI && (ri_count & 1)) join = true;                        // GB12, GB13
        if (!join) break;
        // update the carried state with the character just joined
        if (k & 16) { pict_seq = true; pict_zwj = false; }
        else if (cur == G_EXTEND && pict_seq) { pict_zwj = false; }
        else if (cur == G_ZWJ && pict_seq) { pict_zwj = true; pict_seq = false; }
        else { pict_seq = false; pict_zwj = false; }
        if (cur == G_RI) ++ri_count; else ri_count = 0;
        if (icb == 1) { incb_cons = true; incb_linker = false; }
        else if (incb_cons && icb == 2) incb_linker = true;
        else if (incb_cons && icb == 3) {}
        else { incb_cons = false; incb_linker = false; }
        prev = cur;
        p += len;
    }
    return p;
}

// ---- Precompiled (SentencePiece charsmap) -------------------------------------------------------------------------------
struct Precompiled {
    std::vector<uint32_t> trie;
    std::string normalized;
    bool load(const std::string& blob) {
        if (blob.size() < 4) return false;
        uint32_t tsize;
        std::memcpy(&tsize, blob.data(), 4);
        if (tsize % 4 != 0 || 4 + static_cast<size_t>(tsize) > blob.size()) return false;
        trie.resize(tsize / 4);
        if (tsize) std::memcpy(trie.data(), blob.data() + 4, tsize);
        normalized.assign(blob.data() + 4 + tsize, blob.size() - 4 - tsize);
        return !trie.empty();
    }
    static uint32_t offset(uint32_t u) { return (u >> 10) << ((u & (1u << 9)) >> 6); }
    // The normalised string of the shortest key that is a prefix of s[0, n), or nullptr.
    const char* transform(const unsigned char* s, size_t n, size_t* out_len) const {
        size_t pos = 0;
        uint32_t unit = trie[0];
        pos ^= offset(unit);
        for (size_t i = 0; i < n; ++i) {
            const unsigned char c = s[i];
            if (c == 0) break;
            pos ^= c;
            if (pos >= trie.size()) return nullptr;
            unit = trie[pos];
            if ((unit & ((1u << 31) | 0xFFu)) != c) return nullptr;
            pos ^= offset(unit);
            if ((unit >> 8) & 1) {
                if (pos >= trie.size()) return nullptr;
                const size_t index = trie[pos] & 0x7FFFFFFFu;
                if (index > normalized.size()) return nullptr;
                size_t e = index;
                while (e < normalized.size() && normalized[e] != '\0') ++e;
                *out_len = e - index;
                return normalized.data() + index;
            }
        }
        return nullptr;
    }
    void normalize(const unsigned char* b, const unsigned char* e, std::string& out) const {
        while (b < e) {
            const unsigned char* g = next_grapheme(b, e);
            size_t nl;
            if (g - b < 6) {
                if (const char* norm = transform(b, static_cast<size_t>(g - b), &nl)) {
                    out.append(norm, nl);
                    b = g;
                    continue;
                }
            }
            while (b < g) {
                int len;
                u8_next(b, g, &len);
                if (const char* norm = transform(b, static_cast<size_t>(len), &nl)) out.append(norm, nl);
                else out.append(reinterpret_cast<const char*>(b), static_cast<size_t>(len));
                b += len;
            }
        }
    }
};

inline bool base64_decode(const std::string& in, std::string* out) {
    static int8_t T[256];
    static bool init = false;
    if (!init) {
        for (int i = 0; i < 256; ++i) T[i] = -1;
        const char* a = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 64; ++i) T[static_cast<unsigned char>(a[i])] = static_cast<int8_t>(i);
        init = true;
    }
    out->clear();
    uint32_t acc = 0;
    int bits = 0;
    for (unsigned char c : in) {
        if (c == '=') break;
        if (T[c] < 0) return false;
        acc = (acc << 6) | static_cast<uint32_t>(T[c]);
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            out->push_back(static_cast<char>((acc >> bits) & 0xFF));
        }
    }
    return true;
}

// ---- the Unigram model -----------------------------------------------------------------------------------------------------
struct Model {
    // byte trie over the pieces: node i has edges [first[i], first[i + 1]) sorted by label
    std::vector<uint32_t> first;
    std::vector<uint8_t> label;
    std::vector<uint32_t> child;
    std::vector<int32_t> node_id;  // piece id ending at the node, -1 none
    std::vector<double> score;     // by piece id
    int32_t unk_id = -1;
    double min_score = 0;

    void build(std::vector<std::pair<std::string, int32_t>>& keys) {  // (piece, id); a later duplicate wins, as in a map
        std::stable_sort(keys.begin(), keys.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        struct Item { size_t lo, hi, depth; uint32_t node; };
        std::vector<std::vector<std::pair<uint8_t, uint32_t>>> edges(1);
        node_id.assign(1, -1);
        std::vector<Item> stack{{0, keys.size(), 0, 0}};
        while (!stack.empty()) {
            const Item it = stack.back();
            stack.pop_back();
            size_t i = it.lo;
            while (i < it.hi && keys[i].first.size() == it.depth) { node_id[it.node] = keys[i].second; ++i; }
            while (i < it.hi) {
                const uint8_t c = static_cast<uint8_t>(keys[i].first[it.depth]);
                size_t j = i;
                while (j < it.hi && static_cast<uint8_t>(keys[j].first[it.depth]) == c) ++j;
                const uint32_t nn = static_cast<uint32_t>(edges.size());
                edges.emplace_back();
                node_id.push_back(-1);
                edges[it.node].emplace_back(c, nn);
                stack.push_back({i, j, it.depth + 1, nn});
                i = j;
            }
        }
        first.assign(edges.size() + 1, 0);
        for (size_t n = 0; n < edges.size(); ++n) first[n + 1] = first[n] + static_cast<uint32_t>(edges[n].size());
        label.resize(first.back());
        child.resize(first.back());
        for (size_t n = 0; n < edges.size(); ++n)
            for (size_t k = 0; k < edges[n].size(); ++k) {
                label[first[n] + k] = edges[n][k].first;
                child[first[n] + k] = edges[n][k].second;
            }
    }
    int32_t step(uint32_t node, uint8_t c) const {
        uint32_t lo = first[node], hi = first[node + 1];
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (label[mid] < c) lo = mid + 1; else hi = mid;
        }
        return (lo < first[node + 1] && label[lo] == c) ? static_cast<int32_t>(child[lo]) : -1;
    }
    int32_t find(const unsigned char* s, size_t n) const {
        uint32_t node = 0;
        for (size_t i = 0; i < n; ++i) {
            const int32_t nx = step(node, s[i]);
            if (nx < 0) return -1;
            node = static_cast<uint32_t>(nx);
        }
        return node_id[node];
    }

    struct Node { double best; int32_t start; int32_t id; };
    // One pre-token (UTF-8, already carrying the replacement prefix) -> piece ids appended to out.
    void encode(const unsigned char* s, size_t size, std::vector<Node>& lat, std::vector<int32_t>& out) const {
        if (size == 0) return;
        lat.assign(size + 1, Node{0.0, -1, 0});
        const double unk_score = min_score - 10.0;
        size_t at = 0;
        while (at < size) {
            const double here = lat[at].best;
            int mblen;
            u8_next(s + at, s + size, &mblen);
            if (at + static_cast<size_t>(mblen) > size) mblen = static_cast<int>(size - at);
            bool single = false;
            uint32_t node = 0;
            for (size_t k = at; k < size; ++k) {
                const int32_t nx = step(node, s[k]);
                if (nx < 0) break;
                node = static_cast<uint32_t>(nx);
                const int32_t id = node_id[node];
                if (id < 0) continue;
                Node& t = lat[k + 1];
                const double cand = score[static_cast<size_t>(id)] + here;
                if (t.start < 0 || cand > t.best) { t.best = cand; t.start = static_cast<int32_t>(at); t.id = id; }
                if (!single && k + 1 - at == static_cast<size_t>(mblen)) single = true;
            }
            if (!single) {
                Node& t = lat[at + static_cast<size_t>(mblen)];
                const double cand = unk_score + here;
                if (t.start < 0 || cand > t.best) { t.best = cand; t.start = static_cast<int32_t>(at); t.id = unk_id; }
            }
            at += static_cast<size_t>(mblen);
        }
        // backtrack; runs of unknown pieces are fused and looked up as one string (which is the unknown id unless the
        // fused text happens to be a piece)
        const size_t mark = out.size();
        size_t ends = size;
        size_t fuse_end = 0;
        bool fusing = false;
        auto flush = [&](size_t fuse_begin) {
            const int32_t id = find(s + fuse_begin, fuse_end - fuse_begin);
            out.push_back(id >= 0 ? id : unk_id);
            fusing = false;
        };
        size_t last_start = 0;
        while (ends > 0) {
            const Node& n = lat[ends];
            const size_t st = static_cast<size_t>(n.start);
            if (n.id == unk_id) {
                if (!fusing) { fusing = true; fuse_end = ends; }
                last_start = st;
            } else {
                if (fusing) flush(last_start);
                out.push_back(n.id);
            }
            ends = st;
        }
        if (fusing) flush(last_start);
        std::reverse(out.begin() + static_cast<long>(mark), out.end());
    }
};

}  // namespace unigram
}  // namespace b200
