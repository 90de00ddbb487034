// Host-side weight sources: a safetensors reader/writer and the seeded random init.
//
// Replaces, for this path, the reference's weight loading
// (/root/reference/services/preprocessing_service/src/embedding_generator.rs:106-124:
// VarBuilder::from_mmaped_safetensors(files, F32, device) + BertModel::load). Tensor names are the
// HuggingFace BERT state-dict names candle consumes, with the same fallbacks:
// an optional "bert." / model-type prefix and the legacy LayerNorm "gamma"/"beta" names.
#include "weights_io.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

namespace b200 {

namespace {

// ---- minimal JSON reader for the safetensors header (objects, arrays, strings, numbers) --------
struct JsonCursor {
    const char* p;
    const char* end;
    bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
    bool eat(char c) { ws(); if (p < end && *p == c) { ++p; return true; } return false; }
    std::string str() {
        ws();
        std::string s;
        if (p >= end || *p != '"') { ok = false; return s; }
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) { ++p; }
            s.push_back(*p++);
        }
        if (p < end) ++p; else ok = false;
        return s;
    }
    double num() {
        ws();
        char* e = nullptr;
        double v = strtod(p, &e);
        if (e == p) ok = false;
        p = e;
        return v;
    }
    void skip_value() {
        ws();
        if (p >= end) { ok = false; return; }
        if (*p == '"') { str(); return; }
        if (*p == '{') {
            ++p;
            if (eat('}')) return;
            do { str(); if (!eat(':')) { ok = false; return; } skip_value(); } while (ok && eat(','));
            if (!eat('}')) ok = false;
            return;
        }
        if (*p == '[') {
            ++p;
            if (eat(']')) return;
            do { skip_value(); } while (ok && eat(','));
            if (!eat(']')) ok = false;
            return;
        }
        while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;  // number / true / false / null
    }
};

float half_to_float(uint16_t h) {
    const uint32_t sign = (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ff;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            exp = 127 - 15 + 1;
            while (!(man & 0x400)) { man <<= 1; --exp; }
            bits = sign | (exp << 23) | ((man & 0x3ff) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint64_t fnv1a64(const std::string& s) {
    uint64_t h = 0xCBF29CE484222325ull;
    for (unsigned char c : s) { h ^= c; h *= 0x100000001B3ull; }
    return h;
}
inline uint64_t splitmix64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

}  // namespace

SafetensorsFile::~SafetensorsFile() {
    if (map_ && map_ != MAP_FAILED) munmap(map_, size_);
    if (fd_ >= 0) close(fd_);
}

bool SafetensorsFile::open(const std::string& path, std::string* err) {
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) { *err = "cannot open " + path; return false; }
    struct stat st;
    if (fstat(fd_, &st) != 0 || st.st_size < 8) { *err = "cannot stat / too small: " + path; return false; }
    size_ = static_cast<size_t>(st.st_size);
    map_ = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (map_ == MAP_FAILED) { *err = "mmap failed: " + path; return false; }
    const uint8_t* b = static_cast<const uint8_t*>(map_);
    uint64_t hlen = 0;
    memcpy(&hlen, b, 8);  // little endian
    if (hlen > size_ - 8) { *err = "bad safetensors header length"; return false; }
    data_ = b + 8 + hlen;
    const size_t data_bytes = size_ - 8 - hlen;
    JsonCursor c{reinterpret_cast<const char*>(b + 8), reinterpret_cast<const char*>(b + 8 + hlen)};
    if (!c.eat('{')) { *err = "safetensors header is not a JSON object"; return false; }
    if (!c.eat('}')) {
        do {
            const std::string name = c.str();
            if (!c.eat(':')) { c.ok = false; break; }
            if (name == "__metadata__") { c.skip_value(); continue; }
            Entry e;
            if (!c.eat('{')) { c.ok = false; break; }
            do {
                const std::string key = c.str();
                if (!c.eat(':')) { c.ok = false; break; }
                if (key == "dtype") e.dtype = c.str();
                else if (key == "shape") {
                    if (!c.eat('[')) { c.ok = false; break; }
                    if (!c.eat(']')) { do { e.shape.push_back(static_cast<int64_t>(c.num())); } while (c.ok && c.eat(',')); if (!c.eat(']')) c.ok = false; }
                } else if (key == "data_offsets") {
                    if (!c.eat('[')) { c.ok = false; break; }
                    e.begin = static_cast<uint64_t>(c.num());
                    if (!c.eat(',')) { c.ok = false; break; }
                    e.end = static_cast<uint64_t>(c.num());
                    if (!c.eat(']')) c.ok = false;
                } else c.skip_value();
            } while (c.ok && c.eat(','));
            if (!c.ok || !c.eat('}')) { c.ok = false; break; }
            if (e.end < e.begin || e.end > data_bytes) { *err = "tensor " + name + " lies outside the file"; return false; }
            entries_[name] = e;
        } while (c.ok && c.eat(','));
        if (c.ok && !c.eat('}')) c.ok = false;
    }
    if (!c.ok) { *err = "cannot parse safetensors header of " + path; return false; }
    // candle retries with "<model_type>." in front of every name when the bare names fail
    for (const char* pre : {"", "bert.", "roberta.", "xlm_roberta.", "model."})
        if (entries_.count(std::string(pre) + "embeddings.word_embeddings.weight")) { prefix_ = pre; break; }
    return true;
}

bool SafetensorsFile::read_f32(const std::string& name, const std::vector<int64_t>& shape, float* dst,
                               std::string* err) const {
    auto it = entries_.find(prefix_ + name);
    if (it == entries_.end()) {  // legacy LayerNorm parameter names
        std::string alt = name;
        const size_t n = alt.size();
        if (n > 16 && alt.compare(n - 16, 16, "LayerNorm.weight") == 0) alt.replace(n - 6, 6, "gamma");
        else if (n > 14 && alt.compare(n - 14, 14, "LayerNorm.bias") == 0) alt.replace(n - 4, 4, "beta");
        it = entries_.find(prefix_ + alt);
    }
    if (it == entries_.end()) { *err = "tensor not found in safetensors file: " + name; return false; }
    const Entry& e = it->second;
    if (e.shape != shape) { *err = "tensor " + name + " has an unexpected shape"; return false; }
    size_t n = 1;
    for (int64_t d : shape) n *= static_cast<size_t>(d);
    const uint8_t* src = data_ + e.begin;
    const size_t bytes = e.end - e.begin;
    if (e.dtype == "F32") {
        if (bytes != n * 4) { *err = "tensor " + name + " has an unexpected byte size"; return false; }
        memcpy(dst, src, bytes);
    } else if (e.dtype == "F16" || e.dtype == "BF16") {
        if (bytes != n * 2) { *err = "tensor " + name + " has an unexpected byte size"; return false; }
        const bool bf = e.dtype == "BF16";
        for (size_t i = 0; i < n; ++i) {
            uint16_t h;
            memcpy(&h, src + 2 * i, 2);
            if (bf) { const uint32_t u = static_cast<uint32_t>(h) << 16; memcpy(dst + i, &u, 4); }
            else dst[i] = half_to_float(h);
        }
    } else { *err = "tensor " + name + " has unsupported dtype " + e.dtype; return false; }
    return true;
}

void random_init_tensor(const std::string& name, uint64_t seed, float* dst, size_t n) {
    const uint64_t key = fnv1a64(name) ^ (seed * 0x9E3779B97F4A7C15ull);
    const bool ln_w = name.size() >= 16 && name.compare(name.size() - 16, 16, "LayerNorm.weight") == 0;
    const double mean = ln_w ? 1.0 : 0.0;
    unsigned nt = std::thread::hardware_concurrency();
    if (nt == 0) nt = 1;
    if (nt > 32) nt = 32;
    if (n < (1u << 16)) nt = 1;
    auto work = [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            const uint64_t a = splitmix64(key + 2 * i), b = splitmix64(key + 2 * i + 1);
            const double u1 = (static_cast<double>(a >> 11) + 1.0) * 0x1.0p-53;
            const double u2 = static_cast<double>(b >> 11) * 0x1.0p-53;
            const double z = std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * M_PI * u2);
            dst[i] = static_cast<float>(mean + 0.02 * z);
        }
    };
    if (nt == 1) { work(0, n); return; }
    std::vector<std::thread> th;
    const size_t per = (n + nt - 1) / nt;
    for (unsigned t = 0; t < nt; ++t) {
        const size_t lo = t * per, hi = lo + per < n ? lo + per : n;
        if (lo < hi) th.emplace_back(work, lo, hi);
    }
    for (auto& t : th) t.join();
}

bool write_safetensors(const std::string& path, const std::vector<NamedTensor>& tensors, std::string* err) {
    std::string header = "{";
    uint64_t off = 0;
    for (size_t i = 0; i < tensors.size(); ++i) {
        const NamedTensor& t = tensors[i];
        size_t n = 1;
        std::string shape = "[";
        for (size_t d = 0; d < t.shape.size(); ++d) {
            n *= static_cast<size_t>(t.shape[d]);
            shape += std::to_string(t.shape[d]);
            if (d + 1 < t.shape.size()) shape += ",";
        }
        shape += "]";
        if (i) header += ",";
        header += "\"" + t.name + "\":{\"dtype\":\"F32\",\"shape\":" + shape + ",\"data_offsets\":[" +
                  std::to_string(off) + "," + std::to_string(off + n * 4) + "]}";
        off += n * 4;
    }
    header += "}";
    while (header.size() % 8) header.push_back(' ');
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { *err = "cannot create " + path; return false; }
    const uint64_t hlen = header.size();
    bool ok = fwrite(&hlen, 8, 1, f) == 1 && fwrite(header.data(), 1, header.size(), f) == header.size();
    for (const NamedTensor& t : tensors) {
        size_t n = 1;
        for (int64_t d : t.shape) n *= static_cast<size_t>(d);
        ok = ok && fwrite(t.data, 4, n, f) == n;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) *err = "short write to " + path;
    return ok;
}

}  // namespace b200
