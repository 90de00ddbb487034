// Safetensors reader/writer and the seeded random init (weights_io.cpp).
#pragma once
#include <cstddef>
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace b200 {

class SafetensorsFile {
  public:
    SafetensorsFile() = default;
    ~SafetensorsFile();
    SafetensorsFile(const SafetensorsFile&) = delete;
    SafetensorsFile& operator=(const SafetensorsFile&) = delete;
    bool open(const std::string& path, std::string* err);
    // Copies tensor `name` (HF BERT name without prefix) as f32 into dst; the shape must match.
    bool read_f32(const std::string& name, const std::vector<int64_t>& shape, float* dst, std::string* err) const;

  private:
    struct Entry {
        std::string dtype;
        std::vector<int64_t> shape;
        uint64_t begin = 0, end = 0;
    };
    int fd_ = -1;
    void* map_ = nullptr;
    size_t size_ = 0;
    const uint8_t* data_ = nullptr;
    std::string prefix_;
    std::map<std::string, Entry> entries_;
};

struct NamedTensor {
    std::string name;
    std::vector<int64_t> shape;
    const float* data;
};
bool write_safetensors(const std::string& path, const std::vector<NamedTensor>& tensors, std::string* err);

// value_i = float(mean + 0.02 * N(0,1)) from splitmix64(fnv1a64(name) ^ seed*phi + {2i, 2i+1}) and
// Box-Muller in double; mean = 1 for "...LayerNorm.weight", else 0. Same spec as weights.py.
void random_init_tensor(const std::string& name, uint64_t seed, float* dst, size_t n);

}  // namespace b200
