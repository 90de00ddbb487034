// Exhaustive check of the f32 shortest-digit generator of csrc/json_out.cpp (Schubfach) against the C++ library's
// std::to_chars for every positive finite float, and a timing of both. Host only:
//   g++ -std=c++17 -O2 scripts/probes/f32_digits_check.cpp -o /tmp/f32check -lpthread && /tmp/f32check [stride]
#include "../../codename-symbiont_b200/csrc/json_out.cpp"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
    const uint32_t stride = argc > 1 ? static_cast<uint32_t>(std::atoi(argv[1])) : 1;
    const int nt = static_cast<int>(std::thread::hardware_concurrency() ? std::thread::hardware_concurrency() : 4);
    std::atomic<long> bad{0}, total{0};
    auto work = [&](int tid) {
        long lb = 0, lt = 0;
        for (uint64_t b = 1 + static_cast<uint64_t>(tid) * stride; b < 0x7F800000ull; b += static_cast<uint64_t>(nt) * stride) {
            const uint32_t bits = static_cast<uint32_t>(b);
            float v;
            std::memcpy(&v, &bits, 4);
            char d1[32], d2[32];
            int k1, k2;
            const int n1 = shortest_digits_fast(v, d1, &k1), n2 = shortest_digits(v, d2, &k2);
            ++lt;
            if (n1 != n2 || k1 != k2 || std::memcmp(d1, d2, static_cast<size_t>(n1)) != 0)
                if (++lb <= 3) std::printf("MISMATCH bits %08x (%.9g): %.*s e%d vs %.*s e%d\n", bits, v, n1, d1, k1, n2, d2, k2);
        }
        bad += lb;
        total += lt;
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
    std::printf("checked %ld floats, %ld mismatches\n", total.load(), bad.load());
    std::vector<float> v(4 << 20);
    uint32_t x = 12345;
    for (auto& f : v) { x = x * 1664525u + 1013904223u; f = (static_cast<int>(x >> 8) - (1 << 23)) * (0.3f / (1 << 23)); }
    char buf[64];
    size_t acc = 0;
    int kk;
    auto t0 = std::chrono::steady_clock::now();
    for (float f : v) acc += shortest_digits_fast(std::fabs(f) + 1e-30f, buf, &kk);
    auto t1 = std::chrono::steady_clock::now();
    for (float f : v) acc += shortest_digits(std::fabs(f) + 1e-30f, buf, &kk);
    auto t2 = std::chrono::steady_clock::now();
    for (float f : v) acc += format_f32(f, buf);
    auto t3 = std::chrono::steady_clock::now();
    std::printf("schubfach %.1f ns, to_chars %.1f ns, format_f32 %.1f ns per float (%zu)\n",
                std::chrono::duration<double, std::nano>(t1 - t0).count() / v.size(), std::chrono::duration<double, std::nano>(t2 - t1).count() / v.size(),
                std::chrono::duration<double, std::nano>(t3 - t2).count() / v.size(), acc);
    return bad.load() ? 1 : 0;
}
