// hostbw.cpp — how fast can this box's cores gather small blocks?  (sizing the e2e staging gather)
//   g++ -O2 -pthread tools/hostbw.cpp -o /tmp/hostbw && /tmp/hostbw
#include <emmintrin.h>
#include <immintrin.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <thread>
#include <vector>

static void copy_nt_sse(uint8_t* d, const uint8_t* s, size_t n) {
    for (size_t i = 0; i + 64 <= n; i += 64) {
        __m128i a = _mm_loadu_si128((const __m128i*)(s + i)), b = _mm_loadu_si128((const __m128i*)(s + i + 16)),
                c = _mm_loadu_si128((const __m128i*)(s + i + 32)), e = _mm_loadu_si128((const __m128i*)(s + i + 48));
        _mm_stream_si128((__m128i*)(d + i), a);
        _mm_stream_si128((__m128i*)(d + i + 16), b);
        _mm_stream_si128((__m128i*)(d + i + 32), c);
        _mm_stream_si128((__m128i*)(d + i + 48), e);
    }
}
__attribute__((target("avx512f"))) static void copy_nt_512(uint8_t* d, const uint8_t* s, size_t n) {
    for (size_t i = 0; i + 64 <= n; i += 64) _mm512_stream_si512((__m512i*)(d + i), _mm512_loadu_si512(s + i));
}
__attribute__((target("avx2"))) static void copy_nt_256(uint8_t* d, const uint8_t* s, size_t n) {
    for (size_t i = 0; i + 64 <= n; i += 64) {
        __m256i a = _mm256_loadu_si256((const __m256i*)(s + i)), b = _mm256_loadu_si256((const __m256i*)(s + i + 32));
        _mm256_stream_si256((__m256i*)(d + i), a);
        _mm256_stream_si256((__m256i*)(d + i + 32), b);
    }
}

int main() {
    const size_t N = 1ull << 30, piece = 1024;
    uint8_t* src = (uint8_t*)aligned_alloc(4096, N);
    uint8_t* dst = (uint8_t*)aligned_alloc(4096, N);
    memset(src, 1, N);
    memset(dst, 2, N);
    const bool has512 = __builtin_cpu_supports("avx512f"), has256 = __builtin_cpu_supports("avx2");
    printf("avx2 %d avx512f %d hw threads %u\n", has256, has512, std::thread::hardware_concurrency());
    for (int mode = 0; mode < 5; mode++) {
        if (mode == 2 && !has256) continue;
        if (mode == 3 && !has512) continue;
        for (unsigned T : {4u, 8u, 15u, 16u}) {
            double best = 1e9;
            for (int rep = 0; rep < 3; rep++) {
                auto t0 = std::chrono::steady_clock::now();
                std::vector<std::thread> th;
                for (unsigned t = 0; t < T; t++)
                    th.emplace_back([=] {
                        size_t lo = N / piece * t / T, hi = N / piece * (t + 1) / T;
                        for (size_t i = lo; i < hi; i++) {
                            uint8_t* d = dst + i * piece;
                            const uint8_t* s = src + i * piece;
                            if (mode == 0) memcpy(d, s, piece);
                            else if (mode == 1) copy_nt_sse(d, s, piece);
                            else if (mode == 2) copy_nt_256(d, s, piece);
                            else if (mode == 3) copy_nt_512(d, s, piece);
                            else memcpy(d, s, piece / 2), memcpy(d + piece / 2, s + piece / 2, piece / 2);
                        }
                        _mm_sfence();
                    });
                for (auto& x : th) x.join();
                best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
            }
            const char* names[] = {"memcpy", "nt-sse2", "nt-avx2", "nt-avx512", "memcpy x2"};
            printf("%-10s T=%2u  %.1f GB/s copied\n", names[mode], T, N / best / 1e9);
        }
    }
    return 0;
}
