// host_gather.cpp — the CPU side of the staging gather: one message's IOBuf blocks
// (src/butil/iobuf.h:74-101) concatenated into page-locked staging.
//
// Staging is written once by a core and then only read by the DMA engine, so the bytes go out
// with non-temporal stores, a whole 64-byte line at a time: a line that straddles two blocks is
// first assembled in a stack buffer, so the write-combining buffers only ever see complete
// lines (partial ones fall back to read-for-ownership and cost ~2x).  Plain C++ (no CUDA) so the
// host compiler sees the AVX-512 variant directly; picked at run time.
#include <emmintrin.h>
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "b2c_codec.h"

namespace b2c {

namespace {

typedef void (*LinesFn)(uint8_t* d, const uint8_t* s, size_t lines);  // d 64-byte aligned

void lines_sse2(uint8_t* d, const uint8_t* s, size_t lines) {
    for (size_t i = 0; i < lines; i++, s += 64, d += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 32));
        const __m128i e = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 48));
        _mm_stream_si128(reinterpret_cast<__m128i*>(d), a);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 32), c);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 48), e);
    }
}

__attribute__((target("avx512f"))) void lines_avx512(uint8_t* d, const uint8_t* s, size_t lines) {
    for (size_t i = 0; i < lines; i++, s += 64, d += 64)
        _mm512_stream_si512(reinterpret_cast<__m512i*>(d), _mm512_loadu_si512(s));
}

LinesFn pick_lines() {
    __builtin_cpu_init();
    return __builtin_cpu_supports("avx512f") ? lines_avx512 : lines_sse2;
}

const LinesFn g_lines = pick_lines();

}  // namespace

// Concatenate `nseg` blocks at dst.  `stream`: use non-temporal line stores (the caller issues
// one sfence per batch of messages, see gather_fence).
void gather_message(uint8_t* d, const b2c_iovec* segs, uint32_t nseg, uint64_t total, bool stream) {
    if (!stream || total < 256) {
        for (uint32_t q = 0; q < nseg; q++) {
            memcpy(d, segs[q].iov_base, segs[q].iov_len);
            d += segs[q].iov_len;
        }
        return;
    }
    alignas(64) uint8_t tmp[64];
    size_t fill = 0;                                                  // bytes of the current line waiting in tmp
    size_t head = (64 - (reinterpret_cast<uintptr_t>(d) & 63)) & 63;  // up to the first whole line: plain stores
    for (uint32_t q = 0; q < nseg; q++) {
        const uint8_t* s = static_cast<const uint8_t*>(segs[q].iov_base);
        size_t r = segs[q].iov_len;
        if (head) {
            const size_t t = std::min(head, r);
            memcpy(d, s, t);
            d += t, s += t, r -= t, head -= t;
        }
        if (fill && r) {
            const size_t t = std::min(64 - fill, r);
            memcpy(tmp + fill, s, t);
            fill += t, s += t, r -= t;
            if (fill == 64) {
                g_lines(d, tmp, 1);
                d += 64;
                fill = 0;
            }
        }
        if (r >= 64) {
            const size_t lines = r / 64;
            g_lines(d, s, lines);
            d += lines * 64, s += lines * 64, r -= lines * 64;
        }
        if (r) {  // fill == 0 here
            memcpy(tmp, s, r);
            fill = r;
        }
    }
    if (fill) memcpy(d, tmp, fill);
}

void gather_fence() { _mm_sfence(); }

}  // namespace b2c
