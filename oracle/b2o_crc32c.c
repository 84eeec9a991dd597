/*
 * b2o_crc32c.c — CPU oracle for butil::crc32c (TEST INFRASTRUCTURE ONLY, see
 * b2o.h).  Restates /root/reference/src/butil/crc32c.cc: CRC-32C (Castagnoli),
 * reflected polynomial 0x82F63B78, register pre- and post-inverted
 * (crc32c.cc:386,420).  The reference carries its tables as literals
 * (crc32c.cc:25-288); here they are generated from the polynomial and checked
 * against the reference library by tests/test_oracle.py.
 */
#include "b2o.h"

#include <string.h>
#ifdef __SSE4_2__
#include <nmmintrin.h>
#endif

#define POLY_REFLECTED 0x82F63B78u

static uint32_t T[4][256];
static int t_ready;

__attribute__((constructor)) static void build_tables(void) {
    uint32_t i, k;
    for (i = 0; i < 256; i++) {
        uint32_t c = i;
        for (k = 0; k < 8; k++) c = (c >> 1) ^ ((c & 1) ? POLY_REFLECTED : 0);
        T[0][i] = c;
    }
    /* T[k][b] = T[k-1][b] advanced by one more zero byte: the slice-by-4
     * tables table1_..table3_ of crc32c.cc:90-288 */
    for (i = 0; i < 256; i++)
        for (k = 1; k < 4; k++) T[k][i] = T[0][T[k - 1][i] & 0xff] ^ (T[k - 1][i] >> 8);
    t_ready = 1;
}

static inline uint32_t step1(uint32_t l, uint8_t b) { /* crc32c.cc:390-393 STEP1 */
    return T[0][(l ^ b) & 0xff] ^ (l >> 8);
}

static inline uint32_t step4_tables(uint32_t l, const uint8_t* p) { /* crc32c.cc:333-347 */
    uint32_t w;
    memcpy(&w, p, 4);
    w ^= l;
    return T[3][w & 0xff] ^ T[2][(w >> 8) & 0xff] ^ T[1][(w >> 16) & 0xff] ^ T[0][w >> 24];
}

/* crc32c.cc:379-422 ExtendImpl: bytes up to 16-byte alignment, 16 B and 8 B
 * strides, byte tail.  `fast` selects the SSE4.2 instruction (crc32c.cc:349-363). */
static uint32_t extend_impl(uint32_t crc, const uint8_t* p, size_t n, int fast) {
    const uint8_t* e = p + n;
    uint64_t l = crc ^ 0xffffffffu;
    const uint8_t* x = (const uint8_t*)(((uintptr_t)p + 15) & ~(uintptr_t)15);
    if (!t_ready) build_tables();
    if (x <= e)
        while (p != x) l = step1((uint32_t)l, *p++);
    while ((size_t)(e - p) >= 8) {
#if defined(__SSE4_2__) && defined(__LP64__)
        if (fast) {
            uint64_t w;
            memcpy(&w, p, 8);
            l = _mm_crc32_u64(l, w);
            p += 8;
            continue;
        }
#endif
        l = step4_tables((uint32_t)l, p);
        l = step4_tables((uint32_t)l, p + 4);
        p += 8;
    }
    while (p != e) l = step1((uint32_t)l, *p++);
    (void)fast;
    return (uint32_t)l ^ 0xffffffffu;
}

uint32_t b2o_crc32c_extend(uint32_t init, const uint8_t* data, size_t n) {
    return extend_impl(init, data, n, 1);
}
uint32_t b2o_crc32c_extend_tables(uint32_t init, const uint8_t* data, size_t n) {
    return extend_impl(init, data, n, 0);
}
uint32_t b2o_crc32c_value(const uint8_t* data, size_t n) { /* crc32c.h:27-29 */
    return b2o_crc32c_extend(0, data, n);
}
uint32_t b2o_crc32c_mask(uint32_t crc) { /* crc32c.h:38-41 */
    return ((crc >> 15) | (crc << 17)) + 0xa282ead8u;
}
uint32_t b2o_crc32c_unmask(uint32_t m) { /* crc32c.h:44-47 */
    uint32_t rot = m - 0xa282ead8u;
    return (rot >> 17) | (rot << 15);
}
