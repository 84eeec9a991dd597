/*
 * b2o_snappy.c — CPU oracle: index-based restatement of the reference's snappy
 * 1.1.3 (TEST INFRASTRUCTURE ONLY, see b2o.h).  Each function cites the
 * reference lines whose behaviour it reproduces; file names are relative to
 * /root/reference/src/butil/third_party/snappy/.
 */
#include "b2o.h"

#include <stdlib.h>
#include <string.h>

#define FRAG_BYTES 65536u /* snappy.h:202-203 kBlockSize */
#define MAX_TABLE 16384u  /* snappy.h:205-206 kMaxHashTableSize */
#define INPUT_MARGIN 15u  /* snappy.cc:344 kInputMarginBytes */

static inline uint32_t rd32(const uint8_t* p) {
    uint32_t v;
    memcpy(&v, p, 4);
    return v; /* little-endian host, as the reference's x86 path */
}

/* snappy.cc:47-50 HashBytes */
static inline uint32_t hash32(uint32_t bytes, int shift) {
    return (bytes * 0x1e35a7bdu) >> shift;
}

size_t b2o_snappy_max_compressed_length(size_t n) { return 32 + n + n / 6; }

/* snappy-stubs-internal.h:349-375 Varint::Encode32 */
static size_t put_varint32(uint8_t* dst, uint32_t v) {
    size_t k = 0;
    while (v >= 128) {
        dst[k++] = (uint8_t)(v | 128);
        v >>= 7;
    }
    dst[k++] = (uint8_t)v;
    return k;
}

/* snappy.cc:156-196 EmitLiteral: tag (n=len-1; <60 inline, else 59+count with
 * `count` little-endian length bytes) followed by the bytes. */
static size_t put_literal(uint8_t* dst, const uint8_t* src, uint32_t len) {
    uint32_t n = len - 1;
    size_t k = 0;
    if (n < 60) {
        dst[k++] = (uint8_t)(n << 2);
    } else {
        uint32_t count = 0, t = n;
        k = 1;
        while (t > 0) {
            dst[k++] = (uint8_t)(t & 0xff);
            t >>= 8;
            count++;
        }
        dst[0] = (uint8_t)((59 + count) << 2);
    }
    memcpy(dst + k, src, len);
    return k + len;
}

/* snappy.cc:198-215 EmitCopyLessThan64 */
static size_t put_copy_piece(uint8_t* dst, uint32_t offset, uint32_t len) {
    if (len < 12 && offset < 2048) {
        dst[0] = (uint8_t)(1 + ((len - 4) << 2) + ((offset >> 8) << 5));
        dst[1] = (uint8_t)(offset & 0xff);
        return 2;
    }
    dst[0] = (uint8_t)(2 + ((len - 1) << 2));
    dst[1] = (uint8_t)(offset & 0xff);
    dst[2] = (uint8_t)(offset >> 8);
    return 3;
}

/* snappy.cc:217-232 EmitCopy: 64s while len>=68, one 60 if len>64, the rest. */
static size_t put_copy(uint8_t* dst, uint32_t offset, uint32_t len) {
    size_t k = 0;
    while (len >= 68) {
        k += put_copy_piece(dst + k, offset, 64);
        len -= 64;
    }
    if (len > 64) {
        k += put_copy_piece(dst + k, offset, 60);
        len -= 60;
    }
    k += put_copy_piece(dst + k, offset, len);
    return k;
}

/* snappy-internal.h:86-120 FindMatchLength: common prefix of in[a..] and
 * in[b..], never looking at or beyond in[end] on the b side. */
static uint32_t match_len(const uint8_t* in, uint32_t a, uint32_t b, uint32_t end) {
    uint32_t m = 0;
    while (b + m + 8 <= end) {
        uint64_t x, y;
        memcpy(&x, in + a + m, 8);
        memcpy(&y, in + b + m, 8);
        x ^= y;
        if (x) return m + (uint32_t)(__builtin_ctzll(x) >> 3);
        m += 8;
    }
    while (b + m < end && in[a + m] == in[b + m]) m++;
    return m;
}

/* snappy.cc:247-271 GetHashTable sizing rule */
static uint32_t table_entries(size_t n) {
    uint32_t t = 256;
    while (t < MAX_TABLE && t < n) t <<= 1;
    return t;
}

static int log2_floor(uint32_t v) {
    int r = 0;
    while (v >>= 1) r++;
    return r;
}

/* snappy.cc:329-453 CompressFragment with a caller-provided zeroed table. */
static size_t compress_fragment(const uint8_t* in, uint32_t n, uint8_t* out,
                                uint16_t* table, uint32_t tsize) {
    const int shift = 32 - log2_floor(tsize);
    uint32_t ip = 0, next_emit = 0;
    size_t op = 0;

    if (n >= INPUT_MARGIN) {
        const uint32_t ip_limit = n - INPUT_MARGIN;
        uint32_t next_hash;
        ip = 1;
        next_hash = hash32(rd32(in + ip), shift);
        for (;;) {
            /* probe loop, snappy.cc:377-397 */
            uint32_t skip = 32, next_ip = ip, cand;
            for (;;) {
                uint32_t h = next_hash;
                uint32_t step = skip++ >> 5;
                ip = next_ip;
                next_ip = ip + step;
                if (next_ip > ip_limit) goto remainder;
                next_hash = hash32(rd32(in + next_ip), shift);
                cand = table[h];
                table[h] = (uint16_t)ip;
                if (rd32(in + ip) == rd32(in + cand)) break;
            }
            /* snappy.cc:403 */
            op += put_literal(out + op, in + next_emit, ip - next_emit);
            /* copy loop, snappy.cc:416-439 */
            for (;;) {
                uint32_t base = ip;
                uint32_t matched = 4 + match_len(in, cand + 4, ip + 4, n);
                uint32_t h_prev, h_cur, cur;
                ip += matched;
                op += put_copy(out + op, base - cand, matched);
                next_emit = ip;
                if (ip >= ip_limit) goto remainder;
                h_prev = hash32(rd32(in + ip - 1), shift);
                table[h_prev] = (uint16_t)(ip - 1);
                cur = rd32(in + ip);
                h_cur = hash32(cur, shift);
                cand = table[h_cur];
                table[h_cur] = (uint16_t)ip;
                if (cur != rd32(in + cand)) break;
            }
            /* snappy.cc:441-442 */
            next_hash = hash32(rd32(in + ip + 1), shift);
            ip++;
        }
    }
remainder:
    /* snappy.cc:446-450 */
    if (next_emit < n) op += put_literal(out + op, in + next_emit, n - next_emit);
    return op;
}

size_t b2o_snappy_compress_fragment(const uint8_t* in, size_t n, uint8_t* out) {
    uint16_t* table;
    uint32_t ts = table_entries(n);
    size_t r;
    table = (uint16_t*)calloc(ts, sizeof(uint16_t));
    r = compress_fragment(in, (uint32_t)n, out, table, ts);
    free(table);
    return r;
}

/* snappy.cc:875-954: varint32(N), then each min(N,64 KiB) fragment with a
 * fresh zeroed table.  Fragmentation depends only on N (snappy.cc:892). */
size_t b2o_snappy_compress(const uint8_t* in, size_t n, uint8_t* out) {
    uint16_t* table = (uint16_t*)malloc(MAX_TABLE * sizeof(uint16_t));
    size_t op = put_varint32(out, (uint32_t)n);
    size_t pos = 0;
    while (pos < n) {
        uint32_t take = (uint32_t)((n - pos < FRAG_BYTES) ? (n - pos) : FRAG_BYTES);
        uint32_t ts = table_entries(take);
        memset(table, 0, ts * sizeof(uint16_t));
        op += compress_fragment(in + pos, take, out + op, table, ts);
        pos += take;
    }
    free(table);
    return op;
}

/* snappy.cc:692-711: up to 5 bytes, the 5th byte's high bits silently wrap. */
int b2o_snappy_read_ulen_lenient(const uint8_t* in, size_t n, uint32_t* ulen) {
    uint32_t v = 0, shift = 0;
    size_t i = 0;
    for (;;) {
        uint8_t c;
        if (shift >= 32) return 0;
        if (i >= n) return 0;
        c = in[i++];
        v |= (uint32_t)(c & 0x7f) << shift;
        if (c < 128) break;
        shift += 7;
    }
    *ulen = v;
    return (int)i;
}

/* snappy-stubs-internal.h:325-347 */
int b2o_snappy_read_ulen_strict(const uint8_t* in, size_t n, uint32_t* ulen) {
    uint32_t v = 0;
    int i;
    for (i = 0; i < 5; i++) {
        uint32_t b;
        if ((size_t)i >= n) return 0;
        b = in[i];
        v |= (b & 127u) << (7 * i);
        if (i < 4 ? (b < 128) : (b < 16)) {
            *ulen = v;
            return i + 1;
        }
    }
    return 0;
}

/* Extra bytes that follow a tag byte (bits 11-13 of char_table, snappy.cc:516-549):
 * literal: 0 for len<=60, else 1..4 length bytes; copy-1: 1; copy-2: 2; copy-4: 4. */
static inline uint32_t tag_extra(uint8_t c) {
    switch (c & 3) {
        case 0: return (c >= 0xf0) ? (uint32_t)(c >> 2) - 59 : 0;
        case 1: return 1;
        case 2: return 2;
        default: return 4;
    }
}

static inline uint32_t rd_le(const uint8_t* p, uint32_t k) {
    uint32_t v = 0, i;
    for (i = 0; i < k; i++) v |= (uint32_t)p[i] << (8 * i);
    return v;
}

/* snappy.cc:716-787 DecompressAllTags + RefillTag (snappy.cc:789-847) over one
 * flat buffer, with the writer checks of snappy.cc:1165-1167,1186-1213 (array
 * writer) or 1252-1280 (validator, out == NULL).  Returns 1 iff
 * eof && produced == expected (snappy.cc:867). */
static int decode_tags(const uint8_t* in, size_t n, size_t ip, uint8_t* out,
                       uint32_t expected) {
    size_t produced = 0;
    for (;;) {
        uint8_t c;
        uint32_t extra;
        if (ip == n) return produced == expected; /* RefillTag: clean eof */
        c = in[ip];
        extra = tag_extra(c);
        if (n - ip < (size_t)extra + 1) return 0; /* RefillTag: tag cut short */
        ip++;
        if ((c & 3) == 0) {
            uint32_t len = (uint32_t)(c >> 2) + 1;
            if (len >= 61) len = rd_le(in + ip, extra) + 1u; /* may wrap to 0, snappy.cc:746-750 */
            ip += extra;
            if (n - ip < len) return 0;                   /* premature end, snappy.cc:761 */
            if ((size_t)expected - produced < len) return 0; /* Append overrun */
            if (out) memcpy(out + produced, in + ip, len);
            produced += len;
            ip += len;
        } else {
            uint32_t len, offset;
            if ((c & 3) == 1) {
                len = ((c >> 2) & 7) + 4;
                offset = ((uint32_t)(c >> 5) << 8) | in[ip];
            } else {
                len = (uint32_t)(c >> 2) + 1;
                offset = rd_le(in + ip, extra);
            }
            ip += extra;
            if (produced <= (size_t)offset - 1u) return 0; /* offset==0 or > produced */
            if ((size_t)expected - produced < len) return 0;
            if (out) {
                /* snappy.cc:98-103 IncrementalCopy: byte order matters when len>offset */
                uint8_t* d = out + produced;
                const uint8_t* s = d - offset;
                uint32_t i;
                for (i = 0; i < len; i++) d[i] = s[i];
            }
            produced += len;
        }
    }
}

int b2o_snappy_uncompress(const uint8_t* in, size_t n, uint8_t* out, size_t cap,
                          uint32_t* ulen_out) {
    uint32_t ulen = 0;
    int k = b2o_snappy_read_ulen_lenient(in, n, &ulen);
    if (ulen_out) *ulen_out = ulen;
    if (!k) return 0;
    if ((size_t)ulen > cap) return 0;
    return decode_tags(in, n, (size_t)k, out, ulen);
}

int b2o_snappy_is_valid(const uint8_t* in, size_t n) {
    uint32_t ulen = 0;
    int k = b2o_snappy_read_ulen_lenient(in, n, &ulen);
    if (!k) return 0;
    return decode_tags(in, n, (size_t)k, NULL, ulen);
}
