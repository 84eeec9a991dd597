// snappy_lockstep_dec.cuh — DecompressAllTags (reference snappy.cc:716-787) for streams whose output is at most
// 1 KiB, as a per-lane state machine with one kind of step, so that the 32 lanes of a warp decode 32 different
// streams in lock-step.
//
// snappy_small.cu decodes tag by tag with byte loops, one lane per stream: a lane copying a long element keeps
// the other lanes of its warp waiting, so warps carry 8 streams and issue every instruction for 6 of 32 lanes.
// Here a trip is the same straight-line code for every lane:
//     (if the lane has no element in progress) decode the tag at ip, check everything the reference checks
//     (snappy.cc:731-784, 1200-1213), and then, always, move up to 8 bytes of the element in progress —
//     from the stream (a literal) or from the lane's own output (a copy; offsets below 8 replicate their pattern as
//     IncrementalCopy does, snappy.cc:98-103).
// A `words` body has ~165 elements of 6.6 bytes: one trip each.
//
// The output slot is word-interleaved over the 32 lanes (word w of lane l at index w * 32 + l): lock-step accesses
// never collide on a bank.  Bytes are appended at `op`; the partially filled word is kept in a register, so an
// append is at most three word stores and never a read-modify-write.
//
// Plain C++ on pointers: tests/cpp/lockstep_model.cpp compiles this for the host and tests/test_lockstep_model.py
// checks it against the oracle without a GPU.
#pragma once

#include <stdint.h>

#include "snappy_lockstep.cuh"  // LS_HD, fsr, umin

namespace b2c {
namespace lsd {

constexpr uint32_t kSlots = 32;
constexpr uint32_t kOutMax = 1024;
constexpr uint32_t kSlotWords = kOutMax / 4 + 4;  // an append writes up to three words from the word of `op`

enum : int32_t { kOk = 0, kCorrupt = 1, kTooSmall = 2 };  // B2C_OK / B2C_CORRUPT / B2C_TOO_SMALL

// byte_perm selectors that expand the first o bytes of an 8-byte window into the pattern s[i] = b[i % o], o = 1..7
LS_HD uint32_t pattern_sel(uint32_t o, uint32_t hi_half) {
    // nibble i of the selector = (i + 4 * hi_half) % o
    uint32_t sel = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (uint32_t i = 0; i < 4; i++) sel |= ((i + 4 * hi_half) % o) << (4 * i);
    return sel;
}
LS_HD uint32_t byte_perm(uint32_t lo, uint32_t hi, uint32_t sel) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(lo, hi, sel);
#else
    const uint64_t v = ((uint64_t)hi << 32) | lo;
    uint32_t r = 0;
    for (uint32_t i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
#endif
}

struct DecodeLane {
    const uint8_t* in;  // the stream (global memory); reads touch only aligned words that hold one of its bytes
    uint32_t* slot;     // word w of this lane's output at slot[w * 32]
    uint32_t n, ip, expected, op;
    uint32_t rem;       // bytes of the element in progress still to move
    uint32_t src;       // where they come from: a stream position (literal) or an output position (copy)
    uint32_t is_lit, first;  // first: the next chunk is the first of a copy whose offset is below 8
    uint32_t offset;
    uint32_t cur;       // the bytes of the word that holds `op` which lie below op
    uint32_t live;
    int32_t status;

    // 12 stream bytes from position p (bytes past the end of the stream read as 0)
    LS_HD void load12(uint32_t p, uint32_t& g0, uint32_t& g1, uint32_t& g2) const {
        const uintptr_t a = reinterpret_cast<uintptr_t>(in + p);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
        const uint32_t lead = (uint32_t)(a & 3), sh = lead * 8;
        const uint32_t left = n > p ? n - p : 0u;  // stream bytes from p on
        // word k holds stream bytes iff 4k < lead + left
        const uint32_t have = lead + left;
        const uint32_t w0 = have > 0 ? w[0] : 0u, w1 = have > 4 ? w[1] : 0u, w2 = have > 8 ? w[2] : 0u, w3 = have > 12 ? w[3] : 0u;
        g0 = ls::fsr(w0, w1, sh);
        g1 = ls::fsr(w1, w2, sh);
        g2 = ls::fsr(w2, w3, sh);
    }

    LS_HD void init(const uint8_t* in_, uint32_t n_, uint32_t* slot_, uint32_t cap) {
        in = in_;
        n = n_;
        slot = slot_;
        ip = op = rem = src = is_lit = first = offset = cur = expected = 0;
        live = 1;
        status = kOk;
        // lenient preamble, snappy.cc:692-711
        uint32_t shift = 0;
        for (;;) {
            if (shift >= 32 || ip >= n) {
                status = kCorrupt;
                live = 0;
                return;
            }
            const uint32_t c = in[ip++];
            expected |= (c & 0x7f) << shift;
            if (c < 128) break;
            shift += 7;
        }
        if (expected > cap) {
            status = kTooSmall;
            live = 0;
        }
    }

    LS_HD void trip() {
        // ---- the stream bytes this trip needs: the tag at ip, or the next chunk of a literal in progress ----
        const bool need_tag = rem == 0;
        if (need_tag && ip == n) {  // clean end of input: snappy.cc:867
            if (op != expected) status = kCorrupt;
            live = 0;
            return;
        }
        const uint32_t gp = need_tag ? ip : src;
        uint32_t g0, g1, g2;
        load12(gp, g0, g1, g2);
        uint32_t lo = g0, hi = g1, avail = 8;  // a literal's next bytes
        if (need_tag) {
            const uint32_t c = g0 & 0xff, kind = c & 3;
            uint32_t len = (c >> 2) + 1;             // literal < 61, copy-2, copy-4
            uint32_t used = kind + 1;                // 1, 2, 3 (copy-4 and long literals: below)
            uint32_t off = (g0 >> 8) & 0xffffu;      // copy-2
            if (kind == 1) {
                len = ((c >> 2) & 7) + 4;
                off = ((c >> 5) << 8) | ((g0 >> 8) & 0xff);
            }
            const bool long_lit = kind == 0 && c >= 0xf0u;
            if (long_lit) {  // literal with 1..4 length bytes
                const uint32_t extra = len - 60;
                uint32_t v = (g0 >> 8) | (g1 << 24);
                if (extra < 4) v &= (1u << (8 * extra)) - 1u;
                len = v + 1u;  // may wrap to 0 like the reference
                used = 1 + extra;
            }
            if (kind == 3) {
                off = (g0 >> 8) | (g1 << 24);
                used = 5;
            }
            const uint32_t left = n - ip;
            bool bad = left < used;                              // tag cut short
            if (!bad) {
                if (kind == 0) bad = left - used < len;          // premature end of input
                else bad = off - 1u >= op;                       // offset 0 or beyond what was produced
            }
            if (bad || expected - op < len) {                    // or more output than announced
                status = kCorrupt;
                live = 0;
                return;
            }
            rem = len;
            is_lit = kind == 0 ? 1u : 0u;
            offset = off;
            if (is_lit) {
                src = ip + used;
                ip += used + len;
                // the literal's first bytes came with the tag: bytes used.. of (g0, g1, g2)
                const uint32_t sh = (used & 3) * 8;
                const uint32_t a = used < 4 ? g0 : g1, b = used < 4 ? g1 : g2, d = used < 4 ? g2 : 0u;
                lo = ls::fsr(a, b, sh);
                hi = ls::fsr(b, d, sh);
                avail = 12 - used;  // 7 when the tag had 4 length bytes
            } else {
                ip += used;
                src = op - off;
                first = off < 8 ? 1u : 0u;
            }
        }
        // ---- a copy's next bytes come from the lane's own output ----
        if (!is_lit) {
            const uint32_t* w = slot + (src >> 2) * kSlots;
            const uint32_t sh = (src & 3) * 8;
            const uint32_t w0 = w[0], w1 = w[kSlots], w2 = w[2 * kSlots];
            // the word that holds op is in `cur`, not in the slot yet
            const uint32_t wc = op >> 2, ws = src >> 2;
            const uint32_t x0 = ws == wc ? cur : w0, x1 = ws + 1 == wc ? cur : w1, x2 = ws + 2 == wc ? cur : w2;
            lo = ls::fsr(x0, x1, sh);
            hi = ls::fsr(x1, x2, sh);
            avail = 8;
            if (first) {  // offset o < 8: the 8 bytes are the pattern of the o bytes at src (IncrementalCopy)
                const uint32_t o = offset;
                const uint32_t plo = byte_perm(lo, hi, pattern_sel(o, 0)), phi = byte_perm(lo, hi, pattern_sel(o, 1));
                lo = plo;
                hi = phi;
                // from here on the run continues `o2` bytes back, the smallest multiple of o that is >= 8: those
                // bytes are written by now and no longer overlap the 8 bytes a trip moves
                const uint32_t o2 = o * ((7 + o) / o);
                offset = o2;
                first = 0;
                src = op - o2;  // + k below: the next chunk starts at (op + k) - o2
            }
        }
        // ---- append k bytes at op ----
        const uint32_t k = ls::umin(ls::umin(rem, 8u), avail);
        const uint32_t sh = (op & 3) * 8;
        uint32_t* d = slot + (op >> 2) * kSlots;
        // bytes 0..11 of the three words from the word of op: cur below op, then lo, hi
        const uint32_t d0 = cur | (lo << sh);
        const uint32_t d1 = sh ? (lo >> (32 - sh)) | (hi << sh) : hi;
        const uint32_t d2 = sh ? hi >> (32 - sh) : 0u;
        const uint32_t end = (op & 3) + k;  // bytes of (d0, d1, d2) that are final: 1..11
        if (end >= 4) d[0] = d0;
        if (end >= 8) d[kSlots] = d1;
        // the new partial word: the bytes of the word that holds op + k which lie below op + k
        const uint32_t part = end & 3, wsel = end >> 2;
        const uint32_t dw = wsel == 0 ? d0 : wsel == 1 ? d1 : d2;
        cur = part ? dw & ((1u << (8 * part)) - 1u) : 0u;
        op += k;
        src += k;
        rem -= k;
    }

    // the last partial word goes to the slot; returns the status
    LS_HD void finish() {
        if (op & 3) slot[(op >> 2) * kSlots] = cur;
    }
};

}  // namespace lsd
}  // namespace b2c
