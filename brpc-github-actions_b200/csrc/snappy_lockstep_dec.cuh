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

// For a copy whose offset o is below 8: byte_perm selectors that expand the first o bytes of an 8-byte window into the
// pattern s[i] = b[i % o] (low half in bits 0-15, high half in bits 16-31), and the smallest multiple of o that is
// >= 8 (bits 0-3 of the second table).  Tables, not arithmetic: i % o by a run-time o costs a division each.
LS_HD uint32_t pattern_sel(uint32_t o) {
    const uint32_t t[8] = {0, 0x00000000u, 0x10101010u, 0x10210210u, 0x32103210u, 0x21043210u, 0x10543210u, 0x06543210u};
    return t[o & 7];
}
LS_HD uint32_t pattern_step(uint32_t o) { return (uint32_t)((0xECA89880ull >> (4 * (o & 7))) & 15); }  // 8 8 9 8 10 12 14
LS_HD uint32_t fsl(uint32_t lo, uint32_t hi, uint32_t sh) {  // (hi << sh) | (lo >> (32 - sh)), sh = 0, 8, 16, 24
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, sh);
#else
    return sh ? (hi << sh) | (lo >> (32 - sh)) : hi;
#endif
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
    // The stream is consumed strictly in order (tag, then a literal's bytes), so the lane keeps a 32-byte window of
    // it in registers and asks for the following 16 bytes one refill ahead: the global-memory latency of a lane's
    // private, uncoalesced stream is off the tag-to-tag chain.
    const uint32_t* win;  // aligned word that holds the stream's first byte
    uint32_t lead;        // in - win, in bytes (0..3)
    uint32_t wb;          // word index (from win) of w[0]; lead + ip - 4 * wb is in [0, 16) between trips
    uint32_t w[8], nx[4];
    uint32_t rem;       // bytes of the element in progress still to move
    uint32_t src;       // a copy's source: an output position
    uint32_t is_lit;
    uint32_t offset;
    uint32_t cur;       // the bytes of the word that holds `op` which lie below op
    uint32_t live;
    int32_t status;

    LS_HD uint32_t ldw(uint32_t k) const { return 4 * k < lead + n ? win[k] : 0u; }  // words past the stream read as 0
    // 12 stream bytes from ip, out of the register window
    LS_HD void peek12(uint32_t& g0, uint32_t& g1, uint32_t& g2) const {
        const uint32_t r = lead + ip - 4 * wb, s = r >> 2, sh = (r & 3) * 8;
        const uint32_t x0 = s == 0 ? w[0] : s == 1 ? w[1] : s == 2 ? w[2] : w[3];
        const uint32_t x1 = s == 0 ? w[1] : s == 1 ? w[2] : s == 2 ? w[3] : w[4];
        const uint32_t x2 = s == 0 ? w[2] : s == 1 ? w[3] : s == 2 ? w[4] : w[5];
        const uint32_t x3 = s == 0 ? w[3] : s == 1 ? w[4] : s == 2 ? w[5] : w[6];
        g0 = ls::fsr(x0, x1, sh);
        g1 = ls::fsr(x1, x2, sh);
        g2 = ls::fsr(x2, x3, sh);
    }
    // after ip moved: slide the window by 16 bytes when ip has left its first 16, and ask for the 16 after next
    LS_HD void slide() {
        if (lead + ip - 4 * wb >= 16) {
            w[0] = w[4], w[1] = w[5], w[2] = w[6], w[3] = w[7];
            w[4] = nx[0], w[5] = nx[1], w[6] = nx[2], w[7] = nx[3];
            wb += 4;
            nx[0] = ldw(wb + 8), nx[1] = ldw(wb + 9), nx[2] = ldw(wb + 10), nx[3] = ldw(wb + 11);
        }
    }

    LS_HD void init(const uint8_t* in_, uint32_t n_, uint32_t* slot_, uint32_t cap) {
        in = in_;
        n = n_;
        slot = slot_;
        ip = op = rem = src = is_lit = offset = cur = expected = 0;
        live = 1;
        status = kOk;
        const uintptr_t a0 = reinterpret_cast<uintptr_t>(in);
        win = reinterpret_cast<const uint32_t*>(a0 & ~uintptr_t(3));
        lead = (uint32_t)(a0 & 3);
        wb = 0;
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
            return;
        }
        wb = (lead + ip) >> 2;
        for (uint32_t i = 0; i < 8; i++) w[i] = ldw(wb + i);
        for (uint32_t i = 0; i < 4; i++) nx[i] = ldw(wb + 8 + i);
    }

    // One step.  Straight-line selects on the common path (a short literal, a copy with a 1- or 2-byte offset of at
    // least 8); the rare kinds — a literal with length bytes, a 4-byte offset, a copy that overlaps itself within 8
    // bytes — take a branch of their own.
    LS_HD void trip() {
        const bool need_tag = rem == 0;
        const bool eof = need_tag && ip == n;  // clean end of input: snappy.cc:867
        uint32_t g0, g1, g2;
        peek12(g0, g1, g2);
        // ---- the tag at ip (computed by every lane, applied by those that need one) ----
        const uint32_t c = g0 & 0xff, kind = c & 3;
        uint32_t len = kind == 1 ? ((c >> 2) & 7) + 4 : (c >> 2) + 1;
        uint32_t used = kind + 1;
        uint32_t off = kind == 1 ? ((c >> 5) << 8) | ((g0 >> 8) & 0xff) : (g0 >> 8) & 0xffffu;
        uint32_t lit_lo = need_tag ? ls::fsr(g0, g1, 8) : g0, lit_hi = need_tag ? ls::fsr(g1, g2, 8) : g1, avail = 8;
        if (need_tag && (kind == 3 || (kind == 0 && c >= 0xf0u))) {  // rare: length bytes / 4-byte offset
            if (kind == 0) {
                const uint32_t extra = len - 60;
                uint32_t v = (g0 >> 8) | (g1 << 24);
                if (extra < 4) v &= (1u << (8 * extra)) - 1u;
                len = v + 1u;  // may wrap to 0 like the reference
                used = 1 + extra;
                const uint32_t sh = (used & 3) * 8;
                const uint32_t a = used < 4 ? g0 : g1, b = used < 4 ? g1 : g2, d = used < 4 ? g2 : 0u;
                lit_lo = ls::fsr(a, b, sh);
                lit_hi = ls::fsr(b, d, sh);
                avail = 12 - used;  // 7 when the tag had 4 length bytes
            } else {
                off = (g0 >> 8) | (g1 << 24);
                used = 5;
            }
        }
        const uint32_t left = n - ip;
        const bool lit = kind == 0;
        // tag cut short | premature end of input | offset 0 or beyond what was produced | more output than announced
        const bool bad = need_tag && !eof &&
                         (left < used || (lit ? left - used < len : off - 1u >= op) || expected - op < len);
        const bool stop = eof || bad;
        status = bad || (eof && op != expected) ? kCorrupt : status;
        live = stop ? 0u : 1u;
        const bool take = need_tag && !stop;  // a stopped lane changes nothing below: k == 0, no store, no slide
        rem = take ? len : rem;
        is_lit = take ? (lit ? 1u : 0u) : is_lit;
        offset = take ? off : offset;
        src = take ? (lit ? 0u : op - off) : src;  // always a valid output position
        ip += take ? used : 0u;
        // ---- a copy's next bytes come from the lane's own output (a literal reads, and ignores, a valid word too) ----
        const uint32_t rs = ls::umin(src, kOutMax - 1);  // a finished pattern copy may leave src out of range: read something valid
        const uint32_t* w = slot + (rs >> 2) * kSlots;
        const uint32_t sh2 = (rs & 3) * 8;
        const uint32_t w0 = w[0], w1 = w[kSlots], w2 = w[2 * kSlots];
        const uint32_t dw = (op >> 2) - (rs >> 2);  // the word that holds op is in `cur`, not in the slot yet
        const uint32_t x0 = dw == 0 ? cur : w0, x1 = dw == 1 ? cur : w1, x2 = dw == 2 ? cur : w2;
        uint32_t lo = ls::fsr(x0, x1, sh2), hi = ls::fsr(x1, x2, sh2);
        if (!stop && !is_lit && offset < 8) {  // rare: the 8 bytes are the pattern of the `offset` bytes at src (IncrementalCopy)
            const uint32_t o = offset;
            const uint32_t sel = pattern_sel(o);
            const uint32_t plo = byte_perm(lo, hi, sel & 0xffffu), phi = byte_perm(lo, hi, sel >> 16);
            lo = plo;
            hi = phi;
            // from here on the run continues `o2` bytes back, the smallest multiple of o that is >= 8: those bytes
            // are written by now and no longer overlap the 8 bytes a trip moves
            const uint32_t o2 = pattern_step(o);
            offset = o2;
            src = op - o2;  // + k below: the next chunk starts at (op + k) - o2
        }
        lo = is_lit ? lit_lo : lo;
        hi = is_lit ? lit_hi : hi;
        // ---- append k bytes at op ----
        const uint32_t k = stop ? 0u : ls::umin(ls::umin(rem, 8u), avail);
        const uint32_t sh = (op & 3) * 8;
        uint32_t* d = slot + (op >> 2) * kSlots;
        // bytes 0..11 of the three words from the word of op: cur below op, then lo, hi
        const uint32_t d0 = cur | (lo << sh);
        const uint32_t d1s = fsl(lo, hi, sh);
        const uint32_t d2 = fsl(hi, 0u, sh);
        const uint32_t end = (op & 3) + k;  // bytes of (d0, d1, d2) that are final: 0..11
        if (end >= 4) d[0] = d0;
        if (end >= 8) d[kSlots] = d1s;
        // the new partial word: the bytes of the word that holds op + k which lie below op + k
        const uint32_t part = end & 3, wsel = end >> 2;
        const uint32_t nw = wsel == 0 ? d0 : wsel == 1 ? d1s : d2;
        cur = nw & ~(0xffffffffu << (8 * part));  // part == 0: nothing
        op += k;
        src += is_lit ? 0u : k;  // (a literal leaves src alone: it must stay a valid output position)
        rem -= k;
        ip += is_lit ? k : 0u;   // a literal's bytes are consumed from the stream as they are moved
        slide();
    }

    // the last partial word goes to the slot; returns the status
    LS_HD void finish() {
        if (op & 3) slot[(op >> 2) * kSlots] = cur;
    }
};

}  // namespace lsd
}  // namespace b2c
