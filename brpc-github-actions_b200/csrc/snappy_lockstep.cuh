// snappy_lockstep.cuh — the match-finding half of CompressFragment (reference snappy.cc:329-453) for bodies of
// at most 1 KiB as a per-lane STATE MACHINE with BRANCH-FREE steps, so that the lanes of a warp that carry
// different bodies stay in lock-step: every lane executes the same straight-line code per trip, whatever it is
// doing.
//
// snappy_small.cu runs the reference's loops as loops, one lane per body: a lane that finds a match takes a
// ~60-instruction path while a lane that misses takes an 8-instruction one, and a warp pays for every path any
// of its lanes takes, one after the other.  Here a trip is ONE step for every lane —
//   PROBE  the table step of the probe loop, or of the copy loop's "another match right here?" test
//          (snappy.cc:377-397, 430-439): hash, swap the table slot, compare 8 bytes with the candidate;
//   EXT    the next 8 bytes of FindMatchLength (snappy-internal.h:86-120) of a match that is still running
// — written with selects and predicated stores only.  A finished match is RECORDED as one packed word
// (pack_element) and the warp encodes the records afterwards (snappy_emit.cuh), exactly as the two-phase
// kernel of snappy_small.cu does.
// The elements are the reference's own in the reference's own order: byte-exact.
//
// Shared memory is word-interleaved over 32 slots (word w of slot l at index w * 32 + l), so lanes that work
// on different slots never collide on a bank, whatever their positions are.
//
// The lane code is plain C++ on pointers: tests/cpp/lockstep_model.cpp compiles the same source for the host
// and tests/test_lockstep_model.py checks it against the oracle without a GPU.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define LS_HD __host__ __device__ __forceinline__
#else
#define LS_HD inline
#endif

namespace b2c {
namespace ls {

constexpr uint32_t kSlots = 32;                    // interleave factor of the shared-memory layout
constexpr uint32_t kBodyMax = 1024;
constexpr uint32_t kBodyWords = kBodyMax / 4 + 4;  // 4-word reads may start at the word of the last byte
constexpr uint32_t kTabWords = 512;                // 1024 u16 entries (GetHashTable, snappy.cc:247-271)
constexpr uint32_t kGroupWords = (kTabWords + kBodyWords) * kSlots;  // one group of 32 slots: 98,816 bytes
constexpr uint32_t kMaxElems = kBodyMax / 4;       // copies cover >= 4 input bytes each and do not overlap

LS_HD uint32_t fsr(uint32_t lo, uint32_t hi, uint32_t sh) {  // bytes of hi:lo from bit sh (0, 8, 16, 24)
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
// index of the lowest non-zero byte of x; >= 4 (a large number) when x == 0
LS_HD uint32_t first_diff_byte(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)(__ffs((int)x) - 1) >> 3;  // ffs(0) - 1 = 0xffffffff
#else
    return x ? (uint32_t)__builtin_ctz(x) >> 3 : 0x1fffffffu;
#endif
}
LS_HD uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }
LS_HD uint32_t umax(uint32_t a, uint32_t b) { return a > b ? a : b; }

// One recorded copy: position p (< 1024), distance back (1..1023), length (4..1024 -> stored minus 4).
LS_HD uint32_t pack_element(uint32_t p, uint32_t offset, uint32_t len) { return p | (offset << 10) | ((len - 4) << 20); }
LS_HD void unpack_element(uint32_t e, uint32_t& p, uint32_t& offset, uint32_t& len) {
    p = e & 1023u;
    offset = (e >> 10) & 1023u;
    len = (e >> 20) + 4;
}

struct FindLane {
    uint32_t* tab;         // word k of this slot's table at tab[k * 32]; entry h = half (h & 1) of word h >> 1
    const uint32_t* body;  // word w of this slot's body at body[w * 32]
    uint32_t* list;        // recorded copies, in order
    uint32_t n, ip_limit, shift;
    uint32_t live;         // 0 once the chain has reached emit_remainder (snappy.cc:446)
    uint32_t ext;          // 1 while FindMatchLength of the match at ip is still running
    uint32_t ip, s, cand, m, ne;
    uint32_t b0, b1;       // PROBE: bytes ip..ip+7; EXT: bytes ip+m..ip+m+7

    LS_HD uint16_t* entry(uint32_t h) const { return reinterpret_cast<uint16_t*>(tab + (h >> 1) * kSlots) + (h & 1); }
    LS_HD uint32_t hash(uint32_t bytes) const { return (bytes * 0x1e35a7bdu) >> shift; }  // snappy.cc:47-50

    LS_HD void init(uint32_t* tab_, const uint32_t* body_, uint32_t* list_, uint32_t n_) {
        tab = tab_;
        body = body_;
        list = list_;
        n = n_;
        uint32_t lg = 8;  // GetHashTable sizing, snappy.cc:253-256: 256 entries doubled while < n
        while (lg < 14 && (1u << lg) < n) lg++;
        shift = 32 - lg;
        ext = cand = m = ne = 0;
        ip = 1;
        s = 32;
        b0 = b1 = 0;
        ip_limit = 0;
        live = 0;
        if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
            ip_limit = n - 15;
            live = 1;
            const uint32_t w0 = body[0], w1 = body[kSlots], w2 = body[2 * kSlots];
            b0 = fsr(w0, w1, 8);
            b1 = fsr(w1, w2, 8);
        }
    }

    // One step of a live lane.  `s` is the reference's `skip` (snappy.cc:377-385) with one extra state: s == 31
    // marks the table step the copy loop makes right after a copy (snappy.cc:430-439) — it advances one byte on a
    // miss and is followed by the first probe of a fresh skip = 32 sequence, which is what max(1, s >> 5), s++
    // yields.  Everything below is selects; the only conditional effects are three predicated stores.
    LS_HD void trip() {
        const bool probe = ext == 0;
        const uint32_t nxt = ip + umax(1u, s >> 5);
        const bool stop = probe && nxt > ip_limit;  // snappy.cc:386-388
        // the table step (its load is harmless for a lane that is extending a match)
        uint16_t* e = entry(hash(b0));
        const uint32_t old = *e;
        if (probe && !stop) *e = (uint16_t)ip;
        const uint32_t c = probe ? old : cand;
        // 8 bytes of the candidate side
        const uint32_t pa = c + m;  // m == 0 while probing
        const uint32_t* wa = body + (pa >> 2) * kSlots;
        const uint32_t a_0 = wa[0], a_1 = wa[kSlots], a_2 = wa[2 * kSlots], sa = (pa & 3) * 8;
        const uint32_t x0 = fsr(a_0, a_1, sa) ^ b0, x1 = fsr(a_1, a_2, sa) ^ b1;
        uint32_t nb = umin((x0 ? 0u : 4u) + first_diff_byte(x0 ? x0 : x1), 8u);  // equal leading bytes, 0..8
        const uint32_t rem = n - (ip + m);  // probing: > 15, the 8 bytes are inside the body
        nb = umin(nb, rem);
        const bool hit = !probe || nb >= 4;
        const uint32_t m2 = m + nb;
        const bool more = hit && nb == 8 && rem > 8;  // FindMatchLength goes on
        const bool done = hit && !more && !stop;      // the copy is complete
        if (done) list[ne] = pack_element(ip, ip - c, m2);
        ne += done ? 1u : 0u;
        const uint32_t ip2 = done ? ip + m2 : (hit || stop ? ip : nxt);
        const bool fin = stop || (done && ip2 >= ip_limit);  // snappy.cc:427-429
        // the bytes the next trip starts from: after a copy from ip2 - 1 (its hash is inserted, snappy.cc:430-433)
        const uint32_t lp = done ? ip2 - 1 : (more ? ip + m2 : ip2);
        const uint32_t* wl = body + (lp >> 2) * kSlots;
        const uint32_t l0 = wl[0], l1 = wl[kSlots], l2 = wl[2 * kSlots], l3 = wl[3 * kSlots], sl = (lp & 3) * 8;
        const uint32_t p0 = fsr(l0, l1, sl), p1 = fsr(l1, l2, sl), p2 = fsr(l2, l3, sl);
        if (done && !fin) *entry(hash(p0)) = (uint16_t)lp;
        b0 = done ? fsr(p0, p1, 8) : p0;
        b1 = done ? fsr(p1, p2, 8) : p1;
        s = done ? 31u : s + (probe ? 1u : 0u);
        ip = ip2;
        cand = c;
        m = more ? m2 : 0u;
        ext = more ? 1u : 0u;
        live = fin ? 0u : 1u;
    }
};

}  // namespace ls
}  // namespace b2c
