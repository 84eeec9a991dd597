// snappy_lockstep.cuh — CompressFragment (reference snappy.cc:329-453) for bodies of at most 1 KiB as a
// per-lane STATE MACHINE, so that the 32 lanes of a warp compress 32 different bodies in lock-step.
//
// One lane per body was already the shape of snappy_small.cu, but there the lanes of a warp ran the
// reference's loops as loops: a lane extending a match, pushing a literal or encoding a long copy kept the
// other lanes of its warp waiting, so warps carried 4 bodies and issued every instruction for ~1.4 of them.
// Here every lane advances by ONE step of the same straight-line code per trip, whatever it is doing:
//   PROBE  the table step of the probe loop / of the copy loop's "another match right here?" test
//          (snappy.cc:377-397, 430-439): hash, swap the table slot, compare 8 bytes with the candidate;
//   EXT    the next 8 bytes of FindMatchLength (snappy-internal.h:86-120) of a match that is still running;
//   LIT    the next 4 bytes of the literal in front of the match (EmitLiteral, snappy.cc:156-196);
//   COPY   one tag of EmitCopy (snappy.cc:198-232); the last tag also advances ip, inserts ip-1 and
//          returns to PROBE with skip state 31 (see below) — in the common case (a match shorter than
//          8 bytes with no literal in front) all of this happens in the PROBE trip that found the match.
// A warp-instruction therefore serves 32 bodies, and the elements are the reference's own in the reference's
// own order: byte-exact.
//
// Shared memory is laid out so that lock-step accesses never collide: word w of lane l's body lives at word
// index w * 32 + l (bank l whatever w is), and so do the words of its u16 hash table.
//
// The lane code is plain C++ on pointers, so tests/cpp/lockstep_model.cpp compiles the same source for the
// host and checks it against the oracle without a GPU.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define LS_HD __host__ __device__ __forceinline__
#else
#define LS_HD inline
#endif

namespace b2c {
namespace ls {

constexpr uint32_t kLanes = 32;
constexpr uint32_t kBodyMax = 1024;
constexpr uint32_t kBodyWords = kBodyMax / 4 + 4;  // 4-word reads may start at the word of the last byte
constexpr uint32_t kTabWords = 512;                // 1024 u16 entries (GetHashTable, snappy.cc:247-271)
constexpr uint32_t kWarpWords = (kTabWords + kBodyWords) * kLanes;

LS_HD uint32_t fsr(uint32_t lo, uint32_t hi, uint32_t sh) {  // bytes of hi:lo from bit sh (0, 8, 16, 24)
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
LS_HD uint32_t fsl_hi(uint32_t v, uint32_t sh) {  // v >> (32 - sh), 0 when sh == 0
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(v, 0u, sh);
#else
    return sh ? v >> (32 - sh) : 0u;
#endif
}
LS_HD uint32_t first_diff_byte(uint32_t x) {  // x != 0: index of its lowest non-zero byte
#if defined(__CUDA_ARCH__)
    return (uint32_t)(__ffs((int)x) - 1) >> 3;
#else
    return (uint32_t)__builtin_ctz(x) >> 3;
#endif
}
LS_HD uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }
LS_HD uint32_t umax(uint32_t a, uint32_t b) { return a > b ? a : b; }

enum : uint32_t { kProbe = 0, kExt = 1, kLit = 2, kCopy = 3, kStop = 4 };

struct Lane {
    uint32_t* tab;         // word k of this lane's table at tab[k * 32]; entry h = half (h & 1) of word h >> 1
    const uint32_t* body;  // word w of this lane's body at body[w * 32]
    uint32_t* dst;         // output slot, 4-byte aligned, MaxCompressedLength(n) bytes
    uint32_t n, ip_limit, shift;
    uint32_t mode, ip, s, next_emit, cand, m, lit_pos, lit_rem;
    uint32_t b0, b1;       // PROBE: bytes ip..ip+7; EXT: bytes ip+m..ip+m+7; LIT: bytes lit_pos..lit_pos+3
    uint32_t acc, nacc, wp;  // output funnel: nacc (< 4) pending bytes in acc, wp words stored

    LS_HD uint16_t* entry(uint32_t h) const { return reinterpret_cast<uint16_t*>(tab + (h >> 1) * kLanes) + (h & 1); }
    LS_HD uint32_t hash(uint32_t bytes) const { return (bytes * 0x1e35a7bdu) >> shift; }  // snappy.cc:47-50

    // append the k (0..4) low bytes of v (no bits above 8k) to the output
    LS_HD void push(uint32_t v, uint32_t k) {
        const uint32_t sh = 8 * nacc;
        const uint32_t lo = acc | (v << sh);
        const uint32_t hi = fsl_hi(v, sh);
        nacc += k;
        if (nacc >= 4) {
            dst[wp++] = lo;
            acc = hi;
            nacc -= 4;
        } else {
            acc = lo;
        }
    }
    static LS_HD void literal_header(uint32_t len, uint32_t& v, uint32_t& k) {  // EmitLiteral, snappy.cc:156-196
        const uint32_t n1 = len - 1;
        if (n1 < 60) {
            v = n1 << 2;
            k = 1;
        } else if (n1 < 256) {
            v = (60u << 2) | (n1 << 8);
            k = 2;
        } else {
            v = (61u << 2) | (n1 << 8);
            k = 3;
        }
    }

    // 12 bytes from position p: p0 = bytes p..p+3, p1 = p+4..p+7, p2 = p+8..p+11
    LS_HD void load12(uint32_t p, uint32_t& p0, uint32_t& p1, uint32_t& p2) const {
        const uint32_t* w = body + (p >> 2) * kLanes;
        const uint32_t w0 = w[0], w1 = w[kLanes], w2 = w[2 * kLanes], w3 = w[3 * kLanes], sh = (p & 3) * 8;
        p0 = fsr(w0, w1, sh);
        p1 = fsr(w1, w2, sh);
        p2 = fsr(w2, w3, sh);
    }
    LS_HD void load8(uint32_t p, uint32_t& p0, uint32_t& p1) const {
        const uint32_t* w = body + (p >> 2) * kLanes;
        const uint32_t w0 = w[0], w1 = w[kLanes], w2 = w[2 * kLanes], sh = (p & 3) * 8;
        p0 = fsr(w0, w1, sh);
        p1 = fsr(w1, w2, sh);
    }

    LS_HD void init(uint32_t* tab_, const uint32_t* body_, uint32_t* dst_, uint32_t n_) {
        tab = tab_;
        body = body_;
        dst = dst_;
        n = n_;
        uint32_t lg = 8;  // GetHashTable sizing, snappy.cc:253-256: 256 entries doubled while < n, at most 16384
        while (lg < 14 && (1u << lg) < n) lg++;
        shift = 32 - lg;
        next_emit = 0;
        cand = m = lit_pos = lit_rem = 0;
        acc = nacc = wp = 0;
        ip = 1;
        s = 32;
        b0 = b1 = 0;
        ip_limit = 0;
        mode = kStop;
        if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
            ip_limit = n - 15;
            mode = kProbe;
            load8(1, b0, b1);
        }
    }

    // One step.  `s` is the reference's `skip` (snappy.cc:377-385) with one extra state: s == 31 marks the table
    // step the copy loop makes right after a copy (snappy.cc:430-439) — it advances one byte on a miss and is
    // followed by the first probe of a fresh skip = 32 sequence, which is what max(1, s >> 5), s++ yields.
    LS_HD void trip() {
        const uint32_t mode0 = mode;
        uint32_t pv = 0, pk = 0;  // first push of this trip: a literal's header or 4 of its bytes
        uint32_t cv = 0, ck = 0;  // second push: one copy tag
        bool after_copy = false;
        if (mode0 <= kExt) {
            uint32_t pa = cand + m, nxt = 0;
            bool go = true;
            if (mode0 == kProbe) {
                nxt = ip + umax(1u, s >> 5);
                s++;
                if (nxt > ip_limit) {  // snappy.cc:386-388
                    mode = kStop;
                    go = false;
                } else {
                    uint16_t* e = entry(hash(b0));
                    pa = *e;
                    *e = (uint16_t)ip;
                    cand = pa;
                }
            }
            if (go) {
                uint32_t a0, a1;
                load8(pa, a0, a1);
                const uint32_t x0 = a0 ^ b0, x1 = a1 ^ b1;
                uint32_t nb = x0 ? first_diff_byte(x0) : (x1 ? 4u + first_diff_byte(x1) : 8u);
                if (mode0 == kProbe) {
                    if (nb >= 4) {  // a match (ip + 15 < n: these 8 bytes are inside the body)
                        m = nb;
                        if (ip > next_emit) {  // snappy.cc:403
                            lit_pos = next_emit;
                            lit_rem = ip - next_emit;
                            literal_header(lit_rem, pv, pk);
                        }
                        mode = nb == 8 ? kExt : (lit_rem ? kLit : kCopy);
                    } else {
                        ip = nxt;
                    }
                } else {  // FindMatchLength goes on, bounded by the end of the body
                    const uint32_t rem = n - (ip + m);
                    nb = umin(nb, rem);
                    m += nb;
                    if (!(nb == 8 && rem > 8)) mode = lit_rem ? kLit : kCopy;
                }
            }
        } else if (mode0 == kLit) {
            const uint32_t k = umin(lit_rem, 4u);
            pv = k == 4 ? b0 : b0 & ((1u << (8 * k)) - 1u);
            pk = k;
            lit_pos += k;
            lit_rem -= k;
            if (!lit_rem) mode = kCopy;
        }
        if (mode == kCopy) {  // EmitCopy, snappy.cc:198-232, one tag per trip
            const uint32_t offset = ip - cand;
            ck = 3;
            if (m >= 68) {
                cv = 0xfeu | (offset << 8);
                ip += 64;
                cand += 64;
                m -= 64;
            } else if (m > 64) {
                cv = (2u + (59u << 2)) | (offset << 8);
                ip += 60;
                cand += 60;
                m -= 60;
            } else {
                if (m < 12 && offset < 2048) {
                    cv = (1u + ((m - 4) << 2) + ((offset >> 8) << 5)) | ((offset & 0xff) << 8);
                    ck = 2;
                } else {
                    cv = (2u + ((m - 1) << 2)) | (offset << 8);
                }
                ip += m;
                next_emit = ip;
                if (ip >= ip_limit) {  // snappy.cc:427-429
                    mode = kStop;
                } else {
                    mode = kProbe;
                    s = 31;
                    after_copy = true;
                }
            }
        }
        push(pv, pk);
        push(cv, ck);
        // the bytes the next trip starts from
        if (mode <= kLit) {
            const uint32_t lp = after_copy ? ip - 1 : mode == kProbe ? ip : mode == kExt ? ip + m : lit_pos;
            uint32_t p0, p1, p2;
            load12(lp, p0, p1, p2);
            if (after_copy) {  // snappy.cc:430-433: insert ip - 1; the same words hold bytes ip..ip+7
                *entry(hash(p0)) = (uint16_t)(ip - 1);
                b0 = fsr(p0, p1, 8);
                b1 = fsr(p1, p2, 8);
            } else {
                b0 = p0;
                b1 = p1;
            }
        }
    }

    // emit_remainder (snappy.cc:446-450): a header, then 4 bytes per step while lit_rem != 0
    LS_HD void remainder_begin() {
        lit_pos = next_emit;
        lit_rem = n - next_emit;
        if (lit_rem) {
            uint32_t v, k;
            literal_header(lit_rem, v, k);
            push(v, k);
        }
    }
    LS_HD void remainder_step() {
        const uint32_t k = umin(lit_rem, 4u);
        if (k) {
            const uint32_t* w = body + (lit_pos >> 2) * kLanes;
            const uint32_t v = fsr(w[0], w[kLanes], (lit_pos & 3) * 8);
            push(k == 4 ? v : v & ((1u << (8 * k)) - 1u), k);
            lit_pos += k;
            lit_rem -= k;
        }
    }
    LS_HD uint32_t finish() {  // flush the funnel; the slot has MaxCompressedLength's 32 spare bytes
        if (nacc) dst[wp] = acc;
        return wp * 4 + nacc;
    }
};

}  // namespace ls
}  // namespace b2c
