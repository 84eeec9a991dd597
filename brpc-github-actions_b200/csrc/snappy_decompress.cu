// snappy_decompress.cu — batch snappy decompression for sm_100a.
//
// Replaces butil::snappy::Uncompress(Source*,Sink*) (reference
// src/butil/third_party/snappy/snappy.cc:1526-1552) and its hot loop
// SnappyDecompressor::DecompressAllTags (snappy.cc:716-787) for many messages at
// once, with the reference's accept/reject behaviour: lenient preamble
// (snappy.cc:692-711), tag cut short / premature end (snappy.cc:761,798-801,825),
// offset == 0 or offset > produced (snappy.cc:1200-1202), output overrun
// (snappy.cc:1165-1167,1211-1213) and the final eof && produced == expected
// (snappy.cc:867).
//
// One warp per message.  The warp keeps a 128-byte window of the compressed
// stream in registers (one aligned word per lane) and pulls each tag and its
// trailer out of it with shuffles, so a global load is issued once per ~50
// tags instead of once per tag.  Literal and copy payloads are moved by all 32
// lanes; a copy whose length exceeds its offset replicates the pattern exactly
// like IncrementalCopy (snappy.cc:98-103).  Small messages are decoded into a
// per-warp shared-memory buffer (back-references resolve at shared-memory
// latency) and flushed coalesced; large ones decode straight into HBM.
#include <algorithm>

#include "b2c_internal.cuh"

namespace b2c {

static constexpr uint32_t kFull = 0xffffffffu;

// lenient varint (snappy.cc:692-711).  Returns bytes consumed, 0 on failure.
__device__ __forceinline__ uint32_t read_preamble(const uint8_t* in, uint32_t n, uint32_t* ulen) {
    uint32_t v = 0, shift = 0, i = 0;
    for (;;) {
        if (shift >= 32) return 0;
        if (i >= n) return 0;
        uint32_t c = in[i++];
        v |= (c & 0x7f) << shift;
        if (c < 128) break;
        shift += 7;
    }
    *ulen = v;
    return i;
}

struct Window {
    const uint8_t* base;  // 4-byte aligned address of lane 0's word
    uint32_t word;        // this lane's word
};

__device__ __forceinline__ void window_load(Window& w, const uint8_t* at, const uint8_t* end, uint32_t lane) {
    uintptr_t a = reinterpret_cast<uintptr_t>(at) & ~uintptr_t(3);
    w.base = reinterpret_cast<const uint8_t*>(a);
    const uint8_t* p = w.base + 4 * lane;
    w.word = p < end ? *reinterpret_cast<const uint32_t*>(p) : 0u;
}

// 8 bytes of the stream starting at `at` (at - base <= 116 so that word index + 2 <= 31)
__device__ __forceinline__ uint64_t window_get8(const Window& w, const uint8_t* at) {
    const uint32_t r = (uint32_t)(at - w.base);
    const uint32_t wi = r >> 2, sh = (r & 3) * 8;
    const uint32_t a = __shfl_sync(kFull, w.word, wi);
    const uint32_t b = __shfl_sync(kFull, w.word, wi + 1);
    const uint32_t c = __shfl_sync(kFull, w.word, wi + 2);
    const uint32_t lo = __funnelshift_r(a, b, sh);
    const uint32_t hi = __funnelshift_r(b, c, sh);
    return ((uint64_t)hi << 32) | lo;
}

// Decode one message.  `obuf` is where output bytes are addressed while decoding
// (shared or global memory).  Returns B2C_OK / B2C_CORRUPT.
__device__ int decode_message_warp(const uint8_t* __restrict__ in, const uint32_t n, uint32_t ip, uint8_t* obuf,
                                   const uint32_t expected, const uint32_t lane) {
    const uint8_t* const end = in + n;
    uint32_t produced = 0;
    Window win;
    window_load(win, in + ip, end, lane);
    for (;;) {
        if (ip == n) return produced == expected ? B2C_OK : B2C_CORRUPT;  // RefillTag: clean eof
        const uint8_t* at = in + ip;
        if ((uint32_t)(at - win.base) > 116) window_load(win, at, end, lane);
        const uint64_t bits = window_get8(win, at);
        const uint32_t c = (uint32_t)bits & 0xff;
        const uint32_t trailer = (uint32_t)(bits >> 8);
        const uint32_t kind = c & 3;
        // bytes after the tag byte (char_table bits 11-13, snappy.cc:516-549)
        const uint32_t extra = kind == 0 ? (c >= 0xf0 ? (c >> 2) - 59 : 0) : kind == 1 ? 1 : kind == 2 ? 2 : 4;
        if (n - ip < extra + 1) return B2C_CORRUPT;  // tag cut short
        ip += 1 + extra;
        const uint32_t tmask = extra == 4 ? 0xffffffffu : ((1u << (8 * extra)) - 1u);
        if (kind == 0) {
            uint32_t len = (c >> 2) + 1;
            if (len >= 61) len = (trailer & tmask) + 1u;  // may wrap to 0 like the reference
            if (n - ip < len) return B2C_CORRUPT;          // premature end of input
            if (expected - produced < len) return B2C_CORRUPT;
            warp_copy_bytes(obuf + produced, in + ip, len, lane);
            produced += len;
            ip += len;
        } else {
            uint32_t len, offset;
            if (kind == 1) {
                len = ((c >> 2) & 7) + 4;
                offset = ((c >> 5) << 8) | (trailer & 0xff);
            } else {
                len = (c >> 2) + 1;
                offset = trailer & tmask;
            }
            if (offset == 0 || offset > produced) return B2C_CORRUPT;
            if (expected - produced < len) return B2C_CORRUPT;
            uint8_t* dst = obuf + produced;
            const uint8_t* src = dst - offset;
            // len <= 64: two rounds at most; i % offset replicates short patterns
            if (lane < len) {
                uint32_t i0 = lane, i1 = lane + 32;
                uint8_t v0 = src[offset >= len ? i0 : i0 % offset];
                uint8_t v1 = 0;
                if (i1 < len) v1 = src[offset >= len ? i1 : i1 % offset];
                dst[i0] = v0;
                if (i1 < len) dst[i1] = v1;
            }
            produced += len;
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------
// Batched decoder: up to 32 tags per step.
//
// The serial part of DecompressAllTags (snappy.cc:731-784) is only "where does the next
// tag start"; everything else is independent per tag.  Per step the warp
//   1. keeps a 256-byte window of the stream in shared memory and lets every lane compute,
//      for 8 byte positions each, where the next tag would start IF a tag started there;
//   2. follows that jump table from the current position for up to 32 hops (a chain of
//      broadcast shared-memory loads, ~5 instructions per tag), handing tag t to lane t;
//   3. decodes the 32 tags in parallel, takes a warp prefix sum of their output lengths to
//      place them, and checks every reject condition of the reference at once;
//   4. moves the literals (no dependencies), then the copies in rounds: a copy runs as soon
//      as its source lies entirely before the output position of the first copy that has not
//      run yet (back-references inside the batch resolve in shared memory / L1).
// ---------------------------------------------------------------------------
constexpr uint32_t kWin = 256;
constexpr uint32_t kStop = 0xFFFFu;     // no further tag of this batch from here
constexpr uint32_t kScratchSlot = 3072;  // shared memory per warp for BatchScratch
struct BatchScratch {
    uint32_t win32[kWin / 4 + 4];  // stream bytes [wbase, wbase + 256) + 16 bytes of padding
    // jmp[k][p]: where the chain stands 2^k tags after a tag at p (window-relative), kStop once it meets a tag the
    // batch cannot take: one that is not fully inside the window, lies past the stream's end, or is a literal of more
    // than 1 KiB (it gets a batch of its own)
    uint16_t jmp[5][kWin];
};
static_assert(sizeof(BatchScratch) <= kScratchSlot, "BatchScratch must fit its slot");

__device__ __forceinline__ uint64_t warp_incl_scan_u64(uint64_t v, uint32_t lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t lo = __shfl_up_sync(kFull, (uint32_t)v, d);
        uint32_t hi = __shfl_up_sync(kFull, (uint32_t)(v >> 32), d);
        if (lane >= (uint32_t)d) v += ((uint64_t)hi << 32) | lo;
    }
    return v;
}

// Output policies of the batched decoder.
// LinearOut: the whole output is addressable at base[p] (a per-warp shared-memory buffer for
// small slots, or the final HBM slot).
struct LinearOut {
    static constexpr bool kMovesData = true;
    static constexpr bool kLinear = true;  // byte p lives at base[p]: copy loops may walk raw pointers
    uint8_t* base;
    __device__ __forceinline__ void on_tag(bool, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}
    __device__ __forceinline__ uint8_t ld(uint32_t p) const { return base[p]; }
    __device__ __forceinline__ void st(uint32_t p, uint8_t v) { base[p] = v; }
    // a long literal, by the whole warp: 16-byte stores (incompressible bodies are one literal per 64 KiB fragment)
    __device__ __forceinline__ void copy_in(uint32_t p, const uint8_t* src, uint32_t len, uint32_t lane) {
        warp_copy_bytes(base + p, src, len, lane);
    }
    __device__ __forceinline__ void begin_batch(uint32_t, uint32_t, uint32_t) {}
    __device__ __forceinline__ bool takes_direct(uint32_t) const { return false; }
    __device__ __forceinline__ void direct_literal(const uint8_t*, uint32_t, uint32_t, uint32_t) {}
    __device__ __forceinline__ void finish(uint32_t, uint32_t) {}
};

// RingOut: large outputs.  The most recent 64 KiB of output live in a shared-memory ring
// (byte p at ring[p & 0xffff]) so that back-references — which the reference compressor keeps
// inside one 64 KiB block (snappy.h:193-203) — resolve at shared-memory latency; older bytes
// are read back from HBM, where the ring is flushed in 16-byte stores as it fills.
struct RingOut {
    static constexpr bool kMovesData = true;
    static constexpr bool kLinear = false;
    __device__ __forceinline__ void on_tag(bool, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}
    uint8_t* ring;      // 65536 bytes of shared memory
    uint8_t* gout;      // final slot in HBM
    uint32_t flushed;   // every byte below this is in HBM
    uint32_t ring_lo;   // no byte below this is in the ring
    uint32_t bound;     // during a batch: bytes >= bound are read from the ring
    __device__ __forceinline__ uint8_t ld(uint32_t p) const { return p >= bound ? ring[p & 0xffffu] : gout[p]; }
    __device__ __forceinline__ void st(uint32_t p, uint8_t v) { ring[p & 0xffffu] = v; }
    __device__ __forceinline__ void copy_in(uint32_t p, const uint8_t* src, uint32_t len, uint32_t lane) {
        for (uint32_t i = lane; i < len; i += 32) ring[(p + i) & 0xffffu] = src[i];
    }
    __device__ void flush_to(uint32_t upto, uint32_t lane) {  // warp-cooperative
        uint32_t a = flushed;
        if (upto <= a) return;
        if ((reinterpret_cast<uintptr_t>(gout) & 15) == 0) {
            uint32_t head = min(upto - a, (16u - (a & 15u)) & 15u);
            if (lane < head) gout[a + lane] = ring[(a + lane) & 0xffffu];
            a += head;
            const uint32_t vecs = (upto - a) >> 4;
            for (uint32_t i = lane; i < vecs; i += 32) {
                const uint32_t q = a + 16 * i;
                *reinterpret_cast<uint4*>(gout + q) = *reinterpret_cast<const uint4*>(ring + (q & 0xffffu));
            }
            a += vecs << 4;
        }
        for (uint32_t q = a + lane; q < upto; q += 32) gout[q] = ring[q & 0xffffu];
        flushed = upto;
        __syncwarp();
    }
    // called before a batch that writes [produced, batch_end)
    __device__ __forceinline__ void begin_batch(uint32_t produced, uint32_t batch_end, uint32_t lane) {
        if (batch_end - flushed > 49152u) flush_to(produced, lane);  // keeps flushed >= batch_end - 65536
        const uint32_t lo64 = batch_end > 65536u ? batch_end - 65536u : 0u;
        bound = max(ring_lo, lo64);
    }
    // a literal too long for the ring goes straight from the stream to HBM
    __device__ __forceinline__ bool takes_direct(uint32_t len) const { return len > 8192u; }
    __device__ void direct_literal(const uint8_t* src, uint32_t produced, uint32_t len, uint32_t lane) {
        flush_to(produced, lane);
        warp_copy_bytes(gout + produced, src, len, lane);
        __syncwarp();
        flushed = produced + len;
        ring_lo = produced + len;
    }
    __device__ __forceinline__ void finish(uint32_t produced, uint32_t lane) { flush_to(produced, lane); }
};

// ScanOut: moves no data.  It runs the batched decoder's parse + validation only and records,
// for every 64 KiB output boundary, the stream offset of the tag that starts there — the
// fragment boundaries the raw snappy format does not mark (format_description.txt:18-25).
// A stream is "splittable" when no element straddles such a boundary and no copy reaches back
// across one (always true for the reference compressor, which never matches across its 64 KiB
// blocks, snappy.h:193-203); its blocks can then be decoded independently, one warp each.
struct ScanOut {
    static constexpr bool kMovesData = false;
    static constexpr bool kLinear = false;
    uint32_t* blk_ip;  // [nblk] stream offset where output block k starts
    uint32_t nblk;
    uint32_t bad;      // per-lane: something prevents splitting
    __device__ __forceinline__ void on_tag(bool act, uint32_t at, uint32_t out_pos, uint32_t len, uint32_t kind,
                                           uint32_t offset) {
        if (!act) return;
        const uint32_t k0 = out_pos >> 16;
        if (len && ((out_pos + len - 1) >> 16) != k0) bad = 1;
        if (blk_ip && (out_pos & 0xffffu) == 0 && k0 > 0 && k0 < nblk) blk_ip[k0] = at;
        if (kind != 0 && offset > (out_pos & 0xffffu)) bad = 1;
    }
    __device__ __forceinline__ uint8_t ld(uint32_t) const { return 0; }
    __device__ __forceinline__ void st(uint32_t, uint8_t) {}
    __device__ __forceinline__ void copy_in(uint32_t, const uint8_t*, uint32_t, uint32_t) {}
    __device__ __forceinline__ void begin_batch(uint32_t, uint32_t, uint32_t) {}
    __device__ __forceinline__ bool takes_direct(uint32_t) const { return false; }
    __device__ __forceinline__ void direct_literal(const uint8_t*, uint32_t, uint32_t, uint32_t) {}
    __device__ __forceinline__ void finish(uint32_t, uint32_t) {}
};

template <class Out>
__device__ int decode_message_batched(const uint8_t* __restrict__ in, const uint32_t n, uint32_t ip, Out& out,
                                      const uint32_t expected, const uint32_t lane, BatchScratch& sc) {
    const uint8_t* win = reinterpret_cast<const uint8_t*>(sc.win32);
    uint32_t produced = 0;
    uint32_t wbase = 0;
    bool have_window = false;
    for (;;) {
        if (ip == n) {  // clean eof, snappy.cc:867
            if (produced != expected) return B2C_CORRUPT;
            out.finish(produced, lane);
            return B2C_OK;
        }
        if (!have_window || ip - wbase + 5 > kWin) {
            // ---- 1. refill the window at ip and rebuild the jump table
            __syncwarp();
            wbase = ip;
            have_window = true;
#pragma unroll
            for (uint32_t r = 0; r < 2; r++) {
                const uint32_t wi = lane + 32 * r;
                const uint32_t at = wbase + 4 * wi;
                sc.win32[wi] = at < n ? ld32u(in + at) : 0u;
            }
            if (lane < 4) sc.win32[kWin / 4 + lane] = 0u;
            __syncwarp();
#pragma unroll
            for (uint32_t j = 0; j < 8; j++) {
                const uint32_t p = j * 32 + lane;
                const uint32_t c = win[p];
                const uint32_t kind = c & 3;
                uint64_t L;
                if (kind == 0) {
                    const uint32_t l = c >> 2;
                    if (l < 60) {
                        L = l + 2;
                    } else {
                        const uint32_t ex = l - 59;
                        uint32_t val = win[p + 1] | (win[p + 2] << 8) | (win[p + 3] << 16) | ((uint32_t)win[p + 4] << 24);
                        if (ex < 4) val &= (1u << (8 * ex)) - 1u;
                        L = (uint64_t)1 + ex + (uint32_t)(val + 1u);  // len wraps to 0 like the reference
                    }
                } else {
                    L = kind == 1 ? 2 : kind == 2 ? 3 : 5;
                }
                const uint64_t to = p + L;
                const bool takes = p + 5 <= kWin && wbase + p < n && L <= 1030u;
                sc.jmp[0][p] = (uint16_t)(takes ? to : kStop);
            }
            __syncwarp();
#pragma unroll 1
            for (uint32_t k = 1; k < 5; k++) {  // square the jump table: 2, 4, 8, 16 tags ahead
#pragma unroll
                for (uint32_t j = 0; j < 8; j++) {
                    const uint32_t p = j * 32 + lane;
                    const uint32_t q = sc.jmp[k - 1][p];
                    sc.jmp[k][p] = (uint16_t)(q < kWin ? sc.jmp[k - 1][q] : kStop);
                }
                __syncwarp();
            }
        }
        // ---- 2. follow the chain: tag t of this batch goes to lane t, which composes the powers of the jump table
        //         that make up t (at most 5 dependent shared-memory reads instead of a 32-step walk)
        uint32_t my_pos = ip - wbase;
#pragma unroll
        for (uint32_t k = 0; k < 5; k++)
            if (((lane >> k) & 1u) && my_pos < kWin) my_pos = sc.jmp[k][my_pos];
        // the batch takes tag 0 always (the window was refilled so that it fits; a long literal is taken alone) and tag t
        // while the chain has not met a tag it cannot take
        const bool take = lane == 0 || (my_pos < kWin && sc.jmp[0][my_pos] != kStop);
        const uint32_t taken = __ballot_sync(kFull, take);
        const uint32_t count = taken == kFull ? 32u : (uint32_t)__ffs(~taken) - 1;
        // ---- 3. decode in parallel
        const bool act = lane < count;
        uint32_t kind = 0, len = 0, offset = 0, lit_src = 0, adv = 0;
        bool fail = false;
        if (act) {
            const uint32_t p = my_pos, at = wbase + p;
            const uint32_t c = win[p];
            kind = c & 3;
            const uint32_t extra = kind == 0 ? (c >= 0xf0 ? (c >> 2) - 59 : 0) : kind == 1 ? 1 : kind == 2 ? 2 : 4;
            if (n - at < extra + 1) fail = true;  // tag cut short
            uint32_t trailer = win[p + 1] | (win[p + 2] << 8) | (win[p + 3] << 16) | ((uint32_t)win[p + 4] << 24);
            if (extra < 4) trailer &= (1u << (8 * extra)) - 1u;
            adv = 1 + extra;
            if (kind == 0) {
                len = (c >> 2) + 1;
                if (len >= 61) len = trailer + 1u;
                lit_src = at + 1 + extra;
                if (!fail && n - lit_src < len) fail = true;  // premature end of input
                if (!fail) adv += len;
            } else if (kind == 1) {
                len = ((c >> 2) & 7) + 4;
                offset = ((c >> 5) << 8) | (trailer & 0xff);
            } else {
                len = (c >> 2) + 1;
                offset = trailer;
            }
            if (len > expected) {  // cannot fit whatever came before it
                fail = true;
                len = 0;
            }
        }
        const uint64_t incl = warp_incl_scan_u64(act ? len : 0u, lane);
        const uint64_t pos64 = (uint64_t)produced + incl - len;
        if (act && !fail) {
            if (pos64 + len > expected) fail = true;                       // output overrun
            else if (kind != 0 && (offset == 0 || offset > pos64)) fail = true;  // snappy.cc:1200-1202
        }
        if (__any_sync(kFull, fail)) return B2C_CORRUPT;
        const uint32_t out_pos = (uint32_t)pos64;
        const uint32_t batch_end = __shfl_sync(kFull, (uint32_t)(pos64 + len), count - 1);
        // a literal too long for the output policy's buffer is moved on its own (it is the whole batch)
        const uint32_t kind0 = __shfl_sync(kFull, kind, 0), len0 = __shfl_sync(kFull, len, 0);
        if (Out::kMovesData && count == 1 && kind0 == 0 && out.takes_direct(len0)) {
            const uint32_t l0 = len0, s0 = __shfl_sync(kFull, lit_src, 0);
            out.direct_literal(in + s0, produced, l0, lane);
            produced += l0;
            ip = s0 + l0;
            continue;
        }
        out.on_tag(act, wbase + my_pos, out_pos, len, kind, offset);
        if (!Out::kMovesData) {  // validation / boundary scan only
            produced = batch_end;
            ip = __shfl_sync(kFull, wbase + my_pos + adv, count - 1);
            continue;
        }
        out.begin_batch(produced, batch_end, lane);
        // ---- 4a. literals: short ones per lane, long ones by the whole warp
        const bool is_lit = act && kind == 0;
        if (is_lit && len <= 32) {
            const uint8_t* src = in + lit_src;
            for (uint32_t i = 0; i < len; i++) out.st(out_pos + i, src[i]);
        }
        uint32_t longs = __ballot_sync(kFull, is_lit && len > 32);
        while (longs) {
            const uint32_t l = __ffs(longs) - 1;
            longs &= longs - 1;
            const uint32_t s0 = __shfl_sync(kFull, lit_src, l), d0 = __shfl_sync(kFull, out_pos, l),
                           ln = __shfl_sync(kFull, len, l);
            out.copy_in(d0, in + s0, ln, lane);
        }
        __syncwarp();
        // ---- 4b. copies
        const bool is_copy = act && kind != 0;
        uint32_t pending = __ballot_sync(kFull, is_copy);
        const uint32_t src_pos = out_pos - offset;                         // meaningful for copies
        const uint32_t src_end = offset >= len ? src_pos + len : out_pos;  // bytes read lie in [src_pos, src_end)
        // Which copies of this batch does a copy read from?  Outputs are contiguous in tag order, so the tags whose
        // output overlaps [src_pos, src_end) form a range [a, b] of lanes, found by two binary searches over the lanes'
        // output positions.  A copy may run as soon as no copy of its range is still to run: the rounds below are then
        // as many as the longest chain of copies reading copies (2-3 on text, where a source lies ~15 tags back),
        // not one per copy that happens to precede a source.
        uint32_t deps = 0;
        {
            const uint32_t out_end = out_pos + len;
            uint32_t na = 0, nb = 0;  // tags that end at or before src_pos; tags that begin at or before src_end - 1
#pragma unroll
            for (uint32_t stp = 16; stp; stp >>= 1) {
                const uint32_t ca = na + stp, cb = nb + stp;
                const uint32_t ea = __shfl_sync(kFull, out_end, (ca - 1) & 31), pb = __shfl_sync(kFull, out_pos, (cb - 1) & 31);
                if (ca <= count && ea <= src_pos) na = ca;
                if (cb <= count && pb < src_end) nb = cb;
            }
            if (is_copy && nb > na) deps = (nb >= 32 ? 0xffffffffu : (1u << nb) - 1u) & ~((1u << na) - 1u);
            deps &= pending & ((1u << lane) - 1u);
        }
        // The first copy that has not run yet goes alone, 32 bytes per instruction, when it is long (runs, repeated
        // records); otherwise every short copy none of whose sources is still to be written runs, one lane per tag.
        while (pending) {
            const uint32_t first = __ffs(pending) - 1;
            const uint32_t ln = __shfl_sync(kFull, len, first);
            const uint32_t frontier = __shfl_sync(kFull, out_pos, first);
            if (ln > 16) {
                const uint32_t s0 = __shfl_sync(kFull, src_pos, first), of = __shfl_sync(kFull, offset, first);
                if constexpr (Out::kLinear) {
                    // A run (zero fill, a repeated record) arrives as consecutive copies with one offset: together
                    // they say out[p] = out[p - of] for every p of the run, i.e. ONE copy of their total length.  It
                    // is made by doubling: what is already there is a whole number of periods, so it can be appended
                    // to itself 32 bytes per instruction (source and destination of one step never overlap).
                    const uint32_t same = __ballot_sync(kFull, is_copy && offset == of) >> first;
                    const uint32_t run = same == kFull ? 32u - first : (uint32_t)__ffs(~same) - 1u;  // >= 1
                    const uint32_t lastl = first + run - 1;
                    const uint32_t total = __shfl_sync(kFull, out_pos + len, lastl) - frontier;
                    if (run >= 2 && total >= 128) {
                        uint8_t* d0 = out.base + frontier;
                        const uint8_t* sp = d0 - of;
                        for (uint32_t done = 0; done < total;) {
                            const uint32_t nmove = min(total - done, of + done);
                            for (uint32_t i = lane; i < nmove; i += 32) d0[done + i] = sp[i];
                            __syncwarp();
                            done += nmove;
                        }
                        pending &= ~(((run == 32u ? 0u : (1u << run)) - 1u) << first);
                        continue;
                    }
                }
                pending &= pending - 1;
                const uint32_t i0 = lane, i1 = lane + 32;
                uint8_t v0 = 0, v1 = 0;
                if (of >= ln) {
                    if (i0 < ln) v0 = out.ld(s0 + i0);
                    if (i1 < ln) v1 = out.ld(s0 + i1);
                } else {  // pattern of `of` bytes repeated, as IncrementalCopy (snappy.cc:98-103)
                    if (i0 < ln) v0 = out.ld(s0 + i0 % of);
                    if (i1 < ln) v1 = out.ld(s0 + i1 % of);
                }
                if (i0 < ln) out.st(frontier + i0, v0);
                if (i1 < ln) out.st(frontier + i1, v1);
                __syncwarp();
                continue;
            }
            const bool ready = ((pending >> lane) & 1u) && len <= 16 && (deps & pending) == 0;
            if (ready) {
                if constexpr (Out::kLinear) {
                    // the same three cases as below on raw pointers: one 64-bit address per copy instead of one per byte
                    uint8_t* d = out.base + out_pos;
                    const uint8_t* sp = d - offset;
                    uint8_t* const de = d + len;
                    if (offset >= 4) {
                        for (; d + 4 <= de; d += 4, sp += 4) {
                            const uint8_t b0 = sp[0], b1 = sp[1], b2 = sp[2], b3 = sp[3];
                            d[0] = b0;
                            d[1] = b1;
                            d[2] = b2;
                            d[3] = b3;
                        }
                        const uint32_t r = (uint32_t)(de - d);  // 0-3 bytes left: loaded together, one round trip
                        if (r) {
                            const uint8_t b0 = sp[0], b1 = r > 1 ? sp[1] : (uint8_t)0, b2 = r > 2 ? sp[2] : (uint8_t)0;
                            d[0] = b0;
                            if (r > 1) d[1] = b1;
                            if (r > 2) d[2] = b2;
                        }
                    } else {  // offset 1-3: every byte repeats the one `offset` back, also across the pattern's end
                        for (; d < de; d++, sp++) *d = *sp;
                    }
                } else if (offset >= 4) {
                    // the 4 source bytes of a round precede its first destination byte (also when the copy overlaps
                    // itself: IncrementalCopy reads what it wrote `offset` bytes ago, snappy.cc:98-103), so they are
                    // loaded together: one memory round trip per 4 bytes instead of per byte
                    uint32_t i = 0;
                    for (; i + 4 <= len; i += 4) {
                        const uint8_t b0 = out.ld(src_pos + i), b1 = out.ld(src_pos + i + 1), b2 = out.ld(src_pos + i + 2),
                                      b3 = out.ld(src_pos + i + 3);
                        out.st(out_pos + i, b0);
                        out.st(out_pos + i + 1, b1);
                        out.st(out_pos + i + 2, b2);
                        out.st(out_pos + i + 3, b3);
                    }
                    for (; i < len; i++) out.st(out_pos + i, out.ld(src_pos + i));
                } else if (offset >= len) {
                    for (uint32_t i = 0; i < len; i++) out.st(out_pos + i, out.ld(src_pos + i));
                } else {  // pattern of `offset` bytes repeated, as IncrementalCopy (snappy.cc:98-103)
                    uint32_t j = 0;
                    for (uint32_t i = 0; i < len; i++) {
                        out.st(out_pos + i, out.ld(src_pos + j));
                        if (++j == offset) j = 0;
                    }
                }
            }
            pending &= ~__ballot_sync(kFull, ready);
            __syncwarp();
        }
        produced = batch_end;
        ip = __shfl_sync(kFull, wbase + my_pos + adv, count - 1);
    }
}

struct DecompressParams {
    const uint8_t* in;
    const uint64_t* in_off;
    const uint32_t* in_len;
    const uint64_t* out_off;
    const uint32_t* out_cap;
    const uint32_t* list;  // message ids of this class
    uint8_t* out;
    uint32_t* ulen;        // nullable
    int32_t* status;
    uint32_t n_list;
    uint32_t smem_per_warp;  // 0: decode straight into global memory
    uint32_t batched;        // 1: 32 tags per step (decode_message_batched), 0: one tag per step
    uint32_t ring;           // 1: large slots decode through a 64 KiB shared-memory ring (one warp per CTA)
    // block-parallel decoding of large streams (class 2 only; null otherwise)
    const uint32_t* out_frag_first;  // output layout: first 64 KiB block of message m
    const uint32_t* out_frag_msg;    // output layout: message of block f
    uint32_t* blk_ip;                // [output blocks] stream offset of each block's first tag
    uint32_t* mode;                  // [n] kModeWhole / kModeBlocks / kModeDone
    uint32_t* ulen_tmp;              // [n] uncompressed lengths for the block kernel
    uint32_t n_blocks;
    uint32_t* work_ctr;              // warps take the next stream / block from here (null: one each, by index)
};

enum : uint32_t { kModeWhole = 0, kModeBlocks = 1, kModeDone = 2 };

// Pass 1 for large slots: validate the stream and find its 64 KiB block boundaries.
__global__ void __launch_bounds__(256) snappy_scan_kernel(DecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * (blockDim.x >> 5) + wib;
    if (w >= P.n_list) return;
    const uint32_t m = P.list ? P.list[w] : w;
    const uint32_t cap = P.out_cap[m];
    if (cap <= kFragBytes) {  // single block: nothing to split
        if (lane == 0) P.mode[m] = kModeWhole;
        return;
    }
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    uint32_t expected = 0;
    const uint32_t k = read_preamble(in, n, &expected);
    int st = B2C_OK;
    uint32_t mode = kModeDone;
    if (k == 0) {
        st = B2C_CORRUPT;
    } else if (expected > cap) {
        st = B2C_TOO_SMALL;
    } else if (expected <= kFragBytes) {
        mode = kModeWhole;
    } else {
        const uint32_t f0 = P.out_frag_first[m];
        ScanOut so{P.blk_ip + f0, (expected + kFragBytes - 1) / kFragBytes, 0u};
        if (lane == 0) P.blk_ip[f0] = k;
        BatchScratch& sc = *reinterpret_cast<BatchScratch*>(smem_raw + (size_t)wib * kScratchSlot);
        st = decode_message_batched(in, n, k, so, expected, lane, sc);
        if (st == B2C_OK) mode = __any_sync(kFull, so.bad) ? kModeWhole : kModeBlocks;
    }
    if (lane == 0) {
        P.mode[m] = mode;
        P.ulen_tmp[m] = expected;
        if (mode != kModeWhole) {
            P.status[m] = st;
            if (P.ulen) P.ulen[m] = expected;
        }
    }
}

// ---------------------------------------------------------------------------
// Pass 1, faster: block boundaries by POINTER LIFTING instead of a tag-by-tag walk.
// ---------------------------------------------------------------------------
// Where the next tag starts is a function of the bytes at a position alone: next(p) = p + (bytes the element at p
// occupies in the stream), and so is the output it produces.  So for a 256-byte window of the stream the warp
// evaluates next() and out() at ALL 256 positions at once (as if a tag started at each), then squares the pair
// (next, out-sum) seven times: after level k entry p holds where the chain is 2^k elements on and what they produced
// (an element is >= 2 bytes, so 128 hops leave the window).  One shared-memory read at the window's true entry position
// then yields where the chain leaves the window and how much output the window accounts for: ~7 dependent
// shared-memory rounds per 256 stream bytes instead of ~90 dependent tag steps.  Only a window in which the output
// count crosses a multiple of 64 KiB is walked tag by tag, to find the tag that starts the block (or to learn that an
// element straddles the boundary).  This pass validates nothing beyond "the chain ends exactly at the stream's end with
// the announced output": every block is validated when it is decoded, and a block that fails hands its stream to the
// whole-stream decoder (which alone decides between corrupt and merely unsplittable).
constexpr uint32_t kLiftWin = 256;
constexpr uint32_t kLiftBad = 0xfffffff0u;  // element cut short by the end of the stream (or position past the end)

struct LiftScratch {
    uint32_t N[kLiftWin];
    uint32_t S[kLiftWin];
    uint32_t words[kLiftWin / 4 + 4];
};

// element at window-relative position p: stream bytes it occupies (0 = invalid / cut short) and output bytes
__device__ __forceinline__ uint64_t lift_element(uint32_t b4, uint32_t b5th, uint32_t left /* n - abs, >= 1 */, uint32_t* out) {
    const uint32_t c = b4 & 0xff, kind = c & 3;
    if (kind == 0) {
        uint32_t l = c >> 2, hdr = 1, len = l + 1;
        if (l >= 60) {
            const uint32_t extra = l - 59;
            hdr = 1 + extra;
            if (left < hdr) return 0;
            uint32_t v = b4 >> 8;
            if (extra == 4) v |= b5th << 24;
            else v &= (1u << (8 * extra)) - 1u;
            len = v + 1u;  // may wrap to 0 like the reference
        }
        *out = len;
        const uint64_t size = (uint64_t)hdr + len;
        return size <= left ? size : 0;
    }
    const uint32_t size = kind == 1 ? 2u : kind == 2 ? 3u : 5u;
    *out = kind == 1 ? ((c >> 2) & 7) + 4 : (c >> 2) + 1;
    return size <= left ? size : 0;
}

// next() and out() at all 256 positions of the window at wbase, squared seven times: on return N[e] / S[e] say where
// a chain entering the window at e leaves it (window-relative; kLiftBad: cut short; a fixed point N[e] == e: the
// stream's end) and how much output it produces on the way.  Whole warp; N, S, words are the warp's shared memory.
__device__ __forceinline__ void lift_window(const uint8_t* __restrict__ in, const uint32_t n, const uint32_t wbase,
                                            uint32_t* N, uint32_t* S, uint32_t* words, const uint32_t lane) {
    const uint8_t* wb = reinterpret_cast<const uint8_t*>(words);
    __syncwarp();
    for (uint32_t wi = lane; wi < kLiftWin / 4 + 4; wi += 32) {  // the bytes (+ the look-ahead of a tag at the last position)
        const uint32_t at = wbase + 4 * wi;
        words[wi] = at < n ? ld32u(in + at) : 0u;
    }
    __syncwarp();
    uint32_t Nr[8], Sr[8];
    {
        const uint32_t w0 = words[2 * lane], w1 = words[2 * lane + 1], w2 = words[2 * lane + 2];
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) {
            const uint32_t p = lane * 8 + j, abs = wbase + p;
            const uint32_t lo = j < 4 ? w0 : w1, hi = j < 4 ? w1 : w2;
            const uint32_t b4 = __funnelshift_r(lo, hi, (j & 3) * 8);
            uint32_t o = 0;
            uint64_t size = 0;
            if (abs < n) size = lift_element(b4, wb[p + 4], n - abs, &o);
            // the stream's end is a fixed point (a chain that reaches it stays there)
            Nr[j] = size ? (uint32_t)min((unsigned long long)(p + size), (unsigned long long)(kLiftBad - 1))
                         : (abs == n ? p : kLiftBad);
            Sr[j] = size ? o : 0u;
            N[p] = Nr[j];
            S[p] = Sr[j];
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int lev = 0; lev < 7; lev++) {
        uint32_t nq[8], sq[8];
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) {
            const uint32_t q = Nr[j];
            nq[j] = q < kLiftWin ? N[q] : q;
            sq[j] = q < kLiftWin ? S[q] : 0u;
        }
        __syncwarp();
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) {
            if (Nr[j] < kLiftWin) {
                Nr[j] = nq[j];
                const uint32_t t = Sr[j] + sq[j];
                Sr[j] = t < Sr[j] ? 0xffffffffu : t;  // saturate
                N[lane * 8 + j] = Nr[j];
                S[lane * 8 + j] = Sr[j];
            }
        }
        __syncwarp();
    }
}

// One stream, one warp, window after window from `pos` (with `cum` bytes of output before it) until the chain
// leaves [pos, stop) or ends; returns false when the stream is irregular.  When `want` is a block boundary inside the
// stretch, *found receives the stream offset of the tag that starts that block.
__device__ bool lift_walk(const uint8_t* __restrict__ in, const uint32_t n, const uint32_t expected, uint32_t& pos,
                          uint32_t& cum, const uint64_t stop, const uint32_t want, uint32_t* found, LiftScratch& sc,
                          const uint32_t lane, uint32_t* budget = nullptr) {
    const uint8_t* wb = reinterpret_cast<const uint8_t*>(sc.words);
    while (pos < n && pos < stop) {
        if (budget) {  // a caller that only wants short chains: give up after so many windows
            if (*budget == 0) return false;
            --*budget;
        }
        const uint32_t wbase = pos & ~(kLiftWin - 1);
        lift_window(in, n, wbase, sc.N, sc.S, sc.words, lane);
        const uint32_t e = pos - wbase;
        const uint32_t exit_rel = sc.N[e], add = sc.S[e];
        if (exit_rel == kLiftBad || add > expected - cum) return false;  // cut short, or more output than announced
        if (found && cum <= want && (uint64_t)cum + add >= want) {
            // the block starts inside this window (or right at its exit): find the tag, one hop at a time
            uint32_t p = e, c = cum;
            for (;;) {
                if (c == want) {
                    *found = wbase + p;
                    return true;
                }
                if (p >= kLiftWin) return false;
                const uint32_t b4 = __funnelshift_r(sc.words[p >> 2], sc.words[(p >> 2) + 1], (p & 3) * 8);
                uint32_t o = 0;
                const uint64_t size = lift_element(b4, wb[p + 4], n - (wbase + p), &o);
                if (size == 0 || c + o > want) return false;  // cut short, or an element straddles the boundary
                c += o;
                p += (uint32_t)min((unsigned long long)size, (unsigned long long)kLiftBad);
            }
        }
        cum += add;
        if (exit_rel == e) {  // fixed point: the stream's end
            pos = n;
            break;
        }
        pos = wbase + exit_rel;
    }
    return true;
}

constexpr uint32_t kSegBytes = 2048;  // stream bytes per segment
// A table entry answers, for ONE stream position p of a segment taken as the start of an element: following the chain
// from p, where does it leave the segment and how much output does it produce on the way.
//   bits 0-13  E  >= kSegBytes: the chain leaves the segment and stands at segment-relative position E (< 16384: in
//                 one of the next segments); < kSegBytes: it STOPS in front of position E of this segment, which is
//                 the stream's end, an element cut short, or an element that reaches further than E can say (a long
//                 literal): whoever follows the chain looks at that position itself
//   bits 14-31 S  output bytes of the elements passed (at most 64 bytes per 3 stream bytes, and one last element of
//                 less than 16 KiB: < 2^18)
// Short segments: the sweep is one thread's serial work, a block boundary is then searched element by element inside
// one segment, and the hop count of the chain (stream bytes / kSegBytes) stays in the hundreds for a 4 MiB body.
constexpr uint32_t kDpEBits = 14, kDpEMask = (1u << kDpEBits) - 1u;
constexpr uint32_t kDpRing = 128;  // entries per thread kept in shared memory: the last 64 finished + the 64 in the making
constexpr uint32_t kDpNear = 64;   // a successor at most this far ahead is still in the ring
constexpr uint32_t kDpThreads = 128;

struct SegScanParams {
    DecompressParams D;
    uint32_t* seg_cnt;        // [n_list]    segments of the w-th listed stream (0: not scanned)
    const uint64_t* seg_base; // [n_list+1]  exclusive scan of seg_cnt
    uint32_t* seg_tab;        // [segments][kSegBytes]  one entry per stream position
    uint2* bnd;               // [output blocks]  where to start looking for the block's first tag: (stream pos, output before it)
    uint32_t seg_cap;         // capacity of seg_tab in segments
};

// K0: one warp per large-slot stream.  Reads the preamble and tries the chain window by window with a small budget:
// a stream of few, long elements (incompressible data is one 64 KiB literal per block) is finished here, boundaries
// included, and takes no part in the segment pass, whose cost is per stream byte.  The others get their segment count.

__global__ void __launch_bounds__(256) snappy_seg_count_kernel(SegScanParams Q) {
    __shared__ LiftScratch scratch[8];
    const DecompressParams& P = Q.D;
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * (blockDim.x >> 5) + wib;
    if (w >= P.n_list) return;
    const uint32_t m = P.list ? P.list[w] : w;
    const uint32_t cap = P.out_cap[m];
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    uint32_t expected = 0, cnt = 0, mode = kModeWhole;
    int st = B2C_OK;
    if (cap > kFragBytes) {
        const uint32_t k = read_preamble(in, n, &expected);
        if (k == 0) {
            st = B2C_CORRUPT;
            mode = kModeDone;
        } else if (expected > cap) {
            st = B2C_TOO_SMALL;
            mode = kModeDone;
        } else if (expected > kFragBytes) {
            const uint32_t f0 = P.out_frag_first[m];
            const uint32_t nblk = (expected + kFragBytes - 1) / kFragBytes;
            if (lane == 0) P.blk_ip[f0] = k;
            uint32_t pos = k, cum = 0, budget = n / 32768 + 8;  // enough for one element (and one boundary) per 64 KiB of stream
            bool ok = true;
            for (uint32_t nb = 1; ok && nb < nblk; nb++) {
                uint32_t found = 0xffffffffu;
                ok = lift_walk(in, n, expected, pos, cum, ~0ull, nb * kFragBytes, &found, scratch[wib], lane, &budget) &&
                     found != 0xffffffffu;
                if (ok && lane == 0) {
                    P.blk_ip[f0 + nb] = found;
                    Q.bnd[f0 + nb] = make_uint2(found, nb * kFragBytes);  // the boundary kernel will find it where it is
                }
            }
            if (ok) ok = lift_walk(in, n, expected, pos, cum, ~0ull, 0u, nullptr, scratch[wib], lane, &budget);
            if (ok && pos == n && cum == expected) {
                mode = kModeBlocks;  // done: status below
            } else if (budget == 0) {
                cnt = (n + kSegBytes - 1) / kSegBytes;  // a long chain: the segment pass takes it
            }                                           // else irregular: the whole-stream decoder decides
        }
    }
    if (lane == 0) {
        Q.seg_cnt[w] = cnt;
        P.mode[m] = mode;  // streams in the segment pass: upgraded to kModeBlocks by the chain kernel
        P.ulen_tmp[m] = expected;
        if (mode != kModeWhole) {
            P.status[m] = st;
            if (P.ulen) P.ulen[m] = expected;
        }
    }
}

// KA: one THREAD per segment, from its last position to its first: the entry of p is the entry of p's successor plus
// p's own output (or "leaves here" / "stops here"), so one backward sweep fills the table with a few dozen
// instructions per stream byte and no search.  The successor of p is nearly always within 64 bytes (every element but
// a literal of more than 60 bytes is), where the thread's ring in shared memory still holds its entry (slot-major,
// lanes adjacent: conflict-free); otherwise the entry is read back from the table, to which every finished run of 64
// entries is flushed as 16-byte stores.  Positions are taken four at a time: an element is at least 2 bytes, so the
// entries of p and p - 1 depend on neither each other nor on anything newer, and their lookups go out together.
struct DpElem {
    uint32_t nx;    // successor inside the segment whose entry is to be added (when `look`)
    uint32_t add;   // what to add to it; the whole entry otherwise
    bool look;
};
__device__ __forceinline__ DpElem dp_decode(const uint32_t lo, const uint32_t hi, const uint32_t q, const uint32_t avail) {
    const uint32_t c = lo & 0xffu, kind = c & 3u, l = c >> 2;
    uint32_t size = kind == 1 ? 2u : kind == 2 ? 3u : 5u;
    uint32_t o = kind == 1 ? (l & 7u) + 4u : l + 1u;
    bool ok = q < avail;
    if (kind == 0) {
        size = l + 2u;
        if (l >= 60u) {
            const uint32_t ex = l - 59u;
            uint32_t v = __funnelshift_r(lo, hi, 8);
            if (ex < 4u) v &= (1u << (8u * ex)) - 1u;
            ok = ok && v < kDpEMask;  // else it reaches beyond the next segment (or its length wraps): stop
            o = v + 1u;
            size = 1u + ex + o;
        }
    }
    ok = ok && size <= avail - q;  // else cut short by the stream's end
    const uint32_t nx = q + size;
    ok = ok && nx <= kDpEMask;
    DpElem e;
    e.look = ok && nx < kSegBytes;
    e.nx = nx;
    e.add = !ok ? q : e.look ? (o << kDpEBits) : ((o << kDpEBits) | nx);
    return e;
}

__global__ void __launch_bounds__(kDpThreads) snappy_seg_dp_kernel(SegScanParams Q) {
    extern __shared__ __align__(16) uint32_t dp_ring[];  // [warp][slot][lane]
    const DecompressParams& P = Q.D;
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t rbase = wib * (kDpRing * 32) + lane;
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= Q.seg_base[P.n_list] || g >= Q.seg_cap) return;
    uint32_t lo_w = 0, hi_w = P.n_list;  // the listed stream w with seg_base[w] <= g < seg_base[w+1]
    while (hi_w - lo_w > 1) {
        const uint32_t mid = (lo_w + hi_w) >> 1;
        if (Q.seg_base[mid] <= g) lo_w = mid;
        else hi_w = mid;
    }
    const uint32_t w = lo_w, seg = (uint32_t)(g - Q.seg_base[w]);
    const uint32_t m = P.list ? P.list[w] : w;
    const uint32_t n = P.in_len[m];
    const uint32_t sbase = seg * kSegBytes;
    const uint8_t* sp = P.in + P.in_off[m] + sbase;
    const uint32_t avail = n - sbase;  // > 0; positions >= avail lie at or past the stream's end
    uint32_t* tab = Q.seg_tab + g * kSegBytes;
    // stream bytes [q16, q16 + 16) as four words, whatever the alignment; a word that starts at or past the end is 0
    auto load16 = [&](int q16, uint32_t (&out)[4]) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(sp + q16);
        const uint32_t* a4 = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
        const uint32_t sh = (uint32_t)(a & 3) * 8;
        uint32_t x[5];
#pragma unroll
        for (int k = 0; k < 5; k++) x[k] = q16 + 4 * (k ? k - 1 : 0) < (int)avail ? a4[k] : 0u;  // at most 7 bytes past the end
#pragma unroll
        for (int k = 0; k < 4; k++) out[k] = q16 + 4 * k < (int)avail ? __funnelshift_r(x[k], x[k + 1], sh) : 0u;
    };
    uint32_t cur[4], nxt[4];
    load16((int)kSegBytes, cur);  // the look-ahead of the last positions
    uint32_t prev_lo = cur[0], prev_hi = cur[1];  // stream bytes [q + 4, q + 12) of the group at q
    load16((int)kSegBytes - 16, nxt);
#pragma unroll 1
    for (int q16 = (int)kSegBytes - 16; q16 >= 0; q16 -= 16) {
#pragma unroll
        for (int k = 0; k < 4; k++) cur[k] = nxt[k];
        if (q16 >= 16) load16(q16 - 16, nxt);  // one block ahead: its latency hides behind 16 positions
#pragma unroll
        for (int k = 3; k >= 0; k--) {
            const uint32_t q = (uint32_t)q16 + 4 * k;  // positions q + 3, q + 2, q + 1, q
            const uint32_t w32 = cur[k];
            DpElem e[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t lo = j ? __funnelshift_r(w32, prev_lo, 8 * j) : w32;
                const uint32_t hi = j ? __funnelshift_r(prev_lo, prev_hi, 8 * j) : prev_lo;
                e[j] = dp_decode(lo, hi, q + j, avail);
            }
#pragma unroll
            for (int half = 1; half >= 0; half--) {
                uint32_t t[2];
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const int j = 2 * half + i;
                    t[i] = 0;
                    if (e[j].look)
                        t[i] = e[j].nx - (q + j) <= kDpNear ? dp_ring[rbase + (e[j].nx & (kDpRing - 1)) * 32]
                                                           : __ldcg(tab + e[j].nx);
                }
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const int j = 2 * half + i;
                    dp_ring[rbase + ((q + j) & (kDpRing - 1)) * 32] = t[i] + e[j].add;
                }
            }
            prev_hi = prev_lo;
            prev_lo = w32;
        }
        if ((q16 & 63) == 0) {  // entries [q16, q16 + 64) are final
            uint4* dst = reinterpret_cast<uint4*>(tab + q16);
            const uint32_t r0 = rbase + ((uint32_t)q16 & (kDpRing - 1)) * 32;
#pragma unroll 4
            for (uint32_t i = 0; i < 16; i++)
                dst[i] = make_uint4(dp_ring[r0 + (4 * i) * 32], dp_ring[r0 + (4 * i + 1) * 32],
                                    dp_ring[r0 + (4 * i + 2) * 32], dp_ring[r0 + (4 * i + 3) * 32]);
        }
    }
}

__device__ __forceinline__ uint64_t element_at(const uint8_t* __restrict__ in, uint32_t n, uint32_t pos, uint32_t* o) {
    return lift_element(ld32u(in + pos), pos + 4 < n ? in[pos + 4] : 0u, n - pos, o);
}

// KB: one thread per scanned stream hops from segment to segment through the table (decoding itself the elements the
// table stops at), checks that the chain ends at the stream's end with the announced output, and notes for every block
// boundary the hop it lies in.
__global__ void __launch_bounds__(kDpThreads) snappy_seg_chain_kernel(SegScanParams Q) {
    const DecompressParams& P = Q.D;
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= P.n_list) return;
    const uint64_t g0 = Q.seg_base[w];
    if (Q.seg_base[w + 1] == g0 || Q.seg_base[w + 1] > Q.seg_cap) return;  // not scanned (stays with the whole-stream decoder)
    const uint32_t m = P.list ? P.list[w] : w;
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    const uint32_t expected = P.ulen_tmp[m];
    uint32_t tmp = 0;
    const uint32_t k = read_preamble(in, n, &tmp);
    const uint32_t f0 = P.out_frag_first[m];
    const uint32_t nblk = (expected + kFragBytes - 1) / kFragBytes;
    P.blk_ip[f0] = k;
    uint32_t pos = k, cum = 0, nb = 1;
    bool ok = true;
    while (pos < n) {
        const uint32_t seg = pos / kSegBytes, q = pos - seg * kSegBytes;
        const uint32_t t = Q.seg_tab[(g0 + seg) * kSegBytes + q];
        const uint32_t E = t & kDpEMask, S = t >> kDpEBits;
        const uint32_t pos0 = pos, cum0 = cum;
        if (S > expected - cum) {
            ok = false;
            break;
        }
        cum += S;
        pos = seg * kSegBytes + E;
        if (E < kSegBytes && pos < n) {  // an element the table does not take
            uint32_t o = 0;
            const uint64_t size = element_at(in, n, pos, &o);
            if (size == 0 || o > expected - cum) {
                ok = false;
                break;
            }
            cum += o;
            pos += (uint32_t)size;  // <= n
        }
        while (nb < nblk && (uint64_t)cum >= (uint64_t)nb * kFragBytes) {  // boundary in (cum0, cum]
            Q.bnd[f0 + nb] = make_uint2(pos0, cum0);
            nb++;
        }
    }
    if (ok && pos == n && cum == expected && nb == nblk) {
        P.mode[m] = kModeBlocks;
        P.status[m] = B2C_OK;
        if (P.ulen) P.ulen[m] = expected;
    }
}

// KC: one thread per block boundary walks, element by element, from the noted position to the tag that starts the
// block (at most one segment and one long element away).
__global__ void __launch_bounds__(kDpThreads) snappy_seg_boundary_kernel(SegScanParams Q) {
    const DecompressParams& P = Q.D;
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= P.n_blocks) return;
    const uint32_t m = P.out_frag_msg[f];
    if (P.mode[m] != kModeBlocks) return;
    const uint32_t kb = f - P.out_frag_first[m];
    if (kb == 0) return;
    const uint32_t expected = P.ulen_tmp[m];
    const uint64_t want = (uint64_t)kb * kFragBytes;
    if (want >= expected) return;
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    const uint2 start = Q.bnd[f];
    uint32_t pos = start.x;
    uint64_t cum = start.y;
    bool found = false;
    while (pos < n) {
        if (cum == want) {
            found = true;
            break;
        }
        // 32 lanes walk 32 different streams: without this, some lane misses the L1 at nearly every step
        if (pos + 384 < n) asm volatile("prefetch.global.L1 [%0];" ::"l"(in + pos + 384));
        uint32_t o = 0;
        const uint64_t size = element_at(in, n, pos, &o);
        if (size == 0 || cum + o > want) break;  // cut short, or an element straddles the boundary: not splittable
        cum += o;
        pos += (uint32_t)size;
    }
    if (found) P.blk_ip[f] = pos;
    else P.mode[m] = kModeWhole;
}

// Pass 2 for large slots: every 64 KiB block of a splittable stream is decoded by its own warp.
__global__ void __launch_bounds__(256, 5) snappy_decompress_blocks_kernel(DecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    BatchScratch& sc = *reinterpret_cast<BatchScratch*>(smem_raw + (size_t)wib * kScratchSlot);
    // resident warps pull blocks: the list holds every block of the batch, those of streams decoded elsewhere included
    for (;;) {
        uint32_t f = 0;
        if (lane == 0) f = atomicAdd(P.work_ctr, 1u);
        f = __shfl_sync(kFull, f, 0);
        if (f >= P.n_blocks) return;
        const uint32_t m = P.out_frag_msg[f];
        if (P.mode[m] != kModeBlocks) continue;
        const uint32_t kblk = f - P.out_frag_first[m];
        const uint32_t expected = P.ulen_tmp[m];
        const uint32_t begin = kblk * kFragBytes;
        if (begin >= expected) continue;
        const uint32_t block_len = min(kFragBytes, expected - begin);
        const uint32_t ip0 = P.blk_ip[f];
        const uint32_t ip1 = begin + kFragBytes < expected ? P.blk_ip[f + 1] : P.in_len[m];
        LinearOut lo{P.out + P.out_off[m] + begin};
        __syncwarp();
        const int st = decode_message_batched(P.in + P.in_off[m], ip1, ip0, lo, block_len, lane, sc);
        // a block that does not decode on its own: the stream is corrupt, or merely not splittable (a copy that reaches
        // back across the block's start); the whole-stream decoder, launched next, redoes it and decides
        if (st != B2C_OK && lane == 0) P.mode[m] = kModeWhole;
    }
}


// One instantiation per way of decoding, so that each carries only its own decoder's registers (they are all at the
// 48-register limit of 5 CTAs per SM): kSmemSlot: slots that fit a per-warp shared-memory buffer, batched; kGlobal: batched,
// straight into HBM; kRing: batched through the 64 KiB ring; kPlain: one tag per step (tunable batch_decode = 0).
enum : int { kSmemSlot = 0, kGlobal = 1, kRing = 2, kPlain = 3 };
template <int kPath>
__global__ void __launch_bounds__(256, 5) snappy_decompress_kernel(DecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    // resident warps pull streams (their sizes differ, and those decoded block-parallel are only skipped)
    for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(P.work_ctr, 1u);
    w = __shfl_sync(kFull, w, 0);
    if (w >= P.n_list) return;
    const uint32_t m = P.list ? P.list[w] : w;
    if (P.mode && P.mode[m] != kModeWhole) continue;  // decoded block-parallel, or already rejected by the scan
    __syncwarp();
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    uint8_t* gout = P.out + P.out_off[m];
    const uint32_t cap = P.out_cap[m];
    uint32_t expected = 0;
    const uint32_t k = read_preamble(in, n, &expected);
    int st;
    if (k == 0) {
        st = B2C_CORRUPT;
    } else if (expected > cap) {
        st = B2C_TOO_SMALL;
    } else if ((kPath == kSmemSlot || kPath == kPlain) && P.smem_per_warp && expected <= P.smem_per_warp) {
        uint8_t* sbuf = smem_raw + (size_t)wib * (P.smem_per_warp + kScratchSlot);
        BatchScratch& sc = *reinterpret_cast<BatchScratch*>(sbuf + P.smem_per_warp);
        LinearOut lo{sbuf};
        if constexpr (kPath == kSmemSlot) st = decode_message_batched(in, n, k, lo, expected, lane, sc);
        else st = decode_message_warp(in, n, k, sbuf, expected, lane);
        __syncwarp();
        if (st == B2C_OK) {
            // flush: bytes to 16-byte alignment of the destination, then 16-byte stores
            uint32_t head = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(gout) & 15)) & 15u;
            if (head > expected) head = expected;
            if (lane < head) gout[lane] = sbuf[lane];
            const uint32_t body = (expected - head) >> 4;
            if ((head & 3) == 0) {
                const uint32_t* s32 = reinterpret_cast<const uint32_t*>(sbuf + head);
                uint4* g128 = reinterpret_cast<uint4*>(gout + head);
                for (uint32_t i = lane; i < body; i += 32)
                    g128[i] = make_uint4(s32[4 * i], s32[4 * i + 1], s32[4 * i + 2], s32[4 * i + 3]);
            } else {
                const uint8_t* s = sbuf + head;
                uint4* g128 = reinterpret_cast<uint4*>(gout + head);
                for (uint32_t i = lane; i < body; i += 32) {
                    const uint8_t* q = s + 16 * i;
                    uint32_t x[4];
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        x[j] = q[4 * j] | (q[4 * j + 1] << 8) | (q[4 * j + 2] << 16) | ((uint32_t)q[4 * j + 3] << 24);
                    g128[i] = make_uint4(x[0], x[1], x[2], x[3]);
                }
            }
            const uint32_t done = head + (body << 4);
            if (lane < expected - done) gout[done + lane] = sbuf[done + lane];
        }
    } else {
        if constexpr (kPath == kRing) {
            RingOut ro{smem_raw + (size_t)wib * (65536 + kScratchSlot), gout, 0u, 0u, 0u};
            BatchScratch& rsc = *reinterpret_cast<BatchScratch*>(ro.ring + 65536);
            st = decode_message_batched(in, n, k, ro, expected, lane, rsc);
        } else if constexpr (kPath == kGlobal) {
            BatchScratch& sc = *reinterpret_cast<BatchScratch*>(smem_raw + (size_t)wib * (P.smem_per_warp + kScratchSlot) + P.smem_per_warp);
            LinearOut lo{gout};
            st = decode_message_batched(in, n, k, lo, expected, lane, sc);
        } else {  // kPlain, or a slot larger than its class allows (the lists are built by capacity: does not happen)
            st = decode_message_warp(in, n, k, gout, expected, lane);
        }
    }
    if (lane == 0) {
        P.status[m] = st;
        if (P.ulen) P.ulen[m] = expected;
    }
    }
}

__global__ void snappy_ulen_kernel(const uint8_t* __restrict__ in, const uint64_t* __restrict__ off,
                                   const uint32_t* __restrict__ len, uint32_t n, uint32_t* __restrict__ ulen,
                                   int32_t* __restrict__ status) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n) return;
    uint32_t v = 0;
    const uint32_t k = read_preamble(in + off[m], len[m], &v);
    ulen[m] = k ? v : 0;
    if (status) status[m] = k ? B2C_OK : B2C_CORRUPT;
}

// How far a stream decodes: snappy::UncompressAsMuchAsPossible (snappy.cc:1519-1524).  One thread per stream walks
// the tags exactly as DecompressAllTags does (snappy.cc:716-787) and keeps the books of the writer that function
// uses, SnappyScatteredWriter (snappy.cc:1323-1470) — blocks of min(64 KiB, expected - filled), a literal that does not
// fit fills the current block before it is refused, a literal cut short by the end of the input is appended as far
// as it goes — without moving any data.  Per stream it reports
//   [0] what the reference returns (the writer's Size(): larger than what it wrote when a literal was refused after
//       filling a block, because that block is then counted twice),
//   [1] the bytes the reference's sink receives,
//   [2] the stream offset up to which whole tags were applied,
//   [3] the bytes of one more, partially applied literal, and [4] where in the stream its data starts.
// A rarely used entry point (nothing in bRPC calls it): serial per stream, not a throughput kernel.
__global__ void snappy_prefix_kernel(const uint8_t* __restrict__ in, const uint64_t* __restrict__ off,
                                     const uint32_t* __restrict__ len, uint32_t n_msgs, uint32_t* __restrict__ out5) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_msgs) return;
    const uint8_t* src = in + off[m];
    const uint32_t n = len[m];
    uint32_t expected = 0;
    uint32_t ip = read_preamble(src, n, &expected);
    uint64_t full = 0, blocks = 0;  // full_size_, sum of the sizes of the allocated blocks
    uint32_t bcap = 0, bused = 0;   // the current block: op_limit_ - op_base_, op_ptr_ - op_base_
    uint32_t whole = 0, partial = 0, partial_src = 0, consumed = ip;
    // SnappyScatteredWriter::Append; returns false on refusal, *wrote = bytes that did go in
    auto append = [&](uint32_t l, uint32_t* wrote) -> bool {
        *wrote = 0;
        uint32_t avail = bcap - bused;
        if (l <= avail) {
            bused += l;
            *wrote = l;
            return true;
        }
        while (l > avail) {  // SlowAppend
            bused += avail;
            *wrote += avail;
            full += bused;
            l -= avail;
            if (full + l > expected) return false;  // (bused stays: Size() counts this block twice, like the reference)
            const uint32_t bsize = (uint32_t)min((uint64_t)65536, (uint64_t)expected - full);
            blocks += bsize;
            bcap = bsize;
            bused = 0;
            avail = bsize;
        }
        bused += l;
        *wrote += l;
        return true;
    };
    if (ip) {
        while (ip < n) {
            const uint32_t c = src[ip], kind = c & 3;
            const uint32_t extra = kind == 0 ? (c >= 0xf0 ? (c >> 2) - 59 : 0) : kind == 1 ? 1 : kind == 2 ? 2 : 4;
            if (n - ip < 1 + extra) break;  // RefillTag: the tag is cut short
            uint32_t trailer = 0;
            for (uint32_t i = 0; i < extra; i++) trailer |= (uint32_t)src[ip + 1 + i] << (8 * i);
            uint32_t wrote = 0;
            if (kind == 0) {
                uint32_t l = (c >> 2) + 1;
                if (l >= 61) l = trailer + 1u;
                const uint32_t data = ip + 1 + extra, avail_in = n - data;
                if (avail_in < l) {  // premature end of input: what is there is appended, then the decoder stops
                    append(avail_in, &wrote);
                    partial = wrote;
                    partial_src = data;
                    break;
                }
                if (!append(l, &wrote)) {
                    partial = wrote;
                    partial_src = data;
                    break;
                }
                whole += l;
                ip = data + l;
            } else {
                uint32_t l, offset;
                if (kind == 1) {
                    l = ((c >> 2) & 7) + 4;
                    offset = ((c >> 5) << 8) | trailer;
                } else {
                    l = (c >> 2) + 1;
                    offset = trailer;
                }
                // AppendFromSelf: the fast path needs no check beyond the block; the slow path checks first
                const uint64_t cur = full + bused;
                const bool fast = offset - 1u < bused && bcap - bused >= l + 10;
                if (!fast && ((uint64_t)offset - 1u >= cur || (uint64_t)expected - cur < l)) break;
                append(l, &wrote);  // cannot be refused after the checks above
                whole += l;
                ip += 1 + extra;
            }
            consumed = ip;
        }
    }
    const uint64_t size = full + bused;
    out5[5 * m + 0] = (uint32_t)size;
    out5[5 * m + 1] = (uint32_t)min(size, blocks);
    out5[5 * m + 2] = consumed;
    out5[5 * m + 3] = partial;
    out5[5 * m + 4] = partial_src;
    (void)whole;
}

// IsValidCompressed for a batch: the batched decoder's parse + checks, no data movement.
__global__ void __launch_bounds__(256) snappy_validate_kernel(const uint8_t* __restrict__ in,
                                                              const uint64_t* __restrict__ off,
                                                              const uint32_t* __restrict__ len, uint32_t n_msgs,
                                                              uint32_t* __restrict__ ulen, int32_t* __restrict__ status) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t m = blockIdx.x * (blockDim.x >> 5) + wib;
    if (m >= n_msgs) return;
    const uint8_t* src = in + off[m];
    const uint32_t n = len[m];
    uint32_t expected = 0;
    const uint32_t k = read_preamble(src, n, &expected);
    int st = B2C_CORRUPT;
    if (k) {
        ScanOut so{nullptr, 0u, 0u};
        BatchScratch& sc = *reinterpret_cast<BatchScratch*>(smem_raw + (size_t)wib * kScratchSlot);
        st = decode_message_batched(src, n, k, so, expected, lane, sc);
    }
    if (lane == 0) {
        status[m] = st;
        if (ulen) ulen[m] = k ? expected : 0;
    }
}

// verify step of b2c_snappy_decompress_verify_dev
__global__ void snappy_verify_crc_kernel(const uint32_t* __restrict__ crc, const uint32_t* __restrict__ masked_want,
                                         const uint32_t* __restrict__ ulen, const uint32_t* __restrict__ cap,
                                         int32_t* __restrict__ status, uint32_t n) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n || status[m] != B2C_OK) return;
    const uint32_t c = crc[m];
    const uint32_t masked = ((c >> 15) | (c << 17)) + 0xa282ead8u;  // crc32c::Mask, crc32c.h:38-41
    if (ulen[m] != cap[m] || masked != masked_want[m]) status[m] = B2C_CRC_MISMATCH;
}

int launch_prefix(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_out5) {
    if (L.n == 0) return B2C_OK;
    snappy_prefix_kernel<<<(L.n + 63) / 64, 64, 0, lane.stream>>>(d_in, L.d_off, L.d_len, L.n, d_out5);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

int launch_validate(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_ulen,
                    int32_t* d_status) {
    if (L.n == 0) return B2C_OK;
    snappy_validate_kernel<<<(L.n + 7) / 8, 256, 8 * kScratchSlot, lane.stream>>>(d_in, L.d_off, L.d_len, L.n, d_ulen, d_status);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

int launch_decompress_verify(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in, const LayoutView& out,
                             uint8_t* d_out, const uint32_t* d_masked_crc, uint32_t* d_ulen, int32_t* d_status) {
    if (in.n == 0) return B2C_OK;
    // temporaries after whatever the decoder itself keeps in ws_misc
    // (the decoder keeps block offsets / modes at the start of ws_misc, the CRC launcher its chunk remainders)
    const size_t dec_words = std::max((size_t)out.n_frags + 2 * (size_t)out.n, (size_t)out.n_chunks) + 16;
    int rc = ws_reserve(ctx, lane, lane.ws_misc, (dec_words + 2 * (size_t)in.n + 16) * 4);
    if (rc) return rc;
    uint32_t* tmp = static_cast<uint32_t*>(lane.ws_misc.p) + dec_words;
    uint32_t* d_crc = tmp;
    uint32_t* d_ul = d_ulen ? d_ulen : tmp + in.n;
    if ((rc = launch_decompress(ctx, lane, in, d_in, out, d_out, d_ul, d_status))) return rc;
    // the CRC launcher parks multi-chunk remainders at the start of ws_misc: the decoder is done with it by now
    if ((rc = launch_crc32c(ctx, lane, out, d_out, nullptr, d_crc))) return rc;
    snappy_verify_crc_kernel<<<(in.n + 255) / 256, 256, 0, lane.stream>>>(d_crc, d_masked_crc, d_ul, out.d_len, d_status,
                                                                        in.n);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

int launch_ulen(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_ulen,
                int32_t* d_status) {
    if (L.n == 0) return B2C_OK;
    snappy_ulen_kernel<<<(L.n + 255) / 256, 256, 0, lane.stream>>>(d_in, L.d_off, L.d_len, L.n, d_ulen, d_status);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

int launch_decompress(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in, const LayoutView& out,
                      uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status) {
    if (in.n != out.n) return set_err(ctx, B2C_E_ARG, "decompress: layouts differ in message count", cudaSuccess);
    if (in.n == 0) return B2C_OK;
    static const uint32_t class_smem[3] = {1024, 8192, 0};
    static const uint32_t class_warps[3] = {8, 4, 8};
    prof_mark(ctx, lane, B2C_PROF_DECOMPRESS);
    for (int c = 0; c < 3; c++) {
        if (!out.n_dclass[c]) continue;
        if (c == 0 && ctx->small_kernels && !ctx->batch_decode_small) {  // slots <= 1 KiB: one lane per stream
            int rc = ctx->lockstep_decode ? launch_decompress_lockstep(ctx, lane, in, d_in, out, d_out, d_ulen, d_status)
                                          : launch_decompress_small(ctx, lane, in, d_in, out, d_out, d_ulen, d_status);
            if (rc) return rc;
            continue;
        }
        DecompressParams P;
        P.in = d_in;
        P.in_off = in.d_off;
        P.in_len = in.d_len;
        P.out_off = out.d_off;
        P.out_cap = out.d_len;
        P.list = out.d_dclass_list[c];
        P.out = d_out;
        P.ulen = d_ulen;
        P.status = d_status;
        P.n_list = out.n_dclass[c];
        P.smem_per_warp = class_smem[c];
        P.out_frag_first = nullptr;
        P.out_frag_msg = nullptr;
        P.blk_ip = nullptr;
        P.mode = nullptr;
        P.ulen_tmp = nullptr;
        P.n_blocks = 0;
        P.work_ctr = nullptr;
        P.batched = ctx->batch_decode ? 1u : 0u;
        if (c == 2 && ctx->batch_decode && ctx->block_decode && out.d_frag_msg) {
            // large slots: scan for block boundaries, decode splittable streams one warp per 64 KiB block
            int rc = ws_reserve(ctx, lane, lane.ws_misc, ((size_t)out.n_frags + 2 * (size_t)out.n + 16) * 4);
            if (rc) return rc;
            P.out_frag_first = out.d_frag_first;
            P.out_frag_msg = out.d_frag_msg;
            P.blk_ip = static_cast<uint32_t*>(lane.ws_misc.p);
            P.mode = P.blk_ip + out.n_frags + 1;
            P.ulen_tmp = P.mode + out.n;
            P.n_blocks = out.n_frags;
            // only large-slot messages get a mode from the scan; everything else must read as "whole"
            B2C_CUDA(ctx, cudaMemsetAsync(P.mode, 0, (size_t)out.n * 4, lane.stream));
            const uint32_t sblocks = (P.n_list + 7) / 8;
            if (ctx->lift_scan) {
                // segment tables for every scanned stream, a hop-per-segment chain per stream, a walk per boundary
                SegScanParams Q;
                Q.D = P;
                const size_t seg_cap = (size_t)(in.payload / kSegBytes) + P.n_list + 1;
                const size_t scan_tmp = ((size_t)(P.n_list + 1023) / 1024 + 2) * 8;
                const size_t tab_bytes = seg_cap * kSegBytes * sizeof(uint32_t);  // 4 bytes per stream byte
                const size_t need = tab_bytes + (size_t)out.n_frags * sizeof(uint2) + ((size_t)P.n_list + 1) * 8 +
                                    (size_t)P.n_list * 4 + scan_tmp + 256;
                if ((rc = ws_reserve(ctx, lane, lane.ws_tab, need))) return rc;
                uint8_t* wsb = static_cast<uint8_t*>(lane.ws_tab.p);
                Q.seg_tab = reinterpret_cast<uint32_t*>(wsb);
                Q.bnd = reinterpret_cast<uint2*>(wsb + tab_bytes);
                uint64_t* seg_base = reinterpret_cast<uint64_t*>(wsb + tab_bytes + (size_t)out.n_frags * sizeof(uint2));
                Q.seg_base = seg_base;
                void* tmp = seg_base + P.n_list + 1;
                Q.seg_cnt = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(tmp) + scan_tmp);
                Q.seg_cap = (uint32_t)std::min<size_t>(seg_cap, 0xffffffffu);
                snappy_seg_count_kernel<<<sblocks, 256, 0, lane.stream>>>(Q);
                ctx->launches++;
                if ((rc = launch_scan_u32_u64(ctx, lane, Q.seg_cnt, P.n_list, seg_base, tmp))) return rc;
                const int dp_smem = (int)((kDpThreads / 32) * kDpRing * 32 * sizeof(uint32_t));
                if (!ctx->seg_dp_attr_done) {
                    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_seg_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                       dp_smem));
                    ctx->seg_dp_attr_done = true;
                }
                snappy_seg_dp_kernel<<<(uint32_t)((Q.seg_cap + kDpThreads - 1) / kDpThreads), kDpThreads, dp_smem,
                                       lane.stream>>>(Q);
                snappy_seg_chain_kernel<<<(P.n_list + kDpThreads - 1) / kDpThreads, kDpThreads, 0, lane.stream>>>(Q);
                snappy_seg_boundary_kernel<<<(P.n_blocks + kDpThreads - 1) / kDpThreads, kDpThreads, 0, lane.stream>>>(Q);
                ctx->launches += 2;  // + the chain kernel counted with the blocks kernel below
            } else {
                snappy_scan_kernel<<<sblocks, 256, 8 * kScratchSlot, lane.stream>>>(P);
            }
            if ((rc = next_work_counter(ctx, lane, &P.work_ctr))) return rc;
            const uint32_t bblocks = std::min<uint32_t>((P.n_blocks + 7) / 8, (uint32_t)ctx->num_sms * 5u);
            snappy_decompress_blocks_kernel<<<bblocks, 256, 8 * kScratchSlot, lane.stream>>>(P);
            P.work_ctr = nullptr;
            ctx->launches += 2;
            B2C_CUDA(ctx, cudaGetLastError());
        }
        P.ring = (c == 2 && ctx->batch_decode && ctx->ring_decode) ? 1u : 0u;
        uint32_t wpb = class_warps[c];
        size_t smem = (size_t)wpb * (class_smem[c] + kScratchSlot);
        if (P.ring) {
            wpb = 1;
            smem = 65536 + kScratchSlot;
            if (!ctx->ring_attr_done) {
                B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_decompress_kernel<kRing>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   (int)smem));
                ctx->ring_attr_done = true;
            }
        }
        {
            int rc = next_work_counter(ctx, lane, &P.work_ctr);
            if (rc) return rc;
        }
        // enough CTAs to fill every SM whatever the class's shared memory allows; the others find the list empty
        const uint32_t blocks = std::min<uint32_t>((P.n_list + wpb - 1) / wpb, (uint32_t)ctx->num_sms * (P.ring ? 3u : 8u));
        if (!P.batched) snappy_decompress_kernel<kPlain><<<blocks, wpb * 32, smem, lane.stream>>>(P);
        else if (P.ring) snappy_decompress_kernel<kRing><<<blocks, wpb * 32, smem, lane.stream>>>(P);
        else if (P.smem_per_warp) snappy_decompress_kernel<kSmemSlot><<<blocks, wpb * 32, smem, lane.stream>>>(P);
        else snappy_decompress_kernel<kGlobal><<<blocks, wpb * 32, smem, lane.stream>>>(P);
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    prof_mark(ctx, lane, B2C_PROF_DECOMPRESS);
    return B2C_OK;
}

}  // namespace b2c
