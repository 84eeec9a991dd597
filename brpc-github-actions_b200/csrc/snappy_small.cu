// snappy_small.cu — thread-per-message snappy kernels for bodies of at most 1 KiB
// (BASELINE configs[1]: 2^20 x 1 KiB RPC bodies).
//
// A warp that cooperates on one body spends 32 lanes' issue slots on one serial chain.  For
// small bodies the better shape is ONE LANE PER BODY: every lane runs the reference's scalar
// algorithm (CompressFragment, snappy.cc:329-453 / DecompressAllTags, snappy.cc:716-787) on its
// own body, a few bodies per warp (LPW), all state in shared memory:
//   compress:   u16 hash table [entry][slot] (64 KiB per CTA of 32 bodies) + each body in a
//               contiguous slot fetched by one TMA bulk copy (33 KiB per CTA); two CTAs per SM =
//               64 bodies in flight.  snappy_compress_small2_kernel: lanes FIND matches and record
//               them, warps ENCODE the records (see the comment above that kernel);
//   decompress: each stream's output in a contiguous byte-addressed slot (33 KiB per CTA, 6 CTAs
//               per SM), so back-references resolve at shared-memory latency; flushed with
//               coalesced 16-byte stores.
#include <algorithm>

#include "snappy_emit.cuh"

namespace b2c {

constexpr uint32_t kSmallMax = 1024;

// A CTA always owns 32 bodies ("slots"); LPW of each warp's lanes carry one body each, so a
// CTA has 32/LPW warps.  Fewer bodies per warp = more warps for the same shared memory: the
// serial chains of different bodies then interleave in the schedulers instead of being
// executed in lock-step (ncu on LPW=32: 10 of 32 lanes active, 0.2 IPC per warp, 2 warps/SM).
constexpr uint32_t kSmallInStride = kSmallMax + 32;  // per-slot input buffer: 1 KiB + slack for 16-byte reads past the end
// The hash table of a body holds positions below 1024: 10 bits.  They are kept as a byte (low 8 bits, tlo) and two
// bits packed four entries to a byte (thi), 1.25 KiB per body instead of 2: a CTA of 32 bodies then takes 73 KiB and
// THREE of them fit an SM (96 bodies in flight instead of 64; the kernel waits on its own dependent chains, 56 %
// issue-active with 4 warps per scheduler).  Each body's tables are its own (slot-major): only its lane touches them.
struct SmallCompressSmem {
    uint8_t tlo[1024 * 32];                 // [slot][entry]      low 8 bits
    uint8_t thi[256 * 32];                  // [slot][entry / 4]  bits 8-9, entry & 3 selects the pair
    uint8_t in[kSmallInStride * 32];        // [slot][byte], filled by TMA bulk copies
};

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n\t"
        "@P1 bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ uint32_t hash_bytes_s(uint32_t bytes, uint32_t shift) {  // snappy.cc:47-50
    return (bytes * 0x1e35a7bdu) >> shift;
}

struct SmallCompressParams {
    const uint8_t* in;
    const uint64_t* off;
    const uint32_t* len;
    const uint32_t* frag_list;
    const uint32_t* frag_msg;
    const uint32_t* frag_first;
    const uint64_t* frag_slot;
    uint64_t slot_stride;
    uint8_t* scratch;
    uint32_t* frag_clen;
    uint32_t n_list;
};

// ---------------------------------------------------------------------------
// compress in two phases: FIND (one lane per body) then ENCODE (one warp per body)
// ---------------------------------------------------------------------------
// The chain through the hash table is what has to be serial; encoding what it found is not.  In a
// one-phase kernel (round 1) a lane that finds a match spends ~100 further instructions pushing
// the literal and the copy through its output funnel while the other lanes of its warp wait
// (divergence), so almost every trip through the loop pays for the encode path.  Here a lane only
// RECORDS the match as one packed word (position | offset << 10 | length << 20; all < 1024 for
// a body <= 1 KiB) in a small L2-resident list, and once the warp's bodies are through, the
// whole warp encodes each list 32 elements at a time (emit_elements_warp: one lane per element,
// prefix sum over encoded sizes).  Byte-exact with CompressFragment: the elements and their
// order are the reference's, only the moment they are written changes.
// The table is addressed as a 32-bit shared-memory address kept opaque to the compiler, which
// otherwise re-derives it from the input slot's address on every trip (2 extra instructions per access).
__device__ __forceinline__ uint32_t tab_swap(uint32_t lo_s, uint32_t hi_s, uint32_t h, uint32_t v) {  // returns the old entry, stores v
    uint32_t lo, hb;
    const uint32_t al = lo_s + h, ah = hi_s + (h >> 2), sh = (h & 3u) * 2u;
    asm volatile("ld.shared.u8 %0, [%2];\n\tld.shared.u8 %1, [%3];" : "=r"(lo), "=r"(hb) : "r"(al), "r"(ah) : "memory");
    const uint32_t nhb = (hb & ~(3u << sh)) | ((v >> 8) << sh);  // (bfe / bfi are emulated on sm_100a and measured slower)
    asm volatile("st.shared.u8 [%0], %2;\n\tst.shared.u8 [%1], %3;" ::"r"(al), "r"(ah), "r"(v), "r"(nhb) : "memory");
    return lo | (((hb >> sh) & 3u) << 8);
}
__device__ __forceinline__ void tab_store(uint32_t lo_s, uint32_t hi_s, uint32_t h, uint32_t v) {
    uint32_t hb;
    const uint32_t al = lo_s + h, ah = hi_s + (h >> 2), sh = (h & 3u) * 2u;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(hb) : "r"(ah) : "memory");
    const uint32_t nhb = (hb & ~(3u << sh)) | ((v >> 8) << sh);
    asm volatile("st.shared.u8 [%0], %2;\n\tst.shared.u8 [%1], %3;" ::"r"(al), "r"(ah), "r"(v), "r"(nhb) : "memory");
}

__device__ __forceinline__ uint32_t small_find_body(const uint8_t* in_b, const uint32_t n, uint8_t* tlo, uint8_t* thi,
                                                    uint32_t* __restrict__ elist) {
    uint32_t lo_s = smem_u32(tlo), hi_s = smem_u32(thi);
    asm volatile("" : "+r"(lo_s), "+r"(hi_s));
    uint64_t list = reinterpret_cast<uint64_t>(elist);
    asm volatile("" : "+l"(list));  // one 64-bit base: element k goes to list + 4k with a single wide multiply-add
    uint32_t ne = 0;
    uint32_t tsize = 256;  // GetHashTable sizing, snappy.cc:253-256
    while (tsize < 16384u && tsize < n) tsize <<= 1;
    const uint32_t shift = 32 - (31 - __clz(tsize));
    if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
        const uint32_t ip_limit = n - 15;
        // `s` is the reference's `skip` (snappy.cc:377-385) with one extra state: s == 31 marks the table
        // step the copy loop makes right after a copy (snappy.cc:430-439).  That step advances by one byte
        // on a miss and is followed by the first probe of a fresh `skip = 32` sequence -- exactly what
        // step = max(1, s >> 5), s++ yields -- so one loop without mode flags serves both.
        uint32_t ip = 1, s = 32;
        uint32_t d0, d1;  // bytes ip..ip+7
        {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(in_b);
            d0 = __funnelshift_r(w[0], w[1], 8);
            d1 = __funnelshift_r(w[1], w[2], 8);
        }
        uint32_t h = hash_bytes_s(d0, shift);  // of the bytes at ip, carried like d0 / d1
        for (;;) {
            const uint32_t nxt = ip + max(1u, s >> 5);
            s++;
            if (nxt > ip_limit) break;  // snappy.cc:386-388 (never true right after a copy: ip < ip_limit there)
            // the bytes at nxt are wanted only if this step misses, but asking for them now takes their
            // shared-memory round trip off the dependent chain
            const uint32_t* wn = reinterpret_cast<const uint32_t*>(in_b + (nxt & ~3u));
            const uint32_t n0 = wn[0], n1 = wn[1], n2 = wn[2];
            const uint32_t nd0 = __funnelshift_r(n0, n1, (nxt & 3) * 8);
            const uint32_t nh = hash_bytes_s(nd0, shift);  // likewise: the next step's hash, ready if this one misses
            const uint32_t cand = tab_swap(lo_s, hi_s, h, ip);
            const uint32_t* wa = reinterpret_cast<const uint32_t*>(in_b + (cand & ~3u));
            const uint32_t sa = (cand & 3) * 8;
            const uint32_t a0 = wa[0], a1 = wa[1];
            uint32_t a2 = wa[2];
            if (__funnelshift_r(a0, a1, sa) == d0) {
                // FindMatchLength, snappy-internal.h:86-120.  ip + 15 < n here, so the first 8 bytes are in range.
                uint32_t matched;
                const uint32_t xh = __funnelshift_r(a1, a2, sa) ^ d1;
                if (xh) {
                    matched = 4 + ((uint32_t)(__ffs(xh) - 1) >> 3);
                } else {
                    matched = 8;
                    const uint32_t* wb = reinterpret_cast<const uint32_t*>(in_b + (ip & ~3u)) + 3;
                    const uint32_t sb = (ip & 3) * 8;
                    uint32_t b2 = wb[-1];
                    wa += 3;  // wa / wb: the next unread word of each side; a2 / b2: the last one read
                    for (;;) {
                        const uint32_t rem = n - (ip + matched);
                        if (rem == 0) break;
                        const uint32_t a3 = wa[0], a4 = wa[1], b3 = wb[0], b4 = wb[1];
                        const uint32_t y0 = __funnelshift_r(a2, a3, sa) ^ __funnelshift_r(b2, b3, sb);
                        const uint32_t y1 = __funnelshift_r(a3, a4, sa) ^ __funnelshift_r(b3, b4, sb);
                        const uint32_t y = y0 ? y0 : y1;
                        uint32_t nb = (y0 ? 0u : 4u) + ((uint32_t)(__ffs(y) - 1) >> 3);  // y == 0: ffs - 1 = ~0 -> large
                        nb = min(nb, min(8u, rem));
                        matched += nb;
                        if (nb < 8) break;
                        a2 = a4;
                        b2 = b4;
                        wa += 2;
                        wb += 2;
                    }
                }
                asm volatile("st.global.u32 [%0], %1;" ::"l"(list + 4ull * ne), "r"(ip | ((ip - cand) << 10) | (matched << 20))
                             : "memory");
                ne++;
                ip += matched;
                if (ip >= ip_limit) break;  // snappy.cc:427-429
                {  // snappy.cc:430-433: insert ip-1; one 4-word read yields load32(ip-1) and bytes ip..ip+7
                    const uint32_t q = ip - 1;
                    const uint32_t* w = reinterpret_cast<const uint32_t*>(in_b + (q & ~3u));
                    const uint32_t a = w[0], b = w[1], c = w[2], e = w[3], sh = (q & 3) * 8;
                    const uint32_t p0 = __funnelshift_r(a, b, sh), p1 = __funnelshift_r(b, c, sh),
                                   p2 = __funnelshift_r(c, e, sh);  // bytes q..q+11
                    tab_store(lo_s, hi_s, hash_bytes_s(p0, shift), q);
                    d0 = __funnelshift_r(p0, p1, 8);
                    d1 = __funnelshift_r(p1, p2, 8);
                    h = hash_bytes_s(d0, shift);
                }
                s = 31;
            } else {
                ip = nxt;
                d0 = nd0;
                d1 = __funnelshift_r(n1, n2, (ip & 3) * 8);
                h = nh;
            }
        }
    }
    return ne;
}

constexpr uint32_t kSmallMaxElems = kSmallMax / 4;  // copies cover >= 4 input bytes each and do not overlap

// Persistent CTAs of 32 bodies (LPW per warp); `elists` = gridDim.x * 32 lists of kSmallMaxElems words.
template <int LPW>
__global__ void __launch_bounds__(1024 / LPW, LPW >= 4 ? 3 : 2) snappy_compress_small2_kernel(SmallCompressParams P, uint32_t* elists) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    SmallCompressSmem& sm = *reinterpret_cast<SmallCompressSmem*>(smem_raw);
    __shared__ uint64_t stage_bar;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t slot = warp * LPW + lane;  // meaningful for lane < LPW
    if (threadIdx.x == 0) mbar_init(&stage_bar, blockDim.x);
    const uint32_t groups = (P.n_list + 31) / 32;
    uint32_t parity = 0;
    for (uint32_t g = blockIdx.x; g < groups; g += gridDim.x, parity ^= 1) {
        const uint32_t idx = g * 32 + slot;
        const bool live = lane < LPW && idx < P.n_list;
        __syncthreads();  // previous group's tables and staged bodies are no longer in use (and stage_bar is initialised)
        {
            uint4* t4 = reinterpret_cast<uint4*>(sm.tlo);  // tlo and thi are adjacent
            for (uint32_t i = threadIdx.x; i < (sizeof(sm.tlo) + sizeof(sm.thi)) / 16; i += blockDim.x)
                t4[i] = make_uint4(0, 0, 0, 0);
        }
        uint32_t f = 0, n = 0;
        const uint8_t* gin = nullptr;
        if (live) {
            f = P.frag_list ? P.frag_list[idx] : idx;
            const uint32_t m = P.frag_msg ? P.frag_msg[f] : f;
            const uint32_t k = P.frag_first ? f - P.frag_first[m] : 0u;
            n = min(kFragBytes, P.len[m] - k * kFragBytes);
            gin = P.in + P.off[m] + (size_t)k * kFragBytes;
        }
        uint8_t* in_b = sm.in + (size_t)slot * kSmallInStride;
        if (live && n && (reinterpret_cast<uintptr_t>(gin) & 15) == 0) {  // one TMA bulk copy per body
            const uint32_t bytes = (n + 15u) & ~15u;
            mbar_arrive_expect_tx(&stage_bar, bytes);
            tma_bulk_g2s(in_b, gin, bytes, &stage_bar);
        } else {
            if (live)
                for (uint32_t b = 0; b < n; b += 4) *reinterpret_cast<uint32_t*>(in_b + b) = ld32u(gin + b);
            mbar_arrive(&stage_bar);
        }
        mbar_wait(&stage_bar, parity);
        __syncthreads();  // the zeroed tables (generic stores of other threads) must be visible too
        // ---- phase 1: find ----
        uint32_t* my_list = elists + ((size_t)blockIdx.x * 32 + slot) * kSmallMaxElems;
        uint32_t ne = 0;
        if (live) ne = small_find_body(in_b, n, sm.tlo + slot * 1024, sm.thi + slot * 256, my_list);
        __syncwarp();  // the lists written by this warp's lanes are read by all of its lanes below
        // ---- phase 2: encode, one body at a time by the whole warp ----
        uint8_t* my_out = live ? P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride) : nullptr;
#pragma unroll 1
        for (uint32_t j = 0; j < (uint32_t)LPW; j++) {
            if (!__shfl_sync(0xffffffffu, (int)live, j)) continue;
            const uint32_t nej = __shfl_sync(0xffffffffu, ne, j), nj = __shfl_sync(0xffffffffu, n, j),
                           fj = __shfl_sync(0xffffffffu, f, j);
            const uint64_t oa = reinterpret_cast<uint64_t>(my_out);
            uint8_t* out = reinterpret_cast<uint8_t*>(((uint64_t)__shfl_sync(0xffffffffu, (uint32_t)(oa >> 32), j) << 32) |
                                                      __shfl_sync(0xffffffffu, (uint32_t)oa, j));
            const InShared in{sm.in + (size_t)(warp * LPW + j) * kSmallInStride};
            const uint32_t* el = elists + ((size_t)blockIdx.x * 32 + warp * LPW + j) * kSmallMaxElems;
            uint32_t op = 0, next_emit = 0;
            for (uint32_t base = 0; base < nej; base += 32) {
                const bool has = base + lane < nej;
                const uint32_t e = has ? __ldcg(el + base + lane) : 0u;
                const uint32_t p = e & 1023u, off = (e >> 10) & 1023u, mm = e >> 20;
                const uint32_t end = p + mm;
                uint32_t ls = __shfl_up_sync(0xffffffffu, end, 1);
                if (lane == 0) ls = next_emit;
                next_emit = __shfl_sync(0xffffffffu, end, min(31u, nej - base - 1));
                op = emit_elements_warp(in, has, ls, p, off, mm, out, op, lane);
            }
            op = emit_remainder_warp(in, next_emit, nj, out, op, lane);
            if (lane == 0) P.frag_clen[fj] = op;
        }
    }
}

// ---------------------------------------------------------------------------
// decompress: one lane per stream, output capacity <= 1 KiB
// ---------------------------------------------------------------------------
struct SmallDecompressParams {
    const uint8_t* in;
    const uint64_t* in_off;
    const uint32_t* in_len;
    const uint64_t* out_off;
    const uint32_t* out_cap;
    const uint32_t* list;
    uint8_t* out;
    uint32_t* ulen;
    int32_t* status;
    uint32_t n_list;
};

// Each slot owns a contiguous, byte-addressable output buffer of kSmallSlotBytes in shared memory
// (stride 1040 B: at most LPW lanes of a warp touch it at once, so bank collisions are rare and
// the code needs no interleaved-address arithmetic).  The lane decodes tag by tag with plain byte
// loops — the copy loop IS IncrementalCopy (snappy.cc:98-103), overlap semantics included — and
// when all lanes of the warp are done the warp flushes every slot with coalesced 16-byte stores.
// Back-reference copy inside a slot's shared-memory buffer.  With offset >= 4 the four source bytes
// of a round all precede its first destination byte, so they are loaded together (one shared-memory
// round trip per 4 bytes instead of one per byte); shorter offsets replicate byte by byte exactly as
// IncrementalCopy does (snappy.cc:98-103).
__device__ __forceinline__ void slot_copy(uint8_t* d, const uint8_t* s, uint32_t len, uint32_t offset) {
    uint32_t i = 0;
    if (offset >= 4) {
        for (; i + 4 <= len; i += 4) {  // the 4 source bytes of a round precede its first destination byte
            const uint8_t b0 = s[i], b1 = s[i + 1], b2 = s[i + 2], b3 = s[i + 3];
            d[i] = b0;
            d[i + 1] = b1;
            d[i + 2] = b2;
            d[i + 3] = b3;
        }
    }
    for (; i < len; i++) d[i] = s[i];
}

constexpr uint32_t kSmallSlotBytes = kSmallMax + 16;

template <int LPW>
__global__ void __launch_bounds__(1024 / LPW, 6) snappy_decompress_small_kernel(SmallDecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t slot = warp * LPW + lane;  // meaningful for lane < LPW
    const uint32_t idx = blockIdx.x * 32 + slot;
    const bool live = lane < LPW && idx < P.n_list;
    uint32_t m = 0, expected = 0;
    int st = B2C_OK;
    if (live) {
        m = P.list ? P.list[idx] : idx;
        const uint8_t* in = P.in + P.in_off[m];
        asm volatile("" : "+l"(in));  // one 64-bit register pair, not base + offset re-added on every tag
        const uint32_t n = P.in_len[m];
        const uint32_t cap = P.out_cap[m];
        uint8_t* ob = smem_raw + (size_t)slot * kSmallSlotBytes;
        // lenient preamble, snappy.cc:692-711
        uint32_t ip = 0;
        {
            uint32_t shift = 0;
            for (;;) {
                if (shift >= 32 || ip >= n) {
                    st = B2C_CORRUPT;
                    break;
                }
                const uint32_t c = in[ip++];
                expected |= (c & 0x7f) << shift;
                if (c < 128) break;
                shift += 7;
            }
        }
        if (st == B2C_OK && expected > cap) st = B2C_TOO_SMALL;
        uint32_t op = 0;
        if (st == B2C_OK) {
            // One code path for every tag: the three frequent kinds (short literal, copy with 1-byte offset, copy
            // with 2-byte offset) decode with selects, and a literal is then simply a copy whose source is the
            // stream instead of the slot.  The lanes of a warp decode different streams; separate branches per kind
            // made them take turns (a literal path run for 1.9 of 8 lanes on `words`).
            const uint8_t* obg = ob;  // generic pointer to the slot, for the common source pointer
            asm volatile("" : "+l"(obg));  // converted once, kept
            uint32_t x = ip < n ? ld32u(in + ip) : 0u;  // tag byte + up to 3 trailer bytes
            for (;;) {
                if (ip == n) {  // clean eof: snappy.cc:867
                    if (op != expected) st = B2C_CORRUPT;
                    break;
                }
                const uint32_t c = x & 0xff;
                const uint32_t kind = c & 3;
                const bool is_lit = kind == 0;
                uint32_t len = (c >> 2) + 1;                            // literal < 61, copy-2, copy-4
                uint32_t used = kind + 1;                               // 1, 2, 3 (copy-4: fixed below)
                uint32_t offset = (x >> 8) & 0xffffu;                   // copy-2
                if (kind == 1) {
                    len = ((c >> 2) & 7) + 4;
                    offset = ((c >> 5) << 8) | ((x >> 8) & 0xff);
                }
                if (kind == 3 || (kind == 0 && c >= 0xf0u)) {  // rare: literal with a length trailer, 4-byte offset
                    if (kind == 0) {
                        const uint32_t extra = len - 60;
                        if (n - ip < extra + 1) { st = B2C_CORRUPT; break; }
                        uint32_t v = x >> 8;
                        if (extra == 4) v |= (uint32_t)in[ip + 4] << 24;
                        else v &= (1u << (8 * extra)) - 1u;
                        len = v + 1u;  // may wrap to 0 like the reference
                        used = 1 + extra;
                    } else {
                        if (n - ip < 5) { st = B2C_CORRUPT; break; }
                        offset = (x >> 8) | ((uint32_t)in[ip + 4] << 24);
                        used = 5;
                    }
                }
                if (n - ip < used) { st = B2C_CORRUPT; break; }
                ip += used;
                // literal: the bytes must be there; copy: offset in 1..op (offset - 1 wraps for 0)
                const bool bad = is_lit ? n - ip < len : offset - 1 >= op;
                if (bad || expected - op < len) { st = B2C_CORRUPT; break; }
                const uint8_t* sp = is_lit ? in + ip : obg + op - offset;  // may overlap d: no restrict
                uint8_t* d = ob + op;
                // the next tag's bytes are asked for now: they arrive while this element's bytes are moved
                const uint32_t nip = is_lit ? ip + len : ip;
                const uint32_t xn = nip < n ? ld32u(in + nip) : 0u;
                uint32_t i = 0;
                if (is_lit || offset >= 4) {  // the 4 source bytes of a round precede its first destination byte
                    for (; i + 4 <= len; i += 4) {
                        const uint8_t b0 = sp[i], b1 = sp[i + 1], b2 = sp[i + 2], b3 = sp[i + 3];
                        d[i] = b0;
                        d[i + 1] = b1;
                        d[i + 2] = b2;
                        d[i + 3] = b3;
                    }
                    const uint32_t r = len - i;  // 0-3 bytes left: loaded together, one wait instead of one per byte
                    if (r) {
                        const uint8_t b0 = sp[i], b1 = r > 1 ? sp[i + 1] : (uint8_t)0, b2 = r > 2 ? sp[i + 2] : (uint8_t)0;
                        d[i] = b0;
                        if (r > 1) d[i + 1] = b1;
                        if (r > 2) d[i + 2] = b2;
                    }
                } else {
                    for (; i < len; i++) d[i] = sp[i];  // IncrementalCopy, snappy.cc:98-103
                }
                op += len;
                ip = nip;
                x = xn;
            }
        }
        P.status[m] = st;
        if (P.ulen) P.ulen[m] = expected;
    }
    __syncwarp();
    // warp-cooperative flush of this warp's slots
    const uint32_t flush_len = live && st == B2C_OK ? expected : 0u;
    const uint64_t gaddr = live ? reinterpret_cast<uint64_t>(P.out + P.out_off[m]) : 0ull;
#pragma unroll
    for (int j = 0; j < LPW; j++) {
        const uint32_t len = __shfl_sync(0xffffffffu, flush_len, j);
        if (len == 0) continue;
        const uint64_t ga = ((uint64_t)__shfl_sync(0xffffffffu, (uint32_t)(gaddr >> 32), j) << 32) |
                            __shfl_sync(0xffffffffu, (uint32_t)gaddr, j);
        uint8_t* g = reinterpret_cast<uint8_t*>(ga);
        const uint8_t* sb = smem_raw + (size_t)(warp * LPW + j) * kSmallSlotBytes;
        if ((ga & 15) == 0) {
            const uint32_t vecs = len >> 4;
            for (uint32_t i = lane; i < vecs; i += 32)
                reinterpret_cast<uint4*>(g)[i] = reinterpret_cast<const uint4*>(sb)[i];
            for (uint32_t i = (vecs << 4) + lane; i < len; i += 32) g[i] = sb[i];
        } else {
            for (uint32_t i = lane; i < len; i += 32) g[i] = sb[i];
        }
    }
}

constexpr size_t kSmallDecompressSmem = (size_t)kSmallSlotBytes * 32;

// ---------------------------------------------------------------------------
template <int LPW>
static int configure_one(b2c_ctx* ctx) {
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_small2_kernel<LPW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(SmallCompressSmem)));
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_decompress_small_kernel<LPW>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallDecompressSmem));
    return B2C_OK;
}

int small_kernels_configure(b2c_ctx* ctx) {
    int rc;
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_small2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(SmallCompressSmem)));
    if ((rc = configure_one<4>(ctx))) return rc;
    if ((rc = configure_one<8>(ctx))) return rc;
    if ((rc = configure_one<16>(ctx))) return rc;
    return configure_one<32>(ctx);
}

int launch_compress_small(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* scratch,
                          uint32_t* d_frag_clen) {
    SmallCompressParams P;
    P.in = d_in;
    P.off = L.d_off;
    P.len = L.d_len;
    P.frag_list = L.d_class_list[0];
    P.frag_msg = L.d_frag_msg;
    P.frag_first = L.d_frag_first;
    P.frag_slot = L.d_frag_slot;
    P.slot_stride = L.slot_stride;
    P.scratch = scratch;
    P.frag_clen = d_frag_clen;
    P.n_list = L.n_class[0];
    const uint32_t blocks = (P.n_list + 31) / 32;
    const size_t smem = sizeof(SmallCompressSmem);
    {  // find, then encode; persistent CTAs own their element lists
        const uint32_t grid = std::min<uint32_t>(blocks, (uint32_t)ctx->num_sms * 3);
        int rc;
        if ((rc = ws_reserve(ctx, lane, lane.ws_tab, (size_t)grid * 32 * kSmallMaxElems * 4))) return rc;
        uint32_t* el = static_cast<uint32_t*>(lane.ws_tab.p);
        switch (ctx->small_lpw_compress) {
            case 2: snappy_compress_small2_kernel<2><<<grid, 512, smem, lane.stream>>>(P, el); break;
            case 4: snappy_compress_small2_kernel<4><<<grid, 256, smem, lane.stream>>>(P, el); break;
            case 16: snappy_compress_small2_kernel<16><<<grid, 64, smem, lane.stream>>>(P, el); break;
            case 32: snappy_compress_small2_kernel<32><<<grid, 32, smem, lane.stream>>>(P, el); break;
            default: snappy_compress_small2_kernel<8><<<grid, 128, smem, lane.stream>>>(P, el); break;
        }
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    return B2C_OK;
}

int launch_decompress_small(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in, const LayoutView& out,
                            uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status) {
    SmallDecompressParams P;
    P.in = d_in;
    P.in_off = in.d_off;
    P.in_len = in.d_len;
    P.out_off = out.d_off;
    P.out_cap = out.d_len;
    P.list = out.d_dclass_list[0];
    P.out = d_out;
    P.ulen = d_ulen;
    P.status = d_status;
    P.n_list = out.n_dclass[0];
    const uint32_t blocks = (P.n_list + 31) / 32;
    const size_t smem = kSmallDecompressSmem;
    switch (ctx->small_lpw_decompress) {
        case 4: snappy_decompress_small_kernel<4><<<blocks, 256, smem, lane.stream>>>(P); break;
        case 16: snappy_decompress_small_kernel<16><<<blocks, 64, smem, lane.stream>>>(P); break;
        case 32: snappy_decompress_small_kernel<32><<<blocks, 32, smem, lane.stream>>>(P); break;
        default: snappy_decompress_small_kernel<8><<<blocks, 128, smem, lane.stream>>>(P); break;
    }
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

}  // namespace b2c
