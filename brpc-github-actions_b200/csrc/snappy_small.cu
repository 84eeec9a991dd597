// snappy_small.cu — thread-per-message snappy kernels for bodies of at most 1 KiB
// (BASELINE configs[1]: 2^20 x 1 KiB RPC bodies).
//
// The warp-per-fragment kernels (snappy_compress.cu / snappy_decompress.cu) spend a
// whole warp's issue slots on one serial chain; ncu showed them issue-bound (88-92 %
// issue-active, 25 k / 15 k warp-instructions per 1 KiB body).  For small bodies the
// better shape is ONE LANE PER BODY: every lane runs the reference's scalar algorithm
// (CompressFragment, snappy.cc:329-453 / DecompressAllTags, snappy.cc:716-787) on its
// own body, so one warp instruction advances 32 bodies.
//
// Shared memory is laid out so that lanes never collide on a bank:
//   compress:   u16 hash table [entry][lane] (64 KiB per warp) + the input as
//               32-bit words [word][lane] (33 KiB per warp); both CTAs of an SM = 64 bodies in flight;
//   decompress: the output as words [word][lane] (32 KiB per warp, 6 warps per SM),
//               so back-references resolve at shared-memory latency.
// Output bytes of the compressor are funnelled through a 64-bit register and stored
// as aligned words.
#include "b2c_internal.cuh"

namespace b2c {

constexpr uint32_t kSmallMax = 1024;
constexpr uint32_t kSmallInWords = kSmallMax / 4 + 4;  // spare words: unaligned reads past the end stay in-bounds

// A CTA always owns 32 bodies ("slots"); LPW of each warp's lanes carry one body each, so a
// CTA has 32/LPW warps.  Fewer bodies per warp = more warps for the same shared memory: the
// serial chains of different bodies then interleave in the schedulers instead of being
// executed in lock-step (ncu on LPW=32: 10 of 32 lanes active, 0.2 IPC per warp, 2 warps/SM).
struct SmallCompressSmem {
    uint16_t table[1024 * 32];           // [entry][slot]
    uint32_t in[kSmallInWords * 32];     // [word][slot]
};

__device__ __forceinline__ uint32_t hash_bytes_s(uint32_t bytes, uint32_t shift) {  // snappy.cc:47-50
    return (bytes * 0x1e35a7bdu) >> shift;
}

// Output funnel: bytes are appended above `nacc` pending bytes; full words leave at the bottom.
struct OutFunnel {
    uint32_t* dst;   // 4-byte aligned
    uint32_t acc;
    uint32_t nacc;   // pending bytes (< 4 between pushes)
    uint32_t words;  // words stored so far
    __device__ __forceinline__ void push(uint32_t v, uint32_t k) {  // k in 1..4, v has no bits above 8k
        const uint32_t sh = 8 * nacc;
        const uint32_t lo = acc | (v << sh);
        const uint32_t hi = __funnelshift_l(v, 0u, sh);  // v >> (32 - sh), 0 when sh == 0
        nacc += k;
        if (nacc >= 4) {
            dst[words++] = lo;
            acc = hi;
            nacc -= 4;
        } else {
            acc = lo;
        }
    }
    __device__ __forceinline__ uint32_t finish() {
        if (nacc) dst[words] = acc;  // slot has MaxCompressedLength's 32 spare bytes
        return words * 4 + nacc;
    }
};

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

template <int LPW>
__global__ void __launch_bounds__(1024 / LPW) snappy_compress_small_kernel(SmallCompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    SmallCompressSmem& sm = *reinterpret_cast<SmallCompressSmem*>(smem_raw);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t slot = (threadIdx.x >> 5) * LPW + lane;  // meaningful for lane < LPW
    const uint32_t idx = blockIdx.x * 32 + slot;
    const bool live = lane < LPW && idx < P.n_list;
    // zero all 32 tables together (GetHashTable's memset, snappy.cc:269)
    {
        uint4* t4 = reinterpret_cast<uint4*>(sm.table);
        for (uint32_t i = threadIdx.x; i < sizeof(sm.table) / 16; i += blockDim.x) t4[i] = make_uint4(0, 0, 0, 0);
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
    // stage this lane's body into its word column
    uint32_t* in_s = sm.in + slot;
    if (live) {
        if ((reinterpret_cast<uintptr_t>(gin) & 15) == 0) {
            for (uint32_t b = 0; b < n; b += 16) {
                const uint4 v = *reinterpret_cast<const uint4*>(gin + b);
                const uint32_t w = b >> 2;
                in_s[(w + 0) * 32] = v.x;
                in_s[(w + 1) * 32] = v.y;
                in_s[(w + 2) * 32] = v.z;
                in_s[(w + 3) * 32] = v.w;
            }
        } else {
            for (uint32_t b = 0; b < n; b += 4) in_s[(b >> 2) * 32] = ld32u(gin + b);
        }
    }
    __syncthreads();
    if (!live) return;
    // 8 bytes starting at byte p of this body
    auto ld64s = [&](uint32_t p) -> uint64_t {
        const uint32_t* w = in_s + (p >> 2) * 32;
        const uint32_t a = w[0], b = w[32], c = w[64], sh = (p & 3) * 8;
        return ((uint64_t)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
    };
    auto ld32s = [&](uint32_t p) -> uint32_t {
        const uint32_t* w = in_s + (p >> 2) * 32;
        return __funnelshift_r(w[0], w[32], (p & 3) * 8);
    };
    uint16_t* tab = sm.table + slot;
    uint32_t tsize = 256;  // GetHashTable sizing, snappy.cc:253-256
    while (tsize < 16384u && tsize < n) tsize <<= 1;
    const uint32_t shift = 32 - (31 - __clz(tsize));

    OutFunnel o;
    o.dst = reinterpret_cast<uint32_t*>(P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride));
    o.acc = 0;
    o.nacc = 0;
    o.words = 0;

    auto emit_literal = [&](uint32_t from, uint32_t len) {  // EmitLiteral, snappy.cc:156-196
        const uint32_t n1 = len - 1;
        if (n1 < 60) {
            o.push(n1 << 2, 1);
        } else if (n1 < 256) {
            o.push((60u << 2) | (n1 << 8), 2);
        } else {
            o.push((61u << 2) | (n1 << 8), 3);
        }
        uint32_t i = 0;
        for (; i + 4 <= len; i += 4) o.push(ld32s(from + i), 4);
        if (i < len) o.push(ld32s(from + i) & ((1u << (8 * (len - i))) - 1u), len - i);
    };

    uint32_t next_emit = 0;
    if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
        const uint32_t ip_limit = n - 15;
        uint32_t ip = 1, skip = 32, nxt = 0;
        bool probing = true, have_d = false;
        uint64_t d = 0;  // bytes ip..ip+7 when have_d
        // One loop for both the probe loop (snappy.cc:377-397) and the copy loop's "is there
        // another match right here" test (snappy.cc:430-439): they perform the same table step.
        for (;;) {
            if (probing) {
                nxt = ip + (skip++ >> 5);
                if (nxt > ip_limit) break;  // snappy.cc:386-388
            }
            if (!have_d) d = ld64s(ip);
            have_d = false;
            const uint32_t cur = (uint32_t)d;
            const uint32_t h = hash_bytes_s(cur, shift);
            const uint32_t cand = tab[h * 32];
            tab[h * 32] = (uint16_t)ip;
            const uint64_t x = ld64s(cand) ^ d;
            if ((uint32_t)x == 0) {
                if (ip > next_emit) emit_literal(next_emit, ip - next_emit);  // snappy.cc:403
                // FindMatchLength, snappy-internal.h:86-120.  ip + 15 < n here, so the first 8 bytes are in range.
                uint32_t matched;
                const uint32_t xh = (uint32_t)(x >> 32);
                if (xh) {
                    matched = 4 + ((uint32_t)(__ffs(xh) - 1) >> 3);
                } else {
                    matched = 8;
                    for (;;) {
                        const uint32_t rem = n - (ip + matched);
                        if (rem == 0) break;
                        const uint64_t y = ld64s(cand + matched) ^ ld64s(ip + matched);
                        uint32_t nb = y ? (uint32_t)(__ffsll((long long)y) - 1) >> 3 : 8u;
                        nb = min(nb, rem);
                        matched += nb;
                        if (nb < 8) break;
                    }
                }
                // EmitCopy, snappy.cc:198-232
                {
                    const uint32_t offset = ip - cand;
                    uint32_t len = matched;
                    while (len >= 68) {
                        o.push(0xfeu | (offset << 8), 3);
                        len -= 64;
                    }
                    if (len > 64) {
                        o.push((2u + (59u << 2)) | (offset << 8), 3);
                        len -= 60;
                    }
                    if (len < 12 && offset < 2048)
                        o.push((1u + ((len - 4) << 2) + ((offset >> 8) << 5)) | ((offset & 0xff) << 8), 2);
                    else
                        o.push((2u + ((len - 1) << 2)) | (offset << 8), 3);
                }
                ip += matched;
                next_emit = ip;
                if (ip >= ip_limit) break;  // snappy.cc:427-429
                // snappy.cc:430-433: insert ip-1; one 4-word read yields both load32(ip-1) and bytes ip..ip+7
                {
                    const uint32_t q = ip - 1;
                    const uint32_t* w = in_s + (q >> 2) * 32;
                    const uint32_t a = w[0], b = w[32], c = w[64], e = w[96], sh = (q & 3) * 8;
                    const uint32_t p0 = __funnelshift_r(a, b, sh), p1 = __funnelshift_r(b, c, sh),
                                   p2 = __funnelshift_r(c, e, sh);  // bytes q..q+11
                    tab[hash_bytes_s(p0, shift) * 32] = (uint16_t)q;
                    d = ((uint64_t)__funnelshift_r(p1, p2, 8) << 32) | __funnelshift_r(p0, p1, 8);
                    have_d = true;
                }
                probing = false;
            } else if (probing) {
                ip = nxt;
            } else {  // snappy.cc:441-442: leave the copy loop, resume probing one byte on
                ip += 1;
                probing = true;
                skip = 32;
            }
        }
    }
    if (next_emit < n) emit_literal(next_emit, n - next_emit);  // emit_remainder, snappy.cc:446-450
    P.frag_clen[f] = o.finish();
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

template <int LPW>
__global__ void __launch_bounds__(1024 / LPW) snappy_decompress_small_kernel(SmallDecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t slot = (threadIdx.x >> 5) * LPW + lane;
    const uint32_t idx = blockIdx.x * 32 + slot;
    if (lane >= LPW || idx >= P.n_list) return;
    const uint32_t m = P.list ? P.list[idx] : idx;
    const uint8_t* in = P.in + P.in_off[m];
    const uint32_t n = P.in_len[m];
    const uint32_t cap = P.out_cap[m];
    // this slot's output column: word w lives at col[w * 32]
    uint32_t* col = reinterpret_cast<uint32_t*>(smem_raw) + slot;

    // ---- input funnel: `have` upcoming stream bytes sit in `bits`, refilled one aligned word at a time
    const uint32_t* inext = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(in) & ~uintptr_t(3));
    const uint32_t* const iend =
        reinterpret_cast<const uint32_t*>((reinterpret_cast<uintptr_t>(in) + n + 3) & ~uintptr_t(3));
    uint64_t bits = 0;
    uint32_t have = 0;
    auto refill = [&]() {  // afterwards have >= 5 (or the stream is exhausted and zeros follow)
        while (have <= 4) {
            const uint32_t w = inext < iend ? *inext : 0u;
            inext++;
            bits |= (uint64_t)w << (8 * have);
            have += 4;
        }
    };
    auto consume = [&](uint32_t k) {
        bits >>= 8 * k;
        have -= k;
    };
    {
        const uint32_t lead = (uint32_t)(reinterpret_cast<uintptr_t>(in) & 3);
        refill();
        consume(lead);
        refill();
    }
    // ---- output funnel: the pending partial word is mirrored in shared memory after every push,
    //      so back-references can always read produced bytes from the column
    uint32_t oacc = 0, onacc = 0, owords = 0;
    auto push = [&](uint32_t v, uint32_t k) {  // k in 1..4, v has no bits above 8k
        const uint32_t sh = 8 * onacc;
        const uint32_t lo = oacc | (v << sh);
        const uint32_t hi = __funnelshift_l(v, 0u, sh);
        onacc += k;
        col[owords * 32] = lo;
        if (onacc >= 4) {
            owords++;
            onacc -= 4;
            oacc = hi;
            if (onacc) col[owords * 32] = hi;
        } else {
            oacc = lo;
        }
    };
    auto lds32u = [&](uint32_t p) -> uint32_t {  // 4 produced bytes starting at byte p
        const uint32_t* w = col + (p >> 2) * 32;
        return __funnelshift_r(w[0], w[32], (p & 3) * 8);
    };

    // lenient preamble, snappy.cc:692-711
    uint32_t expected = 0, ip = 0;
    int st = B2C_OK;
    {
        uint32_t shift = 0;
        for (;;) {
            if (shift >= 32 || ip >= n) {
                st = B2C_CORRUPT;
                break;
            }
            const uint32_t c = (uint32_t)bits & 0xff;
            consume(1);
            refill();
            ip++;
            expected |= (c & 0x7f) << shift;
            if (c < 128) break;
            shift += 7;
        }
    }
    if (st == B2C_OK && expected > cap) st = B2C_TOO_SMALL;
    uint32_t produced = 0;
    if (st == B2C_OK) {
        for (;;) {
            if (ip == n) {  // clean eof: snappy.cc:867
                if (produced != expected) st = B2C_CORRUPT;
                break;
            }
            const uint32_t c = (uint32_t)bits & 0xff;
            const uint32_t kind = c & 3;
            const uint32_t extra = kind == 0 ? (c >= 0xf0 ? (c >> 2) - 59 : 0) : kind == 1 ? 1 : kind == 2 ? 2 : 4;
            if (n - ip < extra + 1) {  // tag cut short
                st = B2C_CORRUPT;
                break;
            }
            uint32_t trailer = (uint32_t)(bits >> 8);
            if (extra < 4) trailer &= (1u << (8 * extra)) - 1u;
            consume(1 + extra);
            refill();
            ip += 1 + extra;
            if (kind == 0) {
                uint32_t len = (c >> 2) + 1;
                if (len >= 61) len = trailer + 1u;  // may wrap to 0 like the reference
                if (n - ip < len || expected - produced < len) {
                    st = B2C_CORRUPT;
                    break;
                }
                uint32_t i = 0;
                for (; i + 4 <= len; i += 4) {
                    push((uint32_t)bits, 4);
                    consume(4);
                    refill();
                }
                if (i < len) {
                    const uint32_t k = len - i;
                    push((uint32_t)bits & ((1u << (8 * k)) - 1u), k);
                    consume(k);
                    refill();
                }
                produced += len;
                ip += len;
            } else {
                uint32_t len, offset;
                if (kind == 1) {
                    len = ((c >> 2) & 7) + 4;
                    offset = ((c >> 5) << 8) | (trailer & 0xff);
                } else {
                    len = (c >> 2) + 1;
                    offset = trailer;
                }
                if (offset == 0 || offset > produced || expected - produced < len) {
                    st = B2C_CORRUPT;
                    break;
                }
                const uint32_t src = produced - offset;
                if (offset >= 4) {
                    // 4 bytes per round: the source bytes of a round were all produced before it
                    uint32_t i = 0;
                    for (; i + 4 <= len; i += 4) push(lds32u(src + i), 4);
                    if (i < len) push(lds32u(src + i) & ((1u << (8 * (len - i))) - 1u), len - i);
                } else {
                    for (uint32_t i = 0; i < len; i++) push(lds32u(src + i) & 0xffu, 1);  // IncrementalCopy, snappy.cc:98-103
                }
                produced += len;
            }
        }
    }
    P.status[m] = st;
    if (P.ulen) P.ulen[m] = expected;
    if (st != B2C_OK) return;
    // flush this slot's column
    uint8_t* gout = P.out + P.out_off[m];
    if ((reinterpret_cast<uintptr_t>(gout) & 15) == 0) {
        uint32_t b = 0;
        for (; b + 16 <= expected; b += 16) {
            const uint32_t w = b >> 2;
            *reinterpret_cast<uint4*>(gout + b) =
                make_uint4(col[(w + 0) * 32], col[(w + 1) * 32], col[(w + 2) * 32], col[(w + 3) * 32]);
        }
        for (; b < expected; b++) gout[b] = (uint8_t)(col[(b >> 2) * 32] >> (8 * (b & 3)));
    } else {
        for (uint32_t b = 0; b < expected; b++) gout[b] = (uint8_t)(col[(b >> 2) * 32] >> (8 * (b & 3)));
    }
}

constexpr size_t kSmallDecompressSmem = (size_t)(kSmallMax + 8) * 32;  // 32 columns of 258 words

// ---------------------------------------------------------------------------
template <int LPW>
static int configure_one(b2c_ctx* ctx) {
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_small_kernel<LPW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(SmallCompressSmem)));
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_decompress_small_kernel<LPW>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallDecompressSmem));
    return B2C_OK;
}

int small_kernels_configure(b2c_ctx* ctx) {
    int rc;
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
    switch (ctx->small_lpw_compress) {
        case 4: snappy_compress_small_kernel<4><<<blocks, 256, smem, lane.stream>>>(P); break;
        case 16: snappy_compress_small_kernel<16><<<blocks, 64, smem, lane.stream>>>(P); break;
        case 32: snappy_compress_small_kernel<32><<<blocks, 32, smem, lane.stream>>>(P); break;
        default: snappy_compress_small_kernel<8><<<blocks, 128, smem, lane.stream>>>(P); break;
    }
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
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
