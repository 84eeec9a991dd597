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
            const uint8_t* src = in + ip;
            uint8_t* dst = obuf + produced;
            for (uint32_t i = lane; i < len; i += 32) dst[i] = src[i];
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
};

__global__ void __launch_bounds__(256) snappy_decompress_kernel(DecompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * (blockDim.x >> 5) + wib;
    if (w >= P.n_list) return;
    const uint32_t m = P.list ? P.list[w] : w;
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
    } else if (P.smem_per_warp && expected <= P.smem_per_warp) {
        uint8_t* sbuf = smem_raw + (size_t)wib * P.smem_per_warp;
        st = decode_message_warp(in, n, k, sbuf, expected, lane);
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
        st = decode_message_warp(in, n, k, gout, expected, lane);
    }
    if (lane == 0) {
        P.status[m] = st;
        if (P.ulen) P.ulen[m] = expected;
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
        if (c == 0 && ctx->small_kernels) {  // slots <= 1 KiB: one lane per stream
            int rc = launch_decompress_small(ctx, lane, in, d_in, out, d_out, d_ulen, d_status);
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
        const uint32_t wpb = class_warps[c];
        const uint32_t blocks = (P.n_list + wpb - 1) / wpb;
        snappy_decompress_kernel<<<blocks, wpb * 32, (size_t)wpb * class_smem[c], lane.stream>>>(P);
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    prof_mark(ctx, lane, B2C_PROF_DECOMPRESS);
    return B2C_OK;
}

}  // namespace b2c
