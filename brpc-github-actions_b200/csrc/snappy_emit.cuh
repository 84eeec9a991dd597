// snappy_emit.cuh — warp-cooperative encoding of snappy elements (EmitLiteral / EmitCopy,
// reference snappy.cc:156-232) shared by the compressors that separate FINDING matches from
// WRITING them: one lane per element, placed by a warp prefix sum over the encoded sizes.
#pragma once

#include "b2c_internal.cuh"

namespace b2c {

static constexpr uint32_t kFullMask = 0xffffffffu;
// long literal moves noted for later (emit_elements_warp / emit_remainder_warp): list[0].x = how many, list[1..]
static constexpr uint32_t kDeferMax = 6, kDeferMin = 512, kDeferBytes = (kDeferMax + 2) * 16;

__device__ __forceinline__ uint32_t hash_w(uint32_t bytes, uint32_t shift) {  // snappy.cc:47-50
    return (bytes * 0x1e35a7bdu) >> shift;
}
__device__ __forceinline__ uint32_t prefix8(uint64_t a, uint64_t b) {  // equal leading bytes, 0..8
    const uint64_t x = a ^ b;
    return x ? (uint32_t)(__ffsll((long long)x) - 1) >> 3 : 8u;
}

// input policies -------------------------------------------------------------
struct InShared {  // fragment staged in shared memory, 4-byte aligned base
    const uint8_t* base;
    __device__ __forceinline__ uint64_t ld64(uint32_t p) const {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(base + (p & ~3u));
        const uint32_t a = w[0], b = w[1], c = w[2], sh = (p & 3) * 8;
        return ((uint64_t)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
    }
    __device__ __forceinline__ uint32_t ld32(uint32_t p) const {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(base + (p & ~3u));
        return __funnelshift_r(w[0], w[1], (p & 3) * 8);
    }
    __device__ __forceinline__ uint8_t ld8(uint32_t p) const { return base[p]; }
    static constexpr bool kContiguous = true;
    __device__ __forceinline__ const uint8_t* addr(uint32_t p) const { return base + p; }
};
struct InGlobal {  // fragment read from HBM through L1
    const uint8_t* base;
    __device__ __forceinline__ uint64_t ld64(uint32_t p) const {
        const uintptr_t a = reinterpret_cast<uintptr_t>(base + p);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
        const uint32_t sh = (uint32_t)(a & 3) * 8;
        const uint32_t x = w[0], y = w[1], z = sh ? w[2] : 0u;
        return ((uint64_t)__funnelshift_r(y, z, sh) << 32) | __funnelshift_r(x, y, sh);
    }
    __device__ __forceinline__ uint32_t ld32(uint32_t p) const { return ld32u(base + p); }
    __device__ __forceinline__ uint8_t ld8(uint32_t p) const { return base[p]; }
    static constexpr bool kContiguous = true;
    __device__ __forceinline__ const uint8_t* addr(uint32_t p) const { return base + p; }
};

// literal bytes [from, from + n) of the input to d, by the whole warp
template <class In>
__device__ __forceinline__ void copy_literal_warp(const In& in, uint32_t from, uint8_t* __restrict__ d, uint32_t n,
                                                  uint32_t lane) {
    if constexpr (In::kContiguous) {
        warp_copy_bytes(d, in.addr(from), n, lane);  // 16-byte stores: an incompressible fragment is one 64 KiB literal
    } else {
        for (uint32_t i = lane; i < n; i += 32) d[i] = in.ld8(from + i);
    }
}

__device__ __forceinline__ uint32_t literal_header_bytes(uint32_t len) {  // EmitLiteral, snappy.cc:156-196
    const uint32_t n1 = len - 1;
    return n1 < 60 ? 1u : n1 < (1u << 8) ? 2u : n1 < (1u << 16) ? 3u : n1 < (1u << 24) ? 4u : 5u;
}
__device__ __forceinline__ uint32_t copy_bytes(uint32_t offset, uint32_t len) {  // EmitCopy, snappy.cc:198-232
    uint32_t w = 0;
    if (len >= 68) {
        const uint32_t n64 = (len - 68) / 64 + 1;
        w = 3 * n64;
        len -= 64 * n64;
    }
    if (len > 64) {
        w += 3;
        len -= 60;
    }
    return w + ((len < 12 && offset < 2048) ? 2u : 3u);
}
__device__ __forceinline__ void write_literal_header(uint8_t* o, uint32_t len) {
    const uint32_t n1 = len - 1;
    if (n1 < 60) {
        o[0] = (uint8_t)(n1 << 2);
    } else {
        const uint32_t count = literal_header_bytes(len) - 1;
        o[0] = (uint8_t)((59 + count) << 2);
        for (uint32_t i = 0; i < count; i++) o[1 + i] = (uint8_t)(n1 >> (8 * i));
    }
}
// the final (< 68 bytes) pieces of a copy: an optional 60-byte piece and the rest
__device__ __forceinline__ void write_copy_tail(uint8_t* o, uint32_t offset, uint32_t len) {
    if (len > 64) {
        o[0] = (uint8_t)(2 + (59u << 2));
        o[1] = (uint8_t)(offset & 0xff);
        o[2] = (uint8_t)(offset >> 8);
        o += 3;
        len -= 60;
    }
    if (len < 12 && offset < 2048) {
        o[0] = (uint8_t)(1 + ((len - 4) << 2) + ((offset >> 8) << 5));
        o[1] = (uint8_t)(offset & 0xff);
    } else {
        o[0] = (uint8_t)(2 + ((len - 1) << 2));
        o[1] = (uint8_t)(offset & 0xff);
        o[2] = (uint8_t)(offset >> 8);
    }
}

// Encode up to 32 elements at once.  Lane e (when `has`) holds one element: the literal
// [ls, p) followed by a copy of `mm` bytes from `off` bytes back.  Appends at out + op, returns the new op.
template <class In>
__device__ __forceinline__ uint32_t emit_elements_warp(const In& in, const bool has, const uint32_t e_ls, const uint32_t e_p,
                                                       const uint32_t off, const uint32_t mm, uint8_t* __restrict__ out,
                                                       uint32_t op, const uint32_t lane, uint4* defer = nullptr) {
    const uint32_t ll = has ? e_p - e_ls : 0u;  // literal bytes before the copy
    const uint32_t lit_sz = ll ? literal_header_bytes(ll) + ll : 0u;
    const uint32_t size = has ? lit_sz + copy_bytes(off, mm) : 0u;
    uint32_t incl = size;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(kFullMask, incl, d);
        if (lane >= (uint32_t)d) incl += y;
    }
    uint8_t* o = out + op + incl - size;
    op += __shfl_sync(kFullMask, incl, 31);
    if (has && ll) {
        write_literal_header(o, ll);
        if (ll <= 16) {
            uint8_t* d = o + lit_sz - ll;
            for (uint32_t i = 0; i < ll; i++) d[i] = in.ld8(e_ls + i);
        }
    }
    // long literals and long copies: the whole warp helps, one element at a time
    uint32_t longs = __ballot_sync(kFullMask, has && ll > 16);
    while (longs) {
        const uint32_t l = __ffs(longs) - 1;
        longs &= longs - 1;
        const uint32_t s0 = __shfl_sync(kFullMask, e_ls, l), ln = __shfl_sync(kFullMask, ll, l);
        const uint64_t oa = reinterpret_cast<uint64_t>(o + lit_sz - ll);
        uint8_t* d = reinterpret_cast<uint8_t*>(((uint64_t)__shfl_sync(kFullMask, (uint32_t)(oa >> 32), l) << 32) |
                                                __shfl_sync(kFullMask, (uint32_t)oa, l));
        // Kilobytes of literal in front of a rare match (incompressible data): the vector copy costs registers the
        // orbit kernel has not got inside its loop, so the move is only NOTED (destination, source, length) and made
        // by emit_remainder_warp, which has that copy anyway.  The input does not change and the place is reserved.
        if (In::kContiguous && defer && ln >= kDeferMin && defer[0].x < kDeferMax) {
            __syncwarp();
            if (lane == 0) {
                const uint32_t k = ++defer[0].x;
                const uint64_t da = reinterpret_cast<uint64_t>(d);
                defer[k] = make_uint4((uint32_t)da, (uint32_t)(da >> 32), s0, ln);
            }
            __syncwarp();
            continue;
        }
        for (uint32_t i = lane; i < ln; i += 32) d[i] = in.ld8(s0 + i);
    }
    if (has && mm < 68) write_copy_tail(o + lit_sz, off, mm);
    uint32_t lcopies = __ballot_sync(kFullMask, has && mm >= 68);
    while (lcopies) {
        const uint32_t l = __ffs(lcopies) - 1;
        lcopies &= lcopies - 1;
        const uint32_t of = __shfl_sync(kFullMask, off, l), ln = __shfl_sync(kFullMask, mm, l);
        const uint64_t oa = reinterpret_cast<uint64_t>(o + lit_sz);
        uint8_t* d = reinterpret_cast<uint8_t*>(((uint64_t)__shfl_sync(kFullMask, (uint32_t)(oa >> 32), l) << 32) |
                                                __shfl_sync(kFullMask, (uint32_t)oa, l));
        const uint32_t n64 = (ln - 68) / 64 + 1;
        for (uint32_t k = lane; k < n64; k += 32) {
            d[3 * k] = (uint8_t)(2 + (63u << 2));
            d[3 * k + 1] = (uint8_t)(of & 0xff);
            d[3 * k + 2] = (uint8_t)(of >> 8);
        }
        if (lane == 0) write_copy_tail(d + 3 * n64, of, ln - 64 * n64);
    }
    return op;
}

// emit_remainder (snappy.cc:446-450): the final literal [from, n), written by the whole warp.
template <class In>
__device__ __forceinline__ uint32_t emit_remainder_warp(const In& in, const uint32_t from, const uint32_t n,
                                                        uint8_t* __restrict__ out, uint32_t op, const uint32_t lane,
                                                        uint4* defer = nullptr) {
    uint32_t todo = 0;  // noted moves, the final literal among them: ONE copy loop serves all
    if (defer) {
        __syncwarp();
        todo = defer[0].x;
    }
    if (from < n) {
        const uint32_t ll = n - from;
        uint8_t* o = out + op;
        if (lane == 0) write_literal_header(o, ll);
        uint8_t* d = o + literal_header_bytes(ll);
        op += literal_header_bytes(ll) + ll;
        if (defer) {
            if (lane == 0) {
                const uint64_t da = reinterpret_cast<uint64_t>(d);
                defer[todo + 1] = make_uint4((uint32_t)da, (uint32_t)(da >> 32), from, ll);
            }
            todo++;
            __syncwarp();
        } else {
            copy_literal_warp(in, from, d, ll, lane);
        }
    }
    for (uint32_t k = 1; k <= todo; k++) {
        const uint4 e = defer[k];
        copy_literal_warp(in, e.z, reinterpret_cast<uint8_t*>(((uint64_t)e.y << 32) | e.x), e.w, lane);
    }
    if (defer) {
        __syncwarp();
        if (lane == 0) defer[0].x = 0;
        __syncwarp();
    }
    return op;
}

}  // namespace b2c
