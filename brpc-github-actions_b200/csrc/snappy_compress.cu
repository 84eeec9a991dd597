// snappy_compress.cu — byte-exact batch snappy compression for sm_100a.
//
// Replaces butil::snappy::Compress(Source*,Sink*) (reference
// src/butil/third_party/snappy/snappy.cc:875-954) and its hot loop
// internal::CompressFragment (snappy.cc:329-453) for many messages at once.
// The reference compresses every <=64 KiB fragment independently with a fresh
// zeroed hash table (snappy.cc:887-948), so fragments are the unit of
// parallelism: ONE WARP PER FRAGMENT, its u16 hash table in shared memory.
//
// The warp reproduces the sequential algorithm exactly:
//   * probe phase (snappy.cc:377-397): the next 32 probe positions are a closed
//     form of `skip` (probes 32b..32b+31 all advance by b+1 bytes), so the 32
//     lanes evaluate them together; __match_any_sync resolves hash collisions
//     inside the batch (a lane's candidate is the nearest lower lane with the
//     same hash, else the table), the lowest matching lane wins, and only the
//     lanes the sequential loop would have reached commit their table stores
//     (highest lane wins a slot).
//   * copy phase (snappy.cc:416-439): the 32 lanes compare 128 bytes per step
//     (FindMatchLength, snappy-internal.h:86-120), the copy tags are emitted
//     (EmitCopy, snappy.cc:198-232), lane 0 replays the two table stores and the
//     table load in program order.
// Fragment outputs land in per-fragment scratch slots; a second pass sizes the
// messages, a device scan places them, and a gather kernel writes
// varint32(N) || fragments back to back (snappy.cc:878-881).
#include <algorithm>

#include "b2c_internal.cuh"

namespace b2c {

static constexpr uint32_t kFull = 0xffffffffu;

__device__ __forceinline__ uint32_t hash_bytes(uint32_t bytes, uint32_t shift) {  // snappy.cc:47-50
    return (bytes * 0x1e35a7bdu) >> shift;
}

__device__ __forceinline__ uint32_t varint_len(uint32_t v) {
    return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : v < (1u << 28) ? 4 : 5;
}

// EmitLiteral, snappy.cc:156-196.  Returns bytes written.
__device__ __forceinline__ uint32_t emit_literal(uint8_t* out, const uint8_t* src, uint32_t len, uint32_t lane) {
    uint32_t n = len - 1, hdr;
    if (n < 60) {
        hdr = 1;
        if (lane == 0) out[0] = (uint8_t)(n << 2);
    } else {
        uint32_t count = n < (1u << 8) ? 1 : n < (1u << 16) ? 2 : n < (1u << 24) ? 3 : 4;
        hdr = 1 + count;
        if (lane == 0) {
            out[0] = (uint8_t)((59 + count) << 2);
            for (uint32_t i = 0; i < count; i++) out[1 + i] = (uint8_t)(n >> (8 * i));
        }
    }
    warp_copy_bytes(out + hdr, src, len, lane);
    return hdr + len;
}

// EmitCopy, snappy.cc:198-232: 64-byte pieces while len >= 68, a 60-byte piece if
// len > 64, then the rest (1-byte-offset form when len < 12 and offset < 2048).
__device__ __forceinline__ uint32_t emit_copy(uint8_t* out, uint32_t offset, uint32_t len, uint32_t lane) {
    uint32_t n64 = len >= 68 ? (len - 68) / 64 + 1 : 0;
    for (uint32_t k = lane; k < n64; k += 32) {
        uint8_t* o = out + 3 * k;
        o[0] = (uint8_t)(2 + (63u << 2));
        o[1] = (uint8_t)(offset & 0xff);
        o[2] = (uint8_t)(offset >> 8);
    }
    uint32_t w = 3 * n64;
    len -= 64 * n64;
    if (lane == 0) {
        uint8_t* o = out + w;
        if (len > 64) {
            o[0] = (uint8_t)(2 + (59u << 2));
            o[1] = (uint8_t)(offset & 0xff);
            o[2] = (uint8_t)(offset >> 8);
            o += 3;
        }
        uint32_t rest = len > 64 ? len - 60 : len;
        if (rest < 12 && offset < 2048) {
            o[0] = (uint8_t)(1 + ((rest - 4) << 2) + ((offset >> 8) << 5));
            o[1] = (uint8_t)(offset & 0xff);
        } else {
            o[0] = (uint8_t)(2 + ((rest - 1) << 2));
            o[1] = (uint8_t)(offset & 0xff);
            o[2] = (uint8_t)(offset >> 8);
        }
    }
    if (len > 64) {
        w += 3;
        len -= 60;
    }
    w += (len < 12 && offset < 2048) ? 2 : 3;
    return w;
}

// CompressFragment (snappy.cc:329-453) by one warp.  `table` has `tsize` zeroed
// u16 entries in shared memory.  Returns bytes written to out.
__device__ uint32_t compress_fragment_warp(const uint8_t* __restrict__ in, const uint32_t n, uint8_t* __restrict__ out,
                                           uint16_t* table, const uint32_t tsize, const uint32_t lane) {
    const uint32_t shift = 32 - (31 - __clz(tsize));
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t op = 0, next_emit = 0;
    if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
        const uint32_t ip_limit = n - 15;
        uint32_t ip = 1;  // snappy.cc:350: Hash(++ip)
        for (;;) {
            // ---------------- probe phase ----------------
            uint32_t base = ip, s = 1, cand = 0;
            bool found = false;
            for (;;) {
                const uint32_t pos = base + lane * s;
                const bool valid = pos + s <= ip_limit;  // next_ip <= ip_limit, snappy.cc:386-388
                uint32_t cur = 0, h = 0x80000000u | lane;
                if (valid) {
                    cur = ld32u(in + pos);
                    h = hash_bytes(cur, shift);
                }
                const uint32_t peers = __match_any_sync(kFull, h);
                const uint32_t prev = peers & lt_mask;
                uint32_t c = 0;
                bool match = false;
                if (valid) {
                    c = prev ? base + (31 - __clz(prev)) * s : (uint32_t)table[h];
                    match = ld32u(in + c) == cur;
                }
                const uint32_t mball = __ballot_sync(kFull, match);
                const uint32_t vball = __ballot_sync(kFull, valid);
                const uint32_t fi = vball == kFull ? 32 : __ffs(~vball) - 1;  // first lane past ip_limit
                const uint32_t fm = mball ? __ffs(mball) - 1 : 32;
                // lanes the sequential loop reaches: [0, fm] on a match, [0, fi) otherwise
                const uint32_t reach = fm < fi ? fm + 1 : fi;
                const uint32_t commit = reach >= 32 ? kFull : ((1u << reach) - 1u);
                if (((commit >> lane) & 1u) && (peers & commit & ~lt_mask & ~(1u << lane)) == 0)
                    table[h] = (uint16_t)pos;
                __syncwarp();
                if (fm < fi) {
                    found = true;
                    ip = base + fm * s;
                    cand = __shfl_sync(kFull, c, fm);
                    break;
                }
                if (fi < 32) break;  // ran into ip_limit: emit_remainder
                base += 32 * s;
                s++;
            }
            if (!found) break;
            // ---------------- literal before the match, snappy.cc:403 ----------------
            op += emit_literal(out + op, in + next_emit, ip - next_emit, lane);
            // ---------------- copy phase, snappy.cc:416-439 ----------------
            bool at_end = false;
            for (;;) {
                // common prefix of in[cand..] and in[ip..], bounded by n on the ip side
                uint32_t matched = 0;
                for (;;) {
                    const uint32_t p2 = ip + matched + 4 * lane;
                    const uint32_t p1 = cand + matched + 4 * lane;
                    uint32_t vb = p2 < n ? min(4u, n - p2) : 0u;
                    uint32_t nb = 0;
                    if (vb) {
                        uint32_t x = ld32u(in + p1) ^ ld32u(in + p2);
                        nb = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u;
                        nb = min(nb, vb);
                    }
                    const uint32_t stop = __ballot_sync(kFull, nb < 4);
                    if (stop) {
                        const uint32_t f = __ffs(stop) - 1;
                        matched += 4 * f + __shfl_sync(kFull, nb, f);
                        break;
                    }
                    matched += 128;
                }
                if (matched < 4) break;  // snappy.cc:439: no match at the new ip
                op += emit_copy(out + op, ip - cand, matched, lane);
                ip += matched;
                next_emit = ip;
                if (ip >= ip_limit) {  // snappy.cc:427-429
                    at_end = true;
                    break;
                }
                // snappy.cc:430-437, replayed by lane 0 in program order
                const uint32_t w0 = ld32u(in + ip - 1);
                const uint32_t w1 = ld32u(in + ip);
                const uint32_t hp = hash_bytes(w0, shift), hc = hash_bytes(w1, shift);
                uint32_t c = 0;
                if (lane == 0) {
                    table[hp] = (uint16_t)(ip - 1);
                    c = table[hc];
                    table[hc] = (uint16_t)ip;
                }
                cand = __shfl_sync(kFull, c, 0);
            }
            if (at_end) break;
            ip++;  // snappy.cc:441-442
            __syncwarp();
        }
    }
    // emit_remainder, snappy.cc:446-450
    if (next_emit < n) op += emit_literal(out + op, in + next_emit, n - next_emit, lane);
    return op;
}

struct CompressParams {
    const uint8_t* in;
    const uint64_t* off;
    const uint32_t* len;
    const uint32_t* frag_list;   // fragments of this size class (null: identity)
    const uint32_t* frag_msg;    // null: fragment f belongs to message f
    const uint32_t* frag_first;  // null: first fragment of message m is m
    const uint64_t* frag_slot;   // null: slot = f * slot_stride
    uint64_t slot_stride;
    uint8_t* scratch;
    uint32_t* frag_clen;
    uint32_t n_list;
    uint32_t table_entries;      // per-warp table capacity of this class
};

__global__ void __launch_bounds__(256) snappy_compress_fragments_kernel(CompressParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * (blockDim.x >> 5) + wib;
    if (w >= P.n_list) return;
    uint16_t* table = reinterpret_cast<uint16_t*>(smem_raw) + (size_t)wib * P.table_entries;
    const uint32_t f = P.frag_list ? P.frag_list[w] : w;
    const uint32_t m = P.frag_msg ? P.frag_msg[f] : f;
    const uint32_t k = P.frag_first ? f - P.frag_first[m] : 0u;
    const uint32_t mlen = P.len[m];
    const uint32_t n = min(kFragBytes, mlen - k * kFragBytes);
    // GetHashTable sizing, snappy.cc:253-256
    uint32_t tsize = 256;
    while (tsize < 16384u && tsize < n) tsize <<= 1;
    {
        uint4* t4 = reinterpret_cast<uint4*>(table);
        for (uint32_t i = lane; i < tsize / 8; i += 32) t4[i] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    const uint8_t* in = P.in + P.off[m] + (size_t)k * kFragBytes;
    uint8_t* out = P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride);
    const uint32_t clen = compress_fragment_warp(in, n, out, table, tsize, lane);
    if (lane == 0) P.frag_clen[f] = clen;
}

// Same algorithm with the hash table in HBM/L2 instead of shared memory: used for 64 KiB
// fragments, whose 32 KiB tables allow only 6 resident warps per SM in shared memory.  There the
// warp spends its time waiting on dependent loads of the INPUT (L2 latency), so giving up the
// table's shared-memory latency for 5-10x more resident warps is the better trade (measured on
// C5/words: 82 ms -> see DESIGN.md).  Persistent warps, one private 32 KiB table each.
__global__ void __launch_bounds__(256) snappy_compress_fragments_gtab_kernel(CompressParams P, uint16_t* gtab) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint32_t total = gridDim.x * (blockDim.x >> 5);
    uint16_t* table = gtab + (size_t)gw * 16384;
    for (uint32_t w = gw; w < P.n_list; w += total) {
        const uint32_t f = P.frag_list ? P.frag_list[w] : w;
        const uint32_t m = P.frag_msg ? P.frag_msg[f] : f;
        const uint32_t k = P.frag_first ? f - P.frag_first[m] : 0u;
        const uint32_t mlen = P.len[m];
        const uint32_t n = min(kFragBytes, mlen - k * kFragBytes);
        uint32_t tsize = 256;
        while (tsize < 16384u && tsize < n) tsize <<= 1;
        {
            uint4* t4 = reinterpret_cast<uint4*>(table);
            for (uint32_t i = lane; i < tsize / 8; i += 32) t4[i] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        const uint8_t* in = P.in + P.off[m] + (size_t)k * kFragBytes;
        uint8_t* out = P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride);
        const uint32_t clen = compress_fragment_warp(in, n, out, table, tsize, lane);
        if (lane == 0) P.frag_clen[f] = clen;
        __syncwarp();
    }
}

// message sizes: varint + sum of its fragments; also each fragment's offset
// inside its message stream.
__global__ void snappy_message_sizes_kernel(const uint32_t* __restrict__ len, const uint32_t* __restrict__ frag_first,
                                            const uint32_t* __restrict__ frag_clen, uint32_t* __restrict__ frag_rel,
                                            uint32_t* __restrict__ msg_clen, uint32_t n) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n) return;
    uint32_t acc = varint_len(len[m]);
    const uint32_t f0 = frag_first ? frag_first[m] : m, f1 = frag_first ? frag_first[m + 1] : m + 1;
    for (uint32_t f = f0; f < f1; f++) {
        frag_rel[f] = acc;
        acc += frag_clen[f];
    }
    msg_clen[m] = acc;
}

// one warp per message preamble + one warp per fragment body
__global__ void snappy_gather_kernel(const uint32_t* __restrict__ len, const uint32_t* __restrict__ frag_msg,
                                     const uint64_t* __restrict__ frag_slot, uint64_t slot_stride,
                                     const uint32_t* __restrict__ frag_clen,
                                     const uint32_t* __restrict__ frag_rel, const uint64_t* __restrict__ out_off,
                                     const uint8_t* __restrict__ scratch, uint8_t* __restrict__ out, uint32_t n_frags,
                                     uint32_t n_msgs) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (w < n_frags) {
        const uint32_t m = frag_msg ? frag_msg[w] : w;
        const uint64_t slot = frag_slot ? frag_slot[w] : (uint64_t)w * slot_stride;
        warp_copy_bytes(out + out_off[m] + frag_rel[w], scratch + slot, frag_clen[w], lane);
    } else if (w - n_frags < (n_msgs + 31) / 32) {
        // preambles: one lane per message (Varint::Encode32, snappy-stubs-internal.h:349-375)
        const uint32_t m = (w - n_frags) * 32 + lane;
        if (m < n_msgs) {
            uint8_t* o = out + out_off[m];
            uint32_t v = len[m];
            while (v >= 128) {
                *o++ = (uint8_t)(v | 128);
                v >>= 7;
            }
            *o = (uint8_t)v;
        }
    }
}

// ---------------------------------------------------------------------------
// exclusive scan u32 -> u64 (three passes; n up to 2^32)
// ---------------------------------------------------------------------------
static constexpr uint32_t kScanBlock = 1024;  // elements per block (256 threads x 4)

__global__ void scan_block_sums_kernel(const uint32_t* __restrict__ v, uint32_t n, uint64_t* __restrict__ sums) {
    __shared__ uint64_t red[8];
    const uint32_t base = blockIdx.x * kScanBlock;
    uint64_t s = 0;
    for (uint32_t i = threadIdx.x; i < kScanBlock; i += 256)
        if (base + i < n) s += v[base + i];
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(kFull, s, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t = 0;
        for (int i = 0; i < 8; i++) t += red[i];
        sums[blockIdx.x] = t;
    }
}

__global__ void scan_sums_kernel(uint64_t* sums, uint32_t nblocks) {  // single thread block, serial chunks
    __shared__ uint64_t carry;
    __shared__ uint64_t wsum[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint64_t x = i < nblocks ? sums[i] : 0, incl = x;
        for (int d = 1; d < 32; d <<= 1) {
            uint64_t y = __shfl_up_sync(kFull, incl, d);
            if ((threadIdx.x & 31) >= (uint32_t)d) incl += y;
        }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint64_t ws = wsum[threadIdx.x], wi = ws;
            for (int d = 1; d < 32; d <<= 1) {
                uint64_t y = __shfl_up_sync(kFull, wi, d);
                if (threadIdx.x >= (uint32_t)d) wi += y;
            }
            wsum[threadIdx.x] = wi - ws;  // exclusive warp offsets
        }
        __syncthreads();
        uint64_t excl = carry + wsum[threadIdx.x >> 5] + incl - x;
        if (i < nblocks) sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + x;
        __syncthreads();
    }
}

__global__ void scan_apply_kernel(const uint32_t* __restrict__ v, uint32_t n, const uint64_t* __restrict__ sums,
                                  uint64_t* __restrict__ out) {
    __shared__ uint64_t wsum[8];
    const uint32_t base = blockIdx.x * kScanBlock + threadIdx.x * 4;
    uint32_t a[4];
    uint64_t t = 0;
    for (int j = 0; j < 4; j++) {
        a[j] = base + j < n ? v[base + j] : 0;
        t += a[j];
    }
    uint64_t incl = t;
    for (int d = 1; d < 32; d <<= 1) {
        uint64_t y = __shfl_up_sync(kFull, incl, d);
        if ((threadIdx.x & 31) >= (uint32_t)d) incl += y;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint64_t woff = 0;
    for (uint32_t i = 0; i < (threadIdx.x >> 5); i++) woff += wsum[i];
    uint64_t run = sums[blockIdx.x] + woff + incl - t;
    for (int j = 0; j < 4; j++) {
        if (base + j < n) out[base + j] = run;
        run += a[j];
        if (base + j + 1 == n) out[n] = run;
    }
}

int launch_scan_u32_u64(b2c_ctx* ctx, Lane& lane, const uint32_t* d_vals, uint32_t n, uint64_t* d_out, void* d_tmp) {
    if (n == 0) {
        B2C_CUDA(ctx, cudaMemsetAsync(d_out, 0, 8, lane.stream));
        return B2C_OK;
    }
    const uint32_t nblocks = (n + kScanBlock - 1) / kScanBlock;
    uint64_t* sums = static_cast<uint64_t*>(d_tmp);
    scan_block_sums_kernel<<<nblocks, 256, 0, lane.stream>>>(d_vals, n, sums);
    scan_sums_kernel<<<1, 1024, 0, lane.stream>>>(sums, nblocks);
    scan_apply_kernel<<<nblocks, 256, 0, lane.stream>>>(d_vals, n, sums, d_out);
    ctx->launches += 3;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

// ---------------------------------------------------------------------------
int launch_compress(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* d_out,
                    uint64_t out_cap, uint64_t* d_out_off) {
    if (out_cap < L.max_comp) return set_err(ctx, B2C_E_CAPACITY, "compress output capacity", cudaSuccess);
    if (L.n == 0) {
        B2C_CUDA(ctx, cudaMemsetAsync(d_out_off, 0, 8, lane.stream));
        return B2C_OK;
    }
    int rc = ws_reserve(ctx, lane, lane.ws_frag, L.frag_scratch_bytes + 64);
    if (rc) return rc;
    // misc workspace: frag_clen[F] | frag_rel[F] | msg_clen[n] | scan tmp
    const size_t F = L.n_frags, n = L.n;
    const size_t scan_tmp = ((n + kScanBlock - 1) / kScanBlock + 2) * 8;
    const size_t need = (2 * F + n) * 4 + 16 + scan_tmp;
    rc = ws_reserve(ctx, lane, lane.ws_misc, need);
    if (rc) return rc;
    uint32_t* d_frag_clen = static_cast<uint32_t*>(lane.ws_misc.p);
    uint32_t* d_frag_rel = d_frag_clen + F;
    uint32_t* d_msg_clen = d_frag_rel + F;
    void* d_scan_tmp = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(d_msg_clen + n) + 15) & ~uintptr_t(15));
    uint8_t* scratch = static_cast<uint8_t*>(lane.ws_frag.p);

    static const uint32_t class_entries[3] = {1024, 4096, 16384};
    static const uint32_t class_warps[3] = {8, 4, 1};
    prof_mark(ctx, lane, B2C_PROF_COMPRESS);
    for (int c = 0; c < 3; c++) {
        if (!L.n_class[c]) continue;
        if (ctx->window_mask & (1 << c)) {  // window-parallel warp per fragment
            if ((rc = launch_compress_window(ctx, lane, L, c, d_in, scratch, d_frag_clen))) return rc;
            continue;
        }
        if (c == 0 && ctx->small_kernels && ctx->small_lockstep) {  // bodies <= 1 KiB: 32 per warp in lock-step
            if ((rc = launch_compress_lockstep(ctx, lane, L, d_in, scratch, d_frag_clen))) return rc;
            continue;
        }
        if (c == 0 && ctx->small_kernels) {  // bodies <= 1 KiB: one lane per fragment
            if ((rc = launch_compress_small(ctx, lane, L, d_in, scratch, d_frag_clen))) return rc;
            continue;
        }
        CompressParams P;
        P.in = d_in;
        P.off = L.d_off;
        P.len = L.d_len;
        P.frag_list = L.d_class_list[c];
        P.frag_msg = L.d_frag_msg;
        P.frag_first = L.d_frag_first;
        P.frag_slot = L.d_frag_slot;
        P.slot_stride = L.slot_stride;
        P.scratch = scratch;
        P.frag_clen = d_frag_clen;
        P.n_list = L.n_class[c];
        P.table_entries = class_entries[c];
        if (c == 2 && ctx->gtab_warps_per_sm > 0) {  // 64 KiB fragments: tables in HBM/L2, many resident warps
            const uint32_t want = (uint32_t)ctx->num_sms * (uint32_t)ctx->gtab_warps_per_sm;
            const uint32_t warps = std::min<uint32_t>(want, (P.n_list + 7) / 8 * 8);
            const uint32_t gblocks = (warps + 7) / 8;
            if ((rc = ws_reserve(ctx, lane, lane.ws_tab, (size_t)gblocks * 8 * 32768))) return rc;
            snappy_compress_fragments_gtab_kernel<<<gblocks, 256, 0, lane.stream>>>(
                P, static_cast<uint16_t*>(lane.ws_tab.p));
            ctx->launches++;
            B2C_CUDA(ctx, cudaGetLastError());
            continue;
        }
        const uint32_t wpb = class_warps[c];
        const uint32_t blocks = (P.n_list + wpb - 1) / wpb;
        const size_t smem = (size_t)wpb * class_entries[c] * 2;
        snappy_compress_fragments_kernel<<<blocks, wpb * 32, smem, lane.stream>>>(P);
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    prof_mark(ctx, lane, B2C_PROF_COMPRESS);
    prof_mark(ctx, lane, B2C_PROF_PACK);
    snappy_message_sizes_kernel<<<(L.n + 255) / 256, 256, 0, lane.stream>>>(L.d_len, L.d_frag_first, d_frag_clen,
                                                                            d_frag_rel, d_msg_clen, L.n);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    rc = launch_scan_u32_u64(ctx, lane, d_msg_clen, L.n, d_out_off, d_scan_tmp);
    if (rc) return rc;
    const uint64_t warps = (uint64_t)L.n_frags + (L.n + 31) / 32;
    const uint32_t blocks = (uint32_t)((warps * 32 + 255) / 256);
    snappy_gather_kernel<<<blocks, 256, 0, lane.stream>>>(L.d_len, L.d_frag_msg, L.d_frag_slot, L.slot_stride,
                                                         d_frag_clen, d_frag_rel, d_out_off, scratch, d_out,
                                                         L.n_frags, L.n);
    prof_mark(ctx, lane, B2C_PROF_PACK);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

}  // namespace b2c
