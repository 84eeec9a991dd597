// snappy_window.cu — window-parallel, byte-exact snappy compression: one warp per fragment,
// 32 speculative table steps per round.
//
// CompressFragment (reference snappy.cc:329-453) is a chain: every table step (load 4 bytes at
// ip, hash, look up + replace the table slot, compare with the candidate) decides where the
// next one happens.  What the chain does NOT decide is the work inside a step.  So each round
// the 32 lanes run the table step for the next 32 positions the chain can reach — the next 32
// consecutive bytes after a copy, or the next 32 positions of the `skip` schedule while probing
// (snappy.cc:377-397) — against a snapshot of the table, and the warp then WALKS the chain
// through those precomputed results: one shuffle per visited position instead of a dependent
// sequence of shared-memory round trips.  The walk stays exact because the only thing the
// snapshot cannot know, the slots replaced earlier in the same round, is kept in a short
// in-register list that every lookup checks first (a hit there re-evaluates that one position).
// A round typically resolves ~10 visits; the elements it finds are encoded afterwards by one
// lane each (warp prefix sum over their encoded sizes).  tools/window_model.py is the lane-level
// model this kernel was transcribed from (checked against the oracle on ~500 inputs).
#include "snappy_emit.cuh"

namespace b2c {

// CompressFragment by one warp.  `T` = `tsize` zeroed u16 slots (shared or global memory).
template <class In>
__device__ uint32_t compress_fragment_window(const In in, const uint32_t n, uint8_t* __restrict__ out, uint16_t* T,
                                             const uint32_t tsize, const uint32_t lane) {
    const uint32_t shift = 32 - (31 - __clz(tsize));
    uint32_t op = 0, next_emit = 0;
    if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
        const uint32_t ip_limit = n - 15;
        uint32_t ip = 1, skip = 32;
        bool probing = true, done = false;
        while (!done) {
            // ---------------- round setup: 32 speculative table steps ----------------
            uint32_t pos;
            if (probing) {  // the next 32 positions of the skip schedule (snappy.cc:385)
                const uint32_t st = (skip + lane) >> 5;
                uint32_t incl = st;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t y = __shfl_up_sync(kFullMask, incl, d);
                    if (lane >= (uint32_t)d) incl += y;
                }
                pos = ip + incl - st;
            } else {  // right after a copy the chain can only continue on consecutive bytes
                pos = ip + lane;
            }
            const bool okl = pos <= ip_limit;
            uint64_t d8 = 0;
            uint32_t h = 0, packed = 0;  // packed = snapshot candidate | equal bytes << 16
            if (okl) {
                d8 = in.ld64(pos);
                h = hash_w((uint32_t)d8, shift);
                const uint32_t c0 = T[h];
                packed = c0 | (prefix8(in.ld64(c0), d8) << 16);
            }
            // slots replaced during this round, in order: entry i lives in lane (i & 31), bank (i >> 5)
            uint32_t Lh0 = 0, Lp0 = 0, Lh1 = 0, Lp1 = 0, Lc = 0;
            // elements found during this round: element e lives in lane e
            uint32_t e_ls = 0, e_p = 0, e_cand = 0, e_m = 0, Ec = 0;
            uint32_t p = ip;
            // ---------------- walk the chain through the precomputed steps ----------------
            for (;;) {
                const uint32_t nxt = p + (skip >> 5);
                if (probing && nxt > ip_limit) {  // snappy.cc:386-388
                    done = true;
                    break;
                }
                const uint32_t where = __ballot_sync(kFullMask, okl && pos == p);
                if (!where || Lc >= 62 || Ec >= 32) break;  // not precomputed (or lists full): next round
                const uint32_t j = __ffs(where) - 1;
                if (probing) skip++;
                const uint32_t pk = __shfl_sync(kFullMask, packed, j);
                const uint32_t hh = __shfl_sync(kFullMask, h, j);
                uint32_t cand = pk & 0xffffu, m = pk >> 16;
                // was this slot replaced earlier in the round?  (the most recent replacement wins)
                const uint32_t hit1 = __ballot_sync(kFullMask, lane + 32 < Lc && Lh1 == hh);
                const uint32_t hit0 = __ballot_sync(kFullMask, lane < Lc && Lh0 == hh);
                if (hit1 | hit0) {
                    cand = hit1 ? __shfl_sync(kFullMask, Lp1, 31 - __clz(hit1)) : __shfl_sync(kFullMask, Lp0, 31 - __clz(hit0));
                    m = prefix8(in.ld64(cand), in.ld64(p));
                }
                // table[hash] = ip (snappy.cc:394)
                if (lane == 0) T[hh] = (uint16_t)p;
                if (lane == (Lc & 31)) {
                    if (Lc < 32) { Lh0 = hh; Lp0 = p; } else { Lh1 = hh; Lp1 = p; }
                }
                Lc++;
                if (m >= 4) {  // 4-byte match: snappy.cc:395-396 / 438
                    if (m == 8) {  // FindMatchLength beyond the first 8 bytes, 128 bytes per step
                        for (;;) {
                            const uint32_t p2 = p + m + 4 * lane, p1 = cand + m + 4 * lane;
                            const uint32_t vb = p2 < n ? min(4u, n - p2) : 0u;
                            uint32_t nb = 0;
                            if (vb) {
                                const uint32_t x = in.ld32(p1) ^ in.ld32(p2);
                                nb = min(x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u, vb);
                            }
                            const uint32_t stop = __ballot_sync(kFullMask, nb < 4);
                            if (stop) {
                                const uint32_t f = __ffs(stop) - 1;
                                m += 4 * f + __shfl_sync(kFullMask, nb, f);
                                break;
                            }
                            m += 128;
                        }
                    }
                    if (lane == Ec) {
                        e_ls = next_emit;
                        e_p = p;
                        e_cand = cand;
                        e_m = m;
                    }
                    Ec++;
                    p += m;
                    next_emit = p;
                    if (p >= ip_limit) {  // snappy.cc:427-429
                        done = true;
                        break;
                    }
                    // table[hash(ip - 1)] = ip - 1 (snappy.cc:430-433)
                    const uint32_t hq = hash_w(in.ld32(p - 1), shift);
                    if (lane == 0) T[hq] = (uint16_t)(p - 1);
                    if (lane == (Lc & 31)) {
                        if (Lc < 32) { Lh0 = hq; Lp0 = p - 1; } else { Lh1 = hq; Lp1 = p - 1; }
                    }
                    Lc++;
                    probing = false;
                } else if (probing) {
                    p = nxt;
                } else {  // snappy.cc:441-442: leave the copy loop, resume probing one byte on
                    p += 1;
                    probing = true;
                    skip = 32;
                }
            }
            __syncwarp();  // lane 0's table stores precede the next round's snapshot loads
            ip = p;
            // ---------------- encode this round's elements, one lane each ----------------
            if (Ec) op = emit_elements_warp(in, lane < Ec, e_ls, e_p, e_p - e_cand, e_m, out, op, lane);
        }
    }
    op = emit_remainder_warp(in, next_emit, n, out, op, lane);
    return op;
}

// ---- kernels ------------------------------------------------------------------------------
struct WindowParams {
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
    uint32_t table_entries;  // per-warp table capacity
    uint32_t in_stride;      // per-warp staged-input bytes (0: read the fragment from HBM)
};

__device__ __forceinline__ uint32_t smem_addr32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One warp per fragment.  Shared memory per warp: u16 table [table_entries] (+ the staged fragment).
template <bool kStage>
__global__ void __launch_bounds__(256) snappy_compress_window_kernel(WindowParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ uint64_t bars[8];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * (blockDim.x >> 5) + wib;
    if (w >= P.n_list) return;
    uint8_t* my = smem_raw + (size_t)wib * ((size_t)P.table_entries * 2 + P.in_stride);
    uint16_t* table = reinterpret_cast<uint16_t*>(my);
    const uint32_t f = P.frag_list ? P.frag_list[w] : w;
    const uint32_t m = P.frag_msg ? P.frag_msg[f] : f;
    const uint32_t k = P.frag_first ? f - P.frag_first[m] : 0u;
    const uint32_t n = min(kFragBytes, P.len[m] - k * kFragBytes);
    const uint8_t* gin = P.in + P.off[m] + (size_t)k * kFragBytes;
    uint32_t tsize = 256;  // GetHashTable sizing, snappy.cc:253-256
    while (tsize < 16384u && tsize < n) tsize <<= 1;
    {
        uint4* t4 = reinterpret_cast<uint4*>(table);
        for (uint32_t i = lane; i < tsize / 8; i += 32) t4[i] = make_uint4(0, 0, 0, 0);
    }
    uint8_t* out = P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride);
    uint32_t clen;
    if (kStage) {
        // the fragment comes in by ONE TMA bulk copy per warp (cp.async.bulk + mbarrier) when it is
        // 16-byte aligned, else by word loads
        uint8_t* stage = my + (size_t)P.table_entries * 2;
        uint64_t* bar = &bars[wib];
        const bool bulk = n && (reinterpret_cast<uintptr_t>(gin) & 15) == 0;
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr32(bar)), "r"(1) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            if (bulk) {
                const uint32_t bytes = (n + 15u) & ~15u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr32(bar)), "r"(bytes)
                             : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_addr32(stage)),
                             "l"(gin), "r"(bytes), "r"(smem_addr32(bar))
                             : "memory");
            } else {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr32(bar)) : "memory");
            }
        }
        if (!bulk)
            for (uint32_t b = 4 * lane; b < n; b += 128) *reinterpret_cast<uint32_t*>(stage + b) = ld32u(gin + b);
        __syncwarp();
        asm volatile(
            "{\n\t"
            ".reg .pred P1;\n\t"
            "WWAIT_LOOP:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0, 0x989680;\n\t"
            "@P1 bra WWAIT_DONE;\n\t"
            "bra WWAIT_LOOP;\n\t"
            "WWAIT_DONE:\n\t"
            "}" ::"r"(smem_addr32(bar))
            : "memory");
        __syncwarp();
        clen = compress_fragment_window(InShared{stage}, n, out, table, tsize, lane);
    } else {
        __syncwarp();
        clen = compress_fragment_window(InGlobal{gin}, n, out, table, tsize, lane);
    }
    if (lane == 0) P.frag_clen[f] = clen;
}

int window_kernels_configure(b2c_ctx* ctx) {
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_window_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       64 * 1024));
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_window_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       64 * 1024));
    return B2C_OK;
}

// class c: 0 = fragments <= 1 KiB, 1 = <= 4 KiB, 2 = <= 64 KiB
int launch_compress_window(b2c_ctx* ctx, Lane& lane, const LayoutView& L, int c, const uint8_t* d_in, uint8_t* scratch,
                           uint32_t* d_frag_clen) {
    static const uint32_t entries[3] = {1024, 4096, 16384};
    static const uint32_t warps[3] = {8, 4, 1};
    static const uint32_t stage[3] = {1024 + 32, 4096 + 32, 0};  // staged input bytes per warp (0: from HBM)
    WindowParams P;
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
    P.table_entries = entries[c];
    P.in_stride = stage[c];
    const uint32_t wpb = warps[c];
    const uint32_t blocks = (P.n_list + wpb - 1) / wpb;
    const size_t smem = (size_t)wpb * ((size_t)entries[c] * 2 + stage[c]);
    if (stage[c])
        snappy_compress_window_kernel<true><<<blocks, wpb * 32, smem, lane.stream>>>(P);
    else
        snappy_compress_window_kernel<false><<<blocks, wpb * 32, smem, lane.stream>>>(P);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

}  // namespace b2c
