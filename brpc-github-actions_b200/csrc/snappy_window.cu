// snappy_window.cu — byte-exact snappy compression with one warp per fragment and NO serial walk:
// every round resolves up to 32 steps of CompressFragment's chain (reference snappy.cc:329-453) in
// parallel.
//
// The chain — load 4 bytes at ip, hash, look up + replace the table slot, compare with the
// candidate, advance by the match or by the skip schedule — decides where each step happens, but
// not what a step computes.  So the 32 lanes run the step for the next 32 positions (consecutive
// bytes, the common case; the positions of the skip schedule, snappy.cc:377-385, once 32 probes in a
// row have missed) against a snapshot of the table, each lane extending its own match.  Two facts
// make the snapshot answers final:
//   * a slot a lane reads can only be replaced during this round by an earlier lane of the window
//     with the same hash (all positions a round inserts up to a lane's own are window positions),
//     so the window is CUT at the first lane whose hash repeats an earlier lane's; the next round
//     starts there with a fresh snapshot;
//   * inside the cut, "where does the chain go from lane j" is a function next(j) of lane j's own
//     answer (j + 1 on a miss, the lane the copy lands on after a hit), so the positions the chain
//     really visits are the orbit of lane 0 under next(), found by pointer doubling (5 shuffles).
// Visited lanes then insert themselves, copies insert their last byte's position (snappy.cc:430-433;
// the later position wins a shared slot), and the copies found are queued; whenever 32 are queued
// the warp encodes them at once (snappy_emit.cuh).  tools/orbit_model.py is the lane-level model
// this kernel was transcribed from (byte-identical to the oracle on ~500 inputs).
#include <algorithm>

#include "snappy_emit.cuh"

namespace b2c {

static constexpr uint32_t kCap = 40;    // speculative per-lane match length limit; >= 32, so a capped match leaves the window
static constexpr uint32_t kQueue = 64;  // queued elements per warp (two words each)
static constexpr uint32_t kWarpFixed = kQueue * sizeof(uint2) + kDeferBytes;  // per warp: the queue, then the noted literal moves

// Encode the first `cnt` (<= 32) queued elements; `emitted` = end of the last element already encoded.
template <class In>
__device__ __forceinline__ uint32_t flush_queue(const In& in, const uint2* queue, uint32_t cnt, uint32_t& emitted,
                                                uint8_t* __restrict__ out, uint32_t op, uint32_t lane) {
    const bool has = lane < cnt;
    const uint2 e = has ? queue[lane] : make_uint2(0, 0);
    const uint32_t p = e.x & 0xffffu, off = e.x >> 16, mm = e.y;
    const uint32_t end = p + mm;
    uint32_t ls = __shfl_up_sync(kFullMask, end, 1);
    if (lane == 0) ls = emitted;
    emitted = __shfl_sync(kFullMask, end, cnt - 1);
    return emit_elements_warp(in, has, ls, p, off, mm, out, op, lane,
                              reinterpret_cast<uint4*>(const_cast<uint2*>(queue) + kQueue));
}

// CompressFragment by one warp.  `T` = `tsize` zeroed u16 slots (shared or global memory); `queue` = kQueue
// uint2 of shared memory private to the warp.
template <class In>
__device__ uint32_t compress_fragment_orbit(const In in, const uint32_t n, uint8_t* __restrict__ out, uint16_t* T,
                                            const uint32_t tsize, uint2* queue, const uint32_t lane) {
    const uint32_t shift = 32 - (31 - __clz(tsize));
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t op = 0, next_emit = 0, emitted = 0, cnt = 0;
    if (lane == 0) reinterpret_cast<uint4*>(queue + kQueue)[0].x = 0;  // no literal move noted yet
    __syncwarp();
    if (n >= 15) {  // kInputMarginBytes, snappy.cc:344-347
        const uint32_t ip_limit = n - 15;
        // `s` is the reference's `skip` with one extra state: 31 right after a copy (the copy loop's own table
        // step, snappy.cc:430-439, advances one byte on a miss and is followed by a fresh skip = 32 sequence).
        uint32_t ip = 1, s = 32;
        for (;;) {
            // ---- the window: positions, and whether the chain may stand on them (snappy.cc:386-388) ----
            const uint32_t k1 = s <= 32 ? 32u : (s < 64 ? 64u - s : 0u);  // leading probes that advance one byte
            const bool consecutive = k1 >= 16;
            uint32_t W, step, pos;
            if (consecutive) {
                W = k1;
                step = 1;
                pos = ip + lane;
            } else {
                W = 32;
                step = max(1u, (s + lane) >> 5);
                uint32_t incl = step;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t y = __shfl_up_sync(kFullMask, incl, d);
                    if (lane >= (uint32_t)d) incl += y;
                }
                pos = ip + incl - step;
            }
            const bool vis_ok = lane < W && pos + step <= ip_limit;
            if (!__shfl_sync(kFullMask, (int)vis_ok, 0)) break;  // the chain stops here: emit_remainder
            // ---- every lane's table step against the snapshot ----
            uint32_t h = 0, c0 = 0, m = 0;
            if (vis_ok) {
                const uint64_t d8 = in.ld64(pos);
                h = hash_w((uint32_t)d8, shift);
                c0 = T[h];
                m = prefix8(in.ld64(c0), d8);
                while (m >= 8 && m < kCap) {  // FindMatchLength, 8 bytes per trip (pos + 8 < n: pos <= n - 16)
                    const uint32_t rem = n - (pos + m);
                    if (rem == 0) break;
                    const uint32_t nb = min(prefix8(in.ld64(c0 + m), in.ld64(pos + m)), rem);
                    m += nb;
                    if (nb < 8) break;
                }
            }
            // ---- cut the window at the first repeated hash ----
            const uint32_t same = __match_any_sync(kFullMask, vis_ok ? h : (0x80000000u | lane));
            const uint32_t cuts = __ballot_sync(kFullMask, vis_ok && (same & lt) != 0);
            const uint32_t Wc = cuts ? min(W, (uint32_t)__ffs(cuts) - 1) : W;
            // ---- next(j) ----
            const bool inwin = vis_ok && lane < Wc;
            const bool hit = inwin && m >= 4;
            const uint32_t end = pos + m;
            const bool next_ok = __shfl_down_sync(kFullMask, (int)vis_ok, 1) != 0;
            uint32_t nl = 32;   // 32 = the chain leaves the window after this lane
            bool stop = false;  // ... because the loop-top check fails at the next position
            if (inwin) {
                if (hit) {
                    if (consecutive && m < kCap && end - ip < Wc && end < ip_limit) nl = end - ip;
                } else if (lane + 1 < Wc) {
                    if (next_ok) nl = lane + 1;
                    else stop = true;
                }
            }
            // ---- the orbit of lane 0 by pointer doubling ----
            uint32_t reach = 1u << lane, jump = nl;
#pragma unroll
            for (int it = 0; it < 5; it++) {
                const uint32_t r2 = __shfl_sync(kFullMask, reach, jump & 31);
                const uint32_t j2 = __shfl_sync(kFullMask, jump, jump & 31);
                if (jump < 32) {
                    reach |= r2;
                    jump = j2;
                }
            }
            const uint32_t V = __shfl_sync(kFullMask, reach, 0);
            const bool visited = (V >> lane) & 1u;
            const uint32_t last = 31 - __clz(V);
            // ---- a capped match can only be the last visited lane: finish it, 128 bytes per trip ----
            const bool last_hit = __shfl_sync(kFullMask, (int)hit, last) != 0;
            if (last_hit && __shfl_sync(kFullMask, m, last) >= kCap) {
                const uint32_t pl = __shfl_sync(kFullMask, pos, last), cl = __shfl_sync(kFullMask, c0, last);
                uint32_t ml = __shfl_sync(kFullMask, m, last);
                for (;;) {
                    // a run-heavy fragment is ONE match of up to 64 KiB: compare 512 bytes per trip (four independent
                    // pairs of loads per lane) while that much is left, 128 bytes per trip at the end
                    if (pl + ml + 512 <= n) {
                        uint32_t x = 0;
#pragma unroll
                        for (uint32_t k = 0; k < 4; k++) {
                            const uint32_t q = ml + 128 * k + 4 * lane;
                            x |= in.ld32(cl + q) ^ in.ld32(pl + q);
                        }
                        if (!__any_sync(kFullMask, x != 0)) {
                            ml += 512;
                            continue;
                        }
                    }
                    const uint32_t p2 = pl + ml + 4 * lane, p1 = cl + ml + 4 * lane;
                    const uint32_t vb = p2 < n ? min(4u, n - p2) : 0u;
                    uint32_t nb = 0;
                    if (vb) {
                        const uint32_t x = in.ld32(p1) ^ in.ld32(p2);
                        nb = min(x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u, vb);
                    }
                    const uint32_t brk = __ballot_sync(kFullMask, nb < 4);
                    if (brk) {
                        const uint32_t f = __ffs(brk) - 1;
                        ml += 4 * f + __shfl_sync(kFullMask, nb, f);
                        break;
                    }
                    ml += 128;
                }
                if (lane == last) m = ml;
            }
            const uint32_t endv = pos + m;
            // ---- inserts: visited positions (snappy.cc:394), then each copy's last byte (snappy.cc:430-433) ----
            if (visited) T[h] = (uint16_t)pos;
            __syncwarp();
            const bool vcopy = visited && hit;
            const bool tail = vcopy && endv < ip_limit;
            uint32_t hq = 0;
            if (tail) hq = hash_w(in.ld32(endv - 1), shift);
            const uint32_t grp = __match_any_sync(kFullMask, tail ? hq : (0x80000000u | lane));
            if (tail && lane == 31 - (uint32_t)__clz(grp)) T[hq] = (uint16_t)(endv - 1);  // the later position wins
            __syncwarp();
            // ---- queue the copies, in order ----
            const uint32_t cm = __ballot_sync(kFullMask, vcopy);
            if (vcopy) queue[cnt + __popc(cm & lt)] = make_uint2(pos | ((pos - c0) << 16), m);
            cnt += __popc(cm);
            __syncwarp();
            if (cnt >= 32) {
                op = flush_queue(in, queue, 32, emitted, out, op, lane);
                const uint2 mv = lane + 32 < cnt ? queue[lane + 32] : make_uint2(0, 0);
                __syncwarp();
                if (lane + 32 < cnt) queue[lane] = mv;
                cnt -= 32;
                __syncwarp();
            }
            // ---- where the chain stands now ----
            if (cm) next_emit = __shfl_sync(kFullMask, endv, 31 - __clz(cm));
            if (last_hit) {
                ip = __shfl_sync(kFullMask, endv, last);
                s = 31;
                if (ip >= ip_limit) break;  // snappy.cc:427-429
            } else {
                const uint32_t after = cm ? __popc(V & ~((2u << (31 - __clz(cm))) - 1u)) : 0u;
                s = cm ? 31 + after : s + __popc(V);
                ip = __shfl_sync(kFullMask, pos + step, last);
                if (__shfl_sync(kFullMask, (int)stop, last)) break;
            }
        }
    }
    if (cnt) op = flush_queue(in, queue, cnt, emitted, out, op, lane);
    return emit_remainder_warp(in, next_emit, n, out, op, lane, reinterpret_cast<uint4*>(queue + kQueue));  // snappy.cc:446-450
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
    uint32_t table_entries;  // per-warp table capacity in shared memory (0: tables in `gtab`)
    uint32_t in_stride;      // per-warp staged-input bytes (0: read the fragment from HBM)
    uint16_t* gtab;          // 16384 entries per resident warp when table_entries == 0
    uint32_t* work_ctr;      // warps take the next listed fragment from here (null: every total-th one)
};

__device__ __forceinline__ uint32_t smem_addr32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Persistent warps, one fragment at a time.  Shared memory per warp: element queue, then the u16 table
// (unless it lives in HBM/L2), then the staged fragment (if staged).
template <bool kStage>
// (256, 5): the HBM-table variant lives on resident warps (40 per SM = 48 registers per thread); the vector copies of
// long literals must not cost it a fifth of them
__global__ void __launch_bounds__(256) snappy_compress_orbit_kernel(WindowParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ uint64_t bars[8];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + wib;
    const uint32_t total = gridDim.x * (blockDim.x >> 5);
    const size_t per_warp = kWarpFixed + (size_t)P.table_entries * 2 + P.in_stride;
    uint8_t* my = smem_raw + (size_t)wib * per_warp;
    uint2* queue = reinterpret_cast<uint2*>(my);
    uint16_t* table = P.table_entries ? reinterpret_cast<uint16_t*>(my + kWarpFixed) : P.gtab + (size_t)gw * 16384;
    uint8_t* stage = my + kWarpFixed + (size_t)P.table_entries * 2;
    uint64_t* bar = &bars[wib];
    if (kStage && lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr32(bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    uint32_t parity = 0;
    for (uint32_t w = gw;; w += total) {
        if (P.work_ctr) {  // fragments differ in length (the last one of a body): whoever is free takes the next
            if (lane == 0) w = atomicAdd(P.work_ctr, 1u);
            w = __shfl_sync(kFullMask, w, 0);
        }
        if (w >= P.n_list) break;
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
            const bool bulk = n && (reinterpret_cast<uintptr_t>(gin) & 15) == 0;
            if (lane == 0) {
                if (bulk) {
                    const uint32_t bytes = (n + 15u) & ~15u;
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of the slot
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr32(bar)), "r"(bytes)
                                 : "memory");
                    asm volatile(
                        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
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
                "OWAIT_LOOP:\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n\t"
                "@P1 bra OWAIT_DONE;\n\t"
                "bra OWAIT_LOOP;\n\t"
                "OWAIT_DONE:\n\t"
                "}" ::"r"(smem_addr32(bar)),
                "r"(parity)
                : "memory");
            parity ^= 1;
            __syncwarp();
            clen = compress_fragment_orbit(InShared{stage}, n, out, table, tsize, queue, lane);
        } else {
            __syncwarp();
            clen = compress_fragment_orbit(InGlobal{gin}, n, out, table, tsize, queue, lane);
        }
        if (lane == 0) P.frag_clen[f] = clen;
        __syncwarp();
    }
}

int window_kernels_configure(b2c_ctx* ctx) {
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_orbit_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       100 * 1024));
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_orbit_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
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
    P.in_stride = ctx->window_stage ? stage[c] : 0;
    P.gtab = nullptr;
    uint32_t wpb = warps[c];
    uint32_t resident;  // CTAs per SM the launch is sized for
    if (c == 2 && ctx->window_smem64) {  // 64 KiB fragments: table AND TMA-staged fragment in shared memory, 2 warps/SM
        wpb = 1;
        resident = 2;
        P.in_stride = 65536 + 32;
    } else if (c == 2 && ctx->gtab_warps_per_sm > 0) {  // 64 KiB fragments: tables in HBM/L2, many resident warps
        // A round is a chain of 2-3 L2 round trips, so throughput follows the number of resident warps: measured on
        // C5/words 68.9 / 38.9 / 31.6 / 26.9 / 23.7 ms per GiB at 8 / 16 / 24 / 32 / 40 warps per SM (40 = the
        // register limit; tables in shared memory allow 6: 80 ms).
        wpb = 8;
        resident = std::max(1, ctx->window_warps_per_sm / 8);
        P.table_entries = 0;
    } else {
        const size_t pw = kWarpFixed + (size_t)P.table_entries * 2 + P.in_stride;
        resident = std::max<uint32_t>(1, (uint32_t)((220u << 10) / (pw * wpb + 1024)));
    }
    const size_t per_warp = kWarpFixed + (size_t)P.table_entries * 2 + P.in_stride;
    const uint32_t want_blocks = (uint32_t)ctx->num_sms * resident;
    const uint32_t blocks = std::min<uint32_t>(want_blocks, (P.n_list + wpb - 1) / wpb);
    if (P.table_entries == 0) {
        int rc;
        if ((rc = ws_reserve(ctx, lane, lane.ws_tab, (size_t)blocks * wpb * 32768))) return rc;
        P.gtab = static_cast<uint16_t*>(lane.ws_tab.p);
    }
    {
        int rc = next_work_counter(ctx, lane, &P.work_ctr);
        if (rc) return rc;
    }
    const size_t smem = (size_t)wpb * per_warp;
    if (P.in_stride)
        snappy_compress_orbit_kernel<true><<<blocks, wpb * 32, smem, lane.stream>>>(P);
    else
        snappy_compress_orbit_kernel<false><<<blocks, wpb * 32, smem, lane.stream>>>(P);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

}  // namespace b2c
