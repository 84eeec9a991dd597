// snappy_lockstep.cu — compressor for bodies of at most 1 KiB around the lock-step lane state machine of
// snappy_lockstep.cuh.  Two phases per warp, like snappy_compress_small2_kernel:
//   FIND    LPW lanes of the warp carry one body each and advance one branch-free step per trip
//           (FindLane::trip), recording every copy as one packed word in an L2-resident list;
//   ENCODE  the whole warp encodes each list 32 elements at a time (snappy_emit.cuh).
// A CTA owns 64 slots = two groups of 32; a group's u16 hash tables (64 KiB) and bodies (32.5 KiB) are
// word-interleaved over its slots (word w of slot l at index w * 32 + l), so lanes on different slots never share
// a bank.  Each warp owns LPW slots outright (its table and body columns), so warps never synchronise with each
// other; persistent warps walk the list of bodies LPW at a time.
#include <algorithm>

#include "snappy_emit.cuh"
#include "snappy_lockstep.cuh"
#include "snappy_lockstep_dec.cuh"

namespace b2c {

struct LockstepParams {
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
    uint32_t* elists;  // [resident warps * LPW][ls::kMaxElems]
    uint32_t* gscratch;  // non-null: tables and staged bodies live here (L2 / HBM), one ls::kGroupWords region per resident 64 slots / 2
};

struct InInterleaved {  // a body in the interleaved layout: word w at col[w * 32]
    static constexpr bool kContiguous = false;
    const uint32_t* col;
    __device__ __forceinline__ uint8_t ld8(uint32_t p) const {
        return (uint8_t)(col[(p >> 2) * ls::kSlots] >> ((p & 3) * 8));
    }
};

constexpr uint32_t kLsSlotsPerCta = 64;

// The whole warp copies one body into column `col` (word w at col[w * 32]): coalesced 16-byte reads when the body
// is 16-byte aligned, aligned word pairs otherwise; only aligned units that hold at least one byte of the body are
// read.  (All words of a slot live in one bank by design, so the stores of one body serialise: 256 wavefronts per
// KiB whichever lanes issue them.)
__device__ __forceinline__ void stage_body_warp(uint32_t* __restrict__ col, const uint8_t* __restrict__ gin, uint32_t n,
                                                uint32_t lane) {
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(gin);
    if ((a0 & 15) == 0) {
        const uint4* g = reinterpret_cast<const uint4*>(gin);
        const uint32_t vecs = (n + 15) >> 4;
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (lane < vecs) v0 = __ldg(g + lane);
        if (lane + 32 < vecs) v1 = __ldg(g + lane + 32);
        if (lane < vecs) {
            uint32_t* c = col + (size_t)lane * 4 * ls::kSlots;
            c[0] = v0.x, c[ls::kSlots] = v0.y, c[2 * ls::kSlots] = v0.z, c[3 * ls::kSlots] = v0.w;
        }
        if (lane + 32 < vecs) {
            uint32_t* c = col + (size_t)(lane + 32) * 4 * ls::kSlots;
            c[0] = v1.x, c[ls::kSlots] = v1.y, c[2 * ls::kSlots] = v1.z, c[3 * ls::kSlots] = v1.w;
        }
    } else {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a0 & ~uintptr_t(3));
        const uint32_t sh = (uint32_t)(a0 & 3) * 8;
        const uint32_t have = (uint32_t)(a0 & 3) + n;  // bytes from the first aligned word to the end of the body
        const uint32_t words = (n + 3) >> 2;
        for (uint32_t k = lane; k < words; k += 32) {
            const uint32_t lo = __ldg(w + k);
            const uint32_t hi = (4 * (k + 1) < have) ? __ldg(w + k + 1) : 0u;
            col[(size_t)k * ls::kSlots] = __funnelshift_r(lo, hi, sh);
        }
    }
}

template <int LPW>
__global__ void __launch_bounds__(32 * kLsSlotsPerCta / LPW) snappy_compress_lockstep_kernel(LockstepParams P) {
    extern __shared__ __align__(16) uint32_t ls_smem[];
    constexpr uint32_t kWarps = kLsSlotsPerCta / LPW;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t slot0 = warp * LPW;  // first of this warp's slots within the CTA
    // shared memory (64 slots per SM-resident CTA) or, for the many-more-bodies-in-flight variant, global memory that
    // the L2 keeps: the lane code is the same, only where its table and body columns live differs
    uint32_t* base = P.gscratch ? P.gscratch + (size_t)blockIdx.x * (kLsSlotsPerCta / ls::kSlots) * ls::kGroupWords : ls_smem;
    uint32_t* tab = base + (size_t)(slot0 / ls::kSlots) * ls::kGroupWords + (slot0 % ls::kSlots);  // column of slot0
    uint32_t* body = tab + ls::kTabWords * ls::kSlots;
    const uint32_t gw = blockIdx.x * kWarps + warp;
    uint32_t* my_lists = P.elists + (size_t)gw * LPW * ls::kMaxElems;
    const uint32_t units = (P.n_list + LPW - 1) / LPW;
    for (uint32_t u = gw; u < units; u += gridDim.x * kWarps) {
        const uint32_t idx = u * LPW + lane;
        const bool live = lane < (uint32_t)LPW && idx < P.n_list;
        uint32_t f = 0, n = 0;
        const uint8_t* gin = nullptr;
        uint8_t* out = nullptr;
        if (live) {
            f = P.frag_list ? P.frag_list[idx] : idx;
            const uint32_t m = P.frag_msg ? P.frag_msg[f] : f;
            const uint32_t k = P.frag_first ? f - P.frag_first[m] : 0u;
            n = min(kFragBytes, P.len[m] - k * kFragBytes);
            gin = P.in + P.off[m] + (size_t)k * kFragBytes;
            out = P.scratch + (P.frag_slot ? P.frag_slot[f] : (uint64_t)f * P.slot_stride);
        }
        __syncwarp();  // the previous unit's tables and bodies are no longer in use
        // ---- stage the unit's bodies, zero its tables (GetHashTable's memset, snappy.cc:269) ----
        const uint64_t ga = reinterpret_cast<uint64_t>(gin);
#pragma unroll 1
        for (uint32_t j = 0; j < (uint32_t)LPW; j++) {
            const uint32_t nj = __shfl_sync(kFullMask, n, j);
            const uint8_t* gj = reinterpret_cast<const uint8_t*>(((uint64_t)__shfl_sync(kFullMask, (uint32_t)(ga >> 32), j) << 32) |
                                                                 __shfl_sync(kFullMask, (uint32_t)ga, j));
            if (nj) stage_body_warp(body + j, gj, nj, lane);
        }
        {
            constexpr uint32_t kVecPerRow = LPW / 4;  // 16-byte stores across this warp's LPW consecutive columns
            for (uint32_t i = lane; i < ls::kTabWords * kVecPerRow; i += 32) {
                const uint32_t row = i / kVecPerRow, c4 = i % kVecPerRow;
                *reinterpret_cast<uint4*>(tab + (size_t)row * ls::kSlots + 4 * c4) = make_uint4(0, 0, 0, 0);
            }
        }
        __syncwarp();
        // ---- phase 1: find ----
        ls::FindLane L;
        L.init(tab + lane, body + lane, my_lists + (size_t)lane * ls::kMaxElems, live ? n : 0u);
        while (__any_sync(kFullMask, L.live != 0)) {
            if (L.live) L.trip();
        }
        __syncwarp();  // the lists written by this warp's lanes are read by all of its lanes below
        // ---- phase 2: encode, one body at a time by the whole warp ----
        const uint64_t oa = reinterpret_cast<uint64_t>(out);
#pragma unroll 1
        for (uint32_t j = 0; j < (uint32_t)LPW; j++) {
            if (!__shfl_sync(kFullMask, (int)live, j)) continue;
            const uint32_t nej = __shfl_sync(kFullMask, L.ne, j), nj = __shfl_sync(kFullMask, n, j),
                           fj = __shfl_sync(kFullMask, f, j);
            uint8_t* o = reinterpret_cast<uint8_t*>(((uint64_t)__shfl_sync(kFullMask, (uint32_t)(oa >> 32), j) << 32) |
                                                    __shfl_sync(kFullMask, (uint32_t)oa, j));
            const InInterleaved in{body + j};
            const uint32_t* el = my_lists + (size_t)j * ls::kMaxElems;
            uint32_t op = 0, next_emit = 0;
            for (uint32_t base = 0; base < nej; base += 32) {
                const bool has = base + lane < nej;
                const uint32_t e = has ? __ldcg(el + base + lane) : 0u;
                uint32_t p, off, mm;
                ls::unpack_element(e, p, off, mm);
                const uint32_t end = p + mm;
                uint32_t lsrc = __shfl_up_sync(kFullMask, end, 1);
                if (lane == 0) lsrc = next_emit;
                next_emit = __shfl_sync(kFullMask, end, min(31u, nej - base - 1));
                op = emit_elements_warp(in, has, lsrc, p, off, mm, o, op, lane);
            }
            op = emit_remainder_warp(in, next_emit, nj, o, op, lane);
            if (lane == 0) P.frag_clen[fj] = op;
        }
    }
}

constexpr size_t kLockstepSmem = (size_t)(kLsSlotsPerCta / ls::kSlots) * ls::kGroupWords * 4;

template <int LPW>
static int configure_lpw(b2c_ctx* ctx) {
    B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_compress_lockstep_kernel<LPW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kLockstepSmem));
    return B2C_OK;
}

int lockstep_kernels_configure(b2c_ctx* ctx) {
    int rc;
    if ((rc = configure_lpw<4>(ctx))) return rc;
    if ((rc = configure_lpw<8>(ctx))) return rc;
    if ((rc = configure_lpw<16>(ctx))) return rc;
    return configure_lpw<32>(ctx);
}

int launch_compress_lockstep(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* scratch,
                             uint32_t* d_frag_clen) {
    LockstepParams P;
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
    const uint32_t lpw = (uint32_t)ctx->lockstep_lpw;
    const uint32_t ctas = (P.n_list + kLsSlotsPerCta - 1) / kLsSlotsPerCta;
    const uint32_t per_sm = ctx->lockstep_global > 0 ? (uint32_t)ctx->lockstep_global : 1u;  // resident CTAs per SM
    const uint32_t grid = std::min<uint32_t>(ctas, (uint32_t)ctx->num_sms * per_sm);
    int rc;
    const size_t list_bytes = (size_t)grid * kLsSlotsPerCta * ls::kMaxElems * 4;
    const size_t scratch_bytes = ctx->lockstep_global > 0 ? (size_t)grid * kLockstepSmem : 0;
    if ((rc = ws_reserve(ctx, lane, lane.ws_tab, list_bytes + scratch_bytes + 256))) return rc;
    P.elists = static_cast<uint32_t*>(lane.ws_tab.p);
    P.gscratch = scratch_bytes ? reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(lane.ws_tab.p) + ((list_bytes + 255) & ~size_t(255))) : nullptr;
    const size_t smem_bytes = scratch_bytes ? 0 : kLockstepSmem;
    const uint32_t threads = 32 * kLsSlotsPerCta / lpw;
    switch (lpw) {
        case 4: snappy_compress_lockstep_kernel<4><<<grid, threads, smem_bytes, lane.stream>>>(P); break;
        case 16: snappy_compress_lockstep_kernel<16><<<grid, threads, smem_bytes, lane.stream>>>(P); break;
        case 32: snappy_compress_lockstep_kernel<32><<<grid, threads, smem_bytes, lane.stream>>>(P); break;
        default: snappy_compress_lockstep_kernel<8><<<grid, threads, smem_bytes, lane.stream>>>(P); break;
    }
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// decompress: 32 streams per warp in lock-step (snappy_lockstep_dec.cuh), output slots <= 1 KiB
// ---------------------------------------------------------------------------------------------------------------
struct LockstepDecodeParams {
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

// One lane writes its own decoded body (n bytes, word w at col[w * 32]) to dst: its shared-memory reads stay in the
// lane's bank, and the 16-byte stores of the 32 lanes go to 32 different bodies -- two consecutive stores of a lane
// fill a 32-byte sector, which L2 merges.
__device__ __forceinline__ void flush_slot_lane(const uint32_t* __restrict__ col, uint8_t* __restrict__ dst, uint32_t n) {
    auto byte_at = [&](uint32_t i) -> uint8_t { return (uint8_t)(col[(i >> 2) * lsd::kSlots] >> ((i & 3) * 8)); };
    uint32_t head = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15)) & 15u;
    if (head > n) head = n;
    for (uint32_t i = 0; i < head; i++) dst[i] = byte_at(i);
    const uint32_t vecs = (n - head) >> 4;
    const uint32_t* w = col + (head >> 2) * lsd::kSlots;
    const uint32_t sh = (head & 3) * 8;
    uint4* d4 = reinterpret_cast<uint4*>(dst + head);
    if (sh == 0) {
        for (uint32_t v = 0; v < vecs; v++) {
            const uint32_t* q = w + (size_t)v * 4 * lsd::kSlots;
            d4[v] = make_uint4(q[0], q[lsd::kSlots], q[2 * lsd::kSlots], q[3 * lsd::kSlots]);
        }
    } else {
        for (uint32_t v = 0; v < vecs; v++) {
            const uint32_t* q = w + (size_t)v * 4 * lsd::kSlots;
            const uint32_t a = q[0], b = q[lsd::kSlots], c = q[2 * lsd::kSlots], d = q[3 * lsd::kSlots], e = q[4 * lsd::kSlots];
            d4[v] = make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh),
                               __funnelshift_r(d, e, sh));
        }
    }
    for (uint32_t i = head + (vecs << 4); i < n; i++) dst[i] = byte_at(i);
}

__global__ void __launch_bounds__(32, 6) snappy_decompress_lockstep_kernel(LockstepDecodeParams P) {
    extern __shared__ __align__(16) uint32_t lsd_smem[];
    const uint32_t lane = threadIdx.x;
    const uint32_t groups = (P.n_list + 31) / 32;
    for (uint32_t g = blockIdx.x; g < groups; g += gridDim.x) {
        const uint32_t idx = g * 32 + lane;
        const bool have = idx < P.n_list;
        uint32_t m = 0;
        lsd::DecodeLane L;
        L.live = 0;
        L.status = B2C_OK;
        L.expected = L.op = L.cur = L.rem = 0;
        L.slot = lsd_smem + lane;
        if (have) {
            m = P.list ? P.list[idx] : idx;
            L.init(P.in + P.in_off[m], P.in_len[m], lsd_smem + lane, min(P.out_cap[m], lsd::kOutMax));
        }
        while (__any_sync(kFullMask, L.live != 0)) {
            if (L.live) L.trip();
        }
        if (have) {
            L.finish();
            P.status[m] = L.status;
            if (P.ulen) P.ulen[m] = L.expected;
            if (L.status == B2C_OK) flush_slot_lane(lsd_smem + lane, P.out + P.out_off[m], L.expected);
        }
    }
}

constexpr size_t kLockstepDecodeSmem = (size_t)lsd::kSlotWords * lsd::kSlots * 4;

int launch_decompress_lockstep(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in, const LayoutView& out,
                               uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status) {
    LockstepDecodeParams P;
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
    if (!ctx->lsd_attr_done) {
        B2C_CUDA(ctx, cudaFuncSetAttribute(snappy_decompress_lockstep_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
        ctx->lsd_attr_done = true;
    }
    const uint32_t groups = (P.n_list + 31) / 32;
    const uint32_t grid = std::min<uint32_t>(groups, (uint32_t)ctx->num_sms * (uint32_t)ctx->lockstep_decode_ctas);
    snappy_decompress_lockstep_kernel<<<grid, 32, kLockstepDecodeSmem, lane.stream>>>(P);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

}  // namespace b2c
