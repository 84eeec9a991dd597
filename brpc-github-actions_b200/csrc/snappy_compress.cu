// snappy_compress.cu — byte-exact batch snappy compression for sm_100a: the launcher and the passes around the
// fragment compressors.
//
// Replaces butil::snappy::Compress(Source*,Sink*) (reference src/butil/third_party/snappy/snappy.cc:875-954) for
// many messages at once.  The reference compresses every <= 64 KiB fragment independently with a fresh zeroed hash
// table (snappy.cc:887-948), so fragments are the unit of parallelism; internal::CompressFragment
// (snappy.cc:329-453) itself lives in
//   snappy_small.cu    fragments <= 1 KiB: one lane per fragment finds the matches, the warp encodes them
//   snappy_window.cu   larger fragments: one warp per fragment, up to 32 table steps resolved per round
//   snappy_lockstep.cu fragments <= 1 KiB as lock-step state machines (measured alternative, off by default)
// Fragment outputs land in per-fragment scratch slots; here a second pass sizes the messages, a device scan places
// them, and a gather kernel writes varint32(N) || fragments back to back (snappy.cc:878-881).
#include <algorithm>

#include "b2c_internal.cuh"

namespace b2c {

static constexpr uint32_t kFull = 0xffffffffu;

__device__ __forceinline__ uint32_t varint_len(uint32_t v) {
    return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : v < (1u << 28) ? 4 : 5;
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
        // every fragment class has its kernel above
        if ((rc = launch_compress_window(ctx, lane, L, c, d_in, scratch, d_frag_clen))) return rc;
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
