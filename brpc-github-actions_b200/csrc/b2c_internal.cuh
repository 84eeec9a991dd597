// b2c_internal.cuh — shared declarations of the libb2c.so translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "b2c_codec.h"

namespace b2c {

constexpr uint32_t kFragBytes = 65536;  // snappy.h:202-203 kBlockSize
constexpr uint32_t kCrcChunk = 4096;    // bytes per CRC work item (one quad of lanes)
constexpr int kPipeLanes = 4;           // staging slots of the end-to-end pipeline

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
};

// One stream with its private workspaces: kernels of different lanes may overlap.
struct Lane {
    cudaStream_t stream = nullptr;
    DevBuf ws_frag;   // fragment scratch slots (compress)
    DevBuf ws_misc;   // sizes, scans, chunk remainders
    DevBuf ws_tab;    // hash tables of the 64 KiB-fragment compressor (one 32 KiB table per resident warp)
    DevBuf ws_ctr;    // 64 work counters handed out in turn to kernels whose warps pull their work items
    uint32_t ctr_next = 0;
    DevBuf d_in;      // e2e: device input arena (+ layout blob)
    DevBuf d_out;     // e2e: device output arena (+ result metadata)
    PinBuf h_in;      // e2e: pinned staging, host -> device
    PinBuf h_out;     // e2e: pinned staging, device -> host
    cudaEvent_t ev_meta = nullptr;  // result metadata has reached h_out
    cudaEvent_t ev_done = nullptr;  // everything of this slab has reached h_out
};

// What the kernels consume: device pointers of one batch shape.  Arrays that are the
// identity for this batch are nullptr (fragment f == message f, chunk c == message c, ...).
struct LayoutView {
    uint32_t n = 0;
    uint64_t payload = 0;    // sum len
    uint64_t arena_end = 0;  // max(off+len)
    uint64_t max_comp = 0;   // sum MaxCompressedLength(len)+5
    const uint64_t* d_off = nullptr;  // [n]
    const uint32_t* d_len = nullptr;  // [n]
    // snappy fragments (snappy.cc:887-948): message m owns max(1, ceil(len/65536)) fragments
    uint32_t n_frags = 0;
    const uint32_t* d_frag_msg = nullptr;    // [n_frags]   null: f == m
    const uint32_t* d_frag_first = nullptr;  // [n+1]       null: first[m] == m
    const uint64_t* d_frag_slot = nullptr;   // [n_frags]   null: slot = f * slot_stride
    uint64_t slot_stride = 0;
    uint64_t frag_scratch_bytes = 0;
    // fragments by hash-table size class (<=1024 / <=4096 / <=16384 entries); null list: identity
    uint32_t n_class[3] = {0, 0, 0};
    const uint32_t* d_class_list[3] = {nullptr, nullptr, nullptr};
    // messages by output-slot capacity class (<=1 KiB / <=8 KiB / larger; layout used as decompress OUTPUT)
    uint32_t n_dclass[3] = {0, 0, 0};
    const uint32_t* d_dclass_list[3] = {nullptr, nullptr, nullptr};
    // crc chunks of kCrcChunk bytes: message m owns max(1, ceil(len/4096)) chunks
    uint32_t n_chunks = 0;
    uint32_t n_multi = 0;                      // messages with more than one chunk
    const uint32_t* d_chunk_msg = nullptr;     // [n_chunks]  null: c == m
    const uint32_t* d_chunk_first = nullptr;   // [n+1]       null: first[m] == m
    const uint32_t* d_multi_list = nullptr;    // [n_multi]
};

// What one pass over (off, len) has to know to size a layout; partial sums add up (merge).
struct PlanStats {
    uint64_t payload = 0, arena_end = 0, max_comp = 0, slot_sum = 0, F = 0, C = 0;
    uint32_t max_len = 0, n_multi = 0, n_class[3] = {0, 0, 0}, n_dclass[3] = {0, 0, 0};
    void merge(const PlanStats& p);
};
// accumulate messages [lo, hi); off may be null (arena_end is then left to the caller)
void plan_range(const uint64_t* off, const uint32_t* len, uint32_t lo, uint32_t hi, PlanStats& st);

// Host-side derivation of a LayoutView's arrays into one contiguous blob.
struct LayoutPlan {
    uint32_t n = 0;
    const uint64_t* off = nullptr;
    const uint32_t* len = nullptr;
    LayoutView v;  // counts and scalars filled by plan(); pointers by bind()
    bool has_frag_arrays = false, has_slot_array = false, has_class_lists = false, has_dclass_lists = false,
         has_chunk_arrays = false;
    // blob offsets (bytes)
    size_t o_off = 0, o_slot = 0, o_len = 0, o_frag_msg = 0, o_frag_first = 0, o_class[3] = {0, 0, 0},
           o_dclass[3] = {0, 0, 0}, o_chunk_msg = 0, o_chunk_first = 0, o_multi = 0;
    size_t blob_bytes = 0;

    // `par`: when non-null, large batches are scanned with the ctx's host thread pool
    void plan(uint32_t n, const uint64_t* off, const uint32_t* len, b2c_ctx* par = nullptr);
    void plan_with(uint32_t n, const uint64_t* off, const uint32_t* len, const PlanStats& tot);  // stats made elsewhere
    // upper bound of blob_bytes for any layout of n messages totalling `payload` bytes
    static size_t max_blob_bytes(uint32_t n, uint64_t payload) {
        const uint64_t F = n + payload / kFragBytes, C = n + payload / kCrcChunk;
        return (size_t)(28ull * n + 16 * F + 4 * C + 512);
    }
    void fill(uint8_t* blob) const;                           // blob_bytes of host memory
    void bind(const uint8_t* d_blob, LayoutView* out) const;  // device address of the uploaded blob
};

}  // namespace b2c

struct b2c_ctx {
    int device = 0;
    int num_sms = 148;
    b2c::Lane main;                    // *_dev entry points
    b2c::Lane pipe[b2c::kPipeLanes];   // *_batch entry points (slab pipeline)
    std::string err;
    uint64_t launches = 0;
    bool crc_attr_done = false;
    uint64_t last_h2d = 0, last_d2h = 0;
    uint32_t* d_crc_tabA = nullptr;   // [4][256][32] replicated Z^64 tables (128 KiB)
    uint32_t* d_crc_small = nullptr;  // [3][4][256]: Z^4, Z^16, Z^32 slice tables (12 KiB)
    // optional kernel timing (b2c_prof_*), main lane only
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_ev[B2C_PROF_COUNT];  // start/stop pairs
    uint32_t prof_used[B2C_PROF_COUNT] = {0, 0, 0, 0};
    // reusable host arrays of the *_batch entry points (avoid page-faulting fresh vectors per call)
    std::vector<uint64_t> bp_off, tmp_off;
    std::vector<uint32_t> bp_len, tmp_cap, tmp_cap2;
    int gtab_warps_per_sm = 16;  // B2C_GTAB_WPS: resident warps per SM of the HBM-table compressor (0 = shared-memory tables); 16 keeps all tables L2-resident
    bool small_kernels = true;  // B2C_SMALL_KERNELS=0 routes <=1 KiB bodies through the warp kernels
    int small_lpw_compress = 8, small_lpw_decompress = 8;  // bodies per warp (B2C_SMALL_LPW_C / _D: 2 (compress only), 4, 8, 16, 32); measured with 3 CTAs of 32 bodies per SM: 4: 10.5, 8: 10.0, 16: 12.9 ms per GiB of `words`
    // tunables (environment: B2C_SLAB_MB, B2C_HOST_THREADS, B2C_TRACE)
    size_t slab_bytes = 32u << 20;
    bool slab_fixed = false;  // B2C_SLAB_MB given: no adaptation to the batch's message sizes
    unsigned host_threads = 0;  // 0: cores / ranks on this node, 2..16 (B2C_OPT_HOST_THREADS, B2C_HOST_THREADS)
    void* host_pool = nullptr;  // this ctx's gather / copy-out threads (HostPool, b2c_api.cu), created on first use
    bool trace = false;
    bool input_pinned = false;  // B2C_OPT_INPUT_PINNED: *_batch inputs are page-locked, gather them on the device
    bool ring_decode = false;        // B2C_RING_DECODE=1: unsplittable large slots decode through a 64 KiB smem ring
    bool lift_scan = true;           // B2C_LIFT_SCAN=0: find the 64 KiB block boundaries with the tag-by-tag scan
    bool block_decode = true;        // B2C_BLOCK_DECODE=0: never split large streams into 64 KiB blocks
    bool ring_attr_done = false;
    bool seg_dp_attr_done = false;
    bool gather_attr_done = false;
    bool lsd_attr_done = false;
    bool lockstep_decode = false;  // B2C_LOCKSTEP_D=1: slots <= 1 KiB decode with the lock-step state machine of snappy_lockstep_dec.cuh (measured slower than the lane-per-stream loops of snappy_small.cu: 6.8 vs 5.2 ms per GiB of 1 KiB `words` bodies)
    int lockstep_decode_ctas = 6;  // B2C_LOCKSTEP_D_CTAS: resident one-warp CTAs per SM of the lock-step decoder
    int gather_warps = 8, gather_chunk = 2048;  // B2C_GATHER_WARPS / B2C_GATHER_CHUNK: warps per CTA and bytes per round and warp
    int gather_ctas_per_sm = 1;  // B2C_GATHER_CTAS: CTAs per SM of the device-side gather of pinned input blocks
    bool batch_decode = true;        // B2C_BATCH_DECODE=0: one tag per step in the warp decoder
    bool batch_decode_small = false; // B2C_BATCH_DECODE_SMALL=1: <=1 KiB slots use the warp batch decoder, not one lane per stream
    bool nt_gather = true;  // B2C_NT_GATHER=0: plain memcpy into staging
    bool async_gather = true;  // B2C_ASYNC_GATHER=0: the submitting thread waits for each slab's gather before anything else
    int window_warps_per_sm = 40;  // B2C_WINDOW_WPS: resident warps per SM of the orbit kernel on 64 KiB fragments (tables in HBM/L2; B2C_GTAB_WPS=0 keeps them in shared memory)
    bool window_stage = true;  // B2C_WINDOW_STAGE=0: the window compressor reads <=4 KiB fragments from HBM instead of staging them
    bool window_smem64 = false;  // B2C_WINDOW_SMEM64=1: 64 KiB fragments keep table + TMA-staged fragment in shared memory (one warp per CTA, 2 CTAs per SM); measured 7.6x slower than HBM/L2 tables with 40 warps per SM
    bool small_lockstep = false;  // B2C_LOCKSTEP=1: bodies <= 1 KiB compress with the branch-free lock-step state machines of snappy_lockstep.cu (measured slower than the lane-per-body loops of snappy_small.cu: 15.5 vs 11.0 ms per GiB of 1 KiB `words` bodies)
    int lockstep_global = 0;     // tunable lockstep_global: > 0 = the lock-step compressor keeps tables and staged bodies in global memory (L2), that many 64-body CTAs resident per SM
    int lockstep_lpw = 8;        // B2C_LOCKSTEP_LPW: bodies per warp of the lock-step compressor (4, 8, 16, 32)
    int window_mask = 6;    // B2C_WINDOW: bit c set = fragment class c (0: <=1 KiB, 1: <=4 KiB, 2: <=64 KiB) compresses with the orbit kernel (snappy_window.cu)
};

struct b2c_layout {
    b2c_ctx* ctx = nullptr;
    b2c::LayoutView v;
    void* d_blob = nullptr;
};

struct b2c_event {
    cudaEvent_t ev = nullptr;
};

namespace b2c {

int set_err(b2c_ctx* ctx, int code, const char* what, cudaError_t e);
int ws_reserve(b2c_ctx* ctx, Lane& lane, DevBuf& b, size_t bytes);
// A zeroed u32 in device memory for ONE launch on lane.stream (warps take work items with atomicAdd instead of a fixed
// share: items differ in size, and many are skipped).  64 are in rotation; stream order keeps a reused one safe.
int next_work_counter(b2c_ctx* ctx, Lane& lane, uint32_t** ctr);

#define B2C_CUDA(ctx, call)                                         \
    do {                                                            \
        cudaError_t _e = (call);                                    \
        if (_e != cudaSuccess) return b2c::set_err((ctx), B2C_E_CUDA, #call, _e); \
    } while (0)

// records an event on the main stream when profiling is on (start = even slots, stop = odd)
void prof_mark(b2c_ctx* ctx, Lane& lane, int which);

// kernels' host launchers (defined in the .cu files); all work goes to lane.stream
int crc_tables_init(b2c_ctx* ctx);
int launch_crc32c(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_data, const uint32_t* d_init,
                  uint32_t* d_crc);
int launch_compress(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* d_out,
                    uint64_t out_cap, uint64_t* d_out_off);
int launch_ulen(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_ulen,
                int32_t* d_status);
int launch_decompress(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in, const LayoutView& out,
                      uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status);
int launch_prefix(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_out5);
int launch_validate(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint32_t* d_ulen,
                    int32_t* d_status);
int launch_decompress_verify(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in,
                             const LayoutView& out, uint8_t* d_out, const uint32_t* d_masked_crc, uint32_t* d_ulen,
                             int32_t* d_status);
// thread-per-message kernels for bodies <= 1 KiB (snappy_small.cu)
int small_kernels_configure(b2c_ctx* ctx);
int launch_compress_small(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* scratch,
                          uint32_t* d_frag_clen);
int launch_decompress_small(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in,
                            const LayoutView& out, uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status);
// host_gather.cpp: one message's blocks concatenated into staging (non-temporal whole-line stores when `stream`)
void gather_message(uint8_t* dst, const b2c_iovec* segs, uint32_t nseg, uint64_t total, bool stream);
void gather_fence();
// lock-step state machines, 32 bodies <= 1 KiB per warp (snappy_lockstep.cu)
int lockstep_kernels_configure(b2c_ctx* ctx);
int launch_compress_lockstep(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_in, uint8_t* scratch,
                             uint32_t* d_frag_clen);
int launch_decompress_lockstep(b2c_ctx* ctx, Lane& lane, const LayoutView& in, const uint8_t* d_in,
                               const LayoutView& out, uint8_t* d_out, uint32_t* d_ulen, int32_t* d_status);
// window-parallel warp-per-fragment compressor (snappy_window.cu); c = fragment class 0..2
int window_kernels_configure(b2c_ctx* ctx);
int launch_compress_window(b2c_ctx* ctx, Lane& lane, const LayoutView& L, int c, const uint8_t* d_in, uint8_t* scratch,
                           uint32_t* d_frag_clen);
// exclusive scan of u32 values into u64 offsets: out[0]=0 ... out[n]=total
int launch_scan_u32_u64(b2c_ctx* ctx, Lane& lane, const uint32_t* d_vals, uint32_t n, uint64_t* d_out,
                        void* d_tmp /* >= 8*(n/1024+2) */);

}  // namespace b2c

// ---- device helpers ---------------------------------------------------------
#ifdef __CUDACC__
namespace b2c {

// 4 bytes at arbitrary byte address p (little endian), using aligned word loads.
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t lo = w[0];
    uint32_t hi = sh ? w[1] : 0u;
    return __funnelshift_r(lo, hi, sh);
}

// n bytes from src to dst by one warp, any alignment on either side; nothing outside [dst, dst + n) is written and
// only aligned words that hold at least one byte of [src, src + n) are read:
// bytes up to dst's next 16-byte boundary, then aligned 16-byte stores whose four words are assembled from aligned
// source words (one 16-byte load when source and destination have the same phase), then the tail bytes.  This is
// what moves long literals (incompressible bodies are ONE literal per 64 KiB fragment): 512 bytes per warp
// instruction instead of 32.  The ranges must not overlap.
__device__ __forceinline__ void warp_copy_bytes(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t n,
                                                uint32_t lane) {
    if (n < 64) {
        for (uint32_t i = lane; i < n; i += 32) dst[i] = src[i];
        return;
    }
    const uint32_t head = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15)) & 15u;
    if (lane < head) dst[lane] = src[lane];
    const uint32_t vecs = (n - head) >> 4;
    uint4* d4 = reinterpret_cast<uint4*>(dst + head);
    const uint8_t* s = src + head;
    const uintptr_t sa = reinterpret_cast<uintptr_t>(s);
    if ((sa & 15) == 0) {
        const uint4* s4 = reinterpret_cast<const uint4*>(s);
        for (uint32_t i = lane; i < vecs; i += 32) d4[i] = s4[i];
    } else {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(sa & ~uintptr_t(3));
        const uint32_t sh = (uint32_t)(sa & 3) * 8;
        if (sh == 0) {
            for (uint32_t i = lane; i < vecs; i += 32) {
                const uint32_t* q = w + 4 * i;
                d4[i] = make_uint4(q[0], q[1], q[2], q[3]);
            }
        } else {
            // the fifth word of a vector holds that vector's last bytes in its low part
            for (uint32_t i = lane; i < vecs; i += 32) {
                const uint32_t* q = w + 4 * i;
                const uint32_t a = q[0], b = q[1], c = q[2], d = q[3], e = q[4];
                d4[i] = make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh),
                                   __funnelshift_r(d, e, sh));
            }
        }
    }
    const uint32_t done = head + (vecs << 4);
    if (lane < n - done) dst[done + lane] = src[done + lane];
}

}  // namespace b2c
#endif
