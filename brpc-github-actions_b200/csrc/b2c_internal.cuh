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
constexpr int kNumSM_B200 = 148;

// ---- device workspace -----------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

// ---- layout (host + device mirrors) --------------------------------------
struct Layout {
    b2c_ctx* ctx = nullptr;
    uint32_t n = 0;
    uint64_t payload = 0;      // sum len
    uint64_t arena_end = 0;    // max(off+len)
    uint64_t max_comp = 0;     // sum MaxCompressedLength(len)+5
    std::vector<uint64_t> h_off;
    std::vector<uint32_t> h_len;
    // device copies
    uint64_t* d_off = nullptr;  // [n]
    uint32_t* d_len = nullptr;  // [n]
    // --- snappy fragments (snappy.cc:887-948): fragment f covers message frag_msg[f],
    //     bytes [65536*k, ...) with k = f - frag_first[frag_msg[f]]
    uint32_t n_frags = 0;
    uint32_t* d_frag_msg = nullptr;    // [n_frags]
    uint32_t* d_frag_first = nullptr;  // [n+1]
    uint64_t* d_frag_slot = nullptr;   // [n_frags] byte offset of the fragment's scratch slot
    uint64_t frag_scratch_bytes = 0;
    // fragments grouped by hash-table size class (S: <=1024 entries, M: <=4096, L: <=16384)
    uint32_t n_class[3] = {0, 0, 0};
    uint32_t* d_class_list[3] = {nullptr, nullptr, nullptr};
    // --- decompress classes by slot capacity (this layout used as OUTPUT layout)
    uint32_t n_dclass[3] = {0, 0, 0};
    uint32_t* d_dclass_list[3] = {nullptr, nullptr, nullptr};
    // --- crc chunks: chunk c covers message chunk_msg[c]
    uint32_t n_chunks = 0;
    uint32_t n_multi = 0;              // messages with more than one chunk
    uint32_t* d_chunk_msg = nullptr;   // [n_chunks]
    uint32_t* d_chunk_first = nullptr; // [n+1]
    uint32_t* d_multi_list = nullptr;  // [n_multi] message ids with >1 chunk
    bool all_single_chunk_uniform = false;
};

}  // namespace b2c

struct b2c_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    bool crc_attr_done = false;
    uint64_t last_h2d = 0, last_d2h = 0;
    // CRC tables in device memory
    uint32_t* d_crc_tabA = nullptr;   // [4][256][32] replicated Z^64 tables (128 KiB)
    uint32_t* d_crc_small = nullptr;  // [3][4][256]: Z^4, Z^16, Z^32 slice tables (12 KiB)
    // grow-only device workspace + pinned staging
    b2c::DevBuf ws_frag;      // fragment scratch slots
    b2c::DevBuf ws_misc;      // sizes, scans, chunk remainders
    b2c::DevBuf ws_io_in;     // e2e: device input arena
    b2c::DevBuf ws_io_out;    // e2e: device output arena
    void* h_pin_in = nullptr;
    size_t h_pin_in_cap = 0;
    void* h_pin_out = nullptr;
    size_t h_pin_out_cap = 0;
    // optional kernel timing (b2c_prof_*)
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_ev[B2C_PROF_COUNT];  // start/stop pairs
    uint32_t prof_used[B2C_PROF_COUNT] = {0, 0, 0, 0};
};

struct b2c_layout : b2c::Layout {};

struct b2c_event {
    cudaEvent_t ev = nullptr;
};

namespace b2c {

int set_err(b2c_ctx* ctx, int code, const char* what, cudaError_t e);
int ws_reserve(b2c_ctx* ctx, DevBuf& b, size_t bytes);

#define B2C_CUDA(ctx, call)                                         \
    do {                                                            \
        cudaError_t _e = (call);                                    \
        if (_e != cudaSuccess) return b2c::set_err((ctx), B2C_E_CUDA, #call, _e); \
    } while (0)

// records an event on the ctx stream when profiling is on (start = even slots, stop = odd)
void prof_mark(b2c_ctx* ctx, int which);

// kernels' host launchers (defined in the .cu files)
int crc_tables_init(b2c_ctx* ctx);
int launch_crc32c(b2c_ctx* ctx, const Layout& L, const uint8_t* d_data, const uint32_t* d_init, uint32_t* d_crc);
int launch_compress(b2c_ctx* ctx, const Layout& L, const uint8_t* d_in, uint8_t* d_out, uint64_t out_cap,
                    uint64_t* d_out_off);
int launch_ulen(b2c_ctx* ctx, const Layout& L, const uint8_t* d_in, uint32_t* d_ulen, int32_t* d_status);
int launch_decompress(b2c_ctx* ctx, const Layout& in, const uint8_t* d_in, const Layout& out, uint8_t* d_out,
                      uint32_t* d_ulen, int32_t* d_status);
// exclusive scan of u32 values into u64 offsets: out[0]=0 ... out[n]=total
int launch_scan_u32_u64(b2c_ctx* ctx, const uint32_t* d_vals, uint32_t n, uint64_t* d_out, void* d_tmp /* >= 8*(n/1024+2) */);

}  // namespace b2c

// ---- device helpers ---------------------------------------------------------
#ifdef __CUDACC__
namespace b2c {

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

// 4 bytes at arbitrary byte address p (little endian), using aligned word loads.
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t lo = w[0];
    uint32_t hi = sh ? w[1] : 0u;
    return __funnelshift_r(lo, hi, sh);
}

}  // namespace b2c
#endif
