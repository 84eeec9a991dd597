// b200_block_pool.hpp — page-locked block pool for IOBuf.
//
// bRPC allocates every IOBuf::Block through two function pointers,
//     void* (*butil::iobuf::blockmem_allocate)(size_t)   and   void (*blockmem_deallocate)(void*)
// (src/butil/iobuf.cpp:163-170; default malloc/free, block size 8192: iobuf.h:67), and its RDMA transport already
// swaps them for a registered-memory pool (src/brpc/rdma/rdma_helper.cpp:528-531, rdma/block_pool.cpp).  This is the
// same hook for the GPU codec: blocks come from cudaHostAlloc'ed regions, so message bodies sit in page-locked
// memory from the moment they are received, and a ctx with B2C_OPT_INPUT_PINNED lets the device pull them over
// PCIe itself — no host thread copies them into staging.
//
//     butil::iobuf::blockmem_allocate   = b200::PinnedBlockPool::blockmem_allocate;
//     butil::iobuf::blockmem_deallocate = b200::PinnedBlockPool::blockmem_deallocate;
//
// Regions are never returned to the driver while the pool lives (page-locking is expensive; RDMA's pool does the
// same); blocks are recycled through per-thread caches in front of one shared free list.  Destroy a pool only after the
// threads that used it have ended or moved to another pool.
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <atomic>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "b2c_codec.h"

namespace b200 {

class PinnedBlockPool {
public:
    struct Stats {
        size_t regions = 0, region_bytes = 0;  // page-locked memory held
        size_t blocks_out = 0;                 // pooled blocks outside the shared free list (handed out, or in a thread's cache)
        size_t big_out = 0;                    // oversize allocations currently handed out
        uint64_t allocs = 0, frees = 0, refills = 0;
    };

    // block_bytes: what one Allocate(n <= block_bytes) returns (IOBuf's default block is 8192 bytes);
    // region_bytes: how much is page-locked at a time.
    explicit PinnedBlockPool(b2c_ctx* ctx, size_t block_bytes = 8192, size_t region_bytes = 64u << 20);
    ~PinnedBlockPool();
    PinnedBlockPool(const PinnedBlockPool&) = delete;
    PinnedBlockPool& operator=(const PinnedBlockPool&) = delete;

    void* Allocate(size_t n);   // nullptr when page-locked memory cannot be had
    void Deallocate(void* p);   // p from Allocate of this pool
    bool Owns(const void* p) const;
    Stats stats() const;

    // ---- result slabs ------------------------------------------------------------------------------------------
    // One page-locked buffer that the results of a whole BATCH are DMA'd into; every result is then handed to its
    // owner as a slice of it, which an IOBuf adopts without a copy:
    //     out->append_user_data_with_meta(slice.data, slice.size, slice.deleter, slice.meta);   // iobuf.cpp:1216-1240
    // (meta = the slab's base address, as the RDMA transport stores its registration key there,
    // src/brpc/rdma/rdma_helper.cpp:528-531).  The slab goes back to the pool's slab cache when the batch's creator
    // and every slice have let go of it, in whatever order and on whatever thread.  Slabs are cached and re-used:
    // page-locking memory costs about a millisecond.
    struct Slice {
        char* data = nullptr;
        size_t size = 0;
        void (*deleter)(void*) = nullptr;  // call with `data` when done (IOBuf's user-data deleter signature)
        uint64_t meta = 0;
    };
    char* SlabAcquire(size_t bytes);              // page-locked, >= bytes; nullptr when it cannot be had
    Slice SlabSlice(char* slab, size_t at, size_t n);  // adds a reference
    void SlabRelease(char* slab);                 // the creator's reference
    static void ReleaseSlice(void* data);         // Slice::deleter: finds the slab that holds `data`
    size_t slabs_live() const;                    // acquired and not yet fully released
    void flush_cache(std::vector<void*>& cache, size_t keep);  // thread caches -> shared list (used at thread exit)

    // The process-wide pool behind the two hook functions (created on first use with a context of its own on
    // Context::default_device(), 8192-byte blocks).
    static PinnedBlockPool& global();
    static void* blockmem_allocate(size_t n);
    static void blockmem_deallocate(void* p);

private:
    struct Region {
        char* base;
        size_t bytes;
    };
    bool refill_locked();

    b2c_ctx* ctx_;
    const size_t block_bytes_, region_bytes_;
    mutable std::mutex mu_;
    std::vector<Region> regions_;
    std::vector<void*> free_;                     // shared free list
    std::unordered_map<void*, size_t> big_;       // oversize allocations
    Stats st_;
    std::atomic<uint64_t> allocs_{0}, frees_{0};  // counted without the pool mutex: the thread caches exist to avoid it
    struct SlabRec {
        char* base;
        size_t cap;
    };
    std::vector<SlabRec> slab_free_;              // cached slabs, any sizes
    size_t slabs_live_ = 0;
    void slab_retire(char* base, size_t cap);
    const uint64_t id_;                           // distinguishes pools in the per-thread cache
};

}  // namespace b200
