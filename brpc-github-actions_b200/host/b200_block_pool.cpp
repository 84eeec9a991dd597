// b200_block_pool.cpp — see b200_block_pool.hpp.
#include "b200_block_pool.hpp"

#include <algorithm>
#include <atomic>

#include "b200_codec.hpp"

namespace b200 {

namespace {
constexpr size_t kCacheMax = 64;   // blocks a thread keeps for itself
constexpr size_t kCacheFill = 32;  // blocks moved between a thread's cache and the shared list at a time
std::atomic<uint64_t> g_pool_ids{1};

struct ThreadCache {
    uint64_t pool_id = 0;
    PinnedBlockPool* pool = nullptr;
    std::vector<void*> blocks;
    ~ThreadCache();
};
thread_local ThreadCache t_cache;
}  // namespace

PinnedBlockPool::PinnedBlockPool(b2c_ctx* ctx, size_t block_bytes, size_t region_bytes)
    : ctx_(ctx),
      block_bytes_(std::max<size_t>(64, (block_bytes + 63) / 64 * 64)),
      region_bytes_(std::max(region_bytes, block_bytes_)),
      id_(g_pool_ids.fetch_add(1)) {}

PinnedBlockPool::~PinnedBlockPool() {
    if (t_cache.pool_id == id_) {  // this thread's cache must not outlive the regions
        t_cache.blocks.clear();
        t_cache.pool_id = 0;
        t_cache.pool = nullptr;
    }
    for (const Region& r : regions_) b2c_host_free_pinned(ctx_, r.base);
    for (auto& kv : big_) b2c_host_free_pinned(ctx_, kv.first);
}

bool PinnedBlockPool::refill_locked() {
    void* p = nullptr;
    if (b2c_host_alloc_pinned(ctx_, region_bytes_, &p) != B2C_OK || !p) return false;
    regions_.push_back({static_cast<char*>(p), region_bytes_});
    st_.regions++;
    st_.region_bytes += region_bytes_;
    st_.refills++;
    const size_t nblk = region_bytes_ / block_bytes_;
    free_.reserve(free_.size() + nblk);
    for (size_t i = nblk; i-- > 0;) free_.push_back(static_cast<char*>(p) + i * block_bytes_);  // low addresses first out
    return true;
}

void* PinnedBlockPool::Allocate(size_t n) {
    if (n > block_bytes_) {  // oversize: its own page-locked allocation
        void* p = nullptr;
        if (b2c_host_alloc_pinned(ctx_, n, &p) != B2C_OK || !p) return nullptr;
        std::lock_guard<std::mutex> g(mu_);
        big_[p] = n;
        st_.big_out++;
        st_.allocs++;
        return p;
    }
    ThreadCache& tc = t_cache;
    if (tc.pool_id != id_) {  // the cache serves one pool at a time
        if (tc.pool && !tc.blocks.empty()) tc.pool->flush_cache(tc.blocks, 0);
        tc.blocks.clear();
        tc.pool_id = id_;
        tc.pool = this;
    }
    if (tc.blocks.empty()) {
        std::lock_guard<std::mutex> g(mu_);
        if (free_.empty() && !refill_locked()) return nullptr;
        const size_t take = std::min(kCacheFill, free_.size());
        tc.blocks.assign(free_.end() - take, free_.end());
        free_.resize(free_.size() - take);
    }
    {
        std::lock_guard<std::mutex> g(mu_);
        st_.allocs++;
    }
    void* p = tc.blocks.back();
    tc.blocks.pop_back();
    return p;
}

void PinnedBlockPool::flush_cache(std::vector<void*>& cache, size_t keep) {
    std::lock_guard<std::mutex> g(mu_);
    while (cache.size() > keep) {
        free_.push_back(cache.back());
        cache.pop_back();
    }
}

bool PinnedBlockPool::Owns(const void* p) const {
    std::lock_guard<std::mutex> g(mu_);
    for (const Region& r : regions_)
        if (p >= r.base && p < r.base + r.bytes) return true;
    return big_.count(const_cast<void*>(p)) != 0;
}

void PinnedBlockPool::Deallocate(void* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> g(mu_);
        auto it = big_.find(p);
        if (it != big_.end()) {
            big_.erase(it);
            st_.big_out--;
            st_.frees++;
            b2c_host_free_pinned(ctx_, p);
            return;
        }
    }
    ThreadCache& tc = t_cache;
    if (tc.pool_id != id_) {
        if (tc.pool && !tc.blocks.empty()) tc.pool->flush_cache(tc.blocks, 0);
        tc.blocks.clear();
        tc.pool_id = id_;
        tc.pool = this;
    }
    tc.blocks.push_back(p);
    if (tc.blocks.size() > kCacheMax) flush_cache(tc.blocks, kCacheMax - kCacheFill);
    std::lock_guard<std::mutex> g(mu_);
    st_.frees++;
}

PinnedBlockPool::Stats PinnedBlockPool::stats() const {
    std::lock_guard<std::mutex> g(mu_);
    Stats s = st_;
    s.blocks_out = st_.region_bytes / block_bytes_ - free_.size();
    return s;
}

namespace {
ThreadCache::~ThreadCache() {  // a thread that ends gives its cached blocks back
    if (pool && !blocks.empty()) pool->flush_cache(blocks, 0);
}
}  // namespace

PinnedBlockPool& PinnedBlockPool::global() {
    static PinnedBlockPool* pool = new PinnedBlockPool(Context::local().get());  // lives as long as the process
    return *pool;
}
void* PinnedBlockPool::blockmem_allocate(size_t n) { return global().Allocate(n); }
void PinnedBlockPool::blockmem_deallocate(void* p) { global().Deallocate(p); }

}  // namespace b200
