// b200_block_pool.cpp — see b200_block_pool.hpp.
#include "b200_block_pool.hpp"

#include <algorithm>
#include <atomic>
#include <new>

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
    for (const SlabRec& r : slab_free_) b2c_host_free_pinned(ctx_, r.base);
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
    allocs_.fetch_add(1, std::memory_order_relaxed);
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
    frees_.fetch_add(1, std::memory_order_relaxed);
}

PinnedBlockPool::Stats PinnedBlockPool::stats() const {
    std::lock_guard<std::mutex> g(mu_);
    Stats s = st_;
    s.allocs += allocs_.load(std::memory_order_relaxed);
    s.frees += frees_.load(std::memory_order_relaxed);
    s.blocks_out = st_.region_bytes / block_bytes_ - free_.size();
    return s;
}

// ---- result slabs ---------------------------------------------------------------------------------------------
namespace {
constexpr size_t kSlabHeader = 64;  // keeps the payload 64-byte aligned
constexpr size_t kSlabMin = 256u << 10;
constexpr size_t kSlabCacheMax = 16;
struct SlabHeader {
    std::atomic<long> refs;
    size_t cap;  // bytes after the header
    PinnedBlockPool* pool;
};
static_assert(sizeof(SlabHeader) <= kSlabHeader, "header must fit");
// live slabs of all pools by base address, so that a slice's deleter, which is given nothing but the slice's data
// pointer, finds the slab it belongs to
std::mutex g_slab_mu;
std::vector<std::pair<char*, size_t>> g_slabs;  // sorted by base; (base of payload, cap)
}  // namespace

char* PinnedBlockPool::SlabAcquire(size_t bytes) {
    size_t want = kSlabMin;
    while (want < bytes) want <<= 1;
    char* raw = nullptr;
    size_t cap = 0;
    {
        std::lock_guard<std::mutex> g(mu_);
        size_t best = slab_free_.size();
        for (size_t i = 0; i < slab_free_.size(); i++)
            if (slab_free_[i].cap >= bytes && (best == slab_free_.size() || slab_free_[i].cap < slab_free_[best].cap)) best = i;
        if (best != slab_free_.size()) {
            raw = slab_free_[best].base;
            cap = slab_free_[best].cap;
            slab_free_.erase(slab_free_.begin() + best);
        }
    }
    if (!raw) {
        void* p = nullptr;
        if (b2c_host_alloc_pinned(ctx_, want + kSlabHeader, &p) != B2C_OK || !p) return nullptr;
        raw = static_cast<char*>(p);
        cap = want;
    }
    SlabHeader* h = new (raw) SlabHeader;
    h->refs.store(1, std::memory_order_relaxed);
    h->cap = cap;
    h->pool = this;
    char* payload = raw + kSlabHeader;
    {
        std::lock_guard<std::mutex> g(g_slab_mu);
        g_slabs.insert(std::upper_bound(g_slabs.begin(), g_slabs.end(), std::make_pair(payload, (size_t)0)),
                       std::make_pair(payload, cap));
    }
    std::lock_guard<std::mutex> g(mu_);
    slabs_live_++;
    return payload;
}

void PinnedBlockPool::slab_retire(char* payload, size_t cap) {
    {
        std::lock_guard<std::mutex> g(g_slab_mu);
        auto it = std::lower_bound(g_slabs.begin(), g_slabs.end(), std::make_pair(payload, (size_t)0));
        if (it != g_slabs.end() && it->first == payload) g_slabs.erase(it);
    }
    char* raw = payload - kSlabHeader;
    bool keep;
    {
        std::lock_guard<std::mutex> g(mu_);
        slabs_live_--;
        keep = slab_free_.size() < kSlabCacheMax;
        if (keep) slab_free_.push_back({raw, cap});
    }
    if (!keep) b2c_host_free_pinned(ctx_, raw);
}

PinnedBlockPool::Slice PinnedBlockPool::SlabSlice(char* slab, size_t at, size_t n) {
    SlabHeader* h = reinterpret_cast<SlabHeader*>(slab - kSlabHeader);
    h->refs.fetch_add(1, std::memory_order_relaxed);
    Slice s;
    s.data = slab + at;
    s.size = n;
    s.deleter = &PinnedBlockPool::ReleaseSlice;
    s.meta = reinterpret_cast<uint64_t>(slab);
    return s;
}

void PinnedBlockPool::SlabRelease(char* slab) {
    SlabHeader* h = reinterpret_cast<SlabHeader*>(slab - kSlabHeader);
    if (h->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) h->pool->slab_retire(slab, h->cap);
}

void PinnedBlockPool::ReleaseSlice(void* data) {
    char* p = static_cast<char*>(data);
    char* slab = nullptr;
    {
        std::lock_guard<std::mutex> g(g_slab_mu);
        auto it = std::upper_bound(g_slabs.begin(), g_slabs.end(), std::make_pair(p, ~(size_t)0));
        if (it != g_slabs.begin()) {
            --it;
            if (p >= it->first && p <= it->first + it->second) slab = it->first;  // an empty slice may sit at the very end
        }
    }
    if (!slab) return;  // not ours: nothing sane to do
    SlabHeader* h = reinterpret_cast<SlabHeader*>(slab - kSlabHeader);
    if (h->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) h->pool->slab_retire(slab, h->cap);
}

size_t PinnedBlockPool::slabs_live() const {
    std::lock_guard<std::mutex> g(mu_);
    return slabs_live_;
}

namespace {
ThreadCache::~ThreadCache() {  // a thread that ends gives its cached blocks back
    if (pool && !blocks.empty()) pool->flush_cache(blocks, 0);
}
}  // namespace

PinnedBlockPool& PinnedBlockPool::global() {
    // pool and the context it allocates through live as long as the process (not as long as the first calling thread)
    static PinnedBlockPool* pool = new PinnedBlockPool((new Context(Context::default_device()))->get());
    return *pool;
}
void* PinnedBlockPool::blockmem_allocate(size_t n) { return global().Allocate(n); }
void PinnedBlockPool::blockmem_deallocate(void* p) { global().Deallocate(p); }

}  // namespace b200
