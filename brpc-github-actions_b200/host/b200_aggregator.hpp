// b200_aggregator.hpp — batching aggregator behind the per-message handler API (header-only).
//
// bRPC's CompressHandler is one message per synchronous call (src/brpc/compress.h:28-39), made from one bthread
// per in-flight message (src/brpc/input_messenger.cpp:169-189).  A GPU codec only pays off on batches, so this
// class sits between the two: concurrent callers enqueue their message and wait; one worker thread drains the
// queue into b2c_*_batch calls (the queue/worker shape follows bthread::ExecutionQueue,
// src/bthread/execution_queue.cpp).
//
//  * Results come back as SLICES OF ONE PAGE-LOCKED SLAB per batch (PinnedBlockPool::SlabAcquire): the library
//    DMAs the whole batch's output straight into the slab and every caller adopts its slice without a copy
//    (IOBuf::append_user_data_with_meta, src/butil/iobuf.cpp:1216-1240).
//  * The wait of a caller is a template parameter: std::mutex / std::condition_variable here; inside bRPC
//    instantiate BasicBatchAggregator<BthreadSync> so that a waiting bthread yields its worker
//    pthread (src/bthread/condition_variable.h:44) instead of parking it.
//  * Nothing the GPU path throws or returns leaves the worker: every request of a failed batch completes with
//    `infra_error` set, and the caller falls back to the CPU codec (the bRPC shim does) — a GPU fault must not
//    take the RPC process down or leave callers waiting forever.
#pragma once

#include <string.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <exception>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "b200_block_pool.hpp"
#include "b200_codec.hpp"

namespace b200 {

struct AggregatorOptions {
    int device = 0;              // the worker's codec context binds to this device (no environment variables involved)
    uint32_t max_batch = 8192;   // messages per GPU call
    uint32_t max_wait_us = 100;  // how long the first message of a batch may wait for company
    bool pooled_output = true;   // results as slices of a page-locked slab (false: heap buffers, one copy)
    uint32_t workers = 2;        // batches in flight: each worker thread has its own codec context (own streams and
                                 // staging), so the next batch forms and starts while the last one is on the GPU
};

struct StdSync {  // how a CALLER waits for its result
    using Mutex = std::mutex;
    using CondVar = std::condition_variable;
};

// One result: bytes owned by whoever holds the OutSlice, or by an IOBuf after adopt().
class OutSlice {
public:
    OutSlice() = default;
    ~OutSlice() { reset(); }
    OutSlice(const OutSlice&) = delete;
    OutSlice& operator=(const OutSlice&) = delete;
    const char* data() const { return s_.data; }
    size_t size() const { return s_.size; }
    void (*deleter() const)(void*) { return s_.deleter; }
    uint64_t meta() const { return s_.meta; }
    // ownership moves to the caller, who must call deleter()(data()) once (what an IOBuf does when its last
    // reference to the user data drops)
    PinnedBlockPool::Slice adopt() {
        PinnedBlockPool::Slice s = s_;
        s_ = PinnedBlockPool::Slice();
        return s;
    }
    void reset() {
        if (s_.data && s_.deleter) s_.deleter(s_.data);
        s_ = PinnedBlockPool::Slice();
    }
    void set(const PinnedBlockPool::Slice& s) {
        reset();
        s_ = s;
    }

private:
    PinnedBlockPool::Slice s_;
};

template <class Sync = StdSync>
class BasicBatchAggregator {
public:
    using Options = AggregatorOptions;
    struct Stats {
        uint64_t calls = 0, batches = 0, largest_batch = 0, failed_batches = 0;
    };

    explicit BasicBatchAggregator(const Options& opt = Options()) : opt_(opt) {
        for (uint32_t i = 0; i < (opt_.workers ? opt_.workers : 1u); i++) threads_.emplace_back([this] { worker(); });
    }
    ~BasicBatchAggregator() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        queue_cv_.notify_all();
        for (std::thread& t : threads_) t.join();
    }
    BasicBatchAggregator(const BasicBatchAggregator&) = delete;
    BasicBatchAggregator& operator=(const BasicBatchAggregator&) = delete;

    // Same meaning as brpc::policy::SnappyCompress / SnappyDecompress(const IOBuf&, IOBuf*) and
    // butil::crc32c::Extend, callable from any number of threads.  `infra_error` (nullable) is set when the GPU
    // path itself failed (no device, out of memory, a CUDA error): the result is then false / 0 and says nothing
    // about the data — run the CPU codec instead.
    bool Compress(const BlockList& in, OutSlice* out, bool* infra_error = nullptr) {
        Request r(kCompress, &in);
        submit(&r);
        if (infra_error) *infra_error = r.infra;
        if (r.ok) out->set(r.slice);
        return r.ok;
    }
    bool Decompress(const BlockList& in, OutSlice* out, bool* infra_error = nullptr) {
        Request r(kDecompress, &in);
        submit(&r);
        if (infra_error) *infra_error = r.infra;
        if (r.ok) out->set(r.slice);
        return r.ok;
    }
    uint32_t Crc32c(const BlockList& in, uint32_t init_crc = 0, bool* infra_error = nullptr) {
        Request r(kCrc, &in);
        r.init_crc = init_crc;
        submit(&r);
        if (infra_error) *infra_error = r.infra;
        return r.crc;
    }
    // std::string conveniences (one copy), appending like the reference's IOBuf forms
    bool Compress(const BlockList& in, std::string* out, bool* infra_error = nullptr) {
        OutSlice s;
        if (!Compress(in, &s, infra_error)) return false;
        out->append(s.data(), s.size());
        return true;
    }
    bool Decompress(const BlockList& in, std::string* out, bool* infra_error = nullptr) {
        OutSlice s;
        if (!Decompress(in, &s, infra_error)) return false;
        out->append(s.data(), s.size());
        return true;
    }

    Stats stats() const {
        std::lock_guard<std::mutex> g(mu_);
        return stats_;
    }

private:
    enum Op { kCompress = 0, kDecompress = 1, kCrc = 2 };
    struct Request {
        Op op;
        const BlockList* in;
        uint32_t init_crc = 0, crc = 0;
        bool ok = false, infra = false, done = false;
        PinnedBlockPool::Slice slice;
        typename Sync::Mutex mu;  // the caller waits here, with the caller's kind of wait
        typename Sync::CondVar cv;
        Request(Op o, const BlockList* i) : op(o), in(i) {}
    };

    void submit(Request* r) {
        {
            std::lock_guard<std::mutex> g(mu_);
            queue_.push_back(r);
            stats_.calls++;
        }
        queue_cv_.notify_one();
        std::unique_lock<typename Sync::Mutex> g(r->mu);
        while (!r->done) r->cv.wait(g);
    }

    static void heap_free(void* p) { delete[] static_cast<char*>(p); }

    // the batch's output buffer: a pool slab (page-locked: the library DMAs into it) or, without the pool, heap memory
    struct OutBuf {
        PinnedBlockPool* pool = nullptr;
        char* slab = nullptr;
        std::vector<char> heap;
        char* data = nullptr;
        ~OutBuf() {
            if (slab) pool->SlabRelease(slab);
        }
        bool acquire(PinnedBlockPool* p, size_t bytes) {
            if (p && (slab = p->SlabAcquire(bytes))) {
                pool = p;
                data = slab;
                return true;
            }
            heap.resize(bytes ? bytes : 1);
            data = heap.data();
            return true;
        }
        PinnedBlockPool::Slice slice(size_t at, size_t n) {
            if (slab) return pool->SlabSlice(slab, at, n);
            PinnedBlockPool::Slice s;
            s.data = new char[n ? n : 1];
            memcpy(s.data, data + at, n);
            s.size = n;
            s.deleter = &heap_free;
            return s;
        }
    };

    void run(Context& ctx, PinnedBlockPool* pool, Op op, std::vector<Request*>& reqs) {
        if (reqs.empty()) return;
        const uint32_t n = (uint32_t)reqs.size();
        std::vector<b2c_iovec> segs;
        segs.reserve(reqs.size() + 1);  // never a null array, even for a batch of empty bodies
        std::vector<uint32_t> begin(1, 0u);
        for (Request* r : reqs) {
            segs.insert(segs.end(), r->in->begin(), r->in->end());
            begin.push_back((uint32_t)segs.size());
        }
        auto bytes_of = [](const BlockList& b) {
            size_t t = 0;
            for (const b2c_iovec& v : b) t += v.iov_len;
            return t;
        };
        auto check = [&](const char* what, int rc) {
            if (rc != B2C_OK) throw std::runtime_error(std::string(what) + ": " + b2c_last_error(ctx.get()));
        };
        std::vector<uint64_t> off(n + 1);
        if (op == kCompress) {
            uint64_t cap = 0;
            for (Request* r : reqs) cap += b2c_snappy_max_compressed_length(bytes_of(*r->in)) + 5;
            OutBuf ob;
            ob.acquire(pool, cap);
            check("b2c_snappy_compress_batch",
                  b2c_snappy_compress_batch(ctx.get(), segs.data(), begin.data(), n, ob.data, cap, off.data()));
            for (uint32_t i = 0; i < n; i++) {
                reqs[i]->slice = ob.slice(off[i], off[i + 1] - off[i]);
                reqs[i]->ok = true;  // the reference returns `written`, which is never 0 (snappy_compress.cpp:51-55)
            }
        } else if (op == kDecompress) {
            uint64_t cap = 64;  // what the (lenient) preambles claim, bounded by what streams of these sizes can hold
            for (Request* r : reqs) {
                uint8_t head[5];
                size_t got = 0;
                for (const b2c_iovec& v : *r->in)
                    for (size_t j = 0; j < v.iov_len && got < 5; j++) head[got++] = static_cast<const uint8_t*>(v.iov_base)[j];
                uint32_t ulen = 0;
                if (b2c_snappy_read_preamble(head, got, &ulen)) cap += std::min<uint64_t>(ulen, (uint64_t)bytes_of(*r->in) * 64 / 3 + 64);
            }
            OutBuf ob;
            ob.acquire(pool, cap);
            std::vector<int32_t> st(n);
            check("b2c_snappy_decompress_batch",
                  b2c_snappy_decompress_batch(ctx.get(), segs.data(), begin.data(), n, ob.data, cap, off.data(), st.data()));
            for (uint32_t i = 0; i < n; i++) {
                reqs[i]->ok = st[i] == B2C_OK;
                if (reqs[i]->ok) reqs[i]->slice = ob.slice(off[i], off[i + 1] - off[i]);
            }
        } else {
            std::vector<uint32_t> init(n), crc(n);
            for (uint32_t i = 0; i < n; i++) init[i] = reqs[i]->init_crc;
            check("b2c_crc32c_batch", b2c_crc32c_batch(ctx.get(), segs.data(), begin.data(), n, init.data(), crc.data()));
            for (uint32_t i = 0; i < n; i++) {
                reqs[i]->crc = crc[i];
                reqs[i]->ok = true;
            }
        }
    }

    void worker() {
        std::unique_ptr<Context> ctx;  // created here: a ctx belongs to the thread that uses it
        std::unique_ptr<PinnedBlockPool> pool;
        std::vector<Request*> batch, by_op[3];
        for (;;) {
            {
                std::unique_lock<std::mutex> g(mu_);
                queue_cv_.wait(g, [this] { return stop_ || !queue_.empty(); });
                if (stop_ && queue_.empty()) break;
                // let the batch fill: wait a little for company unless it is already full
                const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(opt_.max_wait_us);
                while (queue_.size() < opt_.max_batch && !stop_) {
                    if (queue_cv_.wait_until(g, deadline) == std::cv_status::timeout) break;
                }
                batch.clear();
                while (!queue_.empty() && batch.size() < opt_.max_batch) {
                    batch.push_back(queue_.front());
                    queue_.pop_front();
                }
                if (batch.empty()) continue;  // another worker took them all
                stats_.batches++;
                if (batch.size() > stats_.largest_batch) stats_.largest_batch = batch.size();
            }
            for (auto& v : by_op) v.clear();
            for (Request* r : batch) by_op[r->op].push_back(r);
            for (int op = 0; op < 3; op++) {
                try {
                    if (!ctx) {
                        ctx.reset(new Context(opt_.device));
                        if (opt_.pooled_output) pool.reset(new PinnedBlockPool(ctx->get()));
                    }
                    run(*ctx, pool.get(), static_cast<Op>(op), by_op[op]);
                } catch (...) {  // no device, out of memory, a CUDA error: the callers fall back, nobody hangs
                    for (Request* r : by_op[op]) {
                        if (r->slice.data && r->slice.deleter) r->slice.deleter(r->slice.data);
                        r->slice = PinnedBlockPool::Slice();
                        r->ok = false;
                        r->crc = 0;
                        r->infra = true;
                    }
                    std::lock_guard<std::mutex> g(mu_);
                    stats_.failed_batches++;
                }
            }
            for (Request* r : batch) {
                // after `done` is set the caller may destroy the request: notify under its lock
                std::lock_guard<typename Sync::Mutex> g(r->mu);
                r->done = true;
                r->cv.notify_one();
            }
        }
        // slices still held by callers keep their slab alive through the pool: give them a moment, and if some are
        // still out (an IOBuf somewhere holds one) let pool and context live on rather than pull memory from under them
        if (pool) {
            for (int i = 0; i < 1000 && pool->slabs_live(); i++) std::this_thread::sleep_for(std::chrono::milliseconds(1));
            if (pool->slabs_live()) {
                pool.release();
                ctx.release();
            }
        }
    }

    Options opt_;
    mutable std::mutex mu_;             // the queue: held for a few instructions at a time
    std::condition_variable queue_cv_;  // the worker waits for work
    std::deque<Request*> queue_;
    bool stop_ = false;
    Stats stats_;
    std::vector<std::thread> threads_;  // last member: started when everything above exists
};

using BatchAggregator = BasicBatchAggregator<StdSync>;

}  // namespace b200
