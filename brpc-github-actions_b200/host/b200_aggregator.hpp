// b200_aggregator.hpp — batching aggregator behind the per-message handler API.
//
// bRPC's CompressHandler is one message per synchronous call (src/brpc/compress.h:28-39), made
// from one bthread per in-flight message (src/brpc/input_messenger.cpp:169-189).  A GPU codec
// only pays off on batches, so this class sits between the two: concurrent callers enqueue
// their message and block; one worker thread drains the queue into b2c_*_batch calls
// (the queue/worker shape follows bthread::ExecutionQueue,
// src/bthread/execution_queue.cpp).  Callers wait on a condition variable; inside bRPC swap
// Waiter for bthread::ConditionVariable so that a waiting bthread yields its worker pthread.
#pragma once

#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "b200_codec.hpp"

namespace b200 {

struct AggregatorOptions {
    int device = 0;
    uint32_t max_batch = 8192;   // messages per GPU call
    uint32_t max_wait_us = 100;  // how long the first message of a batch may wait for company
};

class BatchAggregator {
public:
    using Options = AggregatorOptions;
    struct Stats {
        uint64_t calls = 0, batches = 0, largest_batch = 0;
    };

    explicit BatchAggregator(const Options& opt = Options());
    ~BatchAggregator();
    BatchAggregator(const BatchAggregator&) = delete;
    BatchAggregator& operator=(const BatchAggregator&) = delete;

    // Same meaning as brpc::policy::SnappyCompress / SnappyDecompress(const IOBuf&, IOBuf*) and
    // butil::crc32c::Extend, callable from any number of threads.
    bool Compress(const BlockList& in, std::string* out);
    bool Decompress(const BlockList& in, std::string* out);
    uint32_t Crc32c(const BlockList& in, uint32_t init_crc = 0);

    Stats stats() const;

private:
    enum Op { kCompress = 0, kDecompress = 1, kCrc = 2 };
    struct Request {
        Op op;
        const BlockList* in;
        std::string* out;
        uint32_t init_crc;
        uint32_t crc;
        bool ok;
        bool done;
    };
    void submit(Request* r);
    void worker();
    void run(Op op, std::vector<Request*>& reqs);

    Options opt_;
    mutable std::mutex mu_;
    std::condition_variable queue_cv_;  // worker waits for work
    std::condition_variable done_cv_;   // callers wait for results
    std::deque<Request*> queue_;
    bool stop_ = false;
    Stats stats_;
    std::thread thread_;
};

}  // namespace b200
