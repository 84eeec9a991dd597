// b200_aggregator.cpp — see b200_aggregator.hpp.
#include "b200_aggregator.hpp"

#include <stdlib.h>

#include <chrono>

namespace b200 {

BatchAggregator::BatchAggregator(const Options& opt) : opt_(opt), thread_([this] { worker(); }) {}

BatchAggregator::~BatchAggregator() {
    {
        std::lock_guard<std::mutex> g(mu_);
        stop_ = true;
    }
    queue_cv_.notify_all();
    thread_.join();
}

BatchAggregator::Stats BatchAggregator::stats() const {
    std::lock_guard<std::mutex> g(mu_);
    return stats_;
}

void BatchAggregator::submit(Request* r) {
    std::unique_lock<std::mutex> g(mu_);
    r->done = false;
    queue_.push_back(r);
    stats_.calls++;
    queue_cv_.notify_one();
    done_cv_.wait(g, [r] { return r->done; });
}

bool BatchAggregator::Compress(const BlockList& in, std::string* out) {
    Request r{kCompress, &in, out, 0, 0, false, false};
    submit(&r);
    return r.ok;
}
bool BatchAggregator::Decompress(const BlockList& in, std::string* out) {
    Request r{kDecompress, &in, out, 0, 0, false, false};
    submit(&r);
    return r.ok;
}
uint32_t BatchAggregator::Crc32c(const BlockList& in, uint32_t init_crc) {
    Request r{kCrc, &in, nullptr, init_crc, 0, false, false};
    submit(&r);
    return r.crc;
}

void BatchAggregator::run(Op op, std::vector<Request*>& reqs) {
    if (reqs.empty()) return;
    std::vector<BlockList> in;
    in.reserve(reqs.size());
    for (Request* r : reqs) in.push_back(*r->in);
    if (op == kCompress) {
        std::vector<std::string> out;
        policy::SnappyCompressBatch(in, &out);
        for (size_t i = 0; i < reqs.size(); i++) {
            reqs[i]->out->append(out[i]);
            reqs[i]->ok = !out[i].empty();
        }
    } else if (op == kDecompress) {
        std::vector<std::string> out;
        std::vector<bool> ok;
        policy::SnappyDecompressBatch(in, &out, &ok);
        for (size_t i = 0; i < reqs.size(); i++) {
            reqs[i]->ok = ok[i];
            if (ok[i]) reqs[i]->out->append(out[i]);
        }
    } else {
        std::vector<uint32_t> init(reqs.size()), crc;
        for (size_t i = 0; i < reqs.size(); i++) init[i] = reqs[i]->init_crc;
        policy::Crc32cBatch(in, init.data(), &crc);
        for (size_t i = 0; i < reqs.size(); i++) reqs[i]->crc = crc[i];
    }
}

void BatchAggregator::worker() {
    if (opt_.device != 0) {
        // the codec context of this thread binds to $B200_DEVICE on first use
        setenv("B200_DEVICE", std::to_string(opt_.device).c_str(), 1);
    }
    std::vector<Request*> batch, by_op[3];
    for (;;) {
        {
            std::unique_lock<std::mutex> g(mu_);
            queue_cv_.wait(g, [this] { return stop_ || !queue_.empty(); });
            if (stop_ && queue_.empty()) return;
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
            stats_.batches++;
            if (batch.size() > stats_.largest_batch) stats_.largest_batch = batch.size();
        }
        for (auto& v : by_op) v.clear();
        for (Request* r : batch) by_op[r->op].push_back(r);
        for (int op = 0; op < 3; op++) run(static_cast<Op>(op), by_op[op]);
        {
            std::lock_guard<std::mutex> g(mu_);
            for (Request* r : batch) r->done = true;
        }
        done_cv_.notify_all();
    }
}

}  // namespace b200
