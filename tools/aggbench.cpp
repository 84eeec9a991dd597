// aggbench.cpp — what ONE caller sees behind the CompressHandler binding: T threads (bRPC workers) each compress /
// decompress one body at a time through the process-wide BasicBatchAggregator, exactly as host/brpc_gpu_snappy.cpp
// does, results handed back as slices of the batch's page-locked slab.  Prints calls/s, payload MB/s, mean / p50 / p99
// latency of a call and the mean batch the worker formed, per thread count and body size.
// Build (one line): g++ -std=c++17 -O2 -I include -I brpc-github-actions_b200/host tools/aggbench.cpp
//   brpc-github-actions_b200/host/b200_codec.cpp brpc-github-actions_b200/host/b200_block_pool.cpp
//   -L brpc-github-actions_b200 -lb2c -Wl,-rpath,$PWD/brpc-github-actions_b200 -lpthread -o /tmp/aggbench
// Run: /tmp/aggbench [seconds per point]
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "b200_aggregator.hpp"
#include "b200_codec.hpp"

using b200::BlockList;
using Clock = std::chrono::steady_clock;

// text-like bodies: words of a small vocabulary, the shape of the repository's `words` corpus (not the corpus itself:
// nothing here is compared with the checker; the parity of this path is tests/cpp/test_host_api.cpp)
static std::string make_body(size_t n, uint32_t seed) {
    static const char* const vocab[] = {"user", "id=", "status:ok ", "error ", "0123", "request ", "response ", "name",
                                        "value", "{\"k\":", "\"v\"}, ", "timestamp", "2024-", "GET /", "POST /", "\n"};
    std::string s;
    s.reserve(n + 16);
    uint32_t x = seed * 2654435761u + 12345u;
    while (s.size() < n) {
        x = x * 1664525u + 1013904223u;
        s += vocab[(x >> 24) & 15];
        if (((x >> 8) & 7) == 0) s += std::to_string(x & 0xffff);
    }
    s.resize(n);
    return s;
}

static BlockList blocks_of(const std::string& s, size_t block) {
    BlockList b;
    for (size_t p = 0; p < s.size(); p += block) b.push_back({s.data() + p, std::min(block, s.size() - p)});
    if (s.empty()) b.push_back({s.data(), 0});
    return b;
}

struct Point {
    double calls_s, mbs, mean_us, p50_us, p99_us, mean_batch;
};

static Point run(b200::BatchAggregator& agg, int threads, size_t body_bytes, bool decompress, double seconds) {
    std::atomic<bool> stop(false);
    std::vector<std::vector<float>> lat(threads);
    std::vector<uint64_t> calls(threads, 0);
    const b200::BatchAggregator::Stats before = agg.stats();
    std::vector<std::thread> th;
    const Clock::time_point t0 = Clock::now();
    for (int t = 0; t < threads; t++) {
        th.emplace_back([&, t] {
            std::vector<std::string> bodies, packed;
            for (int i = 0; i < 16; i++) bodies.push_back(make_body(body_bytes, t * 100 + i));
            if (decompress)
                for (const std::string& b : bodies) {
                    std::string z;
                    if (!agg.Compress(blocks_of(b, 8192), &z)) abort();
                    packed.push_back(z);
                }
            lat[t].reserve(1 << 16);
            for (uint64_t i = 0; !stop.load(std::memory_order_relaxed); i++) {
                const std::string& in = decompress ? packed[i & 15] : bodies[i & 15];
                const BlockList bl = blocks_of(in, 8192);
                b200::OutSlice out;
                const Clock::time_point a = Clock::now();
                const bool ok = decompress ? agg.Decompress(bl, &out) : agg.Compress(bl, &out);
                const Clock::time_point b = Clock::now();
                if (!ok) abort();
                lat[t].push_back(std::chrono::duration<float, std::micro>(b - a).count());
                calls[t]++;
            }
        });
    }
    std::this_thread::sleep_for(std::chrono::duration<double>(seconds));
    stop = true;
    for (auto& x : th) x.join();
    const double el = std::chrono::duration<double>(Clock::now() - t0).count();
    const b200::BatchAggregator::Stats after = agg.stats();
    std::vector<float> all;
    uint64_t n = 0;
    for (int t = 0; t < threads; t++) {
        all.insert(all.end(), lat[t].begin(), lat[t].end());
        n += calls[t];
    }
    std::sort(all.begin(), all.end());
    double sum = 0;
    for (float v : all) sum += v;
    Point p;
    p.calls_s = n / el;
    p.mbs = n * (double)body_bytes / el / 1e6;
    p.mean_us = all.empty() ? 0 : sum / all.size();
    p.p50_us = all.empty() ? 0 : all[all.size() / 2];
    p.p99_us = all.empty() ? 0 : all[(size_t)(all.size() * 0.99)];
    p.mean_batch = (double)(after.calls - before.calls) / std::max<uint64_t>(1, after.batches - before.batches);
    return p;
}

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? atof(argv[1]) : 1.5;
    for (uint32_t wait_us : {100u, 20u}) {
        b200::BatchAggregator::Options opt;
        opt.max_wait_us = wait_us;
        if (getenv("AGG_WORKERS")) opt.workers = (uint32_t)atoi(getenv("AGG_WORKERS"));
        b200::BatchAggregator agg(opt);
        for (size_t body : {(size_t)1024, (size_t)16384}) {
            for (int decompress = 0; decompress < 2; decompress++) {
                for (int threads : {1, 8, 32, 128, 512}) {
                    const Point p = run(agg, threads, body, decompress != 0, seconds);
                    printf("{\"workers\": %u, \"max_wait_us\": %u, \"op\": \"%s\", \"body_bytes\": %zu, \"caller_threads\": %d, \"calls_per_s\": %.0f, "
                           "\"payload_MBps\": %.1f, \"latency_us\": {\"mean\": %.0f, \"p50\": %.0f, \"p99\": %.0f}, "
                           "\"mean_batch\": %.1f}\n",
                           opt.workers, wait_us, decompress ? "decompress" : "compress", body, threads, p.calls_s, p.mbs, p.mean_us, p.p50_us,
                           p.p99_us, p.mean_batch);
                    fflush(stdout);
                }
            }
        }
    }
    return 0;
}
