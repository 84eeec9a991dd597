// b2c_multi.cpp — one batch over several GPUs of a box.
//
// Message bodies are independent (reference snappy.cc:887-948; CRC per message), so a batch shards by a host-side
// split and nothing else: the message list is cut into one contiguous range per device, balanced by bytes, every
// range runs through that device's own ctx on that device's own worker thread (the single-device *_batch entry
// points, unchanged), and results land in the caller's buffers in message order.  No collective, no peer traffic.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "b2c_codec.h"

namespace {

struct Worker {
    int device = 0;
    b2c_ctx* ctx = nullptr;
    std::thread thread;
    std::mutex mu;
    std::condition_variable cv;
    std::function<int(b2c_ctx*)> job;
    bool has_job = false, done = true, stop = false;
    int rc = 0;
    int create_rc = B2C_OK;

    void loop(unsigned host_threads) {
        create_rc = b2c_ctx_create(device, &ctx);  // the ctx lives on the thread that uses it
        if (create_rc == B2C_OK && host_threads) b2c_ctx_set_option(ctx, B2C_OPT_HOST_THREADS, (int)host_threads);
        {
            std::lock_guard<std::mutex> g(mu);
            done = true;
        }
        cv.notify_all();
        for (;;) {
            std::unique_lock<std::mutex> g(mu);
            cv.wait(g, [this] { return has_job || stop; });
            if (stop) break;
            std::function<int(b2c_ctx*)> j = std::move(job);
            has_job = false;
            g.unlock();
            const int r = ctx ? j(ctx) : create_rc;
            g.lock();
            rc = r;
            done = true;
            cv.notify_all();
        }
        if (ctx) b2c_ctx_destroy(ctx);
    }
    void start(std::function<int(b2c_ctx*)> j) {
        std::lock_guard<std::mutex> g(mu);
        job = std::move(j);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> g(mu);
        cv.wait(g, [this] { return done; });
        return rc;
    }
};

}  // namespace

struct b2c_multi {
    std::vector<Worker*> workers;
    std::string err;
    std::vector<uint64_t> prefix;           // message bytes, prefix sums (reused between calls)
    std::vector<std::vector<uint32_t>> sb;  // per shard: seg_begin rebased to the shard's first block
};

namespace {

// fn(lo, hi, t) over [0, n) on up to 8 short-lived threads (the per-message host passes of a million-message batch
// would otherwise cost more than the GPUs need for it)
template <class F>
void pfor(uint32_t n, F fn) {
    unsigned hw = std::thread::hardware_concurrency();
    const unsigned T = n < (1u << 15) ? 1u : std::min(8u, std::max(1u, hw));
    if (T == 1) {
        fn(0u, n, 0u);
        return;
    }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < T; t++)
        th.emplace_back([=, &fn] { fn((uint32_t)((uint64_t)n * t / T), (uint32_t)((uint64_t)n * (t + 1) / T), t); });
    for (auto& x : th) x.join();
}

// Cut [0, n) into `parts` contiguous ranges of about equal payload bytes.  cuts[parts + 1].
void split_by_bytes(b2c_multi* m, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, std::vector<uint32_t>& cuts) {
    const size_t parts = m->workers.size();
    m->prefix.resize((size_t)n + 1);
    uint64_t* p = m->prefix.data();
    p[0] = 0;
    uint64_t part[8] = {0};
    pfor(n, [&](uint32_t lo, uint32_t hi, unsigned t) {  // lengths, and each thread's total
        uint64_t sum = 0;
        for (uint32_t i = lo; i < hi; i++) {
            uint64_t l = 0;
            for (uint32_t s = seg_begin[i]; s < seg_begin[i + 1]; s++) l += segs[s].iov_len;
            p[i + 1] = l + 64;  // + a per-message constant: a million empty bodies still split evenly
            sum += l + 64;
        }
        part[t] = sum;
    });
    uint64_t base[9] = {0};
    for (int t = 0; t < 8; t++) base[t + 1] = base[t] + part[t];
    pfor(n, [&](uint32_t lo, uint32_t hi, unsigned t) {  // prefix sums
        uint64_t run = base[t];
        for (uint32_t i = lo; i < hi; i++) {
            run += p[i + 1];
            p[i + 1] = run;
        }
    });
    cuts.assign(parts + 1, n);
    cuts[0] = 0;
    for (size_t r = 1; r < parts; r++) {
        const uint64_t want = p[n] / parts * r;
        uint32_t c = (uint32_t)(std::lower_bound(p, p + n + 1, want) - p);
        cuts[r] = std::min(std::max(c, cuts[r - 1]), n);
    }
}

// Run fn(shard, ctx, m0, m1, segs of the shard, rebased seg_begin) on every non-empty shard; first error wins.
int run_shards(b2c_multi* m, const b2c_iovec* segs, const uint32_t* seg_begin, const std::vector<uint32_t>& cuts,
               const std::function<int(size_t, b2c_ctx*, uint32_t, uint32_t, const b2c_iovec*, const uint32_t*)>& fn) {
    const size_t parts = m->workers.size();
    m->sb.resize(parts);
    for (size_t r = 0; r < parts; r++) {
        const uint32_t m0 = cuts[r], m1 = cuts[r + 1];
        if (m0 == m1) continue;
        // the shard's block list starts at its first block: its worker re-bases seg_begin itself, then runs
        std::vector<uint32_t>* sbp = &m->sb[r];
        const uint32_t s0 = seg_begin[m0];
        const b2c_iovec* shard_segs = segs + s0;
        m->workers[r]->start([=, &fn](b2c_ctx* ctx) {
            sbp->resize((size_t)(m1 - m0) + 1);
            uint32_t* sb = sbp->data();
            for (uint32_t i = m0; i <= m1; i++) sb[i - m0] = seg_begin[i] - s0;
            return fn(r, ctx, m0, m1, shard_segs, sb);
        });
    }
    int rc = B2C_OK;
    for (size_t r = 0; r < parts; r++) {
        if (cuts[r] == cuts[r + 1]) continue;
        const int q = m->workers[r]->wait();
        if (q != B2C_OK && rc == B2C_OK) {
            rc = q;
            m->err = std::string("device ") + std::to_string(m->workers[r]->device) + ": " +
                     (m->workers[r]->ctx ? b2c_last_error(m->workers[r]->ctx) : "no usable CUDA device");
        }
    }
    return rc;
}

}  // namespace

extern "C" {

int b2c_multi_create(const int* devices, int n_devices, b2c_multi** out) {
    if (!out || !devices || n_devices <= 0 || n_devices > 64) return B2C_E_ARG;
    *out = nullptr;
    b2c_multi* m = new b2c_multi();
    unsigned hw = std::thread::hardware_concurrency();
    if (!hw) hw = 4;
    const unsigned per = std::min(16u, std::max(2u, hw / (unsigned)n_devices));  // the devices share the host's cores
    for (int i = 0; i < n_devices; i++) {
        Worker* w = new Worker();
        w->device = devices[i];
        w->done = false;
        w->thread = std::thread([w, per] { w->loop(per); });
        m->workers.push_back(w);
    }
    int rc = B2C_OK;
    for (Worker* w : m->workers) {
        w->wait();
        if (w->create_rc != B2C_OK && rc == B2C_OK) rc = w->create_rc;
    }
    if (rc != B2C_OK) {
        b2c_multi_destroy(m);
        return rc;
    }
    *out = m;
    return B2C_OK;
}

void b2c_multi_destroy(b2c_multi* m) {
    if (!m) return;
    for (Worker* w : m->workers) {
        {
            std::lock_guard<std::mutex> g(w->mu);
            w->stop = true;
        }
        w->cv.notify_all();
        w->thread.join();
        delete w;
    }
    delete m;
}

int b2c_multi_device_count(const b2c_multi* m) { return m ? (int)m->workers.size() : 0; }
const char* b2c_multi_last_error(const b2c_multi* m) { return m ? m->err.c_str() : ""; }

int b2c_multi_crc32c_batch(b2c_multi* m, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, const uint32_t* init,
                           uint32_t* crc) {
    if (!m || (n && (!segs || !seg_begin || !crc))) return B2C_E_ARG;
    if (!n) return B2C_OK;
    std::vector<uint32_t> cuts;
    split_by_bytes(m, segs, seg_begin, n, cuts);
    return run_shards(m, segs, seg_begin, cuts,
                      [&](size_t, b2c_ctx* ctx, uint32_t m0, uint32_t m1, const b2c_iovec* s, const uint32_t* sb) {
                          return b2c_crc32c_batch(ctx, s, sb, m1 - m0, init ? init + m0 : nullptr, crc + m0);
                      });
}

int b2c_multi_snappy_compress_crc_batch(b2c_multi* m, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                                        uint64_t out_cap, uint64_t* out_off, uint32_t* out_len, uint32_t* crc, int masked) {
    if (!m || !out_off || !out_len || (n && (!segs || !seg_begin || !out))) return B2C_E_ARG;
    if (!n) return B2C_OK;
    // every shard gets the part of `out` that its bodies can need at most; inside it the streams are packed
    std::vector<uint64_t> need(m->workers.size() + 1, 0);
    std::vector<uint32_t> cuts;
    split_by_bytes(m, segs, seg_begin, n, cuts);
    for (size_t r = 0; r < m->workers.size(); r++) {
        uint64_t part_sum[8] = {0};
        const uint32_t c0 = cuts[r];
        pfor(cuts[r + 1] - c0, [&](uint32_t lo, uint32_t hi, unsigned t) {
            uint64_t sum = 0;
            for (uint32_t i = c0 + lo; i < c0 + hi; i++)
                sum += b2c_snappy_max_compressed_length((size_t)(m->prefix[i + 1] - m->prefix[i] - 64)) + 5;
            part_sum[t] = sum;
        });
        uint64_t cap = 0;
        for (uint64_t v : part_sum) cap += v;
        need[r + 1] = need[r] + ((cap + 255) & ~255ull);
    }
    if (need.back() > out_cap) {
        m->err = "b2c_multi_snappy_compress_crc_batch: out_cap below the sum of MaxCompressedLength(len) + 5, rounded up to 256 per device";
        return B2C_E_CAPACITY;
    }
    std::vector<std::vector<uint64_t>> offs(m->workers.size());
    int rc = run_shards(m, segs, seg_begin, cuts,
                        [&](size_t r, b2c_ctx* ctx, uint32_t m0, uint32_t m1, const b2c_iovec* s, const uint32_t* sb) {
                            std::vector<uint64_t>& o = offs[r];
                            o.resize((size_t)(m1 - m0) + 1);
                            uint8_t* base = static_cast<uint8_t*>(out) + need[r];
                            int q = crc ? b2c_snappy_compress_crc_batch(ctx, s, sb, m1 - m0, base, need[r + 1] - need[r], o.data(),
                                                                        crc + m0, masked)
                                        : b2c_snappy_compress_batch(ctx, s, sb, m1 - m0, base, need[r + 1] - need[r], o.data());
                            if (q != B2C_OK) return q;
                            for (uint32_t i = m0; i < m1; i++) {
                                out_off[i] = need[r] + o[i - m0];
                                out_len[i] = (uint32_t)(o[i - m0 + 1] - o[i - m0]);
                            }
                            return B2C_OK;
                        });
    return rc;
}

int b2c_multi_snappy_decompress_verify_batch(b2c_multi* m, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                                             const uint32_t* masked_crc, void* out, uint64_t out_cap, uint64_t* out_off,
                                             int32_t* status) {
    if (!m || !out_off || (n && (!segs || !seg_begin || !status))) return B2C_E_ARG;
    out_off[0] = 0;
    if (!n) return B2C_OK;
    // Decoded sizes come from the preambles (a slot is as long as its stream's preamble claims, exactly as in the
    // single-device call), so every shard's place in `out` is known before anything runs and the result is packed.
    std::vector<uint32_t> cuts;
    split_by_bytes(m, segs, seg_begin, n, cuts);
    const size_t parts = m->workers.size();
    std::vector<uint64_t> base(parts + 1, 0);
    for (size_t r = 0; r < parts; r++) {
        uint64_t part_sum[8] = {0};
        const uint32_t c0 = cuts[r];
        pfor(cuts[r + 1] - c0, [&](uint32_t lo, uint32_t hi, unsigned t) {
            uint64_t total = 0;
            for (uint32_t i = c0 + lo; i < c0 + hi; i++) {
                uint8_t head[5];
                size_t got = 0;
                for (uint32_t q = seg_begin[i]; q < seg_begin[i + 1] && got < 5; q++)
                    for (size_t j = 0; j < segs[q].iov_len && got < 5; j++)
                        head[got++] = static_cast<const uint8_t*>(segs[q].iov_base)[j];
                uint32_t ulen = 0;
                const uint64_t len = m->prefix[i + 1] - m->prefix[i] - 64;
                const bool plausible = b2c_snappy_read_preamble(head, got, &ulen) > 0 && (uint64_t)ulen <= len * 64 / 3 + 64;
                total += plausible ? ulen : 0;
            }
            part_sum[t] = total;
        });
        uint64_t total = 0;
        for (uint64_t v : part_sum) total += v;
        base[r + 1] = base[r] + total;
    }
    if (base.back() > out_cap) {
        m->err = "b2c_multi_snappy_decompress_verify_batch: out_cap too small";
        return B2C_E_CAPACITY;
    }
    std::vector<std::vector<uint64_t>> offs(parts);
    return run_shards(m, segs, seg_begin, cuts,
                      [&](size_t r, b2c_ctx* ctx, uint32_t m0, uint32_t m1, const b2c_iovec* s, const uint32_t* sb) {
                          std::vector<uint64_t>& o = offs[r];
                          o.resize((size_t)(m1 - m0) + 1);
                          uint8_t* dst = static_cast<uint8_t*>(out) + base[r];
                          const uint64_t cap = base[r + 1] - base[r];
                          int q = masked_crc ? b2c_snappy_decompress_verify_batch(ctx, s, sb, m1 - m0, masked_crc + m0, dst, cap,
                                                                                   o.data(), status + m0)
                                             : b2c_snappy_decompress_batch(ctx, s, sb, m1 - m0, dst, cap, o.data(), status + m0);
                          if (q != B2C_OK) return q;
                          for (uint32_t i = m0; i <= m1; i++) out_off[i] = base[r] + o[i - m0];
                          return B2C_OK;
                      });
}

}  // extern "C"
