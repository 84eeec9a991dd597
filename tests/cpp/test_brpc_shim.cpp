// brpc_gpu_snappy.cpp (the COMPRESS_TYPE_SNAPPY handler bodies) compiled against the stand-in headers of
// tests/cpp/mock_brpc/ and driven the way bRPC drives a CompressHandler: many threads, one message per call.
// Checked against the CPU oracle: compressed bytes, decoded bytes, accept/reject of damaged streams, and that
// results are adopted zero-copy from the page-locked result slabs (meta != 0, deleters run exactly once).
#ifndef B200_WITH_BRPC
#define B200_WITH_BRPC 1
#endif
#include <stdio.h>
#include <stdlib.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "b2o.h"
#include "brpc/policy/snappy_compress.h"
#include "brpc_gpu_snappy.h"

#define CHECK(c)                                                      \
    do {                                                              \
        if (!(c)) {                                                   \
            fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #c); \
            exit(1);                                                  \
        }                                                             \
    } while (0)

static std::string oracle_compress(const std::string& s) {
    std::string z(b2o_snappy_max_compressed_length(s.size()), '\0');
    z.resize(b2o_snappy_compress(reinterpret_cast<const uint8_t*>(s.data()), s.size(), reinterpret_cast<uint8_t*>(&z[0])));
    return z;
}

// the "stock CPU functions" of the mock tree: the oracle
namespace brpc {
namespace policy {
bool SnappyCompress(const butil::IOBuf& in, butil::IOBuf* out) {
    out->append(oracle_compress(in.to_string()));
    return true;
}
bool SnappyDecompress(const butil::IOBuf& in, butil::IOBuf* out) {
    const std::string z = in.to_string();
    uint32_t ulen = 0;
    if (!b2o_snappy_read_ulen_lenient(reinterpret_cast<const uint8_t*>(z.data()), z.size(), &ulen)) return false;
    if (ulen > z.size() * 22 + 64) return false;
    std::string s(ulen ? ulen : 1, '\0');
    uint32_t got = 0;
    if (!b2o_snappy_uncompress(reinterpret_cast<const uint8_t*>(z.data()), z.size(), reinterpret_cast<uint8_t*>(&s[0]), ulen, &got))
        return false;
    out->append(s.data(), ulen);
    return true;
}
}  // namespace policy
}  // namespace brpc

struct EchoMessage : google::protobuf::Message {
    std::string payload;
    bool SerializeToString(std::string* out) const override {
        *out = payload;
        return true;
    }
    bool ParseFromString(const std::string& in) override {
        payload = in;
        return true;
    }
};

static std::string body(int seed, size_t n) {
    std::string s(n, '\0');
    b2o_corpus_fill(B2O_WORDS, (uint64_t)seed, reinterpret_cast<uint8_t*>(&s[0]), n);
    return s;
}

int main() {
    using namespace brpc::policy;
    SetGpuSnappyMinBytes(0);  // everything through the GPU path
    const int kThreads = 32, kPer = 30;
    std::atomic<int> bad(0), metas(0);
    std::vector<std::thread> th;
    for (int t = 0; t < kThreads; t++)
        th.emplace_back([&, t] {
            for (int i = 0; i < kPer; i++) {
                const std::string s = body(t * 1000 + i, (size_t)((t * 7919 + i * 104729) % 150000));
                butil::IOBuf in, z, back;
                in.append(s);  // 8160-byte blocks, like IOBuf's own
                if (!GpuSnappyCompress(in, &z) || z.to_string() != oracle_compress(s)) bad++;
                if (z.get_first_data_meta() != 0) metas++;  // adopted from a result slab, not copied
                butil::IOBuf zin;
                zin.append(z.to_string());
                if (!GpuSnappyDecompress(zin, &back) || back.to_string() != s) bad++;
                if (i % 6 == 0 && z.size() > 8) {  // a damaged stream fails alone, exactly when the reference rejects it
                    std::string cut = z.to_string();
                    cut.resize(cut.size() - 3);
                    butil::IOBuf cin, junk, want_out;
                    cin.append(cut);
                    const bool want = SnappyDecompress(cin, &want_out);
                    if (GpuSnappyDecompress(cin, &junk) != want) bad++;
                }
                if (i % 10 == 0) {  // the Message overloads (serialize -> compress, decompress -> parse)
                    EchoMessage m, r;
                    m.payload = s;
                    butil::IOBuf zb;
                    if (!GpuSnappyCompress(m, &zb) || zb.to_string() != oracle_compress(s)) bad++;
                    if (!GpuSnappyDecompress(zb, &r) || r.payload != s) bad++;
                }
            }
        });
    for (auto& x : th) x.join();
    CHECK(bad.load() == 0);
    unsigned long long gpu = 0, cpu = 0;
    GpuSnappyCounters(&gpu, &cpu);
    printf("shim: %llu calls on the GPU path, %llu on the CPU path, %d results adopted from result slabs\n", gpu, cpu, metas.load());
    CHECK(cpu == 0 && gpu >= (unsigned long long)kThreads * kPer * 2);
    CHECK(metas.load() >= kThreads * kPer - 32);  // empty bodies aside, every compressed result is a slab slice
    // small bodies stay on the CPU path
    SetGpuSnappyMinBytes(16 * 1024);
    butil::IOBuf in, z;
    in.append(body(5, 1000));
    CHECK(GpuSnappyCompress(in, &z) && z.to_string() == oracle_compress(body(5, 1000)));
    GpuSnappyCounters(&gpu, &cpu);
    CHECK(cpu == 1);
    printf("all checks passed\n");
    return 0;
}
