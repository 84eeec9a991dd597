// test_host_api.cpp — the reference's own unit tests for this path, re-pointed at the GPU
// codec through the C++ host mirror (b200_codec.hpp):
//   test/crc32c_unittest.cc:18-70                      -> crc_*
//   test/brpc_snappy_compress_unittest.cpp:97-104      -> snappy_iobuf
//   test/brpc_snappy_compress_unittest.cpp:137-164     -> snappy_test
//   test/brpc_snappy_compress_unittest.cpp:166-234     -> throughput shapes (round trip + bytes vs oracle)
//   test/brpc_snappy_compress_unittest.cpp:236-255     -> mass_snappy_iobuf
// plus byte-for-byte comparison of every compressed stream with the CPU oracle.
// Exit code 0 = all passed.  (TEST INFRASTRUCTURE: links oracle/libb2oracle.so as the checker.)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <atomic>
#include <thread>

#include "b200_aggregator.hpp"
#include "b200_block_pool.hpp"
#include "b200_codec.hpp"
#include "b200_framing.hpp"
#include "b2o.h"

static int g_fail = 0;
#define CHECK(cond)                                                         \
    do {                                                                    \
        if (!(cond)) {                                                      \
            fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            g_fail++;                                                       \
        }                                                                   \
    } while (0)

using b200::BlockList;

static BlockList blocks_of(const std::string& s, size_t block) {
    BlockList b;
    for (size_t p = 0; p < s.size(); p += block) b.push_back({s.data() + p, std::min(block, s.size() - p)});
    if (s.empty()) b.push_back({s.data(), 0});
    return b;
}

static std::string oracle_compress(const std::string& s) {
    std::string z(b2o_snappy_max_compressed_length(s.size()), '\0');
    size_t k = b2o_snappy_compress(reinterpret_cast<const uint8_t*>(s.data()), s.size(), reinterpret_cast<uint8_t*>(&z[0]));
    z.resize(k);
    return z;
}

static std::string pattern_az09(int len) {  // brpc_snappy_compress_unittest.cpp:110-117
    std::string t;
    for (int j = 0; j < len;) {
        for (int i = 0; i < 26 && j < len; i++, j++) t.push_back('a' + i);
        for (int i = 0; i < 10 && j < len; i++, j++) t.push_back('0' + i);
    }
    return t;
}

static void crc_standard_results() {  // CRC.StandardResults, rfc3720 section B.4
    using namespace b200::crc32c;
    char buf[32];
    memset(buf, 0, sizeof(buf));
    CHECK(0x8a9136aaU == Value(buf, sizeof(buf)));
    memset(buf, 0xff, sizeof(buf));
    CHECK(0x62a8ab43U == Value(buf, sizeof(buf)));
    for (int i = 0; i < 32; i++) buf[i] = i;
    CHECK(0x46dd794eU == Value(buf, sizeof(buf)));
    for (int i = 0; i < 32; i++) buf[i] = 31 - i;
    CHECK(0x113fdb5cU == Value(buf, sizeof(buf)));
    unsigned char data[48] = {
        0x01, 0xc0, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00,
        0x14, 0x00, 0x00, 0x00, 0x00, 0x00, 0x04, 0x00, 0x00, 0x00, 0x00, 0x14, 0x00, 0x00, 0x00, 0x18,
        0x28, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x02, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00,
    };
    CHECK(0xd9963a56 == Value(reinterpret_cast<char*>(data), sizeof(data)));
}

static void crc_values_extend_mask() {  // CRC.Values / Extend / Mask
    using namespace b200::crc32c;
    CHECK(Value("a", 1) != Value("foo", 3));
    CHECK(Value("hello world", 11) == Extend(Value("hello ", 6), "world", 5));
    uint32_t crc = Value("foo", 3);
    CHECK(crc != Mask(crc));
    CHECK(crc != Mask(Mask(crc)));
    CHECK(crc == Unmask(Mask(crc)));
    CHECK(crc == Unmask(Unmask(Mask(Mask(crc)))));
}

static void snappy_iobuf() {  // test_compress_method.snappy_iobuf
    const std::string test = "this is a test";
    std::string output_buf, check_buf;
    CHECK(b200::policy::SnappyCompress(blocks_of(test, 8192), &output_buf));
    CHECK(output_buf == oracle_compress(test));
    CHECK(b200::policy::SnappyDecompress(blocks_of(output_buf, 8192), &check_buf));
    CHECK(check_buf == test);
}

static void snappy_test() {  // test_compress_method.snappy_test: string API, concatenated streams
    std::string text = pattern_az09(200);
    std::string output, append_string;
    CHECK(b200::snappy::Compress(text.data(), text.size(), &output));
    CHECK(output == oracle_compress(text));
    size_t com_len1 = output.size();
    const char* s_text = "123456";
    CHECK(b200::snappy::Compress(s_text, strlen(s_text), &append_string));
    output.append(append_string);
    std::string uncompress_str, uncompress_str_t;
    CHECK(b200::snappy::Uncompress(output.data(), com_len1, &uncompress_str));
    CHECK(uncompress_str == text);
    CHECK(b200::snappy::Uncompress(append_string.data(), append_string.size(), &uncompress_str_t));
    CHECK(uncompress_str_t == s_text);
    // the concatenation as a whole is NOT one valid stream (trailing bytes): reference returns false
    std::string whole;
    CHECK(!b200::snappy::Uncompress(output.data(), output.size(), &whole));
    CHECK(!b200::snappy::IsValidCompressedBuffer(output.data(), output.size()));
    CHECK(b200::snappy::IsValidCompressedBuffer(output.data(), com_len1));
}

static void throughput_shapes() {  // throughput_compare / throughput_compare_complete_random shapes
    const char str_table[] = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789";
    int len_subs[] = {128, 1024, 16 * 1024, 32 * 1024, 512 * 1024};
    srand(1);
    for (int len : len_subs) {
        std::string dup = pattern_az09(len), rnd;
        for (int j = 0; j < len; j++) rnd.push_back(str_table[rand() % 62]);
        for (const std::string* text : {&dup, &rnd}) {
            std::string z, back;
            CHECK(b200::policy::SnappyCompress(blocks_of(*text, 8160), &z));  // 8160-byte IOBuf payload blocks
            CHECK(z == oracle_compress(*text));
            size_t ulen = 0;
            CHECK(b200::snappy::GetUncompressedLength(z.data(), z.size(), &ulen) && ulen == text->size());
            CHECK(b200::policy::SnappyDecompress(blocks_of(z, 100), &back));
            CHECK(back == *text);
            CHECK(b200::crc32c::Value(text->data(), text->size()) ==
                  b2o_crc32c_value(reinterpret_cast<const uint8_t*>(text->data()), text->size()));
        }
    }
}

static void mass_snappy_iobuf() {  // test_compress_method.mass_snappy_iobuf
    std::string text;
    for (int j = 0; j < 782;)
        for (int i = 0; i < 26 && j < 782; i++, j++) text.push_back('a' + i);
    std::string output_buf;
    CHECK(b200::policy::SnappyCompress(blocks_of(text, 8192), &output_buf));
    CHECK(output_buf == oracle_compress(text));
    size_t decompress_length = 0;
    CHECK(b200::snappy::GetUncompressedLength(output_buf.data(), output_buf.size(), &decompress_length));
    std::vector<char> output(decompress_length);
    CHECK(b200::snappy::RawUncompress(output_buf.data(), output_buf.size(), output.data()));
    CHECK(std::string(output.data(), output.size()) == text);
}

static void batch_calls() {
    std::vector<std::string> bodies;
    for (int i = 0; i < 300; i++) {
        std::string s((size_t)(i * 37 % 5000), '\0');
        b2o_corpus_fill(B2O_WORDS, i, reinterpret_cast<uint8_t*>(&s[0]), s.size());
        bodies.push_back(s);
    }
    std::vector<BlockList> in;
    for (const std::string& s : bodies) in.push_back(blocks_of(s, 1000));
    std::vector<std::string> z, back;
    std::vector<bool> ok;
    std::vector<uint32_t> crc;
    b200::policy::SnappyCompressBatch(in, &z);
    b200::policy::Crc32cBatch(in, nullptr, &crc);
    std::vector<BlockList> zin;
    for (const std::string& s : z) zin.push_back(blocks_of(s, 64));
    b200::policy::SnappyDecompressBatch(zin, &back, &ok);
    for (size_t i = 0; i < bodies.size(); i++) {
        CHECK(z[i] == oracle_compress(bodies[i]));
        CHECK(ok[i] && back[i] == bodies[i]);
        CHECK(crc[i] == b2o_crc32c_value(reinterpret_cast<const uint8_t*>(bodies[i].data()), bodies[i].size()));
    }
    // a malformed stream in the middle of a batch fails alone
    std::string bad = z[5];
    bad[bad.size() / 2] ^= 0x55;
    bad.resize(bad.size() - 1);
    zin[5] = blocks_of(bad, 64);
    b200::policy::SnappyDecompressBatch(zin, &back, &ok);
    uint32_t ulen = 0;
    std::vector<uint8_t> tmp(5000);
    int want = b2o_snappy_uncompress(reinterpret_cast<const uint8_t*>(bad.data()), bad.size(), tmp.data(), tmp.size(), &ulen);
    CHECK(ok[5] == (want != 0));
    CHECK(ok[4] && ok[6] && back[6] == bodies[6]);
}

// ---- snappy framing format (framing_format.txt): CPU statement of the spec as the checker ----
static uint32_t cpu_masked_crc(const std::string& s) {
    return b2o_crc32c_mask(b2o_crc32c_value(reinterpret_cast<const uint8_t*>(s.data()), s.size()));
}
static std::string le32(uint32_t v) {
    char b[4] = {(char)(v & 0xff), (char)((v >> 8) & 0xff), (char)((v >> 16) & 0xff), (char)(v >> 24)};
    return std::string(b, 4);
}
static std::string chunk(uint8_t type, const std::string& data) {
    char h[4] = {(char)type, (char)(data.size() & 0xff), (char)((data.size() >> 8) & 0xff), (char)((data.size() >> 16) & 0xff)};
    return std::string(h, 4) + data;
}
static const std::string kSzId = std::string("\xff\x06\x00\x00sNaPpY", 10);
static std::string cpu_frame(const std::string& payload) {
    std::string out = kSzId;
    for (size_t p = 0; p < payload.size(); p += 65536) {
        std::string piece = payload.substr(p, 65536);
        std::string z = oracle_compress(piece);
        if (z.size() < piece.size()) out += chunk(0x00, le32(cpu_masked_crc(piece)) + z);
        else out += chunk(0x01, le32(cpu_masked_crc(piece)) + piece);
    }
    return out;
}

static void framing_format() {
    std::vector<std::string> payloads;
    for (int i = 0; i < 12; i++) {
        static const size_t sizes[] = {0, 1, 100, 65535, 65536, 65537, 200000, 1 << 20, 5000, 131072, 70000, 33};
        std::string s(sizes[i], '\0');
        b2o_corpus_fill(i % 3 == 0 ? B2O_RAND62 : i % 3 == 1 ? B2O_WORDS : B2O_AZ09, i, reinterpret_cast<uint8_t*>(&s[0]), s.size());
        payloads.push_back(s);
    }
    std::vector<BlockList> in;
    for (const std::string& s : payloads) in.push_back(blocks_of(s, 8160));
    std::vector<std::string> framed, back;
    std::vector<bool> ok;
    b200::framing::EncodeBatch(in, &framed);
    for (size_t i = 0; i < payloads.size(); i++) CHECK(framed[i] == cpu_frame(payloads[i]));
    std::vector<BlockList> fin;
    for (const std::string& s : framed) fin.push_back(blocks_of(s, 4000));
    b200::framing::DecodeBatch(fin, &back, &ok);
    for (size_t i = 0; i < payloads.size(); i++) CHECK(ok[i] && back[i] == payloads[i]);
    // hand-built streams: skippable / padding / repeated identifier are ignored; errors are errors
    const std::string body = payloads[2];
    const std::string good_chunk = chunk(0x00, le32(cpu_masked_crc(body)) + oracle_compress(body));
    const std::string raw_chunk = chunk(0x01, le32(cpu_masked_crc(body)) + body);
    struct Case { std::string stream; bool ok; std::string want; };
    std::string bad_crc = good_chunk;
    bad_crc[5] ^= 1;
    std::vector<Case> cases = {
        {kSzId, true, ""},
        {kSzId + good_chunk, true, body},
        {kSzId + raw_chunk + good_chunk, true, body + body},
        {kSzId + chunk(0xfe, std::string(7, '\0')) + good_chunk + chunk(0x80, "future") + kSzId + raw_chunk, true, body + body},
        {good_chunk, false, ""},                                  // no stream identifier
        {kSzId + chunk(0x02, "x") + good_chunk, false, ""},       // reserved unskippable
        {kSzId + bad_crc, false, ""},                             // checksum mismatch
        {kSzId + good_chunk.substr(0, good_chunk.size() - 1), false, ""},  // chunk cut short
        {kSzId + chunk(0xff, "sNaPpX") + good_chunk, false, ""},  // malformed repeated identifier
        {kSzId + chunk(0x00, le32(cpu_masked_crc(body)) + std::string("\x05\x10hel", 5)), false, ""},  // bad snappy data
    };
    std::vector<BlockList> cin;
    for (const Case& c : cases) cin.push_back(blocks_of(c.stream, 64));
    b200::framing::DecodeBatch(cin, &back, &ok);
    for (size_t i = 0; i < cases.size(); i++) {
        CHECK(ok[i] == cases[i].ok);
        CHECK(back[i] == cases[i].want);
    }
}

// Many threads issue one-message calls (the CompressHandler calling pattern, one bthread per message);
// the aggregator must turn them into few GPU batches and hand every caller its own exact result.
static void aggregator_many_threads() {
    b200::BatchAggregator::Options opt;
    opt.max_wait_us = 2000;
    b200::BatchAggregator agg(opt);
    const int kThreads = 32, kPer = 40;
    std::atomic<int> bad(0);
    std::vector<std::thread> th;
    for (int t = 0; t < kThreads; t++) {
        th.emplace_back([&, t] {
            for (int i = 0; i < kPer; i++) {
                std::string body((size_t)((t * 131 + i * 977) % 6000), '\0');
                b2o_corpus_fill(B2O_WORDS, t * 1000 + i, reinterpret_cast<uint8_t*>(&body[0]), body.size());
                std::string z, back;
                if (!agg.Compress(blocks_of(body, 1500), &z) || z != oracle_compress(body)) bad++;
                if (agg.Crc32c(blocks_of(body, 700)) !=
                    b2o_crc32c_value(reinterpret_cast<const uint8_t*>(body.data()), body.size()))
                    bad++;
                if (!agg.Decompress(blocks_of(z, 200), &back) || back != body) bad++;
                if (i % 10 == 0 && z.size() > 4) {  // a corrupt stream fails alone
                    std::string cut = z.substr(0, z.size() - 2), junk;
                    uint32_t ulen = 0;
                    std::vector<uint8_t> tmp(8000);
                    int want = b2o_snappy_uncompress(reinterpret_cast<const uint8_t*>(cut.data()), cut.size(), tmp.data(),
                                                     tmp.size(), &ulen);
                    if (agg.Decompress(blocks_of(cut, 200), &junk) != (want != 0)) bad++;
                }
            }
        });
    }
    for (auto& x : th) x.join();
    CHECK(bad.load() == 0);
    b200::BatchAggregator::Stats st = agg.stats();
    CHECK(st.calls >= (uint64_t)kThreads * kPer * 3);
    CHECK(st.batches < st.calls / 4);  // batching really happened
    CHECK(st.largest_batch > 4);
    printf("aggregator: %llu calls in %llu batches (largest %llu)\n", (unsigned long long)st.calls,
           (unsigned long long)st.batches, (unsigned long long)st.largest_batch);
}

// ---- one-pass sender / receiver calls and RawUncompressToIOVec ----
static void fused_and_iovec_calls() {
    std::vector<std::string> bodies;
    for (int i = 0; i < 200; i++) {
        std::string s((size_t)(i * 53 % 9000), '\0');
        b2o_corpus_fill(B2O_WORDS, 1000 + i, reinterpret_cast<uint8_t*>(&s[0]), s.size());
        bodies.push_back(s);
    }
    std::vector<BlockList> in;
    for (const std::string& s : bodies) in.push_back(blocks_of(s, 8160));
    std::vector<std::string> z, back;
    std::vector<uint32_t> mcrc;
    b200::policy::SnappyCompressCrcBatch(in, &z, &mcrc);
    for (size_t i = 0; i < bodies.size(); i++) {
        CHECK(z[i] == oracle_compress(bodies[i]));
        CHECK(mcrc[i] == cpu_masked_crc(bodies[i]));
    }
    std::vector<BlockList> zin;
    for (const std::string& s : z) zin.push_back(blocks_of(s, 512));
    mcrc[3] ^= 0x10;
    std::vector<b200::policy::Verdict> verdict;
    b200::policy::SnappyDecompressVerifyBatch(zin, mcrc, &back, &verdict);
    for (size_t i = 0; i < bodies.size(); i++) {
        if (i == 3) {
            CHECK(verdict[i] == b200::policy::Verdict::kCrcMismatch);
        } else {
            CHECK(verdict[i] == b200::policy::Verdict::kOk && back[i] == bodies[i]);
        }
    }
    // RawUncompressToIOVec: the reference's unit test shape -- output over several small iovecs
    {
        const std::string& text = bodies[150];
        const std::string& stream = z[150];
        std::string sink(text.size() + 40, '#');
        std::vector<b2c_iovec> iov;
        size_t at = 0, piece = 1;
        while (at < sink.size()) {  // 1, 2, 3, ... byte iovecs, then the rest
            const size_t c = std::min(piece++, sink.size() - at);
            iov.push_back({&sink[at], c});
            at += c;
        }
        CHECK(b200::snappy::RawUncompressToIOVec(stream.data(), stream.size(), iov.data(), iov.size()));
        CHECK(sink.compare(0, text.size(), text) == 0);
        CHECK(sink.substr(text.size()) == std::string(40, '#'));
        // one byte too little room: the reference's writer fails
        std::string tight(text.size() - 1, '#');
        b2c_iovec one = {&tight[0], tight.size()};
        CHECK(!b200::snappy::RawUncompressToIOVec(stream.data(), stream.size(), &one, 1));
    }
    // batch form over 8160-byte output blocks
    std::vector<std::string> sinks;
    std::vector<BlockList> outs;
    for (const std::string& s : bodies) sinks.push_back(std::string(s.size(), '#'));
    for (std::string& s : sinks) outs.push_back(blocks_of(s, 8160));
    std::vector<bool> ok;
    b200::policy::SnappyDecompressToBlocksBatch(zin, outs, &ok);
    for (size_t i = 0; i < bodies.size(); i++) CHECK(ok[i] && sinks[i] == bodies[i]);
}

// ---- Source / Sink (snappy-sinksource.h): the streaming forms, with a Source that yields small pieces and a Sink that
// relies on the base class's defaults, as the reference's own chunked tests do ----
namespace {
class ChunkedSource : public b200::snappy::Source {
public:
    ChunkedSource(const std::string& s, size_t chunk) : s_(s), at_(0), chunk_(chunk) {}
    size_t Available() const override { return s_.size() - at_; }
    const char* Peek(size_t* len) override {
        *len = std::min(chunk_, s_.size() - at_);
        return s_.data() + at_;
    }
    void Skip(size_t n) override { at_ += n; }

private:
    const std::string& s_;
    size_t at_, chunk_;
};
class StringSink : public b200::snappy::Sink {
public:
    std::string got;
    void Append(const char* bytes, size_t n) override { got.append(bytes, n); }
};
}  // namespace

static void source_sink() {
    for (size_t n : {(size_t)0, (size_t)1, (size_t)14, (size_t)1000, (size_t)65536, (size_t)65537, (size_t)300000}) {
        std::string body(n, '\0');
        b2o_corpus_fill(B2O_WORDS, 900 + n, reinterpret_cast<uint8_t*>(&body[0]), n);
        const std::string want = oracle_compress(body);
        for (size_t chunk : {(size_t)7, (size_t)8160, (size_t)70000}) {
            ChunkedSource src(body, chunk);
            StringSink z;
            CHECK(b200::snappy::Compress(&src, &z) == want.size());
            CHECK(z.got == want && src.Available() == 0);
            ChunkedSource zsrc(z.got, chunk);
            StringSink back;
            CHECK(b200::snappy::Uncompress(&zsrc, &back) && back.got == body);
            ChunkedSource vsrc(z.got, chunk);
            CHECK(b200::snappy::IsValidCompressed(&vsrc));
            ChunkedSource lsrc(z.got, chunk);
            uint32_t ulen = 0;
            CHECK(b200::snappy::GetUncompressedLength(&lsrc, &ulen) && ulen == n);
        }
        // flat forms: ByteArraySource -> UncheckedByteArraySink (what RawCompress / RawUncompress are made of)
        std::vector<char> flat(b200::snappy::MaxCompressedLength(n) + 8), out(n + 8);
        b200::snappy::ByteArraySource bsrc(body.data(), n);
        b200::snappy::UncheckedByteArraySink bsink(flat.data());
        CHECK(b200::snappy::Compress(&bsrc, &bsink) == want.size());
        CHECK(bsink.CurrentDestination() == flat.data() + want.size() && memcmp(flat.data(), want.data(), want.size()) == 0);
        b200::snappy::ByteArraySource zsrc(flat.data(), want.size());
        b200::snappy::UncheckedByteArraySink osink(out.data());
        CHECK(b200::snappy::Uncompress(&zsrc, &osink) && memcmp(out.data(), body.data(), n) == 0);
        if (want.size() > 6) {  // a truncated stream is refused
            std::string cut = want.substr(0, want.size() - 2);
            ChunkedSource csrc(cut, 100);
            StringSink junk;
            CHECK(!b200::snappy::Uncompress(&csrc, &junk));
            // ... and UncompressAsMuchAsPossible yields the body's prefix as far as the stream is sound
            // (Python test test_decodable_prefix_* compares the same kernel with the reference's function stream by stream)
            for (size_t keep : {want.size() - 2, want.size() / 2, (size_t)5}) {
                std::string part = want.substr(0, keep);
                ChunkedSource psrc(part, 333);
                StringSink got;
                const size_t produced = b200::snappy::UncompressAsMuchAsPossible(&psrc, &got);
                CHECK(produced == got.got.size() && produced <= n && got.got == body.substr(0, produced));
                if (keep == want.size() - 2) CHECK(produced + 70000 >= n && produced < n);
            }
            ChunkedSource fsrc(want, 4096);
            StringSink all;
            CHECK(b200::snappy::UncompressAsMuchAsPossible(&fsrc, &all) == n && all.got == body);
        }
    }
}

// A GPU path that cannot run (here: a device that does not exist) must neither take the process down nor leave
// callers waiting: every call completes, reports an infrastructure error, and the caller falls back.
static void aggregator_survives_gpu_failure() {
    b200::BatchAggregator::Options opt;
    opt.device = 99;
    b200::BatchAggregator agg(opt);
    std::string body(5000, 'x'), z, back;
    std::vector<std::thread> th;
    std::atomic<int> bad(0);
    for (int t = 0; t < 8; t++)
        th.emplace_back([&] {
            std::string out;
            bool infra = false;
            if (agg.Compress(blocks_of(body, 1000), &out, &infra) || !infra) bad++;
            infra = false;
            if (agg.Decompress(blocks_of(body, 1000), &out, &infra) || !infra) bad++;
            infra = false;
            if (agg.Crc32c(blocks_of(body, 1000), 0, &infra) != 0 || !infra) bad++;
        });
    for (auto& x : th) x.join();
    CHECK(bad.load() == 0);
    CHECK(agg.stats().failed_batches >= 3);
}

// Results of the aggregator are slices of page-locked result slabs; they are released exactly once, on any thread.
static void aggregator_pooled_results() {
    b200::BatchAggregator agg;
    std::string body(30000, '\0');
    b2o_corpus_fill(B2O_WORDS, 77, reinterpret_cast<uint8_t*>(&body[0]), body.size());
    std::vector<b200::PinnedBlockPool::Slice> kept;
    for (int i = 0; i < 20; i++) {
        b200::OutSlice s;
        CHECK(agg.Compress(blocks_of(body, 8160), &s));
        CHECK(std::string(s.data(), s.size()) == oracle_compress(body));
        CHECK(s.meta() != 0 && s.deleter() != nullptr);
        if (i % 2) kept.push_back(s.adopt());  // as an IOBuf would: released later, elsewhere
    }
    std::thread([&] {
        for (auto& k : kept) {
            CHECK(std::string(k.data, k.size) == oracle_compress(body));
            k.deleter(k.data);
        }
    }).join();
}

// ---- page-locked IOBuf block pool (the blockmem_allocate / blockmem_deallocate hook) ----
static void block_pool() {
    b2c_ctx* ctx = b200::Context::local().get();
    {
        b200::PinnedBlockPool pool(ctx, 8192, 1u << 20);  // 128 blocks per region
        std::vector<void*> blocks;
        for (int i = 0; i < 300; i++) blocks.push_back(pool.Allocate(8192));
        std::vector<void*> sorted = blocks;
        std::sort(sorted.begin(), sorted.end());
        CHECK(std::adjacent_find(sorted.begin(), sorted.end()) == sorted.end());
        for (void* b : blocks) CHECK(b && (reinterpret_cast<uintptr_t>(b) & 63) == 0 && pool.Owns(b));
        CHECK(pool.stats().regions == 3);
        // bodies laid into blocks the way IOBuf does it: a 32-byte Block header, then up to 8160 payload bytes
        std::vector<std::string> bodies;
        std::vector<BlockList> in;
        size_t next_block = 0;
        for (int i = 0; i < 120; i++) {
            std::string s((size_t)(i * 197 % 20000), '\0');
            b2o_corpus_fill(B2O_WORDS, 500 + i, reinterpret_cast<uint8_t*>(&s[0]), s.size());
            BlockList bl;
            for (size_t at = 0; at < s.size() || (at == 0 && bl.empty()); at += 8160) {
                if (next_block == blocks.size()) blocks.push_back(pool.Allocate(8192));
                char* blk = static_cast<char*>(blocks[next_block++]);
                const size_t c = std::min<size_t>(8160, s.size() - at);
                memcpy(blk + 32, s.data() + at, c);
                bl.push_back({blk + 32, c});
                if (s.empty()) break;
            }
            bodies.push_back(s);
            in.push_back(bl);
        }
        // page-locked blocks: the device gathers them itself
        CHECK(b2c_ctx_set_option(ctx, B2C_OPT_INPUT_PINNED, 1) == B2C_OK);
        std::vector<std::string> z;
        std::vector<uint32_t> crc;
        b200::policy::SnappyCompressBatch(in, &z);
        b200::policy::Crc32cBatch(in, nullptr, &crc);
        CHECK(b2c_ctx_set_option(ctx, B2C_OPT_INPUT_PINNED, 0) == B2C_OK);
        for (size_t i = 0; i < bodies.size(); i++) {
            CHECK(z[i] == oracle_compress(bodies[i]));
            CHECK(crc[i] == b2o_crc32c_value(reinterpret_cast<const uint8_t*>(bodies[i].data()), bodies[i].size()));
        }
        // recycling: everything back, then out again without new regions
        const size_t regions = pool.stats().regions;
        for (void* b : blocks) pool.Deallocate(b);
        const size_t nblocks = blocks.size();
        blocks.clear();
        for (size_t i = 0; i < nblocks; i++) blocks.push_back(pool.Allocate(100));
        CHECK(pool.stats().regions == regions);
        for (void* b : blocks) pool.Deallocate(b);
        // oversize requests get their own page-locked allocation
        void* big = pool.Allocate(100000);
        CHECK(big && pool.Owns(big) && pool.stats().big_out == 1);
        pool.Deallocate(big);
        CHECK(pool.stats().big_out == 0);
        // many threads
        std::vector<std::thread> th;
        for (int t = 0; t < 4; t++)
            th.emplace_back([&pool, t] {
                std::vector<void*> mine;
                unsigned x = 12345u + t;
                for (int i = 0; i < 20000; i++) {
                    x = x * 1664525u + 1013904223u;
                    if ((x >> 16) % 3 != 0 || mine.empty()) {
                        void* b = pool.Allocate(8192);
                        *static_cast<unsigned*>(b) = x;
                        mine.push_back(b);
                    } else {
                        pool.Deallocate(mine.back());
                        mine.pop_back();
                    }
                    if (mine.size() > 200) {
                        for (void* b : mine) pool.Deallocate(b);
                        mine.clear();
                    }
                }
                for (void* b : mine) pool.Deallocate(b);
            });
        for (auto& x : th) x.join();
        const b200::PinnedBlockPool::Stats st = pool.stats();
        CHECK(st.allocs == st.frees);
        printf("block pool: %zu regions (%zu KiB page-locked), %llu allocations\n", st.regions, st.region_bytes >> 10,
               (unsigned long long)st.allocs);
    }
    // the process-wide hook pair, with the signatures of butil::iobuf::blockmem_allocate / blockmem_deallocate
    void* (*alloc_hook)(size_t) = b200::PinnedBlockPool::blockmem_allocate;
    void (*free_hook)(void*) = b200::PinnedBlockPool::blockmem_deallocate;
    void* b = alloc_hook(8192);
    CHECK(b && b200::PinnedBlockPool::global().Owns(b));
    free_hook(b);
}

int main() {
    crc_standard_results();
    crc_values_extend_mask();
    snappy_iobuf();
    snappy_test();
    throughput_shapes();
    mass_snappy_iobuf();
    batch_calls();
    framing_format();
    fused_and_iovec_calls();
    block_pool();
    source_sink();
    aggregator_many_threads();
    aggregator_survives_gpu_failure();
    aggregator_pooled_results();
    if (g_fail) {
        fprintf(stderr, "%d check(s) failed\n", g_fail);
        return 1;
    }
    printf("host api: all checks passed\n");
    return 0;
}
