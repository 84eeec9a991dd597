// ref_shim.cc — C wrapper over the REFERENCE's own snappy + crc32c, compiled by
// oracle/Makefile together with the reference sources where they lie under
// /root/reference/src (nothing is copied into this repo).  The resulting
// oracle/_ref/libb2ref.so is test infrastructure: it pins the oracle and is the
// "reference" CPU baseline that bench.py times.  Same symbol shapes as b2o.h,
// prefix b2r_.
#include <pthread.h>
#include <stdint.h>
#include <string.h>
#include <time.h>

#include <string>

#include "butil/crc32c.h"
#include "butil/third_party/snappy/snappy.h"
#include "butil/third_party/snappy/snappy-sinksource.h"

namespace sn = butil::snappy;

namespace {
// Source that hands the input out in fixed-size pieces, to mimic IOBuf block
// boundaries (IOBufAsSnappySource::Peek returns the rest of the current block).
class ChunkedSource : public sn::Source {
public:
    ChunkedSource(const char* p, size_t n, size_t chunk) : p_(p), left_(n), chunk_(chunk), in_chunk_(0) {}
    size_t Available() const override { return left_; }
    const char* Peek(size_t* len) override {
        size_t c = chunk_ - in_chunk_;
        *len = c < left_ ? c : left_;
        return p_;
    }
    void Skip(size_t n) override {
        p_ += n;
        left_ -= n;
        in_chunk_ = (in_chunk_ + n) % chunk_;
    }
private:
    const char* p_;
    size_t left_, chunk_, in_chunk_;
};

// Sink with the base-class defaults (like IOBufAsSnappySink, which does not
// override GetAppendBufferVariable): forces the scattered writer path.
class StringSink : public sn::Sink {
public:
    explicit StringSink(std::string* s) : s_(s) {}
    void Append(const char* bytes, size_t n) override { s_->append(bytes, n); }
private:
    std::string* s_;
};
}  // namespace

extern "C" {

size_t b2r_snappy_max_compressed_length(size_t n) { return sn::MaxCompressedLength(n); }

size_t b2r_snappy_compress(const uint8_t* in, size_t n, uint8_t* out) {
    size_t clen = 0;
    sn::RawCompress(reinterpret_cast<const char*>(in), n, reinterpret_cast<char*>(out), &clen);
    return clen;
}

// Compress(Source*, Sink*) with the input delivered in `chunk`-byte pieces.
size_t b2r_snappy_compress_chunked(const uint8_t* in, size_t n, size_t chunk, uint8_t* out) {
    ChunkedSource src(reinterpret_cast<const char*>(in), n, chunk ? chunk : 1);
    sn::UncheckedByteArraySink sink(reinterpret_cast<char*>(out));
    sn::Compress(&src, &sink);
    return sink.CurrentDestination() - reinterpret_cast<char*>(out);
}

int b2r_snappy_read_ulen_strict(const uint8_t* in, size_t n, uint32_t* ulen) {
    size_t r = 0;
    if (!sn::GetUncompressedLength(reinterpret_cast<const char*>(in), n, &r)) return 0;
    *ulen = (uint32_t)r;
    return 1;
}

int b2r_snappy_read_ulen_lenient(const uint8_t* in, size_t n, uint32_t* ulen) {
    sn::ByteArraySource src(reinterpret_cast<const char*>(in), n);
    return sn::GetUncompressedLength(&src, ulen) ? 1 : 0;
}

// RawUncompress(Source*, char*): lenient preamble + array writer.
int b2r_snappy_uncompress(const uint8_t* in, size_t n, uint8_t* out, size_t cap, uint32_t* ulen_out) {
    uint32_t ulen = 0;
    {
        sn::ByteArraySource s0(reinterpret_cast<const char*>(in), n);
        bool ok = sn::GetUncompressedLength(&s0, &ulen);
        if (ulen_out) *ulen_out = ulen;
        if (!ok) return 0;
    }
    if ((size_t)ulen > cap) return 0;
    sn::ByteArraySource src(reinterpret_cast<const char*>(in), n);
    return sn::RawUncompress(&src, reinterpret_cast<char*>(out)) ? 1 : 0;
}

// Uncompress(Source*, Sink*) exactly as brpc::policy::SnappyDecompress drives it
// (scattered writer), input in `chunk`-byte pieces.  Output is copied to `out`
// when it fits.  Returns 1/0; *produced gets the sink's byte count.
int b2r_snappy_uncompress_sink(const uint8_t* in, size_t n, size_t chunk, uint8_t* out, size_t cap,
                               size_t* produced) {
    ChunkedSource src(reinterpret_cast<const char*>(in), n, chunk ? chunk : 1);
    std::string s;
    StringSink sink(&s);
    bool ok = sn::Uncompress(&src, &sink);
    if (produced) *produced = s.size();
    if (s.size() <= cap && !s.empty()) memcpy(out, s.data(), s.size());
    return ok ? 1 : 0;
}

// snappy::UncompressAsMuchAsPossible (snappy.cc:1519-1524): returns what the reference returns; the bytes its
// sink received go to out (up to cap) and their count to *sunk.
size_t b2r_snappy_uncompress_as_much(const uint8_t* in, size_t n, uint8_t* out, size_t cap, size_t* sunk) {
    sn::ByteArraySource src(reinterpret_cast<const char*>(in), n);
    std::string s;
    StringSink sink(&s);
    const size_t r = sn::UncompressAsMuchAsPossible(&src, &sink);
    if (sunk) *sunk = s.size();
    if (!s.empty()) memcpy(out, s.data(), s.size() < cap ? s.size() : cap);
    return r;
}

int b2r_snappy_is_valid(const uint8_t* in, size_t n) {
    sn::ByteArraySource src(reinterpret_cast<const char*>(in), n);
    return sn::IsValidCompressed(&src) ? 1 : 0;
}

// string API (strict preamble): Uncompress(const char*, size_t, std::string*)
int b2r_snappy_uncompress_string(const uint8_t* in, size_t n) {
    std::string s;
    return sn::Uncompress(reinterpret_cast<const char*>(in), n, &s) ? 1 : 0;
}

uint32_t b2r_crc32c_extend(uint32_t init, const uint8_t* data, size_t n) {
    return butil::crc32c::Extend(init, reinterpret_cast<const char*>(data), n);
}
uint32_t b2r_crc32c_value(const uint8_t* data, size_t n) {
    return butil::crc32c::Value(reinterpret_cast<const char*>(data), n);
}
uint32_t b2r_crc32c_mask(uint32_t c) { return butil::crc32c::Mask(c); }
uint32_t b2r_crc32c_unmask(uint32_t c) { return butil::crc32c::Unmask(c); }
int b2r_crc32c_is_fast(void) { return butil::crc32c::IsFastCrc32Supported() ? 1 : 0; }

// ---- batch drivers (same split as b2o_batch.c) -------------------------
enum { OP_CRC = 0, OP_COMP = 1, OP_UNCOMP = 2 };
struct Job {
    int op;
    uint32_t lo, hi;
    const uint8_t* data;
    const uint64_t* off;
    const uint32_t* len;
    uint8_t* out;
    const uint64_t* out_off;
    const uint32_t* out_cap;
    uint32_t* u32_out;
    int32_t* ok;
};

static void* worker(void* arg) {
    Job* j = static_cast<Job*>(arg);
    for (uint32_t i = j->lo; i < j->hi; i++) {
        const uint8_t* src = j->data + j->off[i];
        if (j->op == OP_CRC) {
            j->u32_out[i] = b2r_crc32c_value(src, j->len[i]);
        } else if (j->op == OP_COMP) {
            j->u32_out[i] = (uint32_t)b2r_snappy_compress(src, j->len[i], j->out + j->out_off[i]);
        } else {
            uint32_t ulen = 0;
            j->ok[i] = b2r_snappy_uncompress(src, j->len[i], j->out + j->out_off[i], j->out_cap[i], &ulen);
            j->u32_out[i] = ulen;
        }
    }
    return nullptr;
}

static double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static double run(Job proto, uint32_t n, int threads) {
    pthread_t th[512];
    Job jobs[512];
    if (threads < 1) threads = 1;
    if (threads > 512) threads = 512;
    double t0 = now_s();
    for (int t = 0; t < threads; t++) {
        jobs[t] = proto;
        jobs[t].lo = (uint32_t)((uint64_t)n * t / threads);
        jobs[t].hi = (uint32_t)((uint64_t)n * (t + 1) / threads);
        pthread_create(&th[t], nullptr, worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(th[t], nullptr);
    return now_s() - t0;
}

double b2r_batch_crc32c(const uint8_t* data, const uint64_t* off, const uint32_t* len, uint32_t n,
                        uint32_t* crc, int threads) {
    Job j = {OP_CRC, 0, 0, data, off, len, nullptr, nullptr, nullptr, crc, nullptr};
    return run(j, n, threads);
}
double b2r_batch_compress(const uint8_t* data, const uint64_t* off, const uint32_t* len, uint32_t n,
                          uint8_t* out, const uint64_t* out_off, uint32_t* out_len, int threads) {
    Job j = {OP_COMP, 0, 0, data, off, len, out, out_off, nullptr, out_len, nullptr};
    return run(j, n, threads);
}
double b2r_batch_uncompress(const uint8_t* data, const uint64_t* off, const uint32_t* len, uint32_t n,
                            uint8_t* out, const uint64_t* out_off, const uint32_t* out_cap,
                            uint32_t* ulen, int32_t* ok, int threads) {
    Job j = {OP_UNCOMP, 0, 0, data, off, len, out, out_off, out_cap, ulen, ok};
    return run(j, n, threads);
}

}  // extern "C"
