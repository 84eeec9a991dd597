// b200_codec.cpp — see b200_codec.hpp.  Thin C++ over the C ABI; all arithmetic is in libb2c.so.
#include "b200_codec.hpp"

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <memory>
#include <stdexcept>

namespace b200 {

namespace {
[[noreturn]] void fail(b2c_ctx* ctx, const char* what, int rc) {
    std::string msg = std::string("b200 codec: ") + what + " failed with " + std::to_string(rc);
    if (ctx) msg += std::string(": ") + b2c_last_error(ctx);
    throw std::runtime_error(msg);
}

struct Flat {  // segs + seg_begin for a batch of block lists
    std::vector<b2c_iovec> segs;
    std::vector<uint32_t> begin;
    explicit Flat(const std::vector<BlockList>& in) {
        begin.reserve(in.size() + 1);
        begin.push_back(0);
        for (const BlockList& b : in) {
            segs.insert(segs.end(), b.begin(), b.end());
            begin.push_back((uint32_t)segs.size());
        }
    }
};

size_t total_bytes(const BlockList& b) {
    size_t n = 0;
    for (const b2c_iovec& v : b) n += v.iov_len;
    return n;
}
uint64_t claimed_bytes(const std::vector<BlockList>& in) {  // sum of the (lenient) preambles
    uint64_t claimed = 0;
    for (const BlockList& b : in) {
        uint8_t head[5];
        size_t got = 0;
        for (const b2c_iovec& v : b)
            for (size_t j = 0; j < v.iov_len && got < 5; j++) head[got++] = static_cast<const uint8_t*>(v.iov_base)[j];
        uint32_t ulen = 0;
        if (b2c_snappy_read_preamble(head, got, &ulen)) claimed += ulen;
    }
    return claimed;
}

}  // namespace

Context::Context(int device) : ctx_(nullptr) {
    int rc = b2c_ctx_create(device, &ctx_);
    if (rc != B2C_OK) fail(nullptr, "b2c_ctx_create (no usable CUDA device? there is no CPU fallback)", rc);
}
Context::~Context() { b2c_ctx_destroy(ctx_); }

static std::atomic<int> g_default_device{-1};
void Context::set_default_device(int device) { g_default_device.store(device); }

int Context::default_device() {
    const int dev = g_default_device.load();
    if (dev >= 0) return dev;
    const char* d = getenv("B200_DEVICE");
    return d ? atoi(d) : 0;
}

Context& Context::local() {
    static thread_local std::unique_ptr<Context> c;  // streams, staging and workspaces go when the thread ends
    if (!c) c.reset(new Context(default_device()));
    return *c;
}

namespace policy {

void SnappyCompressBatch(const std::vector<BlockList>& in, std::vector<std::string>* out) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    out->assign(n, std::string());
    if (!n) return;
    Flat f(in);
    uint64_t cap = 0;
    for (const BlockList& b : in) cap += b2c_snappy_max_compressed_length(total_bytes(b)) + 5;
    std::vector<char> buf(cap);
    std::vector<uint64_t> off(n + 1);
    int rc = b2c_snappy_compress_batch(ctx, f.segs.data(), f.begin.data(), n, buf.data(), cap, off.data());
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_compress_batch", rc);
    for (uint32_t i = 0; i < n; i++) (*out)[i].assign(buf.data() + off[i], off[i + 1] - off[i]);
}

void SnappyDecompressBatch(const std::vector<BlockList>& in, std::vector<std::string>* out, std::vector<bool>* ok) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    out->assign(n, std::string());
    ok->assign(n, false);
    if (!n) return;
    Flat f(in);
    // capacity: a stream expands at most 64/3 times; the library sizes every slot from its preamble
    uint64_t cap = 64;
    for (const BlockList& b : in) cap += total_bytes(b) * 64 / 3 + 64;
    // sum of the claimed lengths is usually far smaller: read the preambles first
    cap = std::min<uint64_t>(cap, claimed_bytes(in) + 64);
    std::vector<char> buf(cap);
    std::vector<uint64_t> off(n + 1);
    std::vector<int32_t> st(n);
    int rc = b2c_snappy_decompress_batch(ctx, f.segs.data(), f.begin.data(), n, buf.data(), cap, off.data(), st.data());
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decompress_batch", rc);
    for (uint32_t i = 0; i < n; i++) {
        (*ok)[i] = st[i] == B2C_OK;
        if ((*ok)[i]) (*out)[i].assign(buf.data() + off[i], off[i + 1] - off[i]);
    }
}

void Crc32cBatch(const std::vector<BlockList>& in, const uint32_t* init, std::vector<uint32_t>* crc) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    crc->assign(n, 0);
    if (!n) return;
    Flat f(in);
    int rc = b2c_crc32c_batch(ctx, f.segs.data(), f.begin.data(), n, init, crc->data());
    if (rc != B2C_OK) fail(ctx, "b2c_crc32c_batch", rc);
}

void SnappyCompressCrcBatch(const std::vector<BlockList>& in, std::vector<std::string>* out,
                            std::vector<uint32_t>* masked_crc) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    out->assign(n, std::string());
    masked_crc->assign(n, 0);
    if (!n) return;
    Flat f(in);
    uint64_t cap = 0;
    for (const BlockList& b : in) cap += b2c_snappy_max_compressed_length(total_bytes(b)) + 5;
    std::vector<char> buf(cap);
    std::vector<uint64_t> off(n + 1);
    int rc = b2c_snappy_compress_crc_batch(ctx, f.segs.data(), f.begin.data(), n, buf.data(), cap, off.data(),
                                           masked_crc->data(), /*masked=*/1);
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_compress_crc_batch", rc);
    for (uint32_t i = 0; i < n; i++) (*out)[i].assign(buf.data() + off[i], off[i + 1] - off[i]);
}

void SnappyDecompressVerifyBatch(const std::vector<BlockList>& in, const std::vector<uint32_t>& masked_crc,
                                 std::vector<std::string>* out, std::vector<Verdict>* verdict) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    out->assign(n, std::string());
    verdict->assign(n, Verdict::kCorrupt);
    if (!n) return;
    if (masked_crc.size() != n) fail(ctx, "SnappyDecompressVerifyBatch: one checksum per body", B2C_E_ARG);
    Flat f(in);
    const uint64_t cap = claimed_bytes(in) + 64;
    std::vector<char> buf(cap);
    std::vector<uint64_t> off(n + 1);
    std::vector<int32_t> st(n);
    int rc = b2c_snappy_decompress_verify_batch(ctx, f.segs.data(), f.begin.data(), n, masked_crc.data(), buf.data(), cap,
                                                off.data(), st.data());
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decompress_verify_batch", rc);
    for (uint32_t i = 0; i < n; i++) {
        (*verdict)[i] = st[i] == B2C_OK ? Verdict::kOk : st[i] == B2C_CRC_MISMATCH ? Verdict::kCrcMismatch : Verdict::kCorrupt;
        if (st[i] == B2C_OK) (*out)[i].assign(buf.data() + off[i], off[i + 1] - off[i]);
    }
}

void SnappyDecompressToBlocksBatch(const std::vector<BlockList>& in, const std::vector<BlockList>& out,
                                   std::vector<bool>* ok) {
    b2c_ctx* ctx = Context::local().get();
    const uint32_t n = (uint32_t)in.size();
    ok->assign(n, false);
    if (!n) return;
    if (out.size() != n) fail(ctx, "SnappyDecompressToBlocksBatch: one block list per body", B2C_E_ARG);
    Flat f(in), g(out);
    std::vector<int32_t> st(n);
    int rc = b2c_snappy_decompress_iov_batch(ctx, f.segs.data(), f.begin.data(), n, g.segs.data(), g.begin.data(), st.data());
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decompress_iov_batch", rc);
    for (uint32_t i = 0; i < n; i++) (*ok)[i] = st[i] == B2C_OK;
}

bool SnappyCompress(const BlockList& in, std::string* out) {
    std::vector<std::string> r;
    SnappyCompressBatch({in}, &r);
    out->append(r[0]);
    return !r[0].empty();  // the reference returns `written`, which is never 0 (snappy_compress.cpp:51-55)
}

bool SnappyDecompress(const BlockList& in, std::string* out) {
    std::vector<std::string> r;
    std::vector<bool> ok;
    SnappyDecompressBatch({in}, &r, &ok);
    if (!ok[0]) return false;
    out->append(r[0]);
    return true;
}

}  // namespace policy

namespace snappy {

size_t MaxCompressedLength(size_t n) { return b2c_snappy_max_compressed_length(n); }

bool GetUncompressedLength(const char* compressed, size_t n, size_t* result) {
    uint32_t v = 0;
    if (!b2c_snappy_get_uncompressed_length(compressed, n, &v)) return false;
    *result = v;
    return true;
}

void RawCompress(const char* input, size_t n, char* compressed, size_t* compressed_length) {
    b2c_ctx* ctx = Context::local().get();
    b2c_iovec seg = {input, n};
    const uint32_t begin[2] = {0, 1};
    uint64_t off[2] = {0, 0};
    int rc = b2c_snappy_compress_batch(ctx, &seg, begin, 1, compressed, MaxCompressedLength(n) + 5, off);
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_compress_batch", rc);
    *compressed_length = off[1];
}

size_t Compress(const char* input, size_t n, std::string* output) {
    std::vector<char> buf(MaxCompressedLength(n) + 5);
    size_t clen = 0;
    RawCompress(input, n, buf.data(), &clen);
    output->assign(buf.data(), clen);
    return clen;
}

bool RawUncompress(const char* compressed, size_t n, char* uncompressed) {
    b2c_ctx* ctx = Context::local().get();
    uint32_t ulen = 0;
    if (!b2c_snappy_read_preamble(compressed, n, &ulen)) return false;  // lenient, as RawUncompress(Source*) does
    b2c_iovec seg = {compressed, n};
    const uint32_t begin[2] = {0, 1};
    uint64_t off[2];
    int32_t st = 0;
    int rc = b2c_snappy_decompress_batch(ctx, &seg, begin, 1, uncompressed, ulen, off, &st);
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decompress_batch", rc);
    return st == B2C_OK;
}

bool Uncompress(const char* compressed, size_t n, std::string* uncompressed) {
    size_t ulen = 0;
    if (!GetUncompressedLength(compressed, n, &ulen)) return false;  // strict, snappy.cc:1227-1239
    if (ulen > (uint64_t)n * 64 / 3 + 64) return false;              // cannot be a valid stream (see b2c_api.cu)
    uncompressed->resize(ulen);
    return RawUncompress(compressed, n, ulen ? &(*uncompressed)[0] : nullptr);
}

bool RawUncompressToIOVec(const char* compressed, size_t n, const struct ::b2c_iovec* iov, size_t iov_cnt) {
    b2c_ctx* ctx = Context::local().get();
    b2c_iovec seg = {compressed, n};
    const uint32_t begin[2] = {0, 1}, obegin[2] = {0, (uint32_t)iov_cnt};
    int32_t st = 0;
    int rc = b2c_snappy_decompress_iov_batch(ctx, &seg, begin, 1, iov, obegin, &st);
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decompress_iov_batch", rc);
    return st == B2C_OK;
}

bool IsValidCompressedBuffer(const char* compressed, size_t n) {
    uint32_t ulen = 0;
    if (!b2c_snappy_read_preamble(compressed, n, &ulen)) return false;
    if (ulen > (uint64_t)n * 64 / 3 + 64) return false;
    std::vector<char> tmp(ulen ? ulen : 1);
    return RawUncompress(compressed, n, tmp.data());
}

// ---- Source / Sink ------------------------------------------------------------------------------------------
Sink::~Sink() {}
Source::~Source() {}
char* Sink::GetAppendBuffer(size_t, char* scratch) { return scratch; }
char* Sink::GetAppendBufferVariable(size_t, size_t, char* scratch, size_t scratch_size, size_t* allocated_size) {
    *allocated_size = scratch_size;
    return scratch;
}
void Sink::AppendAndTakeOwnership(char* bytes, size_t n, void (*deleter)(void*, const char*, size_t), void* deleter_arg) {
    Append(bytes, n);
    (*deleter)(deleter_arg, bytes, n);
}
void UncheckedByteArraySink::Append(const char* data, size_t n) {
    if (data != dest_) memcpy(dest_, data, n);  // GetAppendBuffer handed out dest_ itself
    dest_ += n;
}
void UncheckedByteArraySink::AppendAndTakeOwnership(char* bytes, size_t n, void (*deleter)(void*, const char*, size_t),
                                                    void* deleter_arg) {
    if (bytes != dest_) {
        memcpy(dest_, bytes, n);
        (*deleter)(deleter_arg, bytes, n);
    }
    dest_ += n;
}

namespace {
std::string drain(Source* s) {
    std::string all;
    all.reserve(s->Available());
    for (;;) {
        size_t len = 0;
        const char* p = s->Peek(&len);
        if (len == 0) break;
        all.append(p, len);
        s->Skip(len);
    }
    return all;
}
}  // namespace

size_t Compress(Source* source, Sink* sink) {
    const std::string in = drain(source);
    std::string out;
    Compress(in.data(), in.size(), &out);
    sink->Append(out.data(), out.size());
    return out.size();
}

bool GetUncompressedLength(Source* source, uint32_t* result) {
    // SnappyDecompressor::ReadUncompressedLength (snappy.cc:692-711): one byte at a time, lenient about the last byte
    *result = 0;
    uint32_t shift = 0;
    for (;;) {
        if (shift >= 32) return false;
        size_t n = 0;
        const char* ip = source->Peek(&n);
        if (n == 0) return false;
        const unsigned char c = *reinterpret_cast<const unsigned char*>(ip);
        source->Skip(1);
        *result |= static_cast<uint32_t>(c & 0x7f) << shift;
        if (c < 128) return true;
        shift += 7;
    }
}

bool Uncompress(Source* compressed, Sink* uncompressed) {
    const std::string in = drain(compressed);
    uint32_t ulen = 0;
    if (!b2c_snappy_read_preamble(in.data(), in.size(), &ulen)) return false;
    if (ulen > (uint64_t)in.size() * 64 / 3 + 64) return false;  // no valid stream expands more than 64/3 times
    // like the reference, decode into the sink's own buffer when it offers one large enough (snappy.cc:1536-1545)
    char c;
    size_t allocated = 0;
    char* buf = uncompressed->GetAppendBufferVariable(1, ulen, &c, 1, &allocated);
    std::string own;
    if (allocated < ulen) {
        own.resize(ulen ? ulen : 1);
        buf = &own[0];
    }
    if (!RawUncompress(in.data(), in.size(), buf)) return false;
    uncompressed->Append(buf, ulen);
    return true;
}

bool IsValidCompressed(Source* compressed) {
    const std::string in = drain(compressed);
    return IsValidCompressedBuffer(in.data(), in.size());
}

size_t UncompressAsMuchAsPossible(Source* compressed, Sink* uncompressed) {
    const std::string in = drain(compressed);
    if (in.empty()) return 0;
    b2c_ctx* ctx = Context::local().get();
    // how far does it decode?  (b2c_snappy_decodable_prefix_dev keeps the reference writer's books on the device)
    const uint64_t off = 0;
    const uint32_t len = (uint32_t)in.size();
    b2c_layout* lay = nullptr;
    void *d_in = nullptr, *d_out = nullptr;
    uint32_t r[5] = {0, 0, 0, 0, 0};
    int rc = b2c_layout_create(ctx, 1, &off, &len, &lay);
    if (rc == B2C_OK) rc = b2c_dev_alloc(ctx, in.size(), &d_in);
    if (rc == B2C_OK) rc = b2c_dev_alloc(ctx, sizeof(r), &d_out);
    if (rc == B2C_OK) rc = b2c_memcpy_h2d(ctx, d_in, in.data(), in.size());
    if (rc == B2C_OK) rc = b2c_snappy_decodable_prefix_dev(ctx, lay, d_in, static_cast<uint32_t*>(d_out));
    if (rc == B2C_OK) rc = b2c_memcpy_d2h(ctx, r, d_out, sizeof(r));
    if (rc == B2C_OK) rc = b2c_ctx_sync(ctx);
    if (d_in) b2c_dev_free(ctx, d_in);
    if (d_out) b2c_dev_free(ctx, d_out);
    if (lay) b2c_layout_destroy(lay);
    if (rc != B2C_OK) fail(ctx, "b2c_snappy_decodable_prefix_dev", rc);
    const uint32_t sunk = r[1], consumed = r[2], partial = r[3], psrc = r[4];
    if (sunk) {
        // the prefix as a stream of its own: varint(sunk) || the whole tags || the partially applied literal
        size_t k = 0;
        while (k < in.size() && (unsigned char)in[k] >= 128) k++;
        k++;
        std::string z;
        for (uint32_t v = sunk;; v >>= 7) {
            if (v < 128) {
                z.push_back((char)v);
                break;
            }
            z.push_back((char)((v & 0x7f) | 0x80));
        }
        z.append(in, k, consumed - k);
        if (partial) {
            z.push_back((char)(63 << 2));
            const uint32_t n1 = partial - 1;
            for (int i = 0; i < 4; i++) z.push_back((char)(n1 >> (8 * i)));
            z.append(in, psrc, partial);
        }
        std::string out(sunk, '\0');
        if (!RawUncompress(z.data(), z.size(), &out[0])) fail(ctx, "UncompressAsMuchAsPossible: the sound prefix did not decode", B2C_E_ARG);
        uncompressed->Append(out.data(), out.size());
    }
    return r[0];
}

}  // namespace snappy

namespace crc32c {
uint32_t Extend(uint32_t init_crc, const char* data, size_t n) {
    b2c_ctx* ctx = Context::local().get();
    b2c_iovec seg = {data, n};
    const uint32_t begin[2] = {0, 1};
    uint32_t crc = 0;
    int rc = b2c_crc32c_batch(ctx, &seg, begin, 1, &init_crc, &crc);
    if (rc != B2C_OK) fail(ctx, "b2c_crc32c_batch", rc);
    return crc;
}
}  // namespace crc32c

}  // namespace b200
