// b200_codec.hpp — C++ host-side mirror of the reference interfaces on the snappy + CRC32C
// path, implemented over the C ABI (include/b2c_codec.h, libb2c.so).  Names, argument
// meaning and bool/size_t results follow the reference so that callers and tests port 1:1:
//   b200::snappy::*   <- butil::snappy::*  (src/butil/third_party/snappy/snappy.h:63-191)
//   b200::crc32c::*   <- butil::crc32c::*  (src/butil/crc32c.h:19-47)
//   b200::policy::*   <- brpc::policy::SnappyCompress/SnappyDecompress(const IOBuf&, IOBuf*)
//                        (src/brpc/policy/snappy_compress.h:36-39) over IOBuf block lists
// Every call runs on the GPU; there is no CPU fallback (a missing device throws).
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "b2c_codec.h"

namespace b200 {

// One b2c_ctx bound to a device.  Not thread-safe: use one per thread (Context::local()).
class Context {
public:
    explicit Context(int device = 0);
    ~Context();
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    b2c_ctx* get() const { return ctx_; }
    // this thread's default context (created on first use, destroyed when the thread ends), on the device chosen
    // with set_default_device (else $B200_DEVICE, else 0)
    static Context& local();
    static void set_default_device(int device);
    static int default_device();

private:
    b2c_ctx* ctx_;
};

// A message body as IOBuf exposes it: BlockRefs in order (IOBuf::backing_block(i),
// src/butil/iobuf.cpp:1504-1510).  b2c_iovec is layout-identical to butil::const_iovec.
using BlockList = std::vector<b2c_iovec>;

namespace snappy {
size_t MaxCompressedLength(size_t source_bytes);                                     // snappy.h:176
bool GetUncompressedLength(const char* compressed, size_t n, size_t* result);        // snappy.h:182
void RawCompress(const char* input, size_t n, char* compressed, size_t* compressed_length);  // snappy.h:153
size_t Compress(const char* input, size_t n, std::string* output);                   // snappy.h:71
bool RawUncompress(const char* compressed, size_t n, char* uncompressed);            // snappy.h:161
bool Uncompress(const char* compressed, size_t n, std::string* uncompressed);        // snappy.h:79
bool IsValidCompressedBuffer(const char* compressed, size_t n);                      // snappy.h:190
// snappy.h:168-171: the decoded bytes fill iov[0], iov[1], ... in order; false when they do not fit
bool RawUncompressToIOVec(const char* compressed, size_t n, const struct ::b2c_iovec* iov, size_t iov_cnt);

// The streaming abstraction of the reference (src/butil/third_party/snappy/snappy-sinksource.h:38-181), same
// virtual interfaces and default behaviours (snappy-sinksource.cc:36-104), so that a caller that holds a Source* or
// implements a Sink (butil::IOBufAsSnappySource / IOBufAsSnappySink, src/butil/iobuf.h:596-634) ports by renaming
// the namespace.
class Sink {
public:
    Sink() {}
    virtual ~Sink();
    virtual void Append(const char* bytes, size_t n) = 0;
    virtual char* GetAppendBuffer(size_t length, char* scratch);  // default: scratch
    virtual void AppendAndTakeOwnership(char* bytes, size_t n, void (*deleter)(void*, const char*, size_t),
                                        void* deleter_arg);  // default: Append, then the deleter
    virtual char* GetAppendBufferVariable(size_t min_size, size_t desired_size_hint, char* scratch, size_t scratch_size,
                                          size_t* allocated_size);  // default: scratch

private:
    Sink(const Sink&);
    void operator=(const Sink&);
};
class Source {
public:
    Source() {}
    virtual ~Source();
    virtual size_t Available() const = 0;
    virtual const char* Peek(size_t* len) = 0;
    virtual void Skip(size_t n) = 0;

private:
    Source(const Source&);
    void operator=(const Source&);
};
class ByteArraySource : public Source {
public:
    ByteArraySource(const char* p, size_t n) : ptr_(p), left_(n) {}
    size_t Available() const override { return left_; }
    const char* Peek(size_t* len) override {
        *len = left_;
        return ptr_;
    }
    void Skip(size_t n) override {
        left_ -= n;
        ptr_ += n;
    }

private:
    const char* ptr_;
    size_t left_;
};
class UncheckedByteArraySink : public Sink {
public:
    explicit UncheckedByteArraySink(char* dest) : dest_(dest) {}
    void Append(const char* data, size_t n) override;
    char* GetAppendBuffer(size_t len, char* scratch) override { return dest_; }
    char* GetAppendBufferVariable(size_t min_size, size_t desired_size_hint, char* scratch, size_t scratch_size,
                                  size_t* allocated_size) override {
        *allocated_size = desired_size_hint;
        return dest_;
    }
    void AppendAndTakeOwnership(char* bytes, size_t n, void (*deleter)(void*, const char*, size_t), void* deleter_arg) override;
    char* CurrentDestination() const { return dest_; }

private:
    char* dest_;
};
// snappy.h:63-104 on the streaming types.  The Source is consumed to its end, as the reference does; what it
// yields is handed to the GPU as a block list (a piece returned by Peek is copied only when the Source gives no
// promise that it stays valid after Skip, i.e. always here: one host pass, the reference's scratch copies
// included).  Compress returns the bytes written; Uncompress reads the preamble leniently (snappy.cc:692-711).
size_t Compress(Source* source, Sink* sink);
bool Uncompress(Source* compressed, Sink* uncompressed);
bool GetUncompressedLength(Source* source, uint32_t* result);  // consumes the preamble, snappy.cc:870-873
bool IsValidCompressed(Source* compressed);
// snappy.h:104 / snappy.cc:1519-1524: decodes as far as the stream is sound, appends that prefix to the sink and returns
// what the reference returns (its writer's Size(), which exceeds the appended bytes in the one case where the
// reference's writer counts a block twice: a literal refused after it filled the current block).
size_t UncompressAsMuchAsPossible(Source* compressed, Sink* uncompressed);
}  // namespace snappy

namespace crc32c {
uint32_t Extend(uint32_t init_crc, const char* data, size_t n);  // crc32c.h:19-24
inline uint32_t Value(const char* data, size_t n) { return Extend(0, data, n); }
inline uint32_t Mask(uint32_t crc) { return b2c_crc32c_mask(crc); }
inline uint32_t Unmask(uint32_t masked) { return b2c_crc32c_unmask(masked); }
}  // namespace crc32c

namespace policy {
// brpc::policy::SnappyCompress(const IOBuf& in, IOBuf* out): appends the stream to *out.
bool SnappyCompress(const BlockList& in, std::string* out);
// brpc::policy::SnappyDecompress(const IOBuf& in, IOBuf* out): appends the body to *out;
// false on a malformed stream (what the reference's Uncompress(Source*,Sink*) rejects).
bool SnappyDecompress(const BlockList& in, std::string* out);

// The batch forms the GPU path is built for: n bodies per call.
// ok[i] mirrors the per-message bool of the calls above.
void SnappyCompressBatch(const std::vector<BlockList>& in, std::vector<std::string>* out);
void SnappyDecompressBatch(const std::vector<BlockList>& in, std::vector<std::string>* out, std::vector<bool>* ok);
void Crc32cBatch(const std::vector<BlockList>& in, const uint32_t* init /*nullable*/, std::vector<uint32_t>* crc);
// One pass per side of a checksummed, compressed frame (b2c_snappy_compress_crc_batch /
// b2c_snappy_decompress_verify_batch): masked_crc[i] = crc32c::Mask(crc32c::Value(body i)).
void SnappyCompressCrcBatch(const std::vector<BlockList>& in, std::vector<std::string>* out,
                            std::vector<uint32_t>* masked_crc);
enum class Verdict { kOk, kCorrupt, kCrcMismatch };
void SnappyDecompressVerifyBatch(const std::vector<BlockList>& in, const std::vector<uint32_t>& masked_crc,
                                 std::vector<std::string>* out, std::vector<Verdict>* verdict);
// RawUncompressToIOVec for n bodies: body i is laid over out[i]'s blocks.
void SnappyDecompressToBlocksBatch(const std::vector<BlockList>& in, const std::vector<BlockList>& out,
                                   std::vector<bool>* ok);
}  // namespace policy

}  // namespace b200
