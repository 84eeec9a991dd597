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
    // per-thread default context on device $B200_DEVICE (default 0)
    static Context& local();

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
