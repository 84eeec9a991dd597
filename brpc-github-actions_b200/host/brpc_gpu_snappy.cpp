// brpc_gpu_snappy.cpp — drop-in bodies for the COMPRESS_TYPE_SNAPPY handler
// (src/brpc/global.cpp:389-393) on top of libb2c.so.  See INTEGRATION.md.
#ifdef B200_WITH_BRPC

#include "brpc_gpu_snappy.h"

#include <vector>

#include "b200_codec.hpp"
#include "brpc/policy/snappy_compress.h"  // stock CPU functions: fallback on infrastructure errors
#include "brpc/protocol.h"                // ParsePbFromIOBuf
#include "butil/logging.h"

namespace brpc {
namespace policy {

namespace {
const size_t kMinGpuBytes = 16 * 1024;  // below this one launch + two PCIe hops cost more than the CPU codec

void to_blocks(const butil::IOBuf& in, std::vector<b2c_iovec>* segs) {
    const size_t nb = in.backing_block_num();  // src/butil/iobuf.h:360-363
    segs->resize(nb);
    for (size_t i = 0; i < nb; ++i) {
        butil::StringPiece blk = in.backing_block(i);  // src/butil/iobuf.cpp:1504-1510
        (*segs)[i].iov_base = blk.data();
        (*segs)[i].iov_len = blk.size();
    }
}
void free_array(void* p) { delete[] static_cast<char*>(p); }
}  // namespace

bool GpuSnappyCompress(const butil::IOBuf& in, butil::IOBuf* out) {
    if (in.size() < kMinGpuBytes) return SnappyCompress(in, out);
    std::vector<b2c_iovec> segs;
    to_blocks(in, &segs);
    const uint32_t begin[2] = {0, (uint32_t)segs.size()};
    const size_t cap = b2c_snappy_max_compressed_length(in.size()) + 5;
    char* buf = new char[cap];
    uint64_t off[2];
    int rc = b2c_snappy_compress_batch(b200::Context::local().get(), segs.data(), begin, 1, buf, cap, off);
    if (rc != B2C_OK) {  // infrastructure error, never "bad data": keep the RPC alive on the CPU path
        delete[] buf;
        return SnappyCompress(in, out);
    }
    out->append_user_data(buf, off[1], free_array);  // zero-copy adopt, src/butil/iobuf.cpp:1216-1240
    return true;
}

bool GpuSnappyDecompress(const butil::IOBuf& in, butil::IOBuf* out) {
    if (in.size() < kMinGpuBytes) return SnappyDecompress(in, out);
    std::vector<b2c_iovec> segs;
    to_blocks(in, &segs);
    char head[5];
    const size_t got = in.copy_to(head, sizeof(head));
    uint32_t ulen = 0;
    if (!b2c_snappy_read_preamble(head, got, &ulen)) return false;
    if (ulen > in.size() * 64 / 3 + 64) return false;  // no valid stream expands more than 64/3
    const uint32_t begin[2] = {0, (uint32_t)segs.size()};
    char* buf = new char[ulen ? ulen : 1];
    uint64_t off[2];
    int32_t st = 0;
    int rc = b2c_snappy_decompress_batch(b200::Context::local().get(), segs.data(), begin, 1, buf, ulen, off, &st);
    if (rc != B2C_OK) {
        delete[] buf;
        return SnappyDecompress(in, out);
    }
    if (st != B2C_OK) {  // the reference returns false here and the caller sets EREQUEST / ERESPONSE
        delete[] buf;
        return false;
    }
    out->append_user_data(buf, ulen, free_array);
    return true;
}

bool GpuSnappyCompress(const google::protobuf::Message& msg, butil::IOBuf* buf) {
    butil::IOBuf serialized_pb;
    butil::IOBufAsZeroCopyOutputStream wrapper(&serialized_pb);
    if (msg.SerializeToZeroCopyStream(&wrapper)) return GpuSnappyCompress(serialized_pb, buf);
    LOG(WARNING) << "Fail to serialize input pb=" << &msg;
    return false;
}

bool GpuSnappyDecompress(const butil::IOBuf& data, google::protobuf::Message* msg) {
    butil::IOBuf binary_pb;
    if (GpuSnappyDecompress(data, &binary_pb)) return ParsePbFromIOBuf(msg, binary_pb);
    LOG(WARNING) << "Fail to snappy::Uncompress, size=" << data.size();
    return false;
}

}  // namespace policy
}  // namespace brpc

#endif  // B200_WITH_BRPC
