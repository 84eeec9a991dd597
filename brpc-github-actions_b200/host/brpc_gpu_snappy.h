// brpc_gpu_snappy.h — the bRPC-side shim: same four signatures as
// src/brpc/policy/snappy_compress.h:30-39, backed by libb2c.so.  Built only inside a bRPC
// tree (-DB200_WITH_BRPC): it needs butil/iobuf.h and protobuf, which this repo does not carry.
#pragma once
#ifdef B200_WITH_BRPC

#include <google/protobuf/message.h>

#include "butil/iobuf.h"

namespace brpc {
namespace policy {

bool GpuSnappyCompress(const google::protobuf::Message& msg, butil::IOBuf* buf);
bool GpuSnappyDecompress(const butil::IOBuf& data, google::protobuf::Message* msg);
bool GpuSnappyCompress(const butil::IOBuf& in, butil::IOBuf* out);
bool GpuSnappyDecompress(const butil::IOBuf& in, butil::IOBuf* out);

}  // namespace policy
}  // namespace brpc

#endif  // B200_WITH_BRPC
