// TEST DOUBLE of brpc/policy/snappy_compress.h: the stock CPU handlers the shim falls back to
// (src/brpc/policy/snappy_compress.h:30-39).  tests/cpp/test_brpc_shim.cpp defines them over the CPU checker.
#pragma once
#include <google/protobuf/message.h>

#include "butil/iobuf.h"

namespace brpc {
namespace policy {
bool SnappyCompress(const google::protobuf::Message& msg, butil::IOBuf* buf);
bool SnappyDecompress(const butil::IOBuf& data, google::protobuf::Message* msg);
bool SnappyCompress(const butil::IOBuf& in, butil::IOBuf* out);
bool SnappyDecompress(const butil::IOBuf& in, butil::IOBuf* out);
}  // namespace policy
}  // namespace brpc
