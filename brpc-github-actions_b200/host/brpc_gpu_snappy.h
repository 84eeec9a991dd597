// brpc_gpu_snappy.h — the bRPC-side shim: the four signatures of src/brpc/policy/snappy_compress.h:30-39, backed
// by libb2c.so.  It needs butil/iobuf.h, protobuf and the stock brpc::policy::Snappy* functions, so it is built
// inside a bRPC tree with -DB200_WITH_BRPC; this repository compiles and runs it against the small stand-ins of
// those headers under tests/cpp/mock_brpc/ (tests/cpp/test_brpc_shim.cpp).
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

// Bodies below this many bytes go to the stock CPU functions: one queue hop plus two PCIe hops cost more than
// the CPU codec needs for them.  0 sends everything to the GPU (tests).
void SetGpuSnappyMinBytes(size_t n);
// calls served by the GPU path / by the CPU functions (small bodies and fall-backs) so far
void GpuSnappyCounters(unsigned long long* gpu_calls, unsigned long long* cpu_calls);

}  // namespace policy
}  // namespace brpc

#endif  // B200_WITH_BRPC
