// brpc_gpu_snappy.cpp — drop-in bodies for the COMPRESS_TYPE_SNAPPY handler (src/brpc/global.cpp:389-393) on top
// of libb2c.so.  See INTEGRATION.md.
//
// Every call goes through ONE process-wide b200::BasicBatchAggregator: the handler API is one message per call
// (src/brpc/compress.h:31-35), made from one bthread per in-flight message, and the aggregator turns the calls
// that are in flight together into one GPU batch.  Results are slices of the batch's page-locked slab and are
// adopted by the output IOBuf without a copy (IOBuf::append_user_data_with_meta, src/butil/iobuf.cpp:1216-1240;
// meta = the slab, like the RDMA transport's registration key, src/brpc/rdma/rdma_helper.cpp:528-531).
// The GPU path never turns a good RPC into a failed one: any infrastructure error falls back to the stock CPU
// functions; only a stream the reference itself would reject returns false.
#ifdef B200_WITH_BRPC

#include "brpc_gpu_snappy.h"

#include <atomic>
#include <vector>

#include "b200_aggregator.hpp"
#include "brpc/policy/snappy_compress.h"  // stock CPU functions: small bodies and fall-backs
#include "brpc/protocol.h"                // ParsePbFromIOBuf
#include "butil/logging.h"

#ifndef B200_SHIM_SYNC
#define B200_SHIM_SYNC b200::StdSync  // inside bRPC: -DB200_SHIM_SYNC=BthreadSync
#endif

namespace brpc {
namespace policy {

namespace {
std::atomic<size_t> g_min_gpu_bytes{16 * 1024};
std::atomic<unsigned long long> g_gpu_calls{0}, g_cpu_calls{0};

typedef b200::BasicBatchAggregator<B200_SHIM_SYNC> Service;
Service& service() {
    static Service* s = new Service();  // lives as long as the process: IOBufs may hold its slabs until exit
    return *s;
}

void to_blocks(const butil::IOBuf& in, b200::BlockList* segs) {
    const size_t nb = in.backing_block_num();  // src/butil/iobuf.h:360-363
    segs->resize(nb);
    for (size_t i = 0; i < nb; ++i) {
        butil::StringPiece blk = in.backing_block(i);  // src/butil/iobuf.cpp:1504-1510
        (*segs)[i].iov_base = blk.data();
        (*segs)[i].iov_len = blk.size();
    }
}

// hand a result slice to the IOBuf; on failure the slice is released and the caller falls back
bool adopt(b200::OutSlice* r, butil::IOBuf* out) {
    b200::PinnedBlockPool::Slice s = r->adopt();
    if (s.size == 0) {  // append_user_data_with_meta would only run the deleter: do it here
        if (s.deleter) s.deleter(s.data);
        return true;
    }
    if (out->append_user_data_with_meta(s.data, s.size, s.deleter, s.meta) == 0) return true;
    s.deleter(s.data);
    return false;
}
}  // namespace

void SetGpuSnappyMinBytes(size_t n) { g_min_gpu_bytes.store(n); }
void GpuSnappyCounters(unsigned long long* gpu_calls, unsigned long long* cpu_calls) {
    if (gpu_calls) *gpu_calls = g_gpu_calls.load();
    if (cpu_calls) *cpu_calls = g_cpu_calls.load();
}

bool GpuSnappyCompress(const butil::IOBuf& in, butil::IOBuf* out) {
    if (in.size() >= g_min_gpu_bytes.load(std::memory_order_relaxed)) {
        b200::BlockList segs;
        to_blocks(in, &segs);
        b200::OutSlice r;
        bool infra = false;
        if (service().Compress(segs, &r, &infra) && adopt(&r, out)) {
            g_gpu_calls.fetch_add(1, std::memory_order_relaxed);
            return true;  // the reference returns `written`, never 0 (snappy_compress.cpp:51-55)
        }
        // compression cannot fail on data: whatever went wrong is infrastructure — keep the RPC alive on the CPU path
    }
    g_cpu_calls.fetch_add(1, std::memory_order_relaxed);
    return SnappyCompress(in, out);
}

bool GpuSnappyDecompress(const butil::IOBuf& in, butil::IOBuf* out) {
    if (in.size() >= g_min_gpu_bytes.load(std::memory_order_relaxed)) {
        b200::BlockList segs;
        to_blocks(in, &segs);
        b200::OutSlice r;
        bool infra = false;
        if (service().Decompress(segs, &r, &infra)) {
            if (adopt(&r, out)) {
                g_gpu_calls.fetch_add(1, std::memory_order_relaxed);
                return true;
            }
        } else if (!infra) {
            g_gpu_calls.fetch_add(1, std::memory_order_relaxed);
            return false;  // a stream the reference rejects too: the caller sets EREQUEST / ERESPONSE
        }
    }
    g_cpu_calls.fetch_add(1, std::memory_order_relaxed);
    return SnappyDecompress(in, out);
}

bool GpuSnappyCompress(const google::protobuf::Message& msg, butil::IOBuf* buf) {  // snappy_compress.cpp:28-38
    butil::IOBuf serialized_pb;
    butil::IOBufAsZeroCopyOutputStream wrapper(&serialized_pb);
    if (msg.SerializeToZeroCopyStream(&wrapper)) return GpuSnappyCompress(serialized_pb, buf);
    LOG(WARNING) << "Fail to serialize input pb=" << &msg;
    return false;
}

bool GpuSnappyDecompress(const butil::IOBuf& data, google::protobuf::Message* msg) {  // snappy_compress.cpp:40-49
    butil::IOBuf binary_pb;
    if (GpuSnappyDecompress(data, &binary_pb)) return ParsePbFromIOBuf(msg, binary_pb);
    LOG(WARNING) << "Fail to snappy::Uncompress, size=" << data.size();
    return false;
}

}  // namespace policy
}  // namespace brpc

#endif  // B200_WITH_BRPC
