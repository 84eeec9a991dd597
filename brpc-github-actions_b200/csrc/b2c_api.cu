// b2c_api.cu — host side of libb2c.so: contexts, layouts and the extern "C"
// entry points declared in include/b2c_codec.h.  Everything below the C ABI is
// CUDA for sm_100a; there is deliberately no CPU codec here: without a usable
// device b2c_ctx_create fails with B2C_E_NODEVICE.
//
// The *_batch (host memory) entry points run a slab pipeline: the batch is cut
// into slabs of ~32 MiB; host threads gather a slab's IOBuf blocks into pinned
// staging while the previous slab is on the PCIe bus / in the kernels and the one
// before that is being copied out, over kPipeLanes streams with private
// workspaces (replaces the per-fragment gather memcpy of snappy.cc:901-921 and
// the sink copies of src/butil/iobuf.cpp:2066-2093).
#include <emmintrin.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <thread>

#include "b2c_internal.cuh"

namespace b2c {

int set_err(b2c_ctx* ctx, int code, const char* what, cudaError_t e) {
    if (ctx) {
        ctx->err = what ? what : "";
        if (e != cudaSuccess) {
            ctx->err += ": ";
            ctx->err += cudaGetErrorString(e);
        }
    }
    return code;
}

int ws_reserve(b2c_ctx* ctx, Lane& lane, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return B2C_OK;
    if (b.p) {
        // the old buffer may still be in use by queued work
        B2C_CUDA(ctx, cudaStreamSynchronize(lane.stream));
        B2C_CUDA(ctx, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = bytes + bytes / 8 + 4096;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        b.p = nullptr;
        return set_err(ctx, B2C_E_NOMEM, "cudaMalloc(workspace)", e);
    }
    b.cap = want;
    return B2C_OK;
}

int next_work_counter(b2c_ctx* ctx, Lane& lane, uint32_t** ctr) {
    int rc = ws_reserve(ctx, lane, lane.ws_ctr, 64 * sizeof(uint32_t));
    if (rc) return rc;
    uint32_t* p = static_cast<uint32_t*>(lane.ws_ctr.p) + (lane.ctr_next++ & 63u);
    B2C_CUDA(ctx, cudaMemsetAsync(p, 0, sizeof(uint32_t), lane.stream));
    *ctr = p;
    return B2C_OK;
}

// Grow-only pinned buffer; the first `keep` bytes survive a growth.
static int pin_reserve(b2c_ctx* ctx, Lane& lane, PinBuf& b, size_t bytes, size_t keep = 0) {
    if (bytes <= b.cap) return B2C_OK;
    size_t want = bytes + bytes / 8 + 4096;
    void* fresh = nullptr;
    cudaError_t e = cudaHostAlloc(&fresh, want, cudaHostAllocDefault);
    if (e != cudaSuccess) return set_err(ctx, B2C_E_NOMEM, "cudaHostAlloc(staging)", e);
    if (b.p) {
        B2C_CUDA(ctx, cudaStreamSynchronize(lane.stream));
        if (keep) memcpy(fresh, b.p, std::min(keep, b.cap));
        B2C_CUDA(ctx, cudaFreeHost(b.p));
    }
    b.p = fresh;
    b.cap = want;
    return B2C_OK;
}

void prof_mark(b2c_ctx* ctx, Lane& lane, int which) {
    if (!ctx->prof_on || &lane != &ctx->main) return;
    auto& v = ctx->prof_ev[which];
    uint32_t& used = ctx->prof_used[which];
    if (used >= 1024) return;
    if (used >= v.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        v.push_back(e);
    }
    cudaEventRecord(v[used++], lane.stream);
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline size_t max_clen(size_t n) { return 32 + n + n / 6; }

// ---------------------------------------------------------------------------
// persistent host thread pool (gather / copy-out memcpy, preamble parsing)
// ---------------------------------------------------------------------------
class HostPool {
public:
    explicit HostPool(unsigned n) : n_(std::max(1u, n)) {
        for (unsigned t = 1; t < n_; t++) workers_.emplace_back([this, t] { loop(t); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
            gen_++;
        }
        cv_.notify_all();
        for (auto& w : workers_) w.join();
    }
    unsigned size() const { return n_; }
    // run fn(t, T) for t in [0, T) with T = min(size, want)
    void run(unsigned want, const std::function<void(unsigned, unsigned)>& fn) {
        unsigned T = std::max(1u, std::min(n_, want));
        // a section started by this very thread is still running on the workers: do the work here
        if (T == 1 || (async_.load(std::memory_order_acquire) && owner_.load(std::memory_order_relaxed) == std::this_thread::get_id())) {
            fn(0, 1);
            return;
        }
        std::lock_guard<std::mutex> serial(run_mu_);  // one parallel section at a time on this pool
        {
            std::lock_guard<std::mutex> g(mu_);
            fn_ = &fn;
            base_ = 0;
            active_ = T;
            pending_ = T - 1;
            gen_++;
        }
        cv_.notify_all();
        fn(0, T);
        std::unique_lock<std::mutex> g(mu_);
        done_cv_.wait(g, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }
    // Start fn(t, T) on the WORKERS only and return; the caller goes on with other work and
    // must call wait() (same thread) before the next start().  Without workers it runs here.
    void start(unsigned want, std::function<void(unsigned, unsigned)> fn) {
        if (n_ == 1) {
            fn(0, 1);
            return;
        }
        run_mu_.lock();
        owner_.store(std::this_thread::get_id(), std::memory_order_relaxed);
        async_.store(true, std::memory_order_release);
        afn_ = std::move(fn);
        const unsigned T = std::max(1u, std::min(n_ - 1, want));
        {
            std::lock_guard<std::mutex> g(mu_);
            fn_ = &afn_;
            base_ = 1;
            active_ = T;
            pending_ = T;
            gen_++;
        }
        cv_.notify_all();
    }
    bool started_here() const {
        return async_.load(std::memory_order_acquire) && owner_.load(std::memory_order_relaxed) == std::this_thread::get_id();
    }
    void wait() {
        if (!started_here()) return;
        {
            std::unique_lock<std::mutex> g(mu_);
            done_cv_.wait(g, [this] { return pending_ == 0; });
            fn_ = nullptr;
        }
        async_.store(false, std::memory_order_release);
        run_mu_.unlock();
    }

private:
    void loop(unsigned t) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(unsigned, unsigned)>* fn = nullptr;
            unsigned T = 0, t_rel = 0;
            {
                std::unique_lock<std::mutex> g(mu_);
                cv_.wait(g, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                fn = fn_;
                T = active_;
                t_rel = t - base_;  // the caller is share 0 of a synchronous section, not of a started one
            }
            if (fn && t >= base_ && t_rel < T) {
                (*fn)(t_rel, T);
                std::lock_guard<std::mutex> g(mu_);
                if (--pending_ == 0) done_cv_.notify_one();
            }
        }
    }
    unsigned n_;
    std::vector<std::thread> workers_;
    std::mutex mu_, run_mu_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(unsigned, unsigned)>* fn_ = nullptr;
    std::function<void(unsigned, unsigned)> afn_;  // the started section's function (owned)
    std::atomic<bool> async_{false};
    std::atomic<std::thread::id> owner_{};
    unsigned base_ = 0;  // first participating worker id
    unsigned active_ = 0, pending_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

// One pool per ctx (a ctx is used by one thread at a time), created on the first *_batch call.  Default size:
// the host's cores divided among the ranks of this node (torchrun's LOCAL_WORLD_SIZE), between 2 and 16 --
// eight ranks that each start 16 gather threads on a 32-core host only take turns on the cores.
// B2C_OPT_HOST_THREADS / B2C_HOST_THREADS fix it.
static unsigned default_host_threads() {
    unsigned hw = std::thread::hardware_concurrency();
    if (!hw) hw = 4;
    unsigned ranks = 1;
    if (const char* s = getenv("LOCAL_WORLD_SIZE")) ranks = (unsigned)std::max(1L, atol(s));
    return std::min(16u, std::max(2u, hw / ranks));
}

static HostPool& pool(b2c_ctx* ctx) {
    if (!ctx->host_pool) ctx->host_pool = new HostPool(ctx->host_threads ? ctx->host_threads : default_host_threads());
    return *static_cast<HostPool*>(ctx->host_pool);
}

static unsigned threads_for(uint64_t bytes) { return (unsigned)std::min<uint64_t>(64, bytes / (1u << 20) + 1); }

// ---------------------------------------------------------------------------
// LayoutPlan
// ---------------------------------------------------------------------------
static inline int frag_class(uint32_t fl) { return fl <= 1024 ? 0 : fl <= 4096 ? 1 : 2; }
static inline int slot_class(uint32_t cap) { return cap <= 1024 ? 0 : cap <= 8192 ? 1 : 2; }

void PlanStats::merge(const PlanStats& p) {
    payload += p.payload;
    arena_end = std::max(arena_end, p.arena_end);
    max_comp += p.max_comp;
    slot_sum += p.slot_sum;
    F += p.F;
    C += p.C;
    max_len = std::max(max_len, p.max_len);
    n_multi += p.n_multi;
    for (int c = 0; c < 3; c++) {
        n_class[c] += p.n_class[c];
        n_dclass[c] += p.n_dclass[c];
    }
}

void plan_range(const uint64_t* off, const uint32_t* len, uint32_t lo, uint32_t hi, PlanStats& st) {
    const uint64_t full_slot = align_up(max_clen(kFragBytes), 16);
    for (uint32_t m = lo; m < hi; m++) {
        const uint32_t l = len[m];
        st.payload += l;
        if (off) st.arena_end = std::max<uint64_t>(st.arena_end, off[m] + l);
        st.max_comp += max_clen(l) + 5;
        st.max_len = std::max(st.max_len, l);
        const uint32_t nf = l ? (uint32_t)(((uint64_t)l + kFragBytes - 1) / kFragBytes) : 1u;
        const uint32_t last = l - (nf - 1) * kFragBytes;
        st.F += nf;
        st.n_class[2] += nf - 1;
        st.n_class[frag_class(last)] += 1;
        st.slot_sum += (uint64_t)(nf - 1) * full_slot + align_up(max_clen(last), 16);
        const uint32_t nc = l ? (uint32_t)(((uint64_t)l + kCrcChunk - 1) / kCrcChunk) : 1u;
        st.C += nc;
        if (nc > 1) st.n_multi++;
        st.n_dclass[slot_class(l)]++;
    }
}

void LayoutPlan::plan(uint32_t n_, const uint64_t* off_, const uint32_t* len_, b2c_ctx* par) {
    PlanStats tot;
    const uint32_t n = n_;
    const uint64_t* off = off_;
    const uint32_t* len = len_;
    if (par && n >= (1u << 14)) {
        PlanStats part[64];
        unsigned used = 1;
        pool(par).run(threads_for((uint64_t)n * 256), [&](unsigned t, unsigned T) {
            if (t == 0) used = T;
            plan_range(off, len, (uint32_t)((uint64_t)n * t / T), (uint32_t)((uint64_t)n * (t + 1) / T), part[t]);
        });
        for (unsigned t = 0; t < used; t++) tot.merge(part[t]);
    } else {
        plan_range(off, len, 0, n, tot);
    }
    plan_with(n_, off_, len_, tot);
}

void LayoutPlan::plan_with(uint32_t n_, const uint64_t* off_, const uint32_t* len_, const PlanStats& tot) {
    n = n_;
    off = off_;
    len = len_;
    v = LayoutView();
    v.n = n;
    v.payload = tot.payload;
    v.arena_end = tot.arena_end;
    v.max_comp = tot.max_comp;
    v.n_multi = tot.n_multi;
    for (int c = 0; c < 3; c++) {
        v.n_class[c] = tot.n_class[c];
        v.n_dclass[c] = tot.n_dclass[c];
    }
    const uint64_t slot_sum = tot.slot_sum, F = tot.F, C = tot.C;
    const uint32_t max_len = tot.max_len;
    v.n_frags = (uint32_t)F;
    v.n_chunks = (uint32_t)C;
    has_frag_arrays = F != n;
    has_chunk_arrays = C != n;
    const uint64_t stride = align_up(max_clen(std::min(max_len, kFragBytes)), 16);
    const bool uniform = F * stride <= 2 * slot_sum + (1u << 20);
    has_slot_array = !uniform;
    v.slot_stride = uniform ? stride : 0;
    v.frag_scratch_bytes = uniform ? F * stride : slot_sum;
    has_class_lists = (v.n_class[0] != 0) + (v.n_class[1] != 0) + (v.n_class[2] != 0) > 1;
    has_dclass_lists = (v.n_dclass[0] != 0) + (v.n_dclass[1] != 0) + (v.n_dclass[2] != 0) > 1;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        size_t at = o;
        o = align_up(o + bytes, 16);
        return at;
    };
    o_off = take((size_t)n * 8);
    o_slot = has_slot_array ? take(F * 8) : 0;
    o_len = take((size_t)n * 4);
    if (has_frag_arrays) {
        o_frag_msg = take(F * 4);
        o_frag_first = take(((size_t)n + 1) * 4);
    }
    if (has_class_lists)
        for (int c = 0; c < 3; c++) o_class[c] = take((size_t)v.n_class[c] * 4);
    if (has_dclass_lists)
        for (int c = 0; c < 3; c++) o_dclass[c] = take((size_t)v.n_dclass[c] * 4);
    if (has_chunk_arrays) {
        o_chunk_msg = take(C * 4);
        o_chunk_first = take(((size_t)n + 1) * 4);
        o_multi = take((size_t)v.n_multi * 4);
    }
    blob_bytes = std::max<size_t>(o, 16);
}

void LayoutPlan::fill(uint8_t* blob) const {
    if (n) {
        memcpy(blob + o_off, off, (size_t)n * 8);
        memcpy(blob + o_len, len, (size_t)n * 4);
    }
    if (!(has_frag_arrays || has_slot_array || has_class_lists || has_dclass_lists || has_chunk_arrays)) return;
    uint64_t* slot = has_slot_array ? reinterpret_cast<uint64_t*>(blob + o_slot) : nullptr;
    uint32_t* frag_msg = has_frag_arrays ? reinterpret_cast<uint32_t*>(blob + o_frag_msg) : nullptr;
    uint32_t* frag_first = has_frag_arrays ? reinterpret_cast<uint32_t*>(blob + o_frag_first) : nullptr;
    uint32_t* cls[3] = {nullptr, nullptr, nullptr};
    uint32_t* dcls[3] = {nullptr, nullptr, nullptr};
    for (int c = 0; c < 3; c++) {
        if (has_class_lists) cls[c] = reinterpret_cast<uint32_t*>(blob + o_class[c]);
        if (has_dclass_lists) dcls[c] = reinterpret_cast<uint32_t*>(blob + o_dclass[c]);
    }
    uint32_t* chunk_msg = has_chunk_arrays ? reinterpret_cast<uint32_t*>(blob + o_chunk_msg) : nullptr;
    uint32_t* chunk_first = has_chunk_arrays ? reinterpret_cast<uint32_t*>(blob + o_chunk_first) : nullptr;
    uint32_t* multi = has_chunk_arrays ? reinterpret_cast<uint32_t*>(blob + o_multi) : nullptr;
    const uint64_t full_slot = align_up(max_clen(kFragBytes), 16);
    uint64_t slot_at = 0;
    uint32_t f = 0, c = 0, nm = 0, ncls[3] = {0, 0, 0}, ndcls[3] = {0, 0, 0};
    for (uint32_t m = 0; m < n; m++) {
        const uint32_t l = len[m];
        const uint32_t nf = l ? (uint32_t)(((uint64_t)l + kFragBytes - 1) / kFragBytes) : 1u;
        if (frag_first) frag_first[m] = f;
        for (uint32_t k = 0; k < nf; k++, f++) {
            const uint32_t fl = k + 1 < nf ? kFragBytes : l - (nf - 1) * kFragBytes;
            if (frag_msg) frag_msg[f] = m;
            if (slot) {
                slot[f] = slot_at;
                slot_at += k + 1 < nf ? full_slot : align_up(max_clen(fl), 16);
            }
            if (has_class_lists) {
                int q = frag_class(fl);
                cls[q][ncls[q]++] = f;
            }
        }
        if (has_dclass_lists) {
            int q = slot_class(l);
            dcls[q][ndcls[q]++] = m;
        }
        if (chunk_msg) {
            const uint32_t nc = l ? (uint32_t)(((uint64_t)l + kCrcChunk - 1) / kCrcChunk) : 1u;
            chunk_first[m] = c;
            for (uint32_t k = 0; k < nc; k++) chunk_msg[c++] = m;
            if (nc > 1) multi[nm++] = m;
        }
    }
    if (frag_first) frag_first[n] = f;
    if (chunk_first) chunk_first[n] = c;
}

void LayoutPlan::bind(const uint8_t* d, LayoutView* out) const {
    *out = v;
    out->d_off = reinterpret_cast<const uint64_t*>(d + o_off);
    out->d_len = reinterpret_cast<const uint32_t*>(d + o_len);
    out->d_frag_slot = has_slot_array ? reinterpret_cast<const uint64_t*>(d + o_slot) : nullptr;
    out->d_frag_msg = has_frag_arrays ? reinterpret_cast<const uint32_t*>(d + o_frag_msg) : nullptr;
    out->d_frag_first = has_frag_arrays ? reinterpret_cast<const uint32_t*>(d + o_frag_first) : nullptr;
    for (int c = 0; c < 3; c++) {
        out->d_class_list[c] = has_class_lists ? reinterpret_cast<const uint32_t*>(d + o_class[c]) : nullptr;
        out->d_dclass_list[c] = has_dclass_lists ? reinterpret_cast<const uint32_t*>(d + o_dclass[c]) : nullptr;
    }
    out->d_chunk_msg = has_chunk_arrays ? reinterpret_cast<const uint32_t*>(d + o_chunk_msg) : nullptr;
    out->d_chunk_first = has_chunk_arrays ? reinterpret_cast<const uint32_t*>(d + o_chunk_first) : nullptr;
    out->d_multi_list = has_chunk_arrays ? reinterpret_cast<const uint32_t*>(d + o_multi) : nullptr;
}

// Zero-copy input: one warp per message walks the message's blocks in page-locked HOST memory (device-visible
// under UVA) and writes them, concatenated, into the device arena.
//
// What crosses PCIe is decided by the LOADS: a kernel pulls pinned host memory at the copy engines' rate (51 of
// 55 GB/s on this box, tools/sysmem_read_probe.cu) as long as every request is a run of whole aligned 16-byte
// vectors -- 4-byte or byte loads of the same bytes reach 35.  So each block is fetched as the aligned 16-byte
// vectors that cover it (never a vector that holds none of its bytes), parked in the warp's shared-memory
// buffer, and written out from there with aligned 16-byte stores whatever the phase between source and
// destination is (the funnel shifts happen on shared-memory words, not on PCIe reads).
// Small on purpose (16.6 KiB of shared memory per CTA, one CTA per SM): the gather of slab k+1 has to run beside the
// codec kernels of slab k, which fill the SMs' shared memory (2 x 97 KiB for the small-body compressor); with 8 KiB
// buffers the two took turns on the GPU and the sender's call went from 29 to 37 ms.
constexpr uint32_t kGatherSlack = 32;  // per-warp buffer = chunk + the lead-in of an unaligned source + the copy-out's read-ahead

// n bytes from shared memory `s` (any phase) to global `dst` (any phase): bytes up to dst's next 16-byte
// boundary, aligned vectors assembled from 5 shared-memory words each, bytes again for the tail
__device__ __forceinline__ void smem_to_global(uint8_t* __restrict__ dst, const uint8_t* s, uint32_t n, uint32_t lane) {
    uint32_t head = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15)) & 15u;
    if (head > n) head = n;
    if (lane < head) dst[lane] = s[lane];
    const uint32_t vecs = (n - head) >> 4;
    uint4* d4 = reinterpret_cast<uint4*>(dst + head);
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s + head);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(s + head - (sa & 3));  // aligned word holding the first byte
    const uint32_t sh = (sa & 3) * 8;
    for (uint32_t i = lane; i < vecs; i += 32) {
        const uint32_t* q = w + 4 * i;
        const uint32_t a = q[0], b = q[1], c = q[2], d = q[3], e = q[4];
        d4[i] = make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh),
                           __funnelshift_r(d, e, sh));
    }
    const uint32_t done = head + (vecs << 4);
    if (lane < n - done) dst[done + lane] = s[done + lane];
}

__global__ void __launch_bounds__(1024) gather_host_segments_kernel(const b2c_iovec* __restrict__ segs,
                                                                     const uint32_t* __restrict__ seg_begin,
                                                                     const uint64_t* __restrict__ off, uint32_t cnt,
                                                                     uint8_t* __restrict__ d_in, const uint32_t kGatherChunk) {
    extern __shared__ __align__(16) uint8_t gather_smem[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t kGatherWarps = blockDim.x >> 5;
    uint8_t* buf = gather_smem + (size_t)(threadIdx.x >> 5) * (kGatherChunk + kGatherSlack);
    uint4* buf4 = reinterpret_cast<uint4*>(buf);
    const uint32_t warps = gridDim.x * kGatherWarps;
    for (uint32_t m = blockIdx.x * kGatherWarps + (threadIdx.x >> 5); m < cnt; m += warps) {
        uint8_t* dst = d_in + off[m];
        for (uint32_t q = seg_begin[m]; q < seg_begin[m + 1]; q++) {
            const uint8_t* src = static_cast<const uint8_t*>(segs[q].iov_base);
            const uint64_t len = segs[q].iov_len;
            for (uint64_t c0 = 0; c0 < len; c0 += kGatherChunk) {
                const uint8_t* cs = src + c0;
                const uint32_t cl = (uint32_t)(len - c0 < kGatherChunk ? len - c0 : kGatherChunk);
                const uint32_t lead = (uint32_t)(reinterpret_cast<uintptr_t>(cs) & 15);
                const uint4* s4 = reinterpret_cast<const uint4*>(cs - lead);
                const uint32_t nv = (lead + cl + 15) >> 4;  // every vector holds at least one byte of the block
                uint32_t i = lane;
                for (; i + 96 < nv; i += 128) {  // four independent requests per lane in flight
                    const uint4 v0 = __ldg(s4 + i), v1 = __ldg(s4 + i + 32), v2 = __ldg(s4 + i + 64), v3 = __ldg(s4 + i + 96);
                    buf4[i] = v0, buf4[i + 32] = v1, buf4[i + 64] = v2, buf4[i + 96] = v3;
                }
                for (; i < nv; i += 32) buf4[i] = __ldg(s4 + i);
                __syncwarp();
                smem_to_global(dst, buf + lead, cl, lane);
                __syncwarp();
                dst += cl;
            }
        }
    }
}

int launch_gather_host(b2c_ctx* ctx, Lane& lane, const b2c_iovec* d_segs, const uint32_t* d_seg_begin,
                       const uint64_t* d_off, uint32_t cnt, uint8_t* d_in) {
    if (!cnt) return B2C_OK;
    const uint32_t kGatherWarps = (uint32_t)ctx->gather_warps, chunk = (uint32_t)ctx->gather_chunk;
    const size_t smem = (size_t)kGatherWarps * (chunk + kGatherSlack);
    if (!ctx->gather_attr_done) {
        B2C_CUDA(ctx, cudaFuncSetAttribute(gather_host_segments_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        // same L1 / shared-memory split as the codec kernels it has to run beside: an SM does not change its
        // carve-out while CTAs are resident, so a kernel that prefers another split waits for the SM to drain
        B2C_CUDA(ctx, cudaFuncSetAttribute(gather_host_segments_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
        ctx->gather_attr_done = true;
    }
    const uint32_t blocks = std::min<uint32_t>((cnt + kGatherWarps - 1) / kGatherWarps, (uint32_t)ctx->num_sms * (uint32_t)ctx->gather_ctas_per_sm);
    gather_host_segments_kernel<<<blocks, 32 * kGatherWarps, smem, lane.stream>>>(d_segs, d_seg_begin, d_off, cnt, d_in, chunk);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    return B2C_OK;
}

static int lane_create(b2c_ctx* ctx, Lane& l) {
    B2C_CUDA(ctx, cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
    B2C_CUDA(ctx, cudaEventCreateWithFlags(&l.ev_meta, cudaEventDisableTiming));
    B2C_CUDA(ctx, cudaEventCreateWithFlags(&l.ev_done, cudaEventDisableTiming));
    return B2C_OK;
}

static void lane_destroy(Lane& l) {
    if (l.stream) cudaStreamSynchronize(l.stream);
    for (DevBuf* b : {&l.ws_frag, &l.ws_misc, &l.ws_tab, &l.ws_ctr, &l.d_in, &l.d_out})
        if (b->p) cudaFree(b->p);
    for (PinBuf* b : {&l.h_in, &l.h_out})
        if (b->p) cudaFreeHost(b->p);
    if (l.ev_meta) cudaEventDestroy(l.ev_meta);
    if (l.ev_done) cudaEventDestroy(l.ev_done);
    if (l.stream) cudaStreamDestroy(l.stream);
    l = Lane();
}

}  // namespace b2c

using namespace b2c;

// ---------------------------------------------------------------------------
// tunables: which kernel variant runs, pipeline geometry.  The defaults are the fastest measured; the others stay
// selectable (tests/test_gpu_parity.py checks every one for byte-exactness) because which one wins depends on the
// batch shape.  Set before the first call that uses them.
// ---------------------------------------------------------------------------
static const char* const kTunableNames[] = {
    "slab_mb", "host_threads", "trace", "batch_decode", "ring_decode", "block_decode", "lift_scan", "batch_decode_small",
    "nt_gather", "gtab_wps", "small_kernels", "small_lpw_c", "small_lpw_d", "window", "window_stage", "window_smem64",
    "window_wps", "gather_ctas", "gather_warps", "gather_chunk", "lockstep", "lockstep_global", "lockstep_lpw", "lockstep_d",
    "lockstep_d_ctas", "async_gather"};

static int apply_tunable(b2c_ctx* ctx, const char* name, long v) {
    const std::string n(name);
    if (n == "slab_mb") {  // bytes of payload per pipeline slab (fixes the size: no adaptation to the batch)
        ctx->slab_bytes = std::max<size_t>(1, (size_t)v) << 20;
        ctx->slab_fixed = true;
    } else if (n == "host_threads") {
        return b2c_ctx_set_option(ctx, B2C_OPT_HOST_THREADS, (int)std::max(0L, v));
    } else if (n == "trace") ctx->trace = v != 0;
    else if (n == "batch_decode") ctx->batch_decode = v != 0;
    else if (n == "ring_decode") ctx->ring_decode = v != 0;
    else if (n == "block_decode") ctx->block_decode = v != 0;
    else if (n == "lift_scan") ctx->lift_scan = v != 0;
    else if (n == "batch_decode_small") ctx->batch_decode_small = v != 0;
    else if (n == "nt_gather") ctx->nt_gather = v != 0;
    else if (n == "gtab_wps") ctx->gtab_warps_per_sm = (int)v;
    else if (n == "small_kernels") ctx->small_kernels = v != 0;
    else if (n == "small_lpw_c") ctx->small_lpw_compress = (int)v;
    else if (n == "small_lpw_d") ctx->small_lpw_decompress = (int)v;
    else if (n == "window") ctx->window_mask = (int)v;
    else if (n == "window_stage") ctx->window_stage = v != 0;
    else if (n == "window_smem64") ctx->window_smem64 = v != 0;
    else if (n == "window_wps") ctx->window_warps_per_sm = (int)v;
    else if (n == "gather_ctas") ctx->gather_ctas_per_sm = (int)std::max(1L, v);
    else if (n == "gather_warps") ctx->gather_warps = (int)std::min(32L, std::max(1L, v));
    else if (n == "gather_chunk") ctx->gather_chunk = (int)std::min(8192L, std::max(512L, v / 16 * 16));
    else if (n == "lockstep") ctx->small_lockstep = v != 0;
    else if (n == "lockstep_global") ctx->lockstep_global = (int)std::min(32L, std::max(0L, v));
    else if (n == "lockstep_lpw") {
        if (v == 4 || v == 8 || v == 16 || v == 32) ctx->lockstep_lpw = (int)v;
        else return B2C_E_ARG;
    } else if (n == "lockstep_d") ctx->lockstep_decode = v != 0;
    else if (n == "lockstep_d_ctas") ctx->lockstep_decode_ctas = (int)std::min(8L, std::max(1L, v));
    else if (n == "async_gather") ctx->async_gather = v != 0;
    else return B2C_E_ARG;
    return B2C_OK;
}

extern "C" {

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
int b2c_ctx_create(int device, b2c_ctx** out) {
    if (!out) return B2C_E_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0 || device < 0 || device >= count) return B2C_E_NODEVICE;
    b2c_ctx* ctx = new (std::nothrow) b2c_ctx();
    if (!ctx) return B2C_E_NOMEM;
    ctx->device = device;
    // every tunable can also be set through the environment (B2C_<NAME IN CAPITALS>), read once here; the
    // supported way is b2c_ctx_set_tunable
    for (const char* name : kTunableNames) {
        std::string env = "B2C_";
        for (const char* c = name; *c; c++) env += (char)toupper((unsigned char)*c);
        if (const char* v = getenv(env.c_str())) apply_tunable(ctx, name, atol(v));
    }
    int rc = B2C_OK;
    do {
        if ((e = cudaSetDevice(device)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaSetDevice", e); break; }
        cudaDeviceProp prop;
        if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaGetDeviceProperties", e); break; }
        ctx->num_sms = prop.multiProcessorCount;
        if ((rc = lane_create(ctx, ctx->main))) break;
        for (auto& l : ctx->pipe)
            if ((rc = lane_create(ctx, l))) break;
        if (rc) break;
        if ((rc = crc_tables_init(ctx))) break;
        rc = small_kernels_configure(ctx);
        if (!rc) rc = window_kernels_configure(ctx);
        if (!rc) rc = lockstep_kernels_configure(ctx);
    } while (0);
    if (rc != B2C_OK) {
        fprintf(stderr, "b2c_ctx_create failed: %s\n", ctx->err.c_str());
        b2c_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return B2C_OK;
}

void b2c_ctx_destroy(b2c_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    delete static_cast<HostPool*>(ctx->host_pool);
    ctx->host_pool = nullptr;
    lane_destroy(ctx->main);
    for (auto& l : ctx->pipe) lane_destroy(l);
    if (ctx->d_crc_tabA) cudaFree(ctx->d_crc_tabA);
    if (ctx->d_crc_small) cudaFree(ctx->d_crc_small);
    for (auto& v : ctx->prof_ev)
        for (cudaEvent_t e : v) cudaEventDestroy(e);
    delete ctx;
}

const char* b2c_last_error(const b2c_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context"; }
int b2c_ctx_device(const b2c_ctx* ctx) { return ctx ? ctx->device : -1; }
void* b2c_ctx_stream(b2c_ctx* ctx) { return ctx ? (void*)ctx->main.stream : nullptr; }
int b2c_ctx_set_tunable(b2c_ctx* ctx, const char* name, long value) {
    if (!ctx || !name) return B2C_E_ARG;
    return apply_tunable(ctx, name, value);
}

uint64_t b2c_ctx_launch_count(const b2c_ctx* ctx) { return ctx ? ctx->launches : 0; }

int b2c_ctx_set_option(b2c_ctx* ctx, int option, int value) {
    if (!ctx) return B2C_E_ARG;
    switch (option) {
        case B2C_OPT_INPUT_PINNED: ctx->input_pinned = value != 0; return B2C_OK;
        case B2C_OPT_HOST_THREADS:
            if (value < 0 || value > 256) return B2C_E_ARG;
            if (ctx->host_pool && (unsigned)value != ctx->host_threads) {  // takes effect from the next *_batch call
                delete static_cast<HostPool*>(ctx->host_pool);
                ctx->host_pool = nullptr;
            }
            ctx->host_threads = (unsigned)value;
            return B2C_OK;
        default: return set_err(ctx, B2C_E_ARG, "unknown option", cudaSuccess);
    }
}

int b2c_ctx_sync(b2c_ctx* ctx) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->main.stream));
    return B2C_OK;
}

int b2c_dev_alloc(b2c_ctx* ctx, size_t bytes, void** d_ptr) {
    if (!ctx || !d_ptr) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaError_t e = cudaMalloc(d_ptr, bytes + 16);  // room for the aligned unit that holds an arena's last byte
    if (e != cudaSuccess) return set_err(ctx, B2C_E_NOMEM, "cudaMalloc", e);
    return B2C_OK;
}
int b2c_dev_free(b2c_ctx* ctx, void* d_ptr) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaFree(d_ptr));
    return B2C_OK;
}
int b2c_host_alloc_pinned(b2c_ctx* ctx, size_t bytes, void** h_ptr) {
    if (!ctx || !h_ptr) return B2C_E_ARG;
    cudaError_t e = cudaHostAlloc(h_ptr, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) return set_err(ctx, B2C_E_NOMEM, "cudaHostAlloc", e);
    return B2C_OK;
}
int b2c_host_free_pinned(b2c_ctx* ctx, void* h_ptr) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaFreeHost(h_ptr));
    return B2C_OK;
}
int b2c_memcpy_h2d(b2c_ctx* ctx, void* d_dst, const void* h_src, size_t bytes) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, ctx->main.stream));
    return B2C_OK;
}
int b2c_memcpy_d2h(b2c_ctx* ctx, void* h_dst, const void* d_src, size_t bytes) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->main.stream));
    return B2C_OK;
}
int b2c_memset_dev(b2c_ctx* ctx, void* d_dst, int value, size_t bytes) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaMemsetAsync(d_dst, value, bytes, ctx->main.stream));
    return B2C_OK;
}
int b2c_event_create(b2c_ctx* ctx, b2c_event** ev) {
    if (!ctx || !ev) return B2C_E_ARG;
    b2c_event* e = new (std::nothrow) b2c_event();
    if (!e) return B2C_E_NOMEM;
    cudaError_t ce = cudaEventCreate(&e->ev);
    if (ce != cudaSuccess) {
        delete e;
        return set_err(ctx, B2C_E_CUDA, "cudaEventCreate", ce);
    }
    *ev = e;
    return B2C_OK;
}
void b2c_event_destroy(b2c_event* ev) {
    if (!ev) return;
    cudaEventDestroy(ev->ev);
    delete ev;
}
int b2c_event_record(b2c_ctx* ctx, b2c_event* ev) {
    if (!ctx || !ev) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaEventRecord(ev->ev, ctx->main.stream));
    return B2C_OK;
}
int b2c_event_elapsed_ms(b2c_event* start, b2c_event* stop, float* ms) {
    if (!start || !stop || !ms) return B2C_E_ARG;
    if (cudaEventSynchronize(stop->ev) != cudaSuccess) return B2C_E_CUDA;
    if (cudaEventElapsedTime(ms, start->ev, stop->ev) != cudaSuccess) return B2C_E_CUDA;
    return B2C_OK;
}

int b2c_prof_enable(b2c_ctx* ctx, int on) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->main.stream));
    ctx->prof_on = on != 0;
    for (auto& u : ctx->prof_used) u = 0;
    return B2C_OK;
}

int b2c_prof_total_ms(b2c_ctx* ctx, int which, float* total_ms, uint32_t* launches) {
    if (!ctx || which < 0 || which >= B2C_PROF_COUNT || !total_ms) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->main.stream));
    float sum = 0;
    uint32_t pairs = ctx->prof_used[which] / 2;
    for (uint32_t i = 0; i < pairs; i++) {
        float ms = 0;
        B2C_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->prof_ev[which][2 * i], ctx->prof_ev[which][2 * i + 1]));
        sum += ms;
    }
    *total_ms = sum;
    if (launches) *launches = pairs;
    return B2C_OK;
}

// ---------------------------------------------------------------------------
// pure host helpers
// ---------------------------------------------------------------------------
size_t b2c_snappy_max_compressed_length(size_t n) { return 32 + n + n / 6; }

int b2c_snappy_get_uncompressed_length(const void* compressed, size_t n, uint32_t* result) {
    const uint8_t* p = static_cast<const uint8_t*>(compressed);
    uint32_t v = 0;
    for (int i = 0; i < 5; i++) {
        if ((size_t)i >= n) return 0;
        uint32_t b = p[i];
        v |= (b & 127u) << (7 * i);
        if (i < 4 ? (b < 128) : (b < 16)) {
            if (result) *result = v;
            return 1;
        }
    }
    return 0;
}

int b2c_snappy_read_preamble(const void* compressed, size_t n, uint32_t* result) {
    const uint8_t* p = static_cast<const uint8_t*>(compressed);
    uint32_t v = 0, shift = 0;
    size_t i = 0;
    for (;;) {
        if (shift >= 32 || i >= n) return 0;
        uint32_t c = p[i++];
        v |= (c & 0x7f) << shift;
        if (c < 128) break;
        shift += 7;
    }
    if (result) *result = v;
    return (int)i;
}

uint32_t b2c_crc32c_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
uint32_t b2c_crc32c_unmask(uint32_t m) {
    uint32_t rot = m - 0xa282ead8u;
    return (rot >> 17) | (rot << 15);
}

// ---------------------------------------------------------------------------
// layouts
// ---------------------------------------------------------------------------
int b2c_layout_create(b2c_ctx* ctx, uint32_t n, const uint64_t* off, const uint32_t* len, b2c_layout** out) {
    if (!ctx || !out || (n && (!off || !len))) return B2C_E_ARG;
    *out = nullptr;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    b2c_layout* L = new (std::nothrow) b2c_layout();
    if (!L) return B2C_E_NOMEM;
    L->ctx = ctx;
    LayoutPlan plan;
    plan.plan(n, off, len);
    std::vector<uint8_t> blob(plan.blob_bytes);
    plan.fill(blob.data());
    cudaError_t e = cudaMalloc(&L->d_blob, plan.blob_bytes);
    if (e != cudaSuccess) {
        delete L;
        return set_err(ctx, B2C_E_NOMEM, "cudaMalloc(layout)", e);
    }
    e = cudaMemcpy(L->d_blob, blob.data(), plan.blob_bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(L->d_blob);
        delete L;
        return set_err(ctx, B2C_E_CUDA, "layout upload", e);
    }
    plan.bind(static_cast<const uint8_t*>(L->d_blob), &L->v);
    *out = L;
    return B2C_OK;
}

void b2c_layout_destroy(b2c_layout* L) {
    if (!L) return;
    if (L->ctx) {
        cudaSetDevice(L->ctx->device);
        cudaStreamSynchronize(L->ctx->main.stream);
    }
    if (L->d_blob) cudaFree(L->d_blob);
    delete L;
}

uint32_t b2c_layout_count(const b2c_layout* L) { return L ? L->v.n : 0; }
uint64_t b2c_layout_payload_bytes(const b2c_layout* L) { return L ? L->v.payload : 0; }
uint64_t b2c_layout_arena_bytes(const b2c_layout* L) { return L ? L->v.arena_end : 0; }
uint64_t b2c_layout_max_compressed_bytes(const b2c_layout* L) { return L ? L->v.max_comp : 0; }

// ---------------------------------------------------------------------------
// device-resident entry points
// ---------------------------------------------------------------------------
int b2c_crc32c_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_data, const uint32_t* d_init, uint32_t* d_crc) {
    if (!ctx || !L || !d_crc || (L->v.payload && !d_data)) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_crc32c(ctx, ctx->main, L->v, static_cast<const uint8_t*>(d_data), d_init, d_crc);
}

int b2c_snappy_compress_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, void* d_out, uint64_t out_cap,
                            uint64_t* d_out_off) {
    if (!ctx || !L || !d_out || !d_out_off || (L->v.payload && !d_in)) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_compress(ctx, ctx->main, L->v, static_cast<const uint8_t*>(d_in), static_cast<uint8_t*>(d_out),
                           out_cap, d_out_off);
}

int b2c_snappy_uncompressed_length_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, uint32_t* d_ulen,
                                       int32_t* d_status) {
    if (!ctx || !L || !d_ulen) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_ulen(ctx, ctx->main, L->v, static_cast<const uint8_t*>(d_in), d_ulen, d_status);
}

int b2c_snappy_decompress_dev(b2c_ctx* ctx, const b2c_layout* in, const void* d_in, const b2c_layout* out,
                              void* d_out, uint32_t* d_ulen, int32_t* d_status) {
    if (!ctx || !in || !out || !d_status) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_decompress(ctx, ctx->main, in->v, static_cast<const uint8_t*>(d_in), out->v,
                             static_cast<uint8_t*>(d_out), d_ulen, d_status);
}

int b2c_snappy_validate_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, uint32_t* d_ulen, int32_t* d_status) {
    if (!ctx || !L || !d_status) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_validate(ctx, ctx->main, L->v, static_cast<const uint8_t*>(d_in), d_ulen, d_status);
}

int b2c_snappy_decodable_prefix_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, uint32_t* d_out5) {
    if (!ctx || !L || !d_out5) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_prefix(ctx, ctx->main, L->v, static_cast<const uint8_t*>(d_in), d_out5);
}

int b2c_snappy_decompress_verify_dev(b2c_ctx* ctx, const b2c_layout* in, const void* d_in, const b2c_layout* out,
                                     void* d_out, const uint32_t* d_masked_crc, uint32_t* d_ulen, int32_t* d_status) {
    if (!ctx || !in || !out || !d_status || !d_masked_crc) return B2C_E_ARG;
    if (in->v.n != out->v.n) return set_err(ctx, B2C_E_ARG, "decompress_verify: layouts differ in message count", cudaSuccess);
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_decompress_verify(ctx, ctx->main, in->v, static_cast<const uint8_t*>(d_in), out->v,
                                    static_cast<uint8_t*>(d_out), d_masked_crc, d_ulen, d_status);
}

// ---------------------------------------------------------------------------
// host (end-to-end) entry points: slab pipeline
// ---------------------------------------------------------------------------
namespace {

using Clock = std::chrono::steady_clock;
inline double secs(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double>(b - a).count(); }

struct Slab {
    uint32_t m0, m1;   // message range
    uint64_t bytes;    // staging bytes of the payload (messages 16-byte aligned)
    uint64_t payload;  // sum of message lengths
};

struct BatchPlan {
    std::vector<uint64_t>& off;  // per message: offset inside its slab's staging
    std::vector<uint32_t>& len;
    std::vector<Slab> slabs;
    uint64_t payload = 0;
    explicit BatchPlan(b2c_ctx* ctx) : off(ctx->bp_off), len(ctx->bp_len) {}
};

// Is `p` page-locked host memory CUDA can DMA into directly (cudaHostAlloc / cudaHostRegister)?
bool is_pinned_host(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Message lengths, staging offsets and slab cuts.  Staged mode: a message of >= 256 bytes takes a multiple of 64
// bytes of staging (so runs of them start on 64-byte lines and the gather writes whole lines only), a shorter one
// a multiple of 16; sizes are therefore independent of position, the offsets are a prefix sum done by the pool,
// and a slab is the longest run of messages that fits the target (binary search on the prefix sums).
int plan_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, BatchPlan& p) {
    if (p.off.size() < (size_t)n + 1) p.off.resize((size_t)n + 1);
    if (p.len.size() < n) p.len.resize(n);
    std::atomic<bool> too_long(false);
    const bool zc = ctx->input_pinned;
    // zero-copy: a message gets the same phase modulo 16 as its first block has in host memory (so that block moves
    // in whole aligned 16-byte pieces), i.e. `skew` bytes of slack in front of it
    auto skew_of = [&](uint32_t i) -> uint32_t {
        return zc && seg_begin[i] < seg_begin[i + 1] ? (uint32_t)(reinterpret_cast<uintptr_t>(segs[seg_begin[i]].iov_base) & 15) : 0u;
    };
    auto staged_size = [&](uint32_t i, uint32_t l) -> uint64_t {
        if (zc) return align_up((uint64_t)skew_of(i) + l, 16);
        return l >= 256 ? align_up(l, 64) : align_up(l, 16);
    };
    uint64_t part[64] = {0};
    unsigned used = 1;
    std::atomic<bool> pageable(false);
    pool(ctx).run(threads_for((uint64_t)n * 64), [&](unsigned t, unsigned T) {
        if (t == 0) used = T;
        uint32_t lo = (uint32_t)((uint64_t)n * t / T), hi = (uint32_t)((uint64_t)n * (t + 1) / T);
        uint64_t sum = 0;
        // zero-copy: the kernel will dereference every block, and a pageable one is a fatal device fault, so each
        // 2 MiB region a block touches is looked up once per thread (cudaPointerGetAttributes) and remembered
        uint64_t seen[64];
        for (auto& x : seen) x = ~0ull;
        auto pinned_region = [&](uintptr_t a) {
            const uint64_t r = (uint64_t)a >> 21;
            uint64_t& slot = seen[(r * 0x9e3779b97f4a7c15ull) >> 58];
            if (slot == r) return;
            if (!is_pinned_host(reinterpret_cast<const void*>(a))) pageable = true;
            slot = r;
        };
        for (uint32_t i = lo; i < hi; i++) {
            uint64_t l = 0;
            for (uint32_t s = seg_begin[i]; s < seg_begin[i + 1]; s++) l += segs[s].iov_len;
            if (zc)
                for (uint32_t s = seg_begin[i]; s < seg_begin[i + 1]; s++)
                    if (segs[s].iov_len) {
                        const uintptr_t a = reinterpret_cast<uintptr_t>(segs[s].iov_base);
                        pinned_region(a);
                        pinned_region(a + segs[s].iov_len - 1);
                    }
            if (l > 0xffffffffull) {
                too_long = true;
                l = 0;
            }
            p.len[i] = (uint32_t)l;
            sum += staged_size(i, (uint32_t)l);
        }
        part[t] = sum;
    });
    if (too_long) return set_err(ctx, B2C_E_ARG, "message longer than 4 GiB - 1", cudaSuccess);
    if (pageable)
        return set_err(ctx, B2C_E_ARG, "B2C_OPT_INPUT_PINNED is set but an input block is not in page-locked host memory",
                       cudaSuccess);
    // Slab size: ~32 MiB keeps gather, copies and kernels of different slabs overlapped; batches of large messages get
    // up to 128 MiB, because a 64 KiB fragment occupies a warp for milliseconds and a 32 MiB slab holds only 512 of them
    // (measured end to end, 32 -> 128 MiB: C5 10.4 -> 15.2, C4 8.3 -> 13.1 GB/s).  B2C_SLAB_MB fixes the size.
    uint64_t target = ctx->slab_bytes;
    if (!ctx->slab_fixed && n) {
        uint64_t total = 0;
        for (unsigned t = 0; t < used; t++) total += part[t];
        target = std::min<uint64_t>(128ull << 20, std::max<uint64_t>(target, total / n * 8192));
    }
    const uint32_t max_msgs = 1u << 18;
    {
        uint64_t base[65];
        base[0] = 0;
        for (unsigned t = 0; t < used; t++) base[t + 1] = base[t] + part[t];
        uint64_t* g = p.off.data();  // g[i] = staging position of message i if the batch were one slab
        pool(ctx).run(used, [&](unsigned t, unsigned T) {
            uint32_t lo = (uint32_t)((uint64_t)n * t / T), hi = (uint32_t)((uint64_t)n * (t + 1) / T);
            uint64_t pos = base[t];
            for (uint32_t i = lo; i < hi; i++) {
                g[i] = pos;
                pos += staged_size(i, p.len[i]);
            }
        });
        g[n] = base[used];
        for (uint32_t s = 0; s < n;) {
            // the first two slabs are smaller: the pipeline (gather -> H2D -> kernels -> D2H) fills sooner
            const uint64_t tgt = p.slabs.empty() ? target / 4 : p.slabs.size() == 1 ? target / 2 : target;
            const uint32_t lim = (uint32_t)std::min<uint64_t>(n, (uint64_t)s + max_msgs);
            uint32_t e = (uint32_t)(std::upper_bound(g + s + 1, g + lim + 1, g[s] + tgt) - g) - 1;
            if (e < s + 1) e = s + 1;  // a message larger than the target is a slab of its own
            p.slabs.push_back({s, e, g[e] - g[s], 0});
            s = e;
        }
        // offsets relative to each slab's start, and the slab's payload
        std::vector<Slab>& slabs = p.slabs;
        pool(ctx).run((unsigned)std::min<size_t>(slabs.size(), 64), [&](unsigned t, unsigned T) {
            for (size_t k = t; k < slabs.size(); k += T) {
                const uint64_t start = g[slabs[k].m0];
                uint64_t payload = 0;
                for (uint32_t i = slabs[k].m0; i < slabs[k].m1; i++) {
                    g[i] = g[i] - start + skew_of(i);
                    payload += p.len[i];
                }
                slabs[k].payload = payload;
            }
        });
        for (const Slab& sl : slabs) p.payload += sl.payload;
        return B2C_OK;
    }
}

// memcpy every segment of the slab's messages into pinned staging.  When `caps` is given the
// messages are snappy streams: the thread that just copied a stream also reads its (lenient)
// preamble while the bytes are still in its cache and records the plausible output size
// (0 = malformed or impossible claim; a stream cannot expand more than 64/3 times, snappy.cc:867).
// Layout statistics the gathering workers collect on the way (so the submitting thread does not
// have to walk the slab's messages again): one PlanStats per worker, merged by the caller.
struct SlabStats {
    PlanStats in[64];    // of the slab's messages as laid out in staging
    PlanStats caps[64];  // of the preamble-derived output slots (decompress)
    void merged(PlanStats& a, PlanStats& b) const {
        for (int t = 0; t < 64; t++) a.merge(in[t]), b.merge(caps[t]);
    }
};

std::function<void(unsigned, unsigned)> gather_job(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                                                   const BatchPlan& p, const Slab s, uint8_t* stage, uint32_t* caps,
                                                   SlabStats* stats) {
    const uint32_t cnt = s.m1 - s.m0;
    const bool stream = ctx->nt_gather;
    const uint64_t* poff = p.off.data();
    const uint32_t* plen = p.len.data();
    // messages are handed out in chunks of ~256 KiB from a shared counter: with every core busy a
    // static split would wait for whichever worker the OS preempted
    auto next = std::make_shared<std::atomic<uint32_t>>(0u);
    const uint32_t chunk = (uint32_t)std::min<uint64_t>(1024, std::max<uint64_t>(1, (256u << 10) / (s.bytes / std::max(cnt, 1u) + 1)));
    return [=](unsigned t, unsigned) {
        for (;;) {
        const uint32_t c0 = next->fetch_add(chunk, std::memory_order_relaxed);
        if (c0 >= cnt) break;
        const uint32_t lo = s.m0 + c0, hi = s.m0 + std::min(cnt, c0 + chunk);
        if (stats) plan_range(poff, plen, lo, hi, stats->in[t & 63]);
        for (uint32_t i = lo; i < hi; i++) {
            uint8_t* d = stage + poff[i];
            if (caps) {
                uint8_t head[5];
                size_t got = 0;
                for (uint32_t q = seg_begin[i]; q < seg_begin[i + 1] && got < 5; q++)
                    for (size_t j = 0; j < segs[q].iov_len && got < 5; j++)
                        head[got++] = static_cast<const uint8_t*>(segs[q].iov_base)[j];
                uint32_t ulen = 0;
                const int kk = b2c_snappy_read_preamble(head, got, &ulen);
                const bool plausible = kk > 0 && (uint64_t)ulen <= ((uint64_t)plen[i] * 64) / 3 + 64;
                caps[i - s.m0] = plausible ? ulen : 0;
            }
            gather_message(d, segs + seg_begin[i], seg_begin[i + 1] - seg_begin[i], plen[i], stream);
        }
        if (stats && caps) plan_range(nullptr, caps, lo - s.m0, hi - s.m0, stats->caps[t & 63]);
        }
        if (stream) gather_fence();
    };
}

// Start gathering a slab on the pool's workers; the calling thread goes on (layouts, submits,
// copy-outs of other slabs) and calls pool(ctx).wait() before it touches the staging.
void gather_slab_start(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, const BatchPlan& p, const Slab& s,
                       uint8_t* stage, uint32_t* caps, SlabStats* stats) {
    *stats = SlabStats();
    pool(ctx).start(threads_for(s.bytes), gather_job(ctx, segs, seg_begin, p, s, stage, caps, stats));
    if (!ctx->async_gather) pool(ctx).wait();
}

// Ends a started section on every way out of a *_batch call.
struct PoolGuard {
    b2c_ctx* ctx;
    ~PoolGuard() { pool(ctx).wait(); }
};

void parallel_copy(b2c_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    if (!bytes) return;
    pool(ctx).run(threads_for(bytes), [&](unsigned t, unsigned T) {
        uint64_t lo = bytes * t / T, hi = bytes * (t + 1) / T;
        memcpy(static_cast<uint8_t*>(dst) + lo, static_cast<const uint8_t*>(src) + lo, hi - lo);
    });
}

// preamble-derived output slot sizes of a slab of snappy streams, read straight from the segments
void slab_caps(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, const BatchPlan& p, const Slab& s,
               uint32_t* caps) {
    const uint32_t cnt = s.m1 - s.m0;
    pool(ctx).run(threads_for((uint64_t)cnt * 256), [&](unsigned t, unsigned T) {
        uint32_t lo = s.m0 + (uint32_t)((uint64_t)cnt * t / T), hi = s.m0 + (uint32_t)((uint64_t)cnt * (t + 1) / T);
        for (uint32_t i = lo; i < hi; i++) {
            uint8_t head[5];
            size_t got = 0;
            for (uint32_t q = seg_begin[i]; q < seg_begin[i + 1] && got < 5; q++)
                for (size_t j = 0; j < segs[q].iov_len && got < 5; j++)
                    head[got++] = static_cast<const uint8_t*>(segs[q].iov_base)[j];
            uint32_t ulen = 0;
            const int kk = b2c_snappy_read_preamble(head, got, &ulen);
            const bool plausible = kk > 0 && (uint64_t)ulen <= ((uint64_t)p.len[i] * 64) / 3 + 64;
            caps[i - s.m0] = plausible ? ulen : 0;
        }
    });
}

// One slab's way into HBM.  Device image: payload at [0, s.bytes) (messages 16-byte aligned), the
// caller's metadata (layout blobs, ...) at meta_at.  Two modes:
//   staged    (default): host threads gather the blocks into pinned staging, one H2D copy moves
//                        payload + metadata;
//   zero-copy (B2C_OPT_INPUT_PINNED): the blocks already live in page-locked memory (a pinned
//                        IOBuf block pool); only metadata + the block list are copied and a kernel
//                        pulls the payload over PCIe straight from the blocks.
struct SlabIO {
    uint8_t* stage_meta = nullptr;  // host: where the caller writes its metadata
    uint8_t* d_in = nullptr;        // device image base
    size_t meta_at = 0;             // device offset of the metadata
    size_t seg_at = 0;              // zero-copy: offset (from meta start) of the block list
    uint32_t nseg = 0;
};

int slab_begin(b2c_ctx* ctx, Lane& lane, const uint32_t* seg_begin, const Slab& s, size_t meta_cap, SlabIO& io,
               size_t keep = 0) {
    const uint32_t cnt = s.m1 - s.m0;
    io.meta_at = align_up(s.bytes, 256);
    io.nseg = seg_begin[s.m1] - seg_begin[s.m0];
    io.seg_at = align_up(meta_cap, 256);
    const size_t zc_bytes = ctx->input_pinned ? (size_t)io.nseg * sizeof(b2c_iovec) + ((size_t)cnt + 1) * 4 + 256 : 0;
    const size_t meta_total = io.seg_at + zc_bytes;
    int rc;
    if ((rc = ws_reserve(ctx, lane, lane.d_in, io.meta_at + meta_total + 64))) return rc;
    if ((rc = pin_reserve(ctx, lane, lane.h_in, (ctx->input_pinned ? 0 : io.meta_at) + meta_total + 64, keep))) return rc;
    io.d_in = static_cast<uint8_t*>(lane.d_in.p);
    io.stage_meta = static_cast<uint8_t*>(lane.h_in.p) + (ctx->input_pinned ? 0 : io.meta_at);
    return B2C_OK;
}

// `meta_bytes` of metadata are in io.stage_meta; `d_off` = device address of the slab's message offsets
int slab_submit(b2c_ctx* ctx, Lane& lane, const b2c_iovec* segs, const uint32_t* seg_begin, const Slab& s, SlabIO& io,
                size_t meta_bytes, const uint64_t* d_off, uint64_t& h2d) {
    const uint32_t cnt = s.m1 - s.m0;
    if (!ctx->input_pinned) {
        const size_t total = io.meta_at + meta_bytes;
        B2C_CUDA(ctx, cudaMemcpyAsync(io.d_in, lane.h_in.p, total, cudaMemcpyHostToDevice, lane.stream));
        h2d += total;
        return B2C_OK;
    }
    b2c_iovec* h_segs = reinterpret_cast<b2c_iovec*>(io.stage_meta + io.seg_at);
    uint32_t* h_begin = reinterpret_cast<uint32_t*>(io.stage_meta + io.seg_at + (size_t)io.nseg * sizeof(b2c_iovec));
    const uint32_t s0 = seg_begin[s.m0];
    memcpy(h_segs, segs + s0, (size_t)io.nseg * sizeof(b2c_iovec));
    for (uint32_t i = 0; i <= cnt; i++) h_begin[i] = seg_begin[s.m0 + i] - s0;
    const size_t total = io.seg_at + (size_t)io.nseg * sizeof(b2c_iovec) + ((size_t)cnt + 1) * 4;
    B2C_CUDA(ctx, cudaMemcpyAsync(io.d_in + io.meta_at, io.stage_meta, total, cudaMemcpyHostToDevice, lane.stream));
    h2d += total + s.payload;  // the payload crosses PCIe inside the gather kernel
    const uint8_t* d_meta = io.d_in + io.meta_at;
    return launch_gather_host(ctx, lane, reinterpret_cast<const b2c_iovec*>(d_meta + io.seg_at),
                              reinterpret_cast<const uint32_t*>(d_meta + io.seg_at + (size_t)io.nseg * sizeof(b2c_iovec)),
                              d_off, cnt, io.d_in);
}

struct Trace {
    bool on;
    const char* what;
    double plan = 0, gather = 0, layout = 0, submit = 0, wait = 0, copy_out = 0;
    Clock::time_point t0;
    Trace(bool on_, const char* w) : on(on_), what(w), t0(Clock::now()) {}
    void report(size_t slabs, uint64_t payload) {
        if (!on) return;
        double total = secs(t0, Clock::now());
        fprintf(stderr,
                "[b2c] %s: %.2f ms total, %zu slabs, %.1f MB payload | plan %.2f gather %.2f layout %.2f submit %.2f "
                "wait %.2f copy-out %.2f ms\n",
                what, total * 1e3, slabs, payload / 1e6, plan * 1e3, gather * 1e3, layout * 1e3, submit * 1e3,
                wait * 1e3, copy_out * 1e3);
    }
};

#define TIMED(acc, stmt)                       \
    do {                                       \
        auto _t = Clock::now();                \
        stmt;                                  \
        (acc) += secs(_t, Clock::now());       \
    } while (0)

}  // namespace

void b2c_last_transfer_bytes(const b2c_ctx* ctx, uint64_t* h2d, uint64_t* d2h) {
    if (h2d) *h2d = ctx ? ctx->last_h2d : 0;
    if (d2h) *d2h = ctx ? ctx->last_d2h : 0;
}

int b2c_crc32c_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                     const uint32_t* init, uint32_t* crc) {
    if (!ctx || (n && (!segs || !seg_begin || !crc))) return B2C_E_ARG;
    if (n == 0) return B2C_OK;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    Trace tr(ctx->trace, "crc32c_batch");
    BatchPlan bp(ctx);
    int rc;
    TIMED(tr.plan, rc = plan_batch(ctx, segs, seg_begin, n, bp));
    if (rc) return rc;
    uint64_t h2d = 0, d2h = 0;
    const size_t K = bp.slabs.size();
    auto finish = [&](size_t k) -> int {
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        TIMED(tr.wait, B2C_CUDA(ctx, cudaEventSynchronize(lane.ev_done)));
        TIMED(tr.copy_out, memcpy(crc + s.m0, lane.h_out.p, (size_t)(s.m1 - s.m0) * 4));
        return B2C_OK;
    };
    // Slab k+1 is gathered by the pool's workers while this thread lays out, submits and retires
    // the slabs around it.
    // The staging offset of the metadata is known before the layout is (meta_at depends on the payload only), so
    // the gather can start first and the layout is sized afterwards from the statistics the workers collected.
    struct Prep {
        LayoutPlan lp;
        SlabIO io;
        size_t init_rel = 0, meta_bytes = 0;
    } prep[2];
    std::unique_ptr<SlabStats[]> stats(new SlabStats[2]);
    PoolGuard guard{ctx};
    auto layout = [&](size_t k) {  // after the gather: size the layout
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        Prep& P = prep[k & 1];
        if (ctx->input_pinned) {
            P.lp.plan(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, ctx);
        } else {
            PlanStats a, b;
            stats[k & 1].merged(a, b);
            P.lp.plan_with(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, a);
        }
        P.init_rel = align_up(P.lp.blob_bytes, 256);  // metadata: layout blob | init values
        P.meta_bytes = P.init_rel + (init ? (size_t)cnt * 4 : 0);
    };
    auto prepare = [&](size_t k) -> int {  // reserve the lane's input side and start the gather
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        int r;
        Prep& P = prep[k & 1];
        if (ctx->input_pinned) layout(k);  // no gather to hide it behind: the exact size now
        const size_t meta_cap = ctx->input_pinned ? P.meta_bytes
                                                  : LayoutPlan::max_blob_bytes(cnt, s.payload) + 256 + (size_t)cnt * 4;
        if ((r = slab_begin(ctx, lane, seg_begin, s, meta_cap, P.io))) return r;
        if ((r = pin_reserve(ctx, lane, lane.h_out, (size_t)cnt * 4 + 64))) return r;
        if ((r = ws_reserve(ctx, lane, lane.d_out, (size_t)cnt * 4 + 64))) return r;
        if (!ctx->input_pinned)
            gather_slab_start(ctx, segs, seg_begin, bp, s, static_cast<uint8_t*>(lane.h_in.p), nullptr, &stats[k & 1]);
        return B2C_OK;
    };
    if ((rc = prepare(0))) return rc;
    for (size_t k = 0; k < K; k++) {
        TIMED(tr.gather, pool(ctx).wait());
        if (k + 1 >= (size_t)kPipeLanes && (rc = finish(k + 1 - kPipeLanes))) return rc;  // frees slab k+1's lane
        if (k + 1 < K && (rc = prepare(k + 1))) return rc;
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        Prep& P = prep[k & 1];
        if (!ctx->input_pinned) TIMED(tr.layout, layout(k));
        TIMED(tr.layout, P.lp.fill(P.io.stage_meta));
        if (init) memcpy(P.io.stage_meta + P.init_rel, init + s.m0, (size_t)cnt * 4);
        auto ts = Clock::now();
        uint8_t* d_in = P.io.d_in;
        LayoutView lv;
        P.lp.bind(d_in + P.io.meta_at, &lv);
        if ((rc = slab_submit(ctx, lane, segs, seg_begin, s, P.io, P.meta_bytes, lv.d_off, h2d))) return rc;
        const size_t init_at = P.io.meta_at + P.init_rel;
        uint32_t* d_crc = static_cast<uint32_t*>(lane.d_out.p);
        if ((rc = launch_crc32c(ctx, lane, lv, d_in, init ? reinterpret_cast<uint32_t*>(d_in + init_at) : nullptr, d_crc)))
            return rc;
        B2C_CUDA(ctx, cudaMemcpyAsync(lane.h_out.p, d_crc, (size_t)cnt * 4, cudaMemcpyDeviceToHost, lane.stream));
        d2h += (size_t)cnt * 4;
        B2C_CUDA(ctx, cudaEventRecord(lane.ev_done, lane.stream));
        tr.submit += secs(ts, Clock::now());
    }
    for (size_t k = K + 1 > (size_t)kPipeLanes ? K + 1 - kPipeLanes : 0; k < K; k++)
        if ((rc = finish(k))) return rc;
    ctx->last_h2d = h2d;
    ctx->last_d2h = d2h;
    tr.report(K, bp.payload);
    return B2C_OK;
}

// compress (+ optionally checksum the same staged bodies): crc == nullptr is plain b2c_snappy_compress_batch
static int compress_batch_impl(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                               uint64_t out_cap, uint64_t* out_off, uint32_t* crc, bool masked) {
    if (!ctx || !out_off || (n && (!segs || !seg_begin || !out))) return B2C_E_ARG;
    out_off[0] = 0;
    if (n == 0) return B2C_OK;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    Trace tr(ctx->trace, crc ? "compress_crc_batch" : "compress_batch");
    BatchPlan bp(ctx);
    int rc;
    TIMED(tr.plan, rc = plan_batch(ctx, segs, seg_begin, n, bp));
    if (rc) return rc;
    uint64_t h2d = 0, d2h = 0, out_base = 0;
    const size_t K = bp.slabs.size();
    std::vector<uint64_t> slab_total(K, 0), slab_base(K, 0);
    const bool direct = is_pinned_host(out);  // caller's buffer is page-locked: DMA straight into it
    uint64_t fetch_base = 0;
    // stage B: sizes are on the host -> fetch the packed bytes
    auto fetch = [&](size_t k) -> int {
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        TIMED(tr.wait, B2C_CUDA(ctx, cudaEventSynchronize(lane.ev_meta)));
        const uint64_t* meta = static_cast<const uint64_t*>(lane.h_out.p);
        const uint64_t total = meta[cnt];
        slab_total[k] = total;
        slab_base[k] = fetch_base;
        if (fetch_base + total > out_cap)
            return set_err(ctx, B2C_E_CAPACITY, "compress_batch: out_cap too small", cudaSuccess);
        fetch_base += total;
        const size_t data_at = align_up(((size_t)cnt + 1) * 8 + (crc ? (size_t)cnt * 4 : 0), 256);
        uint8_t* dst = direct ? static_cast<uint8_t*>(out) + slab_base[k] : static_cast<uint8_t*>(lane.h_out.p) + data_at;
        B2C_CUDA(ctx, cudaMemcpyAsync(dst, lane.d_out.p, total, cudaMemcpyDeviceToHost, lane.stream));
        d2h += total;
        B2C_CUDA(ctx, cudaEventRecord(lane.ev_done, lane.stream));
        return B2C_OK;
    };
    // stage C: packed bytes are on the host -> caller's buffer
    auto finish = [&](size_t k) -> int {
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        TIMED(tr.wait, B2C_CUDA(ctx, cudaEventSynchronize(lane.ev_done)));
        const uint64_t* meta = static_cast<const uint64_t*>(lane.h_out.p);
        const size_t data_at = align_up(((size_t)cnt + 1) * 8 + (crc ? (size_t)cnt * 4 : 0), 256);
        auto tc = Clock::now();
        if (!direct)
            parallel_copy(ctx, static_cast<uint8_t*>(out) + out_base, static_cast<uint8_t*>(lane.h_out.p) + data_at,
                          slab_total[k]);
        for (uint32_t i = 0; i < cnt; i++) out_off[s.m0 + i + 1] = out_base + meta[i + 1];
        if (crc) {
            const uint32_t* c = reinterpret_cast<const uint32_t*>(meta + cnt + 1);
            if (masked)
                for (uint32_t i = 0; i < cnt; i++) crc[s.m0 + i] = b2c_crc32c_mask(c[i]);
            else
                memcpy(crc + s.m0, c, (size_t)cnt * 4);
        }
        out_base += slab_total[k];
        tr.copy_out += secs(tc, Clock::now());
        return B2C_OK;
    };
    // Software pipeline over the lanes.  While the pool's workers gather slab k+1, this thread retires slab k-3
    // (which frees slab k+1's lane), lays out and submits slab k, and reads slab k-2's sizes to fetch its bytes --
    // so it never waits for the slab it has just submitted.
    struct Prep {
        LayoutPlan lp;
        SlabIO io;
        size_t meta_bytes = 0, out_meta_bytes = 0, off_at = 0;
        uint64_t cap = 0;
    } prep[2];
    const bool timeline = getenv("B2C_TIMELINE") != nullptr;
    std::vector<cudaEvent_t> tl;
    auto tl_mark = [&](cudaStream_t st) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) == cudaSuccess) {
            cudaEventRecord(e, st);
            tl.push_back(e);
        }
    };
    std::unique_ptr<SlabStats[]> stats(new SlabStats[2]);
    PoolGuard guard{ctx};
    auto layout = [&](size_t k) {  // size the layout: from the workers' statistics once the gather is through
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        Prep& P = prep[k & 1];
        if (ctx->input_pinned) {
            P.lp.plan(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, ctx);
        } else {
            PlanStats a, b;
            stats[k & 1].merged(a, b);
            P.lp.plan_with(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, a);
        }
        P.meta_bytes = P.lp.blob_bytes;
    };
    auto prepare = [&](size_t k) -> int {  // reserve the lane's buffers and start the gather
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        Prep& P = prep[k & 1];
        if (ctx->input_pinned) layout(k);
        // sum of MaxCompressedLength(len)+5 without walking the messages: 32 + len + len/6 + 5 each
        P.cap = s.payload + s.payload / 6 + 37ull * cnt;
        P.out_meta_bytes = ((size_t)cnt + 1) * 8 + (crc ? (size_t)cnt * 4 : 0);
        P.off_at = align_up(P.cap, 256);  // device: packed output, then its offsets (then the checksums)
        int r;
        const size_t meta_cap = ctx->input_pinned ? P.meta_bytes : LayoutPlan::max_blob_bytes(cnt, s.payload);
        if ((r = slab_begin(ctx, lane, seg_begin, s, meta_cap, P.io))) return r;
        if ((r = ws_reserve(ctx, lane, lane.d_out, P.off_at + P.out_meta_bytes + 64))) return r;
        if ((r = pin_reserve(ctx, lane, lane.h_out, align_up(P.out_meta_bytes, 256) + (direct ? 0 : P.cap) + 64))) return r;
        if (!ctx->input_pinned)
            gather_slab_start(ctx, segs, seg_begin, bp, s, static_cast<uint8_t*>(lane.h_in.p), nullptr, &stats[k & 1]);
        return B2C_OK;
    };
    // Can slab k's gather start before the slab that used its lane last (k - 4) has been retired?  Its staging is free
    // (that slab's sizes were read two iterations ago, so its H2D copy is long done); what must not happen is a
    // buffer of the lane being re-allocated under the retiring slab, so every capacity has to suffice already.
    auto fits = [&](size_t k) -> bool {
        const Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        const size_t meta_at = align_up(s.bytes, 256);
        const size_t meta_total = align_up(LayoutPlan::max_blob_bytes(cnt, s.payload), 256);
        const uint64_t cap = s.payload + s.payload / 6 + 37ull * cnt;
        const size_t out_meta = ((size_t)cnt + 1) * 8 + (crc ? (size_t)cnt * 4 : 0);
        return meta_at + meta_total + 64 <= lane.d_in.cap && meta_at + meta_total + 64 <= lane.h_in.cap &&
               align_up(cap, 256) + out_meta + 64 <= lane.d_out.cap && align_up(out_meta, 256) + 64 <= lane.h_out.cap;
    };
    if ((rc = prepare(0))) return rc;
    for (size_t k = 0; k < K + 3; k++) {
        if (k < K) TIMED(tr.gather, pool(ctx).wait());
        // keep the gathering threads busy: slab k+1's gather starts before this thread retires slab k-3, when nothing
        // has to grow (outputs DMA'd straight into the caller's page-locked buffer, so retiring needs no pool either)
        const bool early = k + 1 < K && direct && !ctx->input_pinned && fits(k + 1);
        if (early && (rc = prepare(k + 1))) return rc;
        if (k >= 3 && k - 3 < K && (rc = finish(k - 3))) return rc;  // frees slab k+1's lane
        if (k < K) {
            if (k + 1 < K && !early && (rc = prepare(k + 1))) return rc;
            Lane& lane = ctx->pipe[k % kPipeLanes];
            const Slab& s = bp.slabs[k];
            const uint32_t cnt = s.m1 - s.m0;
            Prep& P = prep[k & 1];
            if (!ctx->input_pinned) TIMED(tr.layout, layout(k));
            TIMED(tr.layout, P.lp.fill(P.io.stage_meta));
            auto ts = Clock::now();
            uint8_t* d_in = P.io.d_in;
            LayoutView lv;
            P.lp.bind(d_in + P.io.meta_at, &lv);
            if (timeline) tl_mark(lane.stream);
            if ((rc = slab_submit(ctx, lane, segs, seg_begin, s, P.io, P.meta_bytes, lv.d_off, h2d))) return rc;
            if (timeline) tl_mark(lane.stream);
            uint8_t* d_out = static_cast<uint8_t*>(lane.d_out.p);
            uint64_t* d_out_off = reinterpret_cast<uint64_t*>(d_out + P.off_at);
            if (crc && (rc = launch_crc32c(ctx, lane, lv, d_in, nullptr, reinterpret_cast<uint32_t*>(d_out_off + cnt + 1))))
                return rc;
            if ((rc = launch_compress(ctx, lane, lv, d_in, d_out, P.cap, d_out_off))) return rc;
            if (timeline) tl_mark(lane.stream);
            B2C_CUDA(ctx, cudaMemcpyAsync(lane.h_out.p, d_out_off, P.out_meta_bytes, cudaMemcpyDeviceToHost, lane.stream));
            d2h += P.out_meta_bytes;
            B2C_CUDA(ctx, cudaEventRecord(lane.ev_meta, lane.stream));
            tr.submit += secs(ts, Clock::now());
        }
        if (k >= 2 && k - 2 < K && (rc = fetch(k - 2))) return rc;
    }
    ctx->last_h2d = h2d;
    ctx->last_d2h = d2h;
    tr.report(K, bp.payload);
    if (timeline) {  // B2C_TIMELINE=1: when each slab's input transfer and kernels ran on the GPU, relative to the first mark
        for (size_t i = 0; i + 2 < tl.size(); i += 3) {
            float a = 0, b = 0, c = 0;
            cudaEventSynchronize(tl[i + 2]);
            cudaEventElapsedTime(&a, tl[0], tl[i]);
            cudaEventElapsedTime(&b, tl[0], tl[i + 1]);
            cudaEventElapsedTime(&c, tl[0], tl[i + 2]);
            fprintf(stderr, "[b2c timeline] slab %2zu  input %7.3f .. %7.3f  kernels .. %7.3f ms\n", i / 3, a, b, c);
        }
        for (cudaEvent_t e : tl) cudaEventDestroy(e);
    }
    return B2C_OK;
}

int b2c_snappy_compress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                              uint64_t out_cap, uint64_t* out_off) {
    return compress_batch_impl(ctx, segs, seg_begin, n, out, out_cap, out_off, nullptr, false);
}

int b2c_snappy_compress_crc_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                                  uint64_t out_cap, uint64_t* out_off, uint32_t* crc, int masked) {
    if (n && !crc) return B2C_E_ARG;
    return compress_batch_impl(ctx, segs, seg_begin, n, out, out_cap, out_off, crc, masked != 0);
}

// decompress (+ optionally verify each result against masked_crc): masked_crc == nullptr is plain b2c_snappy_decompress_batch
// Output goes either to one flat buffer (out/out_cap/out_off) or, when oiov is given, message i is scattered over
// oiov[oiov_begin[i] .. oiov_begin[i+1]) the way SnappyIOVecWriter fills its iovecs (snappy.cc:963-1120).
static int decompress_batch_impl(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                                 uint64_t out_cap, uint64_t* out_off, int32_t* status, const uint32_t* masked_crc,
                                 const b2c_iovec* oiov = nullptr, const uint32_t* oiov_begin = nullptr) {
    std::vector<uint64_t> own_off;
    if (oiov_begin) {  // the flat offsets are internal in this mode
        own_off.resize((size_t)n + 1);
        out_off = own_off.data();
        out_cap = ~0ull;
        out = nullptr;
    }
    if (!ctx || !out_off || (n && (!segs || !seg_begin || !status))) return B2C_E_ARG;
    out_off[0] = 0;
    if (n == 0) return B2C_OK;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    Trace tr(ctx->trace, oiov_begin ? "decompress_iov_batch" : masked_crc ? "decompress_verify_batch" : "decompress_batch");
    // scatter mode: per in-flight slab, each message's offset in the slab's staged output and its slot length
    // (0 with too_small set when the iovecs cannot hold what the preamble claims: SnappyIOVecWriter would overrun)
    std::vector<uint64_t> fin_off[kPipeLanes];
    std::vector<uint32_t> fin_len[kPipeLanes];
    std::vector<uint8_t> fin_small[kPipeLanes];
    BatchPlan bp(ctx);
    int rc;
    TIMED(tr.plan, rc = plan_batch(ctx, segs, seg_begin, n, bp));
    if (rc) return rc;
    uint64_t h2d = 0, d2h = 0, out_base = 0;
    const size_t K = bp.slabs.size();
    std::vector<uint64_t> slab_out(K, 0);
    std::vector<uint64_t>& o_off = ctx->tmp_off;
    const bool direct = !oiov_begin && is_pinned_host(out);  // caller's buffer is page-locked: DMA straight into it
    auto finish = [&](size_t k) -> int {
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        TIMED(tr.wait, B2C_CUDA(ctx, cudaEventSynchronize(lane.ev_done)));
        auto tc = Clock::now();
        const uint8_t* h = static_cast<const uint8_t*>(lane.h_out.p);
        const uint64_t base = out_off[s.m0];
        const int32_t* st = reinterpret_cast<const int32_t*>(h + align_up(slab_out[k], 256));
        if (oiov_begin) {  // scatter every decoded message over its iovecs
            const uint64_t* foff = fin_off[k % kPipeLanes].data();
            const uint32_t* flen = fin_len[k % kPipeLanes].data();
            const uint8_t* fsmall = fin_small[k % kPipeLanes].data();
            pool(ctx).run(threads_for(slab_out[k]), [&](unsigned t, unsigned T) {
                const uint32_t lo = (uint32_t)((uint64_t)cnt * t / T), hi = (uint32_t)((uint64_t)cnt * (t + 1) / T);
                for (uint32_t i = lo; i < hi; i++) {
                    if (fsmall[i]) {
                        status[s.m0 + i] = B2C_TOO_SMALL;
                        continue;
                    }
                    status[s.m0 + i] = st[i] == B2C_TOO_SMALL ? B2C_CORRUPT : st[i];
                    if (st[i] != B2C_OK) continue;
                    const uint8_t* src = h + foff[i];
                    uint64_t left = flen[i];
                    for (uint32_t q = oiov_begin[s.m0 + i]; left && q < oiov_begin[s.m0 + i + 1]; q++) {
                        const uint64_t c = std::min<uint64_t>(left, oiov[q].iov_len);
                        memcpy(const_cast<void*>(oiov[q].iov_base), src, c);
                        src += c;
                        left -= c;
                    }
                }
            });
            tr.copy_out += secs(tc, Clock::now());
            return B2C_OK;
        }
        if (!direct) parallel_copy(ctx, static_cast<uint8_t*>(out) + base, h, slab_out[k]);
        for (uint32_t i = 0; i < cnt; i++)  // TOO_SMALL can only mean "implausible preamble" here
            status[s.m0 + i] = st[i] == B2C_TOO_SMALL ? B2C_CORRUPT : st[i];
        tr.copy_out += secs(tc, Clock::now());
        return B2C_OK;
    };
    // While the pool's workers gather slab k+1 (and read its preambles), this thread sizes slab k's output
    // slots from ITS preambles, lays it out, submits it and retires slab k-3.
    LayoutPlan lin_ring[2];
    std::vector<uint32_t>* cap_ring[2] = {&ctx->tmp_cap, &ctx->tmp_cap2};
    std::unique_ptr<SlabStats[]> stats(new SlabStats[2]);
    PoolGuard guard{ctx};
    auto prepare = [&](size_t k) -> int {
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        std::vector<uint32_t>& caps = *cap_ring[k & 1];
        if (caps.size() < cnt) caps.resize(cnt);
        // output slot sizes come from the (lenient) preambles
        if (ctx->input_pinned) {
            TIMED(tr.layout, lin_ring[k & 1].plan(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, ctx));
            TIMED(tr.gather, slab_caps(ctx, segs, seg_begin, bp, s, caps.data()));
        } else {  // read by the gathering threads while the bytes are in their cache
            int r;
            if ((r = pin_reserve(ctx, lane, lane.h_in,
                                 align_up(s.bytes, 256) + 2 * LayoutPlan::max_blob_bytes(cnt, s.payload) + (size_t)cnt * 16 + 4096)))
                return r;
            gather_slab_start(ctx, segs, seg_begin, bp, s, static_cast<uint8_t*>(lane.h_in.p), caps.data(), &stats[k & 1]);
        }
        return B2C_OK;
    };
    // as in the compress pipeline: slab k+1's gather may start before slab k-3 is retired when its staging need not grow
    auto fits = [&](size_t k) -> bool {
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        return align_up(s.bytes, 256) + 2 * LayoutPlan::max_blob_bytes(cnt, s.payload) + (size_t)cnt * 16 + 4096 <=
               ctx->pipe[k % kPipeLanes].h_in.cap;
    };
    if ((rc = prepare(0))) return rc;
    for (size_t k = 0; k < K; k++) {
        TIMED(tr.gather, pool(ctx).wait());
        const bool early = k + 1 < K && direct && !ctx->input_pinned && fits(k + 1);
        if (early) {
            // the lane's staging is free once the H2D copy of the slab that used it last has completed (ev_meta)
            if (k + 1 >= (size_t)kPipeLanes) B2C_CUDA(ctx, cudaEventSynchronize(ctx->pipe[(k + 1) % kPipeLanes].ev_meta));
            if ((rc = prepare(k + 1))) return rc;
        }
        if (k + 1 >= (size_t)kPipeLanes && (rc = finish(k + 1 - kPipeLanes))) return rc;  // frees slab k+1's lane
        if (k + 1 < K && !early && (rc = prepare(k + 1))) return rc;
        Lane& lane = ctx->pipe[k % kPipeLanes];
        const Slab& s = bp.slabs[k];
        const uint32_t cnt = s.m1 - s.m0;
        LayoutPlan& lin = lin_ring[k & 1];
        std::vector<uint32_t>& o_cap = *cap_ring[k & 1];
        if (o_off.size() < cnt) o_off.resize(cnt);
        auto tl = Clock::now();
        PlanStats st_in, st_out;
        if (!ctx->input_pinned) {
            stats[k & 1].merged(st_in, st_out);
            lin.plan_with(cnt, bp.off.data() + s.m0, bp.len.data() + s.m0, st_in);
        }
        bool clamped = false;
        if (oiov_begin) {  // a message whose iovecs cannot hold what its preamble claims is not decoded at all
            std::vector<uint8_t>& small = fin_small[k % kPipeLanes];
            small.assign(cnt, 0);
            for (uint32_t i = 0; i < cnt; i++) {
                uint64_t room = 0;
                for (uint32_t q = oiov_begin[s.m0 + i]; q < oiov_begin[s.m0 + i + 1]; q++) room += oiov[q].iov_len;
                if (o_cap[i] > room) {
                    o_cap[i] = 0;
                    small[i] = 1;
                    clamped = true;
                }
            }
        }
        uint64_t pos = 0;
        for (uint32_t i = 0; i < cnt; i++) {
            o_off[i] = pos;
            out_off[s.m0 + i] = out_base + pos;
            pos += o_cap[i];
        }
        slab_out[k] = pos;
        if (out_base + pos > out_cap) return set_err(ctx, B2C_E_CAPACITY, "decompress_batch: out_cap too small", cudaSuccess);
        out_base += pos;
        out_off[s.m1] = out_base;
        if (oiov_begin) {
            fin_off[k % kPipeLanes].assign(o_off.begin(), o_off.begin() + cnt);
            fin_len[k % kPipeLanes].assign(o_cap.begin(), o_cap.begin() + cnt);
        }
        LayoutPlan lout;
        if (ctx->input_pinned || clamped) {
            lout.plan(cnt, o_off.data(), o_cap.data(), ctx);
        } else {  // the workers summed everything but the end of the arena, which is the running total
            st_out.arena_end = pos;
            lout.plan_with(cnt, o_off.data(), o_cap.data(), st_out);
        }
        const size_t blob_out_rel = align_up(lin.blob_bytes, 256);  // metadata: input layout blob | output layout blob | crcs
        const size_t crc_rel = align_up(blob_out_rel + lout.blob_bytes, 256);
        const size_t meta_bytes = masked_crc ? crc_rel + (size_t)cnt * 4 : blob_out_rel + lout.blob_bytes;
        SlabIO io;
        if ((rc = slab_begin(ctx, lane, seg_begin, s, meta_bytes, io, /*keep=*/s.bytes))) return rc;
        lin.fill(io.stage_meta);
        lout.fill(io.stage_meta + blob_out_rel);
        if (masked_crc) memcpy(io.stage_meta + crc_rel, masked_crc + s.m0, (size_t)cnt * 4);
        tr.layout += secs(tl, Clock::now());
        const size_t st_at = align_up(pos, 256);
        if ((rc = ws_reserve(ctx, lane, lane.d_out, st_at + (size_t)cnt * 4 + 64))) return rc;
        if ((rc = pin_reserve(ctx, lane, lane.h_out, (direct ? 0 : st_at) + st_at + (size_t)cnt * 4 + 64))) return rc;
        auto ts = Clock::now();
        uint8_t* d_in = io.d_in;
        LayoutView vin, vout;
        lin.bind(d_in + io.meta_at, &vin);
        lout.bind(d_in + io.meta_at + blob_out_rel, &vout);
        if ((rc = slab_submit(ctx, lane, segs, seg_begin, s, io, meta_bytes, vin.d_off, h2d))) return rc;
        B2C_CUDA(ctx, cudaEventRecord(lane.ev_meta, lane.stream));  // the staging has been read
        uint8_t* d_out = static_cast<uint8_t*>(lane.d_out.p);
        int32_t* d_status = reinterpret_cast<int32_t*>(d_out + st_at);
        if (masked_crc) {
            const uint32_t* d_expect = reinterpret_cast<const uint32_t*>(d_in + io.meta_at + crc_rel);
            if ((rc = launch_decompress_verify(ctx, lane, vin, d_in, vout, d_out, d_expect, nullptr, d_status))) return rc;
        } else if ((rc = launch_decompress(ctx, lane, vin, d_in, vout, d_out, nullptr, d_status))) {
            return rc;
        }
        if (direct) {
            const uint64_t base = out_off[s.m0];
            if (pos) B2C_CUDA(ctx, cudaMemcpyAsync(static_cast<uint8_t*>(out) + base, d_out, pos, cudaMemcpyDeviceToHost, lane.stream));
            B2C_CUDA(ctx, cudaMemcpyAsync(static_cast<uint8_t*>(lane.h_out.p) + st_at, d_status, (size_t)cnt * 4,
                                          cudaMemcpyDeviceToHost, lane.stream));
            d2h += pos + (size_t)cnt * 4;
        } else {
            B2C_CUDA(ctx, cudaMemcpyAsync(lane.h_out.p, d_out, st_at + (size_t)cnt * 4, cudaMemcpyDeviceToHost, lane.stream));
            d2h += st_at + (size_t)cnt * 4;
        }
        B2C_CUDA(ctx, cudaEventRecord(lane.ev_done, lane.stream));
        tr.submit += secs(ts, Clock::now());
    }
    for (size_t k = K + 1 > (size_t)kPipeLanes ? K + 1 - kPipeLanes : 0; k < K; k++)
        if ((rc = finish(k))) return rc;
    ctx->last_h2d = h2d;
    ctx->last_d2h = d2h;
    tr.report(K, bp.payload);
    return B2C_OK;
}

int b2c_snappy_decompress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                                uint64_t out_cap, uint64_t* out_off, int32_t* status) {
    return decompress_batch_impl(ctx, segs, seg_begin, n, out, out_cap, out_off, status, nullptr);
}

int b2c_snappy_decompress_iov_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                                    const b2c_iovec* out_iov, const uint32_t* out_iov_begin, int32_t* status) {
    if (n && (!out_iov_begin || (!out_iov && out_iov_begin[n] != out_iov_begin[0]))) return B2C_E_ARG;
    if (n == 0) return ctx ? B2C_OK : B2C_E_ARG;
    return decompress_batch_impl(ctx, segs, seg_begin, n, nullptr, 0, nullptr, status, nullptr, out_iov, out_iov_begin);
}

int b2c_snappy_decompress_verify_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                                       const uint32_t* masked_crc, void* out, uint64_t out_cap, uint64_t* out_off,
                                       int32_t* status) {
    if (n && !masked_crc) return B2C_E_ARG;
    return decompress_batch_impl(ctx, segs, seg_begin, n, out, out_cap, out_off, status, masked_crc);
}

}  // extern "C"
