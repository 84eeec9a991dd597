// b2c_api.cu — host side of libb2c.so: contexts, layouts and the extern "C"
// entry points declared in include/b2c_codec.h.  Everything below the C ABI is
// CUDA for sm_100a; there is deliberately no CPU codec here: without a usable
// device b2c_ctx_create fails with B2C_E_NODEVICE.
#include <string.h>

#include <algorithm>
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

int ws_reserve(b2c_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return B2C_OK;
    if (b.p) {
        // the old buffer may still be in use by queued work
        B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
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

static int pin_reserve(b2c_ctx* ctx, void** p, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return B2C_OK;
    if (*p) {
        B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        B2C_CUDA(ctx, cudaFreeHost(*p));
        *p = nullptr;
        *cap = 0;
    }
    size_t want = bytes + bytes / 8 + 4096;
    cudaError_t e = cudaHostAlloc(p, want, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        *p = nullptr;
        return set_err(ctx, B2C_E_NOMEM, "cudaHostAlloc(staging)", e);
    }
    *cap = want;
    return B2C_OK;
}

void prof_mark(b2c_ctx* ctx, int which) {
    if (!ctx->prof_on) return;
    auto& v = ctx->prof_ev[which];
    uint32_t& used = ctx->prof_used[which];
    if (used >= 1024) return;
    if (used >= v.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        v.push_back(e);
    }
    cudaEventRecord(v[used++], ctx->stream);
}

template <typename T>
static int upload(b2c_ctx* ctx, T** d, const std::vector<T>& h) {
    *d = nullptr;
    size_t bytes = std::max<size_t>(h.size(), 1) * sizeof(T);
    B2C_CUDA(ctx, cudaMalloc(d, bytes));
    if (!h.empty()) B2C_CUDA(ctx, cudaMemcpyAsync(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return B2C_OK;
}

// Run fn(t, T) on T host threads (t = 0..T-1).
template <typename F>
static void parallel_for(unsigned T, F fn) {
    if (T <= 1) {
        fn(0u, 1u);
        return;
    }
    std::vector<std::thread> th;
    th.reserve(T - 1);
    for (unsigned t = 1; t < T; t++) th.emplace_back([=] { fn(t, T); });
    fn(0u, T);
    for (auto& x : th) x.join();
}

static unsigned host_threads(uint64_t bytes) {
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 4;
    unsigned want = (unsigned)std::min<uint64_t>(16, bytes / (4u << 20) + 1);
    return std::max(1u, std::min(hw, want));
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace b2c

using namespace b2c;

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
    int rc = B2C_OK;
    do {
        if ((e = cudaSetDevice(device)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaSetDevice", e); break; }
        cudaDeviceProp prop;
        if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaGetDeviceProperties", e); break; }
        ctx->num_sms = prop.multiProcessorCount;
        if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaStreamCreate", e); break; }
        if ((e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = set_err(ctx, B2C_E_CUDA, "cudaStreamCreate", e); break; }
        rc = crc_tables_init(ctx);
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
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (DevBuf* b : {&ctx->ws_frag, &ctx->ws_misc, &ctx->ws_io_in, &ctx->ws_io_out})
        if (b->p) cudaFree(b->p);
    if (ctx->h_pin_in) cudaFreeHost(ctx->h_pin_in);
    if (ctx->h_pin_out) cudaFreeHost(ctx->h_pin_out);
    if (ctx->d_crc_tabA) cudaFree(ctx->d_crc_tabA);
    if (ctx->d_crc_small) cudaFree(ctx->d_crc_small);
    for (auto& v : ctx->prof_ev)
        for (cudaEvent_t e : v) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
}

const char* b2c_last_error(const b2c_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context"; }
int b2c_ctx_device(const b2c_ctx* ctx) { return ctx ? ctx->device : -1; }
void* b2c_ctx_stream(b2c_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
uint64_t b2c_ctx_launch_count(const b2c_ctx* ctx) { return ctx ? ctx->launches : 0; }

int b2c_ctx_sync(b2c_ctx* ctx) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B2C_OK;
}

int b2c_dev_alloc(b2c_ctx* ctx, size_t bytes, void** d_ptr) {
    if (!ctx || !d_ptr) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaError_t e = cudaMalloc(d_ptr, bytes ? bytes : 1);
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
    B2C_CUDA(ctx, cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return B2C_OK;
}
int b2c_memcpy_d2h(b2c_ctx* ctx, void* h_dst, const void* d_src, size_t bytes) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return B2C_OK;
}
int b2c_memset_dev(b2c_ctx* ctx, void* d_dst, int value, size_t bytes) {
    if (!ctx) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaMemsetAsync(d_dst, value, bytes, ctx->stream));
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
    B2C_CUDA(ctx, cudaEventRecord(ev->ev, ctx->stream));
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
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->prof_on = on != 0;
    for (auto& u : ctx->prof_used) u = 0;
    return B2C_OK;
}

int b2c_prof_total_ms(b2c_ctx* ctx, int which, float* total_ms, uint32_t* launches) {
    if (!ctx || which < 0 || which >= B2C_PROF_COUNT || !total_ms) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
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
    L->n = n;
    L->h_off.assign(off, off + n);
    L->h_len.assign(len, len + n);
    std::vector<uint32_t> frag_msg, frag_first(n + 1), chunk_msg, chunk_first(n + 1), multi;
    std::vector<uint64_t> frag_slot;
    std::vector<uint32_t> cls[3], dcls[3];
    uint64_t slot = 0;
    bool single = true;
    for (uint32_t m = 0; m < n; m++) {
        const uint32_t l = len[m];
        L->payload += l;
        L->arena_end = std::max<uint64_t>(L->arena_end, off[m] + l);
        L->max_comp += b2c_snappy_max_compressed_length(l) + 5;
        frag_first[m] = (uint32_t)frag_msg.size();
        for (uint64_t pos = 0; pos < l; pos += kFragBytes) {
            const uint32_t fl = (uint32_t)std::min<uint64_t>(kFragBytes, l - pos);
            const uint32_t f = (uint32_t)frag_msg.size();
            frag_msg.push_back(m);
            frag_slot.push_back(slot);
            slot += align_up(b2c_snappy_max_compressed_length(fl), 16);
            cls[fl <= 1024 ? 0 : fl <= 4096 ? 1 : 2].push_back(f);
        }
        chunk_first[m] = (uint32_t)chunk_msg.size();
        const uint32_t nc = std::max<uint32_t>(1, (uint32_t)(((uint64_t)l + kCrcChunk - 1) / kCrcChunk));
        for (uint32_t k = 0; k < nc; k++) chunk_msg.push_back(m);
        if (nc > 1) {
            multi.push_back(m);
            single = false;
        }
        dcls[l <= 2048 ? 0 : l <= 8192 ? 1 : 2].push_back(m);
    }
    frag_first[n] = (uint32_t)frag_msg.size();
    chunk_first[n] = (uint32_t)chunk_msg.size();
    L->n_frags = (uint32_t)frag_msg.size();
    L->frag_scratch_bytes = slot;
    L->n_chunks = (uint32_t)chunk_msg.size();
    L->n_multi = (uint32_t)multi.size();
    L->all_single_chunk_uniform = single;
    int rc = B2C_OK;
    do {
        if ((rc = upload(ctx, &L->d_off, L->h_off))) break;
        if ((rc = upload(ctx, &L->d_len, L->h_len))) break;
        if ((rc = upload(ctx, &L->d_frag_msg, frag_msg))) break;
        if ((rc = upload(ctx, &L->d_frag_first, frag_first))) break;
        if ((rc = upload(ctx, &L->d_frag_slot, frag_slot))) break;
        if ((rc = upload(ctx, &L->d_chunk_msg, chunk_msg))) break;
        if ((rc = upload(ctx, &L->d_chunk_first, chunk_first))) break;
        if ((rc = upload(ctx, &L->d_multi_list, multi))) break;
        for (int c = 0; c < 3 && rc == B2C_OK; c++) {
            L->n_class[c] = (uint32_t)cls[c].size();
            L->n_dclass[c] = (uint32_t)dcls[c].size();
            if ((rc = upload(ctx, &L->d_class_list[c], cls[c]))) break;
            if ((rc = upload(ctx, &L->d_dclass_list[c], dcls[c]))) break;
        }
    } while (0);
    if (rc == B2C_OK) {
        // host vectors above are pageable: make sure the staged copies are done before they die
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = set_err(ctx, B2C_E_CUDA, "layout upload", e);
    }
    if (rc != B2C_OK) {
        b2c_layout_destroy(L);
        return rc;
    }
    *out = L;
    return B2C_OK;
}

void b2c_layout_destroy(b2c_layout* L) {
    if (!L) return;
    if (L->ctx) {
        cudaSetDevice(L->ctx->device);
        cudaStreamSynchronize(L->ctx->stream);
    }
    void* ptrs[] = {L->d_off, L->d_len, L->d_frag_msg, L->d_frag_first, L->d_frag_slot, L->d_chunk_msg,
                    L->d_chunk_first, L->d_multi_list, L->d_class_list[0], L->d_class_list[1], L->d_class_list[2],
                    L->d_dclass_list[0], L->d_dclass_list[1], L->d_dclass_list[2]};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    delete L;
}

uint32_t b2c_layout_count(const b2c_layout* L) { return L ? L->n : 0; }
uint64_t b2c_layout_payload_bytes(const b2c_layout* L) { return L ? L->payload : 0; }
uint64_t b2c_layout_arena_bytes(const b2c_layout* L) { return L ? L->arena_end : 0; }
uint64_t b2c_layout_max_compressed_bytes(const b2c_layout* L) { return L ? L->max_comp : 0; }

// ---------------------------------------------------------------------------
// device-resident entry points
// ---------------------------------------------------------------------------
int b2c_crc32c_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_data, const uint32_t* d_init, uint32_t* d_crc) {
    if (!ctx || !L || !d_crc || (L->payload && !d_data)) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_crc32c(ctx, *L, static_cast<const uint8_t*>(d_data), d_init, d_crc);
}

int b2c_snappy_compress_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, void* d_out, uint64_t out_cap,
                            uint64_t* d_out_off) {
    if (!ctx || !L || !d_out || !d_out_off || (L->payload && !d_in)) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_compress(ctx, *L, static_cast<const uint8_t*>(d_in), static_cast<uint8_t*>(d_out), out_cap,
                           d_out_off);
}

int b2c_snappy_uncompressed_length_dev(b2c_ctx* ctx, const b2c_layout* L, const void* d_in, uint32_t* d_ulen,
                                       int32_t* d_status) {
    if (!ctx || !L || !d_ulen) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_ulen(ctx, *L, static_cast<const uint8_t*>(d_in), d_ulen, d_status);
}

int b2c_snappy_decompress_dev(b2c_ctx* ctx, const b2c_layout* in, const void* d_in, const b2c_layout* out,
                              void* d_out, uint32_t* d_ulen, int32_t* d_status) {
    if (!ctx || !in || !out || !d_status) return B2C_E_ARG;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_decompress(ctx, *in, static_cast<const uint8_t*>(d_in), *out, static_cast<uint8_t*>(d_out), d_ulen,
                             d_status);
}

// ---------------------------------------------------------------------------
// host (end-to-end) entry points
// ---------------------------------------------------------------------------
namespace {

struct Gathered {
    std::vector<uint64_t> off;  // staging offsets, 16-byte aligned
    std::vector<uint32_t> len;
    uint64_t total = 0;         // staging bytes
    uint64_t payload = 0;
};

// Sum segment lengths per message and lay the messages out in staging.
int plan_gather(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, Gathered& g) {
    g.off.resize(n);
    g.len.resize(n);
    uint64_t pos = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint64_t l = 0;
        for (uint32_t s = seg_begin[i]; s < seg_begin[i + 1]; s++) l += segs[s].iov_len;
        if (l > 0xffffffffull) return set_err(ctx, B2C_E_ARG, "message longer than 4 GiB - 1", cudaSuccess);
        g.off[i] = pos;
        g.len[i] = (uint32_t)l;
        g.payload += l;
        pos += align_up(l, 16);
    }
    g.total = pos;
    return B2C_OK;
}

// memcpy every segment into pinned staging, split over host threads by message range.
void run_gather(const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, const Gathered& g, uint8_t* stage) {
    parallel_for(host_threads(g.payload), [&](unsigned t, unsigned T) {
        uint32_t lo = (uint32_t)((uint64_t)n * t / T), hi = (uint32_t)((uint64_t)n * (t + 1) / T);
        for (uint32_t i = lo; i < hi; i++) {
            uint8_t* d = stage + g.off[i];
            for (uint32_t s = seg_begin[i]; s < seg_begin[i + 1]; s++) {
                memcpy(d, segs[s].iov_base, segs[s].iov_len);
                d += segs[s].iov_len;
            }
        }
    });
}

void parallel_copy(void* dst, const void* src, uint64_t bytes) {
    parallel_for(host_threads(bytes), [&](unsigned t, unsigned T) {
        uint64_t lo = bytes * t / T, hi = bytes * (t + 1) / T;
        memcpy(static_cast<uint8_t*>(dst) + lo, static_cast<const uint8_t*>(src) + lo, hi - lo);
    });
}

struct LayoutGuard {
    b2c_layout* L = nullptr;
    ~LayoutGuard() { b2c_layout_destroy(L); }
};

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
    Gathered g;
    int rc = plan_gather(ctx, segs, seg_begin, n, g);
    if (rc) return rc;
    const size_t meta = (size_t)n * 4;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_in, &ctx->h_pin_in_cap, g.total + meta + 64))) return rc;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_out, &ctx->h_pin_out_cap, meta + 64))) return rc;
    if ((rc = ws_reserve(ctx, ctx->ws_io_in, g.total + meta + 64))) return rc;
    if ((rc = ws_reserve(ctx, ctx->ws_io_out, meta + 64))) return rc;
    uint8_t* stage = static_cast<uint8_t*>(ctx->h_pin_in);
    run_gather(segs, seg_begin, n, g, stage);
    uint8_t* d_in = static_cast<uint8_t*>(ctx->ws_io_in.p);
    uint32_t* d_init = nullptr;
    uint64_t h2d = g.total;
    B2C_CUDA(ctx, cudaMemcpyAsync(d_in, stage, g.total, cudaMemcpyHostToDevice, ctx->stream));
    if (init) {
        const size_t init_at = align_up(g.total, 16);
        memcpy(stage + init_at, init, meta);
        d_init = reinterpret_cast<uint32_t*>(d_in + init_at);
        B2C_CUDA(ctx, cudaMemcpyAsync(d_init, stage + init_at, meta, cudaMemcpyHostToDevice, ctx->stream));
        h2d += meta;
    }
    LayoutGuard lg;
    if ((rc = b2c_layout_create(ctx, n, g.off.data(), g.len.data(), &lg.L))) return rc;
    uint32_t* d_crc = static_cast<uint32_t*>(ctx->ws_io_out.p);
    if ((rc = launch_crc32c(ctx, *lg.L, d_in, d_init, d_crc))) return rc;
    B2C_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin_out, d_crc, meta, cudaMemcpyDeviceToHost, ctx->stream));
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(crc, ctx->h_pin_out, meta);
    ctx->last_h2d = h2d + (uint64_t)n * 12;
    ctx->last_d2h = meta;
    return B2C_OK;
}

int b2c_snappy_compress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                              uint64_t out_cap, uint64_t* out_off) {
    if (!ctx || !out_off || (n && (!segs || !seg_begin || !out))) return B2C_E_ARG;
    out_off[0] = 0;
    if (n == 0) return B2C_OK;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    Gathered g;
    int rc = plan_gather(ctx, segs, seg_begin, n, g);
    if (rc) return rc;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_in, &ctx->h_pin_in_cap, g.total + 64))) return rc;
    if ((rc = ws_reserve(ctx, ctx->ws_io_in, g.total + 64))) return rc;
    uint8_t* stage = static_cast<uint8_t*>(ctx->h_pin_in);
    run_gather(segs, seg_begin, n, g, stage);
    uint8_t* d_in = static_cast<uint8_t*>(ctx->ws_io_in.p);
    B2C_CUDA(ctx, cudaMemcpyAsync(d_in, stage, g.total, cudaMemcpyHostToDevice, ctx->stream));
    LayoutGuard lg;
    if ((rc = b2c_layout_create(ctx, n, g.off.data(), g.len.data(), &lg.L))) return rc;
    const uint64_t cap = lg.L->max_comp;
    const size_t off_bytes = ((size_t)n + 1) * 8;
    if ((rc = ws_reserve(ctx, ctx->ws_io_out, cap + off_bytes + 64))) return rc;
    uint8_t* d_out = static_cast<uint8_t*>(ctx->ws_io_out.p);
    uint64_t* d_out_off = reinterpret_cast<uint64_t*>(d_out + align_up(cap, 16));
    if ((rc = launch_compress(ctx, *lg.L, d_in, d_out, cap, d_out_off))) return rc;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_out, &ctx->h_pin_out_cap, off_bytes + 64))) return rc;
    B2C_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin_out, d_out_off, off_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(out_off, ctx->h_pin_out, off_bytes);
    const uint64_t total = out_off[n];
    if (total > out_cap) return set_err(ctx, B2C_E_CAPACITY, "compress_batch: out_cap too small", cudaSuccess);
    if ((rc = pin_reserve(ctx, &ctx->h_pin_out, &ctx->h_pin_out_cap, total + 64))) return rc;
    B2C_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin_out, d_out, total, cudaMemcpyDeviceToHost, ctx->stream));
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    parallel_copy(out, ctx->h_pin_out, total);
    ctx->last_h2d = g.total + (uint64_t)n * 12;
    ctx->last_d2h = total + off_bytes;
    return B2C_OK;
}

int b2c_snappy_decompress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n, void* out,
                                uint64_t out_cap, uint64_t* out_off, int32_t* status) {
    if (!ctx || !out_off || (n && (!segs || !seg_begin || !status))) return B2C_E_ARG;
    out_off[0] = 0;
    if (n == 0) return B2C_OK;
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    Gathered g;
    int rc = plan_gather(ctx, segs, seg_begin, n, g);
    if (rc) return rc;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_in, &ctx->h_pin_in_cap, g.total + 64))) return rc;
    if ((rc = ws_reserve(ctx, ctx->ws_io_in, g.total + 64))) return rc;
    uint8_t* stage = static_cast<uint8_t*>(ctx->h_pin_in);
    run_gather(segs, seg_begin, n, g, stage);
    uint8_t* d_in = static_cast<uint8_t*>(ctx->ws_io_in.p);
    B2C_CUDA(ctx, cudaMemcpyAsync(d_in, stage, g.total, cudaMemcpyHostToDevice, ctx->stream));
    // Output slots from the (lenient) preambles.  A stream cannot expand by more than
    // 64/3 (a 3-byte copy yields at most 64 bytes), so a larger claim is malformed: the
    // reference would run out of input before producing it (snappy.cc:867).
    std::vector<uint64_t> o_off(n);
    std::vector<uint32_t> o_cap(n);
    uint64_t pos = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t ulen = 0;
        int k = b2c_snappy_read_preamble(stage + g.off[i], g.len[i], &ulen);
        bool plausible = k > 0 && (uint64_t)ulen <= ((uint64_t)g.len[i] * 64) / 3 + 64;
        o_off[i] = pos;
        o_cap[i] = plausible ? ulen : 0;
        out_off[i] = pos;
        pos += o_cap[i];
    }
    out_off[n] = pos;
    if (pos > out_cap) return set_err(ctx, B2C_E_CAPACITY, "decompress_batch: out_cap too small", cudaSuccess);
    LayoutGuard lin, lout;
    if ((rc = b2c_layout_create(ctx, n, g.off.data(), g.len.data(), &lin.L))) return rc;
    if ((rc = b2c_layout_create(ctx, n, o_off.data(), o_cap.data(), &lout.L))) return rc;
    const size_t st_bytes = (size_t)n * 4;
    if ((rc = ws_reserve(ctx, ctx->ws_io_out, pos + st_bytes + 64))) return rc;
    uint8_t* d_out = static_cast<uint8_t*>(ctx->ws_io_out.p);
    int32_t* d_status = reinterpret_cast<int32_t*>(d_out + align_up(pos, 16));
    if ((rc = launch_decompress(ctx, *lin.L, d_in, *lout.L, d_out, nullptr, d_status))) return rc;
    if ((rc = pin_reserve(ctx, &ctx->h_pin_out, &ctx->h_pin_out_cap, align_up(pos, 16) + st_bytes + 64))) return rc;
    uint8_t* h_out = static_cast<uint8_t*>(ctx->h_pin_out);
    if (pos) B2C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, pos, cudaMemcpyDeviceToHost, ctx->stream));
    B2C_CUDA(ctx, cudaMemcpyAsync(h_out + align_up(pos, 16), d_status, st_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (pos) parallel_copy(out, h_out, pos);
    memcpy(status, h_out + align_up(pos, 16), st_bytes);
    for (uint32_t i = 0; i < n; i++)  // TOO_SMALL here only means "implausible preamble"
        if (status[i] == B2C_TOO_SMALL) status[i] = B2C_CORRUPT;
    ctx->last_h2d = g.total + (uint64_t)n * 24;
    ctx->last_d2h = pos + st_bytes;
    return B2C_OK;
}

}  // extern "C"
