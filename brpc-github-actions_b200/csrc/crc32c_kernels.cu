// crc32c_kernels.cu — batch CRC-32C (Castagnoli) for sm_100a.
//
// Replaces butil::crc32c::Extend / Value (reference src/butil/crc32c.cc:379-454,
// crc32c.h:19-29) for many messages at once.  The reference walks one buffer
// with a serial 8-bytes-per-instruction chain (crc32c.cc:407-410); a GPU has no
// CRC instruction, so the work is re-shaped around GF(2) linearity:
//
//   * a message is cut into 4 KiB chunks; 4 lanes (a "quad") own one chunk;
//   * each step the quad reads one 64-byte row (16 B per lane, full 32-byte
//     sectors) and every lane keeps FOUR independent Horner accumulators, one per
//     32-bit column of its uint4:  H <- Z64(H) ^ word, where Z64 advances a CRC
//     register across 64 bytes.  Z64 is applied with 4 table lookups (one per
//     state byte) — 1 lookup per payload byte, the same cost as slicing, but
//     with no serial dependency between the 16 lookups of a row;
//   * the Z64 tables are replicated once per lane ([byte][lane]) in shared
//     memory (128 KiB), so a warp-wide lookup is bank-conflict free;
//   * columns, lanes, the ragged tail and (for multi-chunk messages) chunks are
//     then combined with the shift operators Z4/Z16/Z32 (small tables) and a
//     GF(2) multiply by x^(8*len) (crc32c_combine_kernel).
//
// Register convention is the reference's: l = crc ^ 0xffffffff, one byte step is
// l = table0[(l ^ b) & 0xff] ^ (l >> 8) (crc32c.cc:390-393), result l ^ 0xffffffff.
#include <mutex>

#include "b2c_internal.cuh"

namespace b2c {

static constexpr uint32_t kPolyReflected = 0x82F63B78u;

// ---------------------------------------------------------------------------
// host-side table construction
// ---------------------------------------------------------------------------
namespace {

struct HostTables {
    uint32_t t0[256];
    HostTables() {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c >> 1) ^ ((c & 1) ? kPolyReflected : 0);
            t0[i] = c;
        }
    }
    uint32_t zbytes(uint32_t l, uint32_t k) const {  // advance register across k zero bytes
        for (uint32_t i = 0; i < k; i++) l = t0[l & 0xff] ^ (l >> 8);
        return l;
    }
};

uint32_t host_mulmod(uint32_t a, uint32_t b) {  // reflected domain: bit 31 is x^0
    uint32_t p = 0;
    for (int i = 0; i < 32; i++) {
        if ((a >> (31 - i)) & 1) p ^= b;
        b = (b >> 1) ^ ((b & 1) ? kPolyReflected : 0);
    }
    return p;
}

}  // namespace

__constant__ uint32_t c_x2n[64];  // x^(2^j) mod P, reflected

int crc_tables_init(b2c_ctx* ctx) {
    HostTables ht;
    std::vector<uint32_t> tabA(4 * 256 * 32);
    std::vector<uint32_t> small(3 * 4 * 256);
    const uint32_t dist[3] = {4, 16, 32};
    for (uint32_t k = 0; k < 4; k++)
        for (uint32_t b = 0; b < 256; b++) {
            uint32_t v = ht.zbytes(b << (8 * k), 64);
            for (uint32_t l = 0; l < 32; l++) tabA[(k * 256 + b) * 32 + l] = v;
            for (uint32_t d = 0; d < 3; d++) small[(d * 4 + k) * 256 + b] = ht.zbytes(b << (8 * k), dist[d]);
        }
    uint32_t x2n[64];
    x2n[0] = 1u << 30;  // x^1
    for (int j = 1; j < 64; j++) x2n[j] = host_mulmod(x2n[j - 1], x2n[j - 1]);
    B2C_CUDA(ctx, cudaMalloc(&ctx->d_crc_tabA, tabA.size() * 4));
    B2C_CUDA(ctx, cudaMalloc(&ctx->d_crc_small, small.size() * 4));
    B2C_CUDA(ctx, cudaMemcpy(ctx->d_crc_tabA, tabA.data(), tabA.size() * 4, cudaMemcpyHostToDevice));
    B2C_CUDA(ctx, cudaMemcpy(ctx->d_crc_small, small.data(), small.size() * 4, cudaMemcpyHostToDevice));
    B2C_CUDA(ctx, cudaMemcpyToSymbol(c_x2n, x2n, sizeof(x2n)));
    return B2C_OK;
}

// ---------------------------------------------------------------------------
// device code
// ---------------------------------------------------------------------------
// Shared-memory image.  The replicated Z^64 tables sit in two 64 KiB regions whose SHARED
// address is 64 KiB-aligned: entry (table k, byte b, lane l) lives at
//     region(k >> 1) + (b << 8) + (k & 1) * 128 + l * 4
// so one PRMT builds the whole address — it drops state byte k into address bits 8..15 next to a
// per-lane constant (region base | lane * 4) — and the table's +128 is the LDS immediate:
// 2 instructions per lookup instead of the 4-5 of shift/mask/scale/add indexing.
struct CrcSmem {
    uint32_t S[3][4][256];   // Z^4, Z^16, Z^32: 12 KiB (plain layout, rarely used)
};
constexpr uint32_t kCrcRegion = 65536;
// dynamic smem: S tables | up to 64 KiB of alignment slack | 2 regions
constexpr size_t kCrcSmemBytes = sizeof(CrcSmem) + kCrcRegion + 2 * kCrcRegion;

__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u32_p128(uint32_t saddr) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+128];" : "=r"(v) : "r"(saddr));
    return v;
}

struct CrcLane {
    uint32_t reg0;   // shared address of region 0 | lane * 4   (bits 8..15 are zero)
    uint32_t reg1;   // same for region 1
    uint32_t stage;  // shared address of this thread's first staging slot (0: no room, load directly)
};
constexpr uint32_t kCrcStages = 3;
constexpr uint32_t kCrcStageBytes = 1024 * 16;  // one 16-byte slot per thread of the CTA

__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
    return r;
}

// Z^64 applied to a 32-bit state: 4 conflict-free lookups
__device__ __forceinline__ uint32_t apply_A(const CrcLane& cl, uint32_t x) {
    const uint32_t t0 = lds_u32(__byte_perm(x, cl.reg0, 0x7604));
    const uint32_t t1 = lds_u32_p128(__byte_perm(x, cl.reg0, 0x7614));
    const uint32_t t2 = lds_u32(__byte_perm(x, cl.reg1, 0x7624));
    const uint32_t t3 = lds_u32_p128(__byte_perm(x, cl.reg1, 0x7634));
    return t0 ^ t1 ^ t2 ^ t3;
}
template <int D>
__device__ __forceinline__ uint32_t apply_S(const CrcSmem& sm, uint32_t x) {
    return sm.S[D][0][x & 0xff] ^ sm.S[D][1][(x >> 8) & 0xff] ^ sm.S[D][2][(x >> 16) & 0xff] ^ sm.S[D][3][x >> 24];
}
__device__ __forceinline__ uint32_t byte_step(const CrcSmem& sm, uint32_t l, uint32_t b) {
    return sm.S[0][3][(l ^ b) & 0xff] ^ (l >> 8);  // Z^4 of a top byte == table0
}

__device__ __forceinline__ uint4 ldg128_stream(const uint8_t* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// S(l, p[0..L)) computed by the 4 lanes of one quad; result valid on all 4 lanes.
__device__ uint32_t crc_chunk_quad(const CrcSmem& sm, const CrcLane& cl, const uint8_t* p, uint32_t L, uint32_t l,
                                   uint32_t lane) {
    const uint32_t q = lane & 3;
    const uint32_t qbase = lane & ~3u;
    const uint32_t qmask = 0xFu << qbase;
    // head: bytes up to 16-byte alignment
    uint32_t hb = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(p) & 15)) & 15u;
    if (hb > L) hb = L;
    for (uint32_t i = 0; i < hb; i++) l = byte_step(sm, l, p[i]);
    p += hb;
    L -= hb;
    const uint32_t rows = L >> 6;
    if (rows) {
        const uint8_t* lp = p + q * 16;
        uint32_t h0, h1, h2, h3;
        if (cl.stage) {
            // The rows come through a private 3-deep cp.async ring in shared memory (16 bytes per
            // stage per thread, L2 -> smem, no L1 allocation and no registers held while in flight):
            // with 192 KiB of tables only ~20 KiB of L1 is left, too little to keep enough plain
            // loads outstanding for HBM speed.
            const uint32_t s0 = cl.stage;
#pragma unroll
            for (uint32_t s = 0; s < kCrcStages; s++) {
                if (s < rows) cp_async16(s0 + s * kCrcStageBytes, lp + (size_t)s * 64);
                cp_async_commit();
            }
            cp_async_wait<kCrcStages - 1>();
            uint4 w = lds128(s0);
            h0 = w.x; h1 = w.y; h2 = w.z; h3 = w.w;
            if (q == 0) h0 ^= l;
            if (kCrcStages < rows) cp_async16(s0, lp + (size_t)kCrcStages * 64);
            cp_async_commit();
            uint32_t slot = 1;
            for (uint32_t it = 1; it < rows; it++) {
                cp_async_wait<kCrcStages - 1>();
                const uint32_t sa = s0 + slot * kCrcStageBytes;
                w = lds128(sa);
                h0 = apply_A(cl, h0) ^ w.x;
                h1 = apply_A(cl, h1) ^ w.y;
                h2 = apply_A(cl, h2) ^ w.z;
                h3 = apply_A(cl, h3) ^ w.w;
                if (it + kCrcStages < rows) cp_async16(sa, lp + (size_t)(it + kCrcStages) * 64);
                cp_async_commit();
                slot = slot + 1 == kCrcStages ? 0 : slot + 1;
            }
            cp_async_wait<0>();
        } else {
            uint4 w = ldg128_stream(lp);
            h0 = w.x; h1 = w.y; h2 = w.z; h3 = w.w;
            if (q == 0) h0 ^= l;
#pragma unroll 4
            for (uint32_t it = 1; it < rows; it++) {
                w = ldg128_stream(lp + (size_t)it * 64);
                h0 = apply_A(cl, h0) ^ w.x;
                h1 = apply_A(cl, h1) ^ w.y;
                h2 = apply_A(cl, h2) ^ w.z;
                h3 = apply_A(cl, h3) ^ w.w;
            }
        }
        // fold the 4 columns of this lane: g = S(0, 16 raw bytes)
        uint32_t g = apply_S<0>(sm, h0);
        g = apply_S<0>(sm, g ^ h1);
        g = apply_S<0>(sm, g ^ h2);
        g = apply_S<0>(sm, g ^ h3);
        uint32_t g0 = __shfl_sync(qmask, g, qbase + 0);
        uint32_t g1 = __shfl_sync(qmask, g, qbase + 1);
        uint32_t g2 = __shfl_sync(qmask, g, qbase + 2);
        uint32_t g3 = __shfl_sync(qmask, g, qbase + 3);
        l = apply_S<2>(sm, apply_S<1>(sm, g0) ^ g1) ^ (apply_S<1>(sm, g2) ^ g3);
        p += (size_t)rows * 64;
        L -= rows * 64;
    }
    if (L == 0) return l;
    // ragged tail: L < 64 bytes, p is 16-byte aligned
    const uint32_t tw = L >> 2, tb = L & 3;
    const uint32_t a = tw >> 2, r = tw & 3;  // full lanes, words in the partial lane
    uint4 w = make_uint4(0, 0, 0, 0);
    if (q * 16 < L) w = ldg128_stream(p + q * 16);
    uint32_t g = apply_S<0>(sm, w.x);
    g = apply_S<0>(sm, g ^ w.y);
    g = apply_S<0>(sm, g ^ w.z);
    g = apply_S<0>(sm, g ^ w.w);
    uint32_t g0 = __shfl_sync(qmask, g, qbase + 0);
    uint32_t g1 = __shfl_sync(qmask, g, qbase + 1);
    uint32_t g2 = __shfl_sync(qmask, g, qbase + 2);
    if (a > 0) l = apply_S<1>(sm, l) ^ g0;
    if (a > 1) l = apply_S<1>(sm, l) ^ g1;
    if (a > 2) l = apply_S<1>(sm, l) ^ g2;
    // partial lane `a` finishes: r whole words then tb bytes
    uint32_t wx = __shfl_sync(qmask, w.x, qbase + a);
    uint32_t wy = __shfl_sync(qmask, w.y, qbase + a);
    uint32_t wz = __shfl_sync(qmask, w.z, qbase + a);
    uint32_t ww = __shfl_sync(qmask, w.w, qbase + a);
    uint32_t last = wx;
    if (r > 0) { l = apply_S<0>(sm, l ^ wx); last = wy; }
    if (r > 1) { l = apply_S<0>(sm, l ^ wy); last = wz; }
    if (r > 2) { l = apply_S<0>(sm, l ^ wz); last = ww; }
    for (uint32_t i = 0; i < tb; i++) l = byte_step(sm, l, (last >> (8 * i)) & 0xff);
    return l;
}

struct CrcParams {
    const uint8_t* data;
    const uint64_t* off;
    const uint32_t* len;
    const uint32_t* init;        // nullable
    const uint32_t* chunk_msg;   // nullable when uniform single chunk
    const uint32_t* chunk_first; // nullable when uniform single chunk
    const uint32_t* tabA;
    const uint32_t* small;
    uint32_t* rem;               // [n_chunks] raw remainders of multi-chunk messages
    uint32_t* crc;               // [n]
    uint32_t n_chunks;
};

__global__ void __launch_bounds__(1024, 1) crc32c_chunks_kernel(CrcParams P) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    CrcSmem& sm = *reinterpret_cast<CrcSmem*>(smem_raw);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t region0 = (sbase + (uint32_t)sizeof(CrcSmem) + kCrcRegion - 1) & ~(kCrcRegion - 1);
    {
        // P.tabA is [k][b][lane]; scatter 16 bytes (4 lanes) at a time into the interleaved regions
        const uint4* srcA = reinterpret_cast<const uint4*>(P.tabA);
        uint8_t* gen0 = smem_raw + (region0 - sbase);
        for (uint32_t i = threadIdx.x; i < 4 * 256 * 8; i += blockDim.x) {
            const uint32_t k = i >> 11, b = (i >> 3) & 255, l4 = i & 7;
            *reinterpret_cast<uint4*>(gen0 + (k >> 1) * kCrcRegion + (b << 8) + (k & 1) * 128 + l4 * 16) = srcA[i];
        }
        const uint4* srcS = reinterpret_cast<const uint4*>(P.small);
        uint4* dstS = reinterpret_cast<uint4*>(&sm.S[0][0][0]);
        for (uint32_t i = threadIdx.x; i < sizeof(sm.S) / 16; i += blockDim.x) dstS[i] = srcS[i];
    }
    __syncthreads();
    CrcLane cl;
    cl.reg0 = region0 | ((threadIdx.x & 31) << 2);
    cl.reg1 = cl.reg0 + kCrcRegion;
    {   // staging ring in the alignment slack between the small tables and region 0
        const uint32_t slack0 = (sbase + (uint32_t)sizeof(CrcSmem) + 15u) & ~15u;
        cl.stage = region0 - slack0 >= kCrcStages * kCrcStageBytes ? slack0 + threadIdx.x * 16 : 0u;
    }
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t quads_per_cta = blockDim.x >> 2;
    const uint32_t total_quads = gridDim.x * quads_per_cta;
    for (uint32_t c = blockIdx.x * quads_per_cta + (threadIdx.x >> 2); c < P.n_chunks; c += total_quads) {
        uint32_t m = c, k = 0;
        if (P.chunk_msg) {
            m = P.chunk_msg[c];
            k = c - P.chunk_first[m];
        }
        const uint32_t len = P.len[m];
        const uint32_t begin = k * kCrcChunk;
        const uint32_t L = min(kCrcChunk, len - begin);
        const bool single = len <= kCrcChunk;
        uint32_t l = 0;
        if (k == 0) l = (P.init ? P.init[m] : 0u) ^ 0xffffffffu;
        l = crc_chunk_quad(sm, cl, P.data + P.off[m] + begin, L, l, lane);
        if ((lane & 3) == 0) {
            if (single) P.crc[m] = l ^ 0xffffffffu;
            else P.rem[c] = l;
        }
    }
}

// GF(2)[x] multiply mod P in the reflected domain (bit 31 is x^0).
__device__ __forceinline__ uint32_t mulmod(uint32_t a, uint32_t b) {
    uint32_t p = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) {
        p ^= b & (0u - ((a >> (31 - i)) & 1u));
        b = (b >> 1) ^ (kPolyReflected & (0u - (b & 1u)));
    }
    return p;
}

// x^(8*bytes) mod P
__device__ uint32_t xpow_bytes(uint64_t bytes) {
    uint64_t e = bytes << 3;
    uint32_t r = 0x80000000u;  // x^0
    for (int j = 0; e; j++, e >>= 1)
        if (e & 1) r = mulmod(r, c_x2n[j]);
    return r;
}

// One warp per multi-chunk message: message remainder = sum over chunks of
// rem_k * x^(8 * bytes after chunk k).
__global__ void crc32c_combine_kernel(const uint32_t* __restrict__ multi_list, uint32_t n_multi,
                                      const uint32_t* __restrict__ chunk_first, const uint32_t* __restrict__ len,
                                      const uint32_t* __restrict__ rem, uint32_t* __restrict__ crc) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (warp >= n_multi) return;
    const uint32_t m = multi_list[warp];
    const uint32_t first = chunk_first[m];
    const uint32_t K = chunk_first[m + 1] - first;
    const uint32_t n = len[m];
    const uint32_t per = (K + 31) / 32;
    const uint32_t a = min(K, lane * per), b = min(K, a + per);
    uint32_t acc = 0;
    if (a < b) {
        const uint32_t xch = xpow_bytes(kCrcChunk);
        acc = rem[first + a];
        for (uint32_t k = a + 1; k < b; k++) {
            uint32_t lk = min(kCrcChunk, n - k * kCrcChunk);
            acc = mulmod(acc, lk == kCrcChunk ? xch : xpow_bytes(lk)) ^ rem[first + k];
        }
        uint64_t covered = min((uint64_t)n, (uint64_t)b * kCrcChunk);
        uint64_t after = (uint64_t)n - covered;
        if (after) acc = mulmod(acc, xpow_bytes(after));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) crc[m] = acc ^ 0xffffffffu;
}

int launch_crc32c(b2c_ctx* ctx, Lane& lane, const LayoutView& L, const uint8_t* d_data, const uint32_t* d_init,
                  uint32_t* d_crc) {
    if (L.n == 0) return B2C_OK;
    if (!ctx->crc_attr_done) {  // per device: a ctx is bound to one
        B2C_CUDA(ctx, cudaFuncSetAttribute(crc32c_chunks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)kCrcSmemBytes));
        ctx->crc_attr_done = true;
    }
    uint32_t* d_rem = nullptr;
    if (L.n_multi) {
        int rc = ws_reserve(ctx, lane, lane.ws_misc, (size_t)L.n_chunks * 4);
        if (rc) return rc;
        d_rem = static_cast<uint32_t*>(lane.ws_misc.p);
    }
    CrcParams P;
    P.data = d_data;
    P.off = L.d_off;
    P.len = L.d_len;
    P.init = d_init;
    P.chunk_msg = L.d_chunk_msg;
    P.chunk_first = L.d_chunk_first;
    P.tabA = ctx->d_crc_tabA;
    P.small = ctx->d_crc_small;
    P.rem = d_rem;
    P.crc = d_crc;
    P.n_chunks = L.n_chunks;
    const uint32_t quads_per_cta = 256;
    uint32_t grid = (L.n_chunks + quads_per_cta - 1) / quads_per_cta;
    if (grid > (uint32_t)ctx->num_sms) grid = ctx->num_sms;
    prof_mark(ctx, lane, B2C_PROF_CRC32C);
    crc32c_chunks_kernel<<<grid, 1024, kCrcSmemBytes, lane.stream>>>(P);
    prof_mark(ctx, lane, B2C_PROF_CRC32C);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    if (L.n_multi) {
        uint32_t blocks = (L.n_multi * 32 + 255) / 256;
        crc32c_combine_kernel<<<blocks, 256, 0, lane.stream>>>(L.d_multi_list, L.n_multi, L.d_chunk_first, L.d_len,
                                                               d_rem, d_crc);
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    return B2C_OK;
}

// ---------------------------------------------------------------------------
// Allocation-free variant (b2c_crc32c_noalloc_dev).
//
// The chunk kernel needs one 32-bit slot per 4 KiB chunk to park the partial CRCs
// of multi-chunk messages until the combine kernel folds them.  launch_crc32c
// takes those slots from a workspace that grows with cudaMalloc; callers that
// cannot allocate on their call path (see include/b2c_codec.h) get the same slots
// from this fixed array instead.  Inside host code `g_noalloc_rem` is only a
// handle for the runtime, not an address a kernel could use, so the real device
// address is looked up once per device and cached.
// ---------------------------------------------------------------------------
static __device__ __align__(256) uint32_t g_noalloc_rem[B2C_CRC32C_NOALLOC_MAX_CHUNKS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_addr[kMaxDevices] = {};  // device address of g_noalloc_rem, per device
}  // namespace

// First use on a device: resolve the array's address with that device current.
static int noalloc_scratch(b2c_ctx* ctx, uint32_t** out) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= kMaxDevices) return set_err(ctx, B2C_E_ARG, "noalloc: device index out of range", cudaSuccess);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_addr[dev]) {
        B2C_CUDA(ctx, cudaSetDevice(dev));
        void* p = nullptr;
        B2C_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_rem));
        if (!p) return set_err(ctx, B2C_E_CUDA, "noalloc: cudaGetSymbolAddress returned null", cudaSuccess);
        g_noalloc_addr[dev] = static_cast<uint32_t*>(p);
    }
    *out = g_noalloc_addr[dev];
    return B2C_OK;
}

}  // namespace b2c

extern "C" int b2c_crc32c_noalloc_dev(b2c_ctx* ctx, const b2c_layout* layout, const void* d_data,
                                      const uint32_t* d_init, uint32_t* d_crc) {
    using namespace b2c;
    if (!ctx || !layout || !d_crc || (layout->v.payload && !d_data)) return B2C_E_ARG;
    const LayoutView& L = layout->v;
    if (L.n_chunks > B2C_CRC32C_NOALLOC_MAX_CHUNKS)
        return set_err(ctx, B2C_E_CAPACITY,
                       "b2c_crc32c_noalloc_dev: layout has more than B2C_CRC32C_NOALLOC_MAX_CHUNKS (2^20) 4 KiB chunks; "
                       "split the batch or use b2c_crc32c_dev",
                       cudaSuccess);
    B2C_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t* d_rem = nullptr;
    int rc = noalloc_scratch(ctx, &d_rem);
    if (rc) return rc;
    if (L.n == 0) return B2C_OK;
    if (!ctx->crc_attr_done) {
        B2C_CUDA(ctx, cudaFuncSetAttribute(crc32c_chunks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)kCrcSmemBytes));
        ctx->crc_attr_done = true;
    }
    CrcParams P;
    P.data = static_cast<const uint8_t*>(d_data);
    P.off = L.d_off;
    P.len = L.d_len;
    P.init = d_init;
    P.chunk_msg = L.d_chunk_msg;
    P.chunk_first = L.d_chunk_first;
    P.tabA = ctx->d_crc_tabA;
    P.small = ctx->d_crc_small;
    P.rem = d_rem;  // slot c is written by chunk c before the combine kernel reads it
    P.crc = d_crc;
    P.n_chunks = L.n_chunks;
    uint32_t grid = (L.n_chunks + 255) / 256;
    if (grid > (uint32_t)ctx->num_sms) grid = ctx->num_sms;
    crc32c_chunks_kernel<<<grid, 1024, kCrcSmemBytes, ctx->main.stream>>>(P);
    ctx->launches++;
    B2C_CUDA(ctx, cudaGetLastError());
    if (L.n_multi) {
        const uint32_t blocks = (L.n_multi * 32 + 255) / 256;
        crc32c_combine_kernel<<<blocks, 256, 0, ctx->main.stream>>>(L.d_multi_list, L.n_multi, L.d_chunk_first, L.d_len,
                                                                d_rem, d_crc);
        ctx->launches++;
        B2C_CUDA(ctx, cudaGetLastError());
    }
    return B2C_OK;
}
