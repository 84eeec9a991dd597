// How fast can a KERNEL pull page-locked host memory over PCIe, and with which instructions?  (The copy engines
// reach ~55 GB/s on this box.)  Variants: plain 16-byte loads, cp.async (LDGSTS) into shared memory, and TMA
// bulk copies (cp.async.bulk, UBLKCP) of several sizes.  Tuning aid for the device-side gather of pinned IOBuf
// blocks (B2C_OPT_INPUT_PINNED); build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o sysmem_read_probe ...
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// piece = bytes one warp moves per step (512 = one 16-byte load per lane)
__global__ void plain_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t vecs, int unroll) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (unroll == 4) {
        for (; i + 3 * stride < vecs; i += 4 * stride) {
            uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
            dst[i] = a, dst[i + stride] = b, dst[i + 2 * stride] = c, dst[i + 3 * stride] = d;
        }
    }
    for (; i < vecs; i += stride) dst[i] = src[i];
}

// one CTA moves `piece` bytes per step through shared memory with ONE bulk copy in and one bulk copy out,
// `depth` steps in flight
template <int DEPTH>
__global__ void __launch_bounds__(32) bulk_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t bytes,
                                                  uint32_t piece) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ uint64_t bar[DEPTH];
    const size_t pieces = bytes / piece;
    if (threadIdx.x == 0) {
        for (int d = 0; d < DEPTH; d++)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[d])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint32_t phase[DEPTH];
        for (int d = 0; d < DEPTH; d++) phase[d] = 0;
        size_t issued = 0, p = blockIdx.x;
        // prologue
        size_t q = p;
        for (int d = 0; d < DEPTH && q < pieces; d++, q += gridDim.x) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[d])), "r"(piece) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             s32(sm + (size_t)d * piece)), "l"(src + q * piece), "r"(piece), "r"(s32(&bar[d])) : "memory");
            issued++;
        }
        int d = 0;
        for (; p < pieces; p += gridDim.x) {
            asm volatile(
                "{\n\t.reg .pred P1;\n\tW1:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 0x989680;\n\t@P1 bra D1;\n\tbra W1;\n\tD1:\n\t}" ::"r"(
                    s32(&bar[d])), "r"(phase[d]) : "memory");
            phase[d] ^= 1;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + p * piece),
                         "r"(s32(sm + (size_t)d * piece)), "r"(piece) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the slot may be refilled
            if (q < pieces) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[d])), "r"(piece) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 s32(sm + (size_t)d * piece)), "l"(src + q * piece), "r"(piece), "r"(s32(&bar[d])) : "memory");
                q += gridDim.x;
            }
            d = (d + 1) % DEPTH;
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

int main() {
    const size_t bytes = 512ull << 20;
    uint8_t *h, *d;
    CK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
    CK(cudaMalloc(&d, bytes));
    for (size_t i = 0; i < bytes; i += 4096) h[i] = (uint8_t)(i >> 12);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    auto report = [&](const char* name, float ms) { printf("%-44s %8.2f ms  %6.1f GB/s\n", name, ms, bytes / (ms * 1e-3) / 1e9); };
    float ms;
    for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(e0));
        CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    report("copy engine (cudaMemcpyAsync)", ms);
    const int grids[] = {148, 592, 2368, 9472};
    for (int unroll : {1, 4})
        for (int g : grids) {
            for (int rep = 0; rep < 2; rep++) {
                CK(cudaEventRecord(e0));
                plain_kernel<<<g, 256>>>(reinterpret_cast<const uint4*>(h), reinterpret_cast<uint4*>(d), bytes / 16, unroll);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
            }
            char nm[96];
            snprintf(nm, sizeof nm, "plain 16-byte loads, grid %d x 256, unroll %d", g, unroll);
            report(nm, ms);
        }
    CK(cudaFuncSetAttribute(bulk_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    CK(cudaFuncSetAttribute(bulk_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    for (uint32_t piece : {512u, 1024u, 2048u, 4096u, 8192u, 16384u})
        for (int g : {148, 592, 1184}) {
            for (int rep = 0; rep < 2; rep++) {
                CK(cudaEventRecord(e0));
                if (piece * 8 <= 65536)
                    bulk_kernel<8><<<g, 32, piece * 8>>>(h, d, bytes, piece);
                else
                    bulk_kernel<4><<<g, 32, piece * 4>>>(h, d, bytes, piece);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
            }
            char nm[96];
            snprintf(nm, sizeof nm, "TMA bulk copies of %u B, %d CTAs, depth %d", piece, g, piece * 8 <= 65536 ? 8 : 4);
            report(nm, ms);
        }
    // the same reads while the copy engine writes results back to the host (what the sender's pipeline does)
    {
        uint8_t *h2, *d2;
        CK(cudaHostAlloc(&h2, bytes, cudaHostAllocDefault));
        CK(cudaMalloc(&d2, bytes));
        cudaStream_t s1, s2;
        CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
        for (int mode = 0; mode < 2; mode++) {
            for (int rep = 0; rep < 2; rep++) {
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpyAsync(h2, d2, bytes, cudaMemcpyDeviceToHost, s2));
                CK(cudaEventRecord(e0, s1));
                if (mode == 0)
                    plain_kernel<<<148, 256, 0, s1>>>(reinterpret_cast<const uint4*>(h), reinterpret_cast<uint4*>(d), bytes / 16, 1);
                else
                    CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s1));
                CK(cudaEventRecord(e1, s1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
            }
            report(mode == 0 ? "plain loads WHILE a D2H copy runs" : "copy engine H2D WHILE a D2H copy runs", ms);
        }
    }
    // spot check
    uint8_t probe[4];
    CK(cudaMemcpy(probe, d + (123ull << 12), 1, cudaMemcpyDeviceToHost));
    printf("check %s\n", probe[0] == 123 ? "ok" : "BAD");
    return 0;
}
