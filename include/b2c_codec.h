/*
 * b2c_codec.h — C ABI of the B200 batch payload codec (libb2c.so).
 *
 * Drop-in boundary for bRPC's one data-parallel hot path: snappy compress /
 * decompress and CRC32C over many independent RPC message bodies.  Plain
 * pointers and sizes only; no C++ or torch types.  Each entry point cites the
 * reference interface (bRPC 1.4.0, paths relative to the reference root) it
 * replaces.  INTEGRATION.md shows the bRPC-side binding.
 *
 * Results are bit-exact with the reference: compressed bytes equal
 * butil::snappy::RawCompress, decompressed bytes / accept-reject decisions
 * equal butil::snappy::Uncompress(Source*,Sink*), CRCs equal
 * butil::crc32c::Extend.
 *
 * Conventions
 *   - every function returning int returns B2C_OK (0) or a negative B2C_E_*
 *     infrastructure error; b2c_last_error(ctx) has the text.
 *   - per-message results live in status arrays: B2C_OK, B2C_CORRUPT (the
 *     reference would return false for this message) or B2C_TOO_SMALL.
 *   - a "layout" describes n messages inside one arena: message i occupies
 *     bytes [off[i], off[i]+len[i]).
 *   - *_dev entry points take DEVICE pointers and enqueue work on the ctx
 *     stream; they do not synchronise (call b2c_ctx_sync).  *_batch entry
 *     points take HOST memory (iovec block lists in, flat buffers out) and
 *     return when the results are in host memory.
 *   - thread safety: one ctx may be used by one thread at a time; create one
 *     ctx per worker thread (or guard it) — all ctxs on a device share it.
 *   - device input arenas of the *_dev entry points (d_in, d_data) must be
 *     readable for 16 bytes past b2c_layout_arena_bytes(layout): the kernels
 *     fetch bodies in aligned 16-byte / 4-byte units and the unit that holds a
 *     message's last byte may extend past it (the bytes are never used).
 *     b2c_dev_alloc always leaves that room; with your own allocator add 16.
 *     The *_batch entry points stage into their own padded buffers.
 */
#ifndef B2C_CODEC_H_
#define B2C_CODEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2C_OK 0
#define B2C_CORRUPT 1   /* malformed snappy stream: reference returns false (snappy.cc:867) */
#define B2C_TOO_SMALL 2 /* output slot smaller than the stream's uncompressed length */
#define B2C_CRC_MISMATCH 3 /* b2c_snappy_decompress_verify_dev: decoded fine, checksum differs */

#define B2C_E_CUDA (-1)     /* CUDA runtime/driver failure (text in b2c_last_error) */
#define B2C_E_ARG (-2)      /* invalid argument */
#define B2C_E_NOMEM (-3)    /* host or device allocation failed */
#define B2C_E_CAPACITY (-4) /* caller-provided output buffer too small */
#define B2C_E_NODEVICE (-5) /* no usable CUDA device: the codec has no CPU fallback */

typedef struct b2c_ctx b2c_ctx;
typedef struct b2c_layout b2c_layout;

/* Layout-identical to butil's `struct const_iovec` (src/butil/iobuf.h:40-44);
 * one entry per IOBuf BlockRef as returned by IOBuf::backing_block(i)
 * (src/butil/iobuf.cpp:1504-1510). */
typedef struct b2c_iovec {
    const void* iov_base;
    size_t iov_len;
} b2c_iovec;

/* ---- context ----------------------------------------------------------- */

/* Bind a codec context to CUDA device `device` (streams, lookup tables, pinned
 * staging, device workspace).  Fails with B2C_E_NODEVICE when there is no GPU.
 * Replaces nothing in the reference (its codec is stateless); it is the state
 * behind the handlers registered at src/brpc/global.cpp:389-393. */
int b2c_ctx_create(int device, b2c_ctx** ctx);
void b2c_ctx_destroy(b2c_ctx* ctx);
const char* b2c_last_error(const b2c_ctx* ctx);
int b2c_ctx_device(const b2c_ctx* ctx);
/* cudaStream_t of the ctx, as an opaque pointer */
void* b2c_ctx_stream(b2c_ctx* ctx);
int b2c_ctx_sync(b2c_ctx* ctx);
/* Options.
 *   B2C_OPT_INPUT_PINNED (default 0): promise that EVERY input block handed to the *_batch
 *     entry points lies in page-locked host memory that the device can address with the same
 *     pointer (cudaHostAlloc / b2c_host_alloc_pinned; the bRPC-side equivalent is a pinned IOBuf
 *     block pool installed through iobuf::blockmem_allocate, src/butil/iobuf.cpp:163-170, as
 *     src/brpc/rdma/block_pool.cpp does for RDMA).  The library then skips the host-side gather:
 *     only the block list is copied and a kernel pulls the blocks over PCIe itself.  Passing a
 *     pageable block with this option set is a fatal device fault. */
#define B2C_OPT_INPUT_PINNED 1
/*   B2C_OPT_HOST_THREADS (default 0 = automatic): threads of this ctx's host pool, which gathers IOBuf
 *     blocks into pinned staging and copies results out for the *_batch entry points.  Automatic is
 *     the host's cores divided by the ranks on this node (LOCAL_WORLD_SIZE when set), at least 2 and
 *     at most 16.  Every ctx owns its pool, so ctxs used from different threads (a sender and a
 *     receiver, or one per device) never wait for each other; give each its share of the cores. */
#define B2C_OPT_HOST_THREADS 2
int b2c_ctx_set_option(b2c_ctx* ctx, int option, int value);
/* Implementation tunables by name (kernel variants, pipeline geometry): "window_wps", "small_lpw_c", "slab_mb",
 * "lockstep", ... -- the list and the meaning of each value are in csrc/b2c_api.cu (kTunableNames /
 * apply_tunable).  The defaults are the fastest measured; every value is checked for byte-exact results by the
 * GPU tests.  Each name can also be given through the environment as B2C_<NAME IN CAPITALS>, read once by
 * b2c_ctx_create (a convenience for experiments).  Unknown names / invalid values: B2C_E_ARG.  No counterpart in
 * the reference. */
int b2c_ctx_set_tunable(b2c_ctx* ctx, const char* name, long value);
/* Number of kernels this ctx has launched so far (bench.py's gpu_launches). */
uint64_t b2c_ctx_launch_count(const b2c_ctx* ctx);

/* Device memory / transfer / timing helpers so that hosts without their own
 * CUDA binding can drive the *_dev entry points. */
int b2c_dev_alloc(b2c_ctx* ctx, size_t bytes, void** d_ptr);
int b2c_dev_free(b2c_ctx* ctx, void* d_ptr);
int b2c_host_alloc_pinned(b2c_ctx* ctx, size_t bytes, void** h_ptr);
int b2c_host_free_pinned(b2c_ctx* ctx, void* h_ptr);
int b2c_memcpy_h2d(b2c_ctx* ctx, void* d_dst, const void* h_src, size_t bytes); /* async on ctx stream */
int b2c_memcpy_d2h(b2c_ctx* ctx, void* h_dst, const void* d_src, size_t bytes); /* async on ctx stream */
int b2c_memset_dev(b2c_ctx* ctx, void* d_dst, int value, size_t bytes);
typedef struct b2c_event b2c_event;
int b2c_event_create(b2c_ctx* ctx, b2c_event** ev);
void b2c_event_destroy(b2c_event* ev);
int b2c_event_record(b2c_ctx* ctx, b2c_event* ev);                /* on the ctx stream */
int b2c_event_elapsed_ms(b2c_event* start, b2c_event* stop, float* ms); /* syncs on stop */

/* Kernel timing for bench.py's roofline line (instrumentation; no counterpart in
 * the reference).  While enabled, the launchers bracket their dominant kernels
 * with CUDA events on the ctx stream; b2c_prof_total_ms waits for them and
 * returns the summed duration and the number of timed launches since the last
 * enable/reset. */
#define B2C_PROF_CRC32C 0      /* crc32c_chunks_kernel */
#define B2C_PROF_COMPRESS 1    /* snappy_compress_fragments_kernel (all size classes) */
#define B2C_PROF_DECOMPRESS 2  /* snappy_decompress_kernel (all size classes) */
#define B2C_PROF_PACK 3        /* compress post-pass: sizes + scan + gather */
#define B2C_PROF_COUNT 4
int b2c_prof_enable(b2c_ctx* ctx, int on);
int b2c_prof_total_ms(b2c_ctx* ctx, int which, float* total_ms, uint32_t* launches);

/* ---- pure host helpers ------------------------------------------------- */

/* snappy::MaxCompressedLength — src/butil/third_party/snappy/snappy.cc:55-77 */
size_t b2c_snappy_max_compressed_length(size_t source_len);
/* snappy::GetUncompressedLength(const char*, size_t, size_t*) — snappy.cc:235-244
 * (strict varint, snappy-stubs-internal.h:325-347).  Returns 1/0 like the reference's bool. */
int b2c_snappy_get_uncompressed_length(const void* compressed, size_t n, uint32_t* result);
/* SnappyDecompressor::ReadUncompressedLength — snappy.cc:692-711: the lenient
 * preamble parse that Uncompress(Source*,Sink*) (and so brpc's SnappyDecompress)
 * uses.  Returns the number of preamble bytes, 0 on failure. */
int b2c_snappy_read_preamble(const void* compressed, size_t n, uint32_t* result);
/* butil::crc32c::Mask / Unmask — src/butil/crc32c.h:38-47 */
uint32_t b2c_crc32c_mask(uint32_t crc);
uint32_t b2c_crc32c_unmask(uint32_t masked_crc);

/* ---- layouts ----------------------------------------------------------- */

/* Describe n messages in an arena: host arrays off[n] (byte offsets) and len[n].
 * The arrays are copied; derived work lists (64 KiB snappy fragments —
 * snappy.cc:887-948 — and CRC chunks) are precomputed and uploaded.  The
 * reference has no such object: it walks one IOBuf at a time
 * (IOBufAsSnappySource, src/butil/iobuf.cpp:2062-2090). */
int b2c_layout_create(b2c_ctx* ctx, uint32_t n, const uint64_t* off, const uint32_t* len,
                      b2c_layout** layout);
void b2c_layout_destroy(b2c_layout* layout);
uint32_t b2c_layout_count(const b2c_layout* layout);
/* sum of len[] */
uint64_t b2c_layout_payload_bytes(const b2c_layout* layout);
/* one past the last arena byte any message touches */
uint64_t b2c_layout_arena_bytes(const b2c_layout* layout);
/* sum over messages of MaxCompressedLength(len[i]) + 5: capacity that always
 * suffices for b2c_snappy_compress_dev's packed output */
uint64_t b2c_layout_max_compressed_bytes(const b2c_layout* layout);

/* ---- device-resident entry points -------------------------------------- */

/* d_crc[i] = butil::crc32c::Extend(d_init ? d_init[i] : 0, msg_i, len_i)
 * — src/butil/crc32c.cc:451-454 / crc32c.h:19-29 (Value == Extend(0,...)). */
int b2c_crc32c_dev(b2c_ctx* ctx, const b2c_layout* layout, const void* d_data,
                   const uint32_t* d_init, uint32_t* d_crc);

/* Same result as b2c_crc32c_dev, for callers that must not allocate device memory
 * at call time.  No counterpart in the reference (its CRC needs no scratch at all).
 *
 * Who needs it: an RPC worker that checksums bodies on the request path cannot
 * afford a cudaMalloc there — the allocator call serialises the whole device and
 * every other in-flight batch waits behind it — and a service that handed the rest
 * of HBM to another engine at start-up has nothing left to allocate from, so a
 * late workspace growth inside b2c_crc32c_dev would simply fail.  b2c_crc32c_dev
 * keeps the per-chunk partial CRCs of messages longer than 4 KiB in a workspace it
 * grows on demand; this variant keeps them in a fixed array that is part of the
 * library image (4 bytes per 4 KiB chunk, B2C_CRC32C_NOALLOC_MAX_CHUNKS entries =
 * 4 MiB, resident on every GPU that loads the library until the process exits).
 *
 * Limits and rules:
 *   - layouts with more than B2C_CRC32C_NOALLOC_MAX_CHUNKS chunks (sum over
 *     messages of max(1, ceil(len/4096))) are refused with B2C_E_CAPACITY;
 *   - the first call on each device looks up the array's device address; make
 *     that first call before starting any stream capture on that device, and
 *     let it carry at least one message longer than 4 KiB: the CUDA runtime
 *     loads a kernel's code on its first launch (which allocates), and only a
 *     multi-chunk message launches both kernels of this entry point;
 *   - the looked-up address is remembered per device ordinal for the life of
 *     the process; it does not survive cudaDeviceReset;
 *   - all calls on a device share the one array: two calls whose kernels can
 *     overlap on the GPU (different contexts or streams) race — serialise them;
 *   - the array keeps whatever the previous call left; every slot a call reads
 *     is written earlier in that same call, so nothing has to be cleared;
 *   - it calls neither cudaMalloc nor cudaFree nor any other allocating routine;
 *     the layout (created earlier, at set-up time) and the caller's input and
 *     output buffers are the only device memory it touches besides the array. */
#define B2C_CRC32C_NOALLOC_MAX_CHUNKS (1u << 20)
int b2c_crc32c_noalloc_dev(b2c_ctx* ctx, const b2c_layout* layout, const void* d_data,
                           const uint32_t* d_init, uint32_t* d_crc);

/* For every message: the exact byte stream butil::snappy::Compress(Source*,Sink*)
 * produces (snappy.cc:875-954; what brpc::policy::SnappyCompress(const IOBuf&,
 * IOBuf*) appends, src/brpc/policy/snappy_compress.cpp:51-55).  Streams are
 * packed back to back in d_out; d_out_off[n+1] (device) receives their offsets
 * (stream i = [d_out_off[i], d_out_off[i+1])).  d_out must hold out_cap >=
 * b2c_layout_max_compressed_bytes(layout) bytes. */
int b2c_snappy_compress_dev(b2c_ctx* ctx, const b2c_layout* layout, const void* d_in,
                            void* d_out, uint64_t out_cap, uint64_t* d_out_off);

/* d_ulen[i] / d_status[i]: lenient preamble of stream i (snappy.cc:692-711);
 * status B2C_CORRUPT where the reference's ReadUncompressedLength fails. */
int b2c_snappy_uncompressed_length_dev(b2c_ctx* ctx, const b2c_layout* in_layout, const void* d_in,
                                       uint32_t* d_ulen, int32_t* d_status);

/* butil::snappy::Uncompress(Source*,Sink*) per message (snappy.cc:1526-1552,
 * tag loop 716-787; what brpc::policy::SnappyDecompress(const IOBuf&, IOBuf*)
 * runs, snappy_compress.cpp:57-61).  Stream i is decoded into
 * d_out + out_layout.off[i]; out_layout.len[i] is the slot capacity.
 * d_status[i] = B2C_OK iff the reference returns true (then d_ulen[i] bytes were
 * written), B2C_CORRUPT iff it returns false, B2C_TOO_SMALL when the preamble
 * asks for more than the slot.  d_ulen may be NULL.
 * Workspace (owned by the ctx, grow-only): when the batch holds slots larger than 64 KiB, their streams are cut into
 * independently decoded 64 KiB blocks behind a table of 4 bytes per compressed byte of the batch (so about four
 * times in_layout's payload of device memory, on top of input and output). */
int b2c_snappy_decompress_dev(b2c_ctx* ctx, const b2c_layout* in_layout, const void* d_in,
                              const b2c_layout* out_layout, void* d_out, uint32_t* d_ulen,
                              int32_t* d_status);

/* butil::snappy::IsValidCompressed(Source*) per stream (snappy.cc:1252-1291, the
 * SnappyDecompressionValidator writer): d_status[i] = B2C_OK iff the reference would
 * accept stream i, B2C_CORRUPT otherwise.  Nothing is decoded; d_ulen (nullable)
 * receives the preamble's length. */
int b2c_snappy_validate_dev(b2c_ctx* ctx, const b2c_layout* in_layout, const void* d_in, uint32_t* d_ulen,
                            int32_t* d_status);

/* How far each stream decodes — the bookkeeping of snappy::UncompressAsMuchAsPossible (snappy.cc:1519-1524, writer
 * snappy.cc:1323-1470) without moving data.  d_out5[5*i + ...]: [0] the value the reference returns, [1] the number
 * of bytes its sink receives, [2] the stream offset up to which whole tags were applied, [3] the bytes of one more,
 * partially applied literal and [4] where in the stream that literal's data starts.  The decoded prefix itself is
 * then the decode of varint(d[1]) || stream[preamble, d[2]) || literal(stream[d[4], d[4] + d[3])), which the host
 * mirror b200::snappy::UncompressAsMuchAsPossible builds and runs through b2c_snappy_decompress_batch.  One thread
 * per stream: a rarely used entry point, not a throughput path. */
int b2c_snappy_decodable_prefix_dev(b2c_ctx* ctx, const b2c_layout* in_layout, const void* d_in, uint32_t* d_out5);

/* Decompress and verify in one call — the unit of the snappy framing format's
 * compressed chunk (src/butil/third_party/snappy/framing_format.txt:39-59,82-92):
 * d_status[i] = B2C_OK iff stream i decodes (as b2c_snappy_decompress_dev) AND
 * butil::crc32c::Mask(crc32c::Value(decoded bytes)) == d_masked_crc[i]; a stream that
 * decodes but whose checksum differs gets B2C_CRC_MISMATCH.  Output slots must be sized
 * exactly (out_layout.len[i] == uncompressed length, see
 * b2c_snappy_uncompressed_length_dev); a stream shorter than its slot also reports
 * B2C_CRC_MISMATCH. */
int b2c_snappy_decompress_verify_dev(b2c_ctx* ctx, const b2c_layout* in_layout, const void* d_in,
                                     const b2c_layout* out_layout, void* d_out, const uint32_t* d_masked_crc,
                                     uint32_t* d_ulen, int32_t* d_status);

/* ---- host (end-to-end) entry points ------------------------------------- */
/* Message i is the concatenation of segs[seg_begin[i] .. seg_begin[i+1]) — the
 * BlockRefs of one IOBuf (src/butil/iobuf.h:74-101).  Inputs are gathered into
 * pinned staging, moved with async H2D copies, processed, and results copied
 * back into the caller's flat buffers. */

/* brpc::policy::SnappyCompress(const IOBuf&, IOBuf*) for a batch
 * (snappy_compress.cpp:51-55).  out_off[n+1]: stream i = out[out_off[i], out_off[i+1]). */
int b2c_snappy_compress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                              uint32_t n, void* out, uint64_t out_cap, uint64_t* out_off);

/* brpc::policy::SnappyDecompress(const IOBuf&, IOBuf*) for a batch
 * (snappy_compress.cpp:57-61).  Decoded message i = out[out_off[i], out_off[i+1]);
 * status[i] as for b2c_snappy_decompress_dev.  Slot i is as long as stream i's preamble
 * claims (0 when the preamble is malformed or claims more than a stream of that size can
 * expand to), so a stream that fails later still occupies its slot, contents unspecified. */
int b2c_snappy_decompress_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                                uint32_t n, void* out, uint64_t out_cap, uint64_t* out_off,
                                int32_t* status);

/* snappy::RawUncompressToIOVec (snappy.cc:1122-1132, SnappyIOVecWriter :963-1120) for a batch:
 * decoded message i is laid over out_iov[out_iov_begin[i] .. out_iov_begin[i+1]) in order, each
 * iovec filled before the next (the blocks an IOBuf would hand out).  status[i] = B2C_OK iff the
 * reference returns true; B2C_TOO_SMALL when the iovecs cannot hold the length the preamble
 * claims (the reference's writer overruns and returns false; nothing is written here);
 * B2C_CORRUPT otherwise.  Unused tail space of the iovecs is left untouched. */
int b2c_snappy_decompress_iov_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                                    uint32_t n, const b2c_iovec* out_iov, const uint32_t* out_iov_begin,
                                    int32_t* status);

/* The sender's two passes over a body in one: b2c_snappy_compress_batch plus
 * crc[i] = butil::crc32c::Value(body i) (crc32c.h:31-33), or Mask() of it when masked != 0
 * (crc32c.h:44-48) -- the pair a checksummed, snappy-compressed frame carries
 * (framing_format.txt:39-59).  The bodies cross PCIe once for both results.  No single
 * counterpart in the reference: it replaces the SnappyCompress call and the crc32c::Value
 * call a sender makes on the same IOBuf. */
int b2c_snappy_compress_crc_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                                  uint32_t n, void* out, uint64_t out_cap, uint64_t* out_off,
                                  uint32_t* crc, int masked);

/* The receiver's counterpart: b2c_snappy_decompress_batch plus the check of
 * b2c_snappy_decompress_verify_dev against masked_crc[i]; status[i] additionally takes
 * B2C_CRC_MISMATCH.  The decoded bytes are checksummed in HBM before they leave the GPU. */
int b2c_snappy_decompress_verify_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin,
                                       uint32_t n, const uint32_t* masked_crc, void* out,
                                       uint64_t out_cap, uint64_t* out_off, int32_t* status);

/* butil::crc32c::Extend for a batch (crc32c.h:19-29); init may be NULL (== Value). */
int b2c_crc32c_batch(b2c_ctx* ctx, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                     const uint32_t* init, uint32_t* crc);

/* ---- one batch over several GPUs ----------------------------------------------- */
/* Message bodies are independent (snappy.cc:887-948; CRC per message), so a batch shards over the GPUs of a box
 * by a host-side split and nothing else: the message list is cut into one contiguous range per device, balanced
 * by bytes; every range runs through its device's own ctx on that device's own worker thread; results land in
 * the caller's buffers in message order.  No collective, no peer traffic.  The reference has no counterpart (it
 * has no device at all); this is the proposed `ctx_create(const int* devices, int n, ...)` of SURVEY.md 8(b).
 * A b2c_multi is used by one thread at a time.  The host's cores are divided among the devices' gather pools. */
typedef struct b2c_multi b2c_multi;
int b2c_multi_create(const int* devices, int n_devices, b2c_multi** multi);
void b2c_multi_destroy(b2c_multi* multi);
int b2c_multi_device_count(const b2c_multi* multi);
const char* b2c_multi_last_error(const b2c_multi* multi);
/* b2c_crc32c_batch over all devices */
int b2c_multi_crc32c_batch(b2c_multi* multi, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                           const uint32_t* init, uint32_t* crc);
/* b2c_snappy_compress_crc_batch (crc == NULL: b2c_snappy_compress_batch) over all devices.  Compressed sizes are not
 * known before the devices have run, so every device packs its streams into its own part of `out`: stream i =
 * out[out_off[i], out_off[i] + out_len[i]) (both arrays have n entries); out_cap must be at least the sum of
 * MaxCompressedLength(len) + 5 over the batch plus 256 bytes per device. */
int b2c_multi_snappy_compress_crc_batch(b2c_multi* multi, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                                        void* out, uint64_t out_cap, uint64_t* out_off, uint32_t* out_len, uint32_t* crc,
                                        int masked);
/* b2c_snappy_decompress_verify_batch (masked_crc == NULL: b2c_snappy_decompress_batch) over all devices; the decoded
 * sizes come from the preambles, so the result is packed exactly as by the single-device call: message i =
 * out[out_off[i], out_off[i+1]). */
int b2c_multi_snappy_decompress_verify_batch(b2c_multi* multi, const b2c_iovec* segs, const uint32_t* seg_begin, uint32_t n,
                                             const uint32_t* masked_crc, void* out, uint64_t out_cap, uint64_t* out_off,
                                             int32_t* status);

/* Bytes the last *_batch call moved over PCIe (bench.py's e2e accounting). */
void b2c_last_transfer_bytes(const b2c_ctx* ctx, uint64_t* h2d, uint64_t* d2h);

#ifdef __cplusplus
}
#endif
#endif /* B2C_CODEC_H_ */
