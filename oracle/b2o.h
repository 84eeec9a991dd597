/*
 * b2o — CPU oracle for the snappy + CRC32C payload path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of the reference's
 * algorithms (bRPC 1.4.0: src/butil/third_party/snappy/snappy.cc,
 * src/butil/crc32c.cc).  Nothing under oracle/ is part of the shipped codec:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it, and only as the checker or the timed
 * CPU baseline.  The product (libb2c.so) never links or calls it.
 *
 * Parity status: PINNED.  oracle/Makefile also compiles the reference's own
 * sources from /root/reference into oracle/_ref/libb2ref.so; tests/test_oracle.py
 * checks this restatement against that library on every corpus, against the
 * reference's known-answer vectors (test/crc32c_unittest.cc:18-53) and against
 * the committed fixtures in tests/golden/ that were generated from it.
 */
#ifndef B2O_H_
#define B2O_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- snappy (snappy.cc) ------------------------------------------------ */

/* snappy.cc:55-77 MaxCompressedLength */
size_t b2o_snappy_max_compressed_length(size_t n);

/* snappy.cc:875-954 Compress(Source*,Sink*) over a flat byte string
 * (== RawCompress, snappy.cc:1293-1304).  Returns compressed length. */
size_t b2o_snappy_compress(const uint8_t* in, size_t n, uint8_t* out);

/* snappy.cc:329-453 CompressFragment on one <=64 KiB fragment, table sized as
 * snappy.cc:247-271.  Returns bytes written. */
size_t b2o_snappy_compress_fragment(const uint8_t* in, size_t n, uint8_t* out);

/* snappy.cc:692-711 ReadUncompressedLength (lenient varint, what
 * Uncompress(Source*,Sink*) and therefore brpc's SnappyDecompress use).
 * Returns number of preamble bytes consumed, 0 on failure. */
int b2o_snappy_read_ulen_lenient(const uint8_t* in, size_t n, uint32_t* ulen);

/* snappy-stubs-internal.h:325-347 Varint::Parse32WithLimit (strict; used by
 * GetUncompressedLength(const char*,...), snappy.cc:235-244). */
int b2o_snappy_read_ulen_strict(const uint8_t* in, size_t n, uint32_t* ulen);

/* snappy.cc:716-787 DecompressAllTags + SnappyArrayWriter (snappy.cc:1141-1225)
 * behind the lenient preamble: RawUncompress(Source*, char*).  `out` must hold
 * `cap` bytes; a stream whose preamble exceeds `cap` fails.  Returns 1 on
 * success (eof && produced == expected, snappy.cc:867), else 0. */
int b2o_snappy_uncompress(const uint8_t* in, size_t n, uint8_t* out, size_t cap,
                          uint32_t* ulen_out);

/* snappy.cc:1252-1291 IsValidCompressed (Source-based, lenient preamble). */
int b2o_snappy_is_valid(const uint8_t* in, size_t n);

/* ---- crc32c (crc32c.cc / crc32c.h) ------------------------------------ */

/* crc32c.cc:379-422,451-454 Extend */
uint32_t b2o_crc32c_extend(uint32_t init, const uint8_t* data, size_t n);
/* same arithmetic, table path only (crc32c.cc:333-347 Slow_CRC32) */
uint32_t b2o_crc32c_extend_tables(uint32_t init, const uint8_t* data, size_t n);
/* crc32c.h:27-47 */
uint32_t b2o_crc32c_value(const uint8_t* data, size_t n);
uint32_t b2o_crc32c_mask(uint32_t crc);
uint32_t b2o_crc32c_unmask(uint32_t masked);

/* ---- deterministic corpora ---------------------------- */
enum { B2O_AZ09 = 0, B2O_RAND62 = 1, B2O_WORDS = 2 };
/* Fill out[0..n) with message `index` of corpus `kind`. */
void b2o_corpus_fill(int kind, uint64_t index, uint8_t* out, size_t n);
/* Fill n_msgs messages; message i goes to out + off[i], length len[i]. */
void b2o_corpus_fill_batch(int kind, uint64_t first_index, uint32_t n_msgs,
                           const uint64_t* off, const uint32_t* len,
                           uint8_t* out, int threads);

/* ---- batch drivers (static contiguous split over `threads` pthreads) ---- */
/* Each returns wall seconds spent inside the parallel region. */
double b2o_batch_crc32c(const uint8_t* data, const uint64_t* off,
                        const uint32_t* len, uint32_t n, uint32_t* crc,
                        int threads);
/* out slots: message i is written at out + out_off[i]; out_len[i] receives its size */
double b2o_batch_compress(const uint8_t* data, const uint64_t* off,
                          const uint32_t* len, uint32_t n, uint8_t* out,
                          const uint64_t* out_off, uint32_t* out_len,
                          int threads);
/* ok[i] = 1/0 ; ulen[i] = decoded length */
double b2o_batch_uncompress(const uint8_t* data, const uint64_t* off,
                            const uint32_t* len, uint32_t n, uint8_t* out,
                            const uint64_t* out_off, const uint32_t* out_cap,
                            uint32_t* ulen, int32_t* ok, int threads);
/* message-by-message memcmp of two arenas; returns how many messages differ */
uint32_t b2o_batch_equal(const uint8_t* a, const uint64_t* a_off, const uint8_t* b,
                         const uint64_t* b_off, const uint32_t* len, uint32_t n,
                         int32_t* scratch_ok, uint32_t* first_bad, int threads);

#ifdef __cplusplus
}
#endif
#endif
