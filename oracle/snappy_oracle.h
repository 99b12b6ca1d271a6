/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement ("oracle") of the reference's
 * host Snappy codec in the PIM-compression multi-block format.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may build, load or call this.  The shipped product
 * (pim-compression_b200/) never links or falls back to it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement
 *   (a) against the reference's six golden pairs (tests/golden/, taken from
 *       /root/reference/test: decompress(.snappy) == .txt and
 *       compress(.txt, 32768) == .snappy, byte for byte), and
 *   (b) against the reference's own host codec compiled unmodified from
 *       the .c files of /root/reference/snappy into oracle/_ref/ (see oracle/Makefile) on
 *       seeded synthetic inputs over block sizes 1 ... 65536.
 *
 * Reference citations are relative to /root/reference/snappy/.
 */
#ifndef SNAPPY_ORACLE_H
#define SNAPPY_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORACLE_OK = 0, ORACLE_INVALID_INPUT = 1, ORACLE_BUFFER_TOO_SMALL = 2 };

/* snappy_compress.c:55-60 -- the reference's (per-file) output bound. */
size_t oracle_max_compressed_length(size_t input_length);

/* Safe capacity for a whole multi-block file: header + 4 B per block + the
 * per-block literal-only worst case.  (The reference bound above ignores the
 * 4-byte block headers; tests allocate with this one.) */
size_t oracle_file_bound(size_t input_length, uint32_t block_size);

/* snappy_compress.c:284-413 (compress_block) with :139-146 (table sizing).
 * Compresses in[0..n) as ONE block; writes the 4-byte LE size followed by the
 * elements at out; returns total bytes written (4 + compressed size). */
size_t oracle_compress_block(const uint8_t *in, uint32_t n, uint8_t *out);

/* snappy_compress.c:455-485 (snappy_compress_host). */
int oracle_compress(const uint8_t *in, size_t n, uint32_t block_size,
                    uint8_t *out, size_t out_cap, size_t *out_len);

/* snappy_decompress.c:187-215 + :218-289 (setup_decompression +
 * snappy_decompress_host), reference semantics INCLUDING its quirks: silent
 * clamping at buffer ends (:145-147,174-175), the uint16_t element length
 * (:233), copies may reach into earlier blocks (:167-173).
 * *out_len receives the header's decompressed length. */
int oracle_decompress(const uint8_t *in, size_t n, uint8_t *out, size_t out_cap,
                      size_t *out_len);

/* Independent strict decoder of the same format: every element must fit its
 * block exactly, copies may not reach before their own block, no 16-bit
 * length wrap.  This is the behaviour the CUDA decoder implements (SURVEY.md
 * 8(a) row Q "GPU contract"); it agrees with oracle_decompress on every
 * stream the reference compressor can produce with block_size <= 65535. */
int oracle_decompress_strict(const uint8_t *in, size_t n, uint8_t *out,
                             size_t out_cap, size_t *out_len);

/* Reads the two header varints (snappy_decompress.c:23-37); returns bytes
 * consumed or 0 on a malformed varint. */
size_t oracle_read_header(const uint8_t *in, size_t n, uint32_t *dlen,
                          uint32_t *block_size);

#ifdef __cplusplus
}
#endif
#endif
