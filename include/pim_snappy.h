/*
 * pim_snappy.h -- C ABI of the B200-native backend for PIM-compression's
 * block-parallel Snappy codec (multi-block format, snappy/README.md:19-33 of
 * the reference).
 *
 * This header is the drop-in boundary.  Part 1 re-declares, layout-identical,
 * the types and entry points that the reference's CLI (snappy/dpu_snappy.c)
 * binds for its UPMEM-DPU backend; the library implements them with sm_100a
 * CUDA kernels.  Part 2 is the device-resident batched API the benchmarks and
 * callers with data already in HBM use.  Plain pointers and sizes only.
 *
 * There is NO CPU fallback: every codec entry point fails (SNAPPY_INVALID_INPUT
 * and a message on stderr) when no sm_100 device / CUDA driver is usable.
 *
 * All reference citations are relative to /root/reference/.
 */
#ifndef PIM_SNAPPY_H
#define PIM_SNAPPY_H

#include <stddef.h>
#include <stdint.h>
#include <sys/time.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------
 * Part 1 -- reference-compatible surface
 * ---------------------------------------------------------------------- */

/* snappy/dpu_snappy.h:21-25 */
typedef enum {
	SNAPPY_OK = 0,           /* success */
	SNAPPY_INVALID_INPUT,    /* malformed stream, or the device path failed */
	SNAPPY_BUFFER_TOO_SMALL  /* a length exceeds what the buffers/format allow */
} snappy_status;

/* snappy/dpu_snappy.h:37-44 -- same member order and types */
typedef struct host_buffer_context {
	const char *file_name;   /* file name (informational) */
	uint8_t *buffer;         /* whole buffer */
	uint8_t *curr;           /* cursor */
	unsigned long length;    /* valid bytes */
	unsigned long max;       /* largest length the caller allows */
} host_buffer_context;

/* snappy/dpu_snappy.h:47-55 -- seconds; printed by the CLI in this order */
struct program_runtime {
	double pre;       /* host pre-processing (header parse, block-offset walk) */
	double d_alloc;   /* device + pinned staging allocation */
	double load;      /* context / module initialisation ("program load") */
	double copy_in;   /* host -> device */
	double run;       /* kernels (CUDA-event time, max over GPUs) */
	double copy_out;  /* device -> host */
	double d_free;    /* release */
};

/* snappy/dpu_snappy.h:60 (defined in dpu_snappy.c:87-91) */
double get_runtime(struct timeval *start, struct timeval *end);

/* snappy/snappy_compress.h:16 (snappy_compress.c:418-453).  Allocates
 * output->buffer (sized for the worst case INCLUDING the 4-byte block headers
 * the reference's bound forgets; page-locked when the call will be streamed,
 * i.e. beyond two chunks, plain malloc below that -- release either with
 * pim_snappy_host_free()), output->curr = buffer, output->length = 0,
 * runtime->pre = elapsed. */
void setup_compression(struct host_buffer_context *input, struct host_buffer_context *output,
                       struct program_runtime *runtime);

/* snappy/snappy_decompress.h:15 (snappy_decompress.c:187-215).  Consumes the
 * first header varint (decompressed length) from input->curr, checks it
 * against output->max, allocates output->buffer, sets output->length. */
snappy_status setup_decompression(struct host_buffer_context *input, struct host_buffer_context *output,
                                  struct program_runtime *runtime);

/* snappy/snappy_compress.h:37 (snappy_compress.c:487-714) -- the DPU entry
 * point, now served by the GPU(s).  Pre: setup_compression() ran.  Post:
 * output->buffer[0 .. output->length) holds the complete file image, byte
 * identical to the reference's snappy_compress_host(); runtime fields filled
 * (pre is accumulated with +=, as in the reference :528).  Unlike the DPU
 * path (:626-627,697-703) it does not fwrite() the file itself unless
 * pim_snappy_set_write_file(1) was called -- the CLI writes output->buffer
 * (see INTEGRATION.md).  The host entry points are serialised internally:
 * they may be called from several threads, one runs at a time. */
snappy_status snappy_compress_dpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                  uint32_t block_size, struct program_runtime *runtime);

/* snappy/snappy_decompress.h:34 (snappy_decompress.c:292-493).  Pre:
 * setup_decompression() ran (input->curr at the block-size varint).  Post:
 * output->buffer[0 .. output->length) holds the decoded bytes. */
snappy_status snappy_decompress_dpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                    struct program_runtime *runtime);

/* The reference's DPU compressor writes output->file_name itself (snappy_compress.c:626-627,697-703) and its
 * main() skips write_output_host for that case (dpu_snappy.c:209).  With enable != 0 (or $PIM_SNAPPY_WRITE_FILE=1
 * in the environment) snappy_compress_dpu() does the same after filling output->buffer, so the reference's
 * unmodified main() links against this library and still produces the file.  Default: off (the CLI shipped
 * here writes the buffer, like the host path). */
void pim_snappy_set_write_file(int enable);

/* Preferred names for the same two functions (they never write files). */
snappy_status snappy_compress_gpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                  uint32_t block_size, struct program_runtime *runtime);
snappy_status snappy_decompress_gpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                    struct program_runtime *runtime);

/* Page-locked host memory for buffers handed to the entry points above (what
 * read_input_host's malloc, dpu_snappy.c:41, becomes).  Pageable memory is
 * accepted too, it is just copied at pageable-memcpy speed. */
void *pim_snappy_host_alloc(size_t bytes);
void pim_snappy_host_free(void *p);

/* Which GPUs the entry points shard over (contiguous block ranges, one host
 * thread per GPU; replaces the compile-time NR_DPUS, snappy/Makefile:10-12).
 * Default: all visible devices, or the first $PIM_SNAPPY_NR_GPUS of them.
 * Returns the number of devices that will be used, <= 0 if none is usable. */
int pim_snappy_set_devices(const int *device_ids, int n);
int pim_snappy_device_count(void);

/* The contiguous block range [*block_lo, *block_hi) that shard `shard` of
 * `n_shards` codes: ceil(nblocks / n_shards) blocks each, the last ones
 * possibly fewer or none (snappy_compress.c:495, snappy_decompress.c:307).  The
 * entry points use it with one shard per GPU; multi-process callers (one
 * process per GPU) use it to pick their rank's blocks.  Pure arithmetic. */
void pim_snappy_shard_range(uint64_t nblocks, int n_shards, int shard, uint64_t *block_lo, uint64_t *block_hi);

/* The host entry points stream shards larger than two chunks: chunk i+1 is
 * copied to the device while chunk i is coded and chunk i-1 is copied back
 * (what the reference's separate copy_in / run / copy_out phases,
 * snappy_compress.c:560-670, become when they overlap; compress groups go round
 * robin over $PIM_SNAPPY_STREAMS (default 8) compute streams).  Default 32 MiB, or
 * $PIM_SNAPPY_CHUNK_MIB; 0 processes every shard in one piece. */
void pim_snappy_set_chunk_bytes(uint64_t bytes);

/* Largest block size of the format's valid domain. */
#define PIM_SNAPPY_MAX_BLOCK_SIZE 65536u

/* Extended-domain mode (BASELINE.json config 3 sweeps block sizes up to 1 MB).  Above 65,536 bytes
 * per block the reference's host compressor still runs, but its 16-bit hash-table entries (snappy_compress.c:139,347,
 * 392,397) and 16-bit copy offsets (:241-243) make it emit streams that NO decoder reproduces the input from -- its own
 * included.  With enable != 0 (or $PIM_SNAPPY_EXTENDED_BLOCKS=1) the COMPRESSOR accepts block sizes up to
 * PIM_SNAPPY_MAX_EXTENDED_BLOCK_SIZE and reproduces the reference's bytes there too, truncations included, so that
 * compressed sizes / ratios of the reference's sweep can be reproduced.  The decoder is untouched: it keeps rejecting
 * block sizes above 65,536 (SNAPPY_BUFFER_TOO_SMALL).  Default: off. */
#define PIM_SNAPPY_MAX_EXTENDED_BLOCK_SIZE 1048576u
void pim_snappy_set_extended_blocks(int enable);

/* ------------------------------------------------------------------------
 * Part 2 -- device-resident batched API (current device, caller's stream)
 *
 * A "block stream" is the body of the file format without the two header
 * varints:  ( u32le(compressed_size) element* )*  -- one record per block.
 * ---------------------------------------------------------------------- */

uint64_t pim_snappy_num_blocks(uint64_t input_len, uint32_t block_size);
/* capacity needed for the block stream of input_len bytes (worst case) */
uint64_t pim_snappy_compress_bound(uint64_t input_len, uint32_t block_size);
/* bytes of device scratch pim_snappy_compress_device() needs */
uint64_t pim_snappy_compress_scratch_bytes(uint64_t input_len, uint32_t block_size);

/* Compress d_in[0..input_len) (input_len < 2^32 * block_size) as independent
 * blocks.  Writes the compacted block stream to d_out (capacity >=
 * pim_snappy_compress_bound) and d_block_off[0..nblocks] (device, u64):
 * d_block_off[i] = offset in d_out of block i's 4-byte size field,
 * d_block_off[nblocks] = total bytes.  Asynchronous on `stream`
 * (a cudaStream_t).  Returns SNAPPY_OK if everything was enqueued. */
snappy_status pim_snappy_compress_device(const uint8_t *d_in, uint64_t input_len, uint32_t block_size,
                                         uint8_t *d_scratch, uint64_t scratch_bytes,
                                         uint8_t *d_out, uint64_t *d_block_off, void *stream);

/* pim_snappy_compress_device() for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs,
 * outputs and bytes as pim_snappy_compress_device(); the scratch is a device array built into the library
 * (PIM_SNAPPY_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library until exit), so the call makes no
 * cudaMalloc / cudaFree or any other allocation.  Inputs whose pim_snappy_compress_scratch_bytes() exceeds the array
 * (about 18 MiB of input at 32 KiB blocks: 42 MB of the array are hash tables) are rejected with
 * SNAPPY_BUFFER_TOO_SMALL and a message.  All calls on one GPU share the array: it keeps the last call's contents
 * (every call clears what must start at zero), and calls that OVERLAP on the GPU -- different streams, or several host
 * threads -- race; serialise them, or use pim_snappy_compress_device() with scratch of your own. */
#define PIM_SNAPPY_NOALLOC_SCRATCH_BYTES (64u << 20)
snappy_status pim_snappy_compress_device_noalloc(const uint8_t *d_in, uint64_t input_len, uint32_t block_size,
                                                 uint8_t *d_out, uint64_t *d_block_off, void *stream);

/* Decompress nblocks blocks.  d_block_off[0..nblocks] as produced above (or by
 * the host header walk): block i occupies d_comp[d_block_off[i] ..
 * d_block_off[i+1]) = 4-byte size + elements.  Block i decodes to exactly
 * min(block_size, out_len - i*block_size) bytes at d_out + i*block_size.
 * *d_status (device u32, zeroed by this call) is OR-ed with 1 for every block
 * that is malformed (SNAPPY_INVALID_INPUT).  Asynchronous on `stream`. */
snappy_status pim_snappy_decompress_device(const uint8_t *d_comp, const uint64_t *d_block_off,
                                           uint64_t nblocks, uint32_t block_size, uint64_t out_len,
                                           uint8_t *d_out, uint32_t *d_status, void *stream);

/* The same for n_streams independent streams at once (callers whose data already lives in HBM
 * hand over a batch instead of one slurped file, cf. read_input_host, dpu_snappy.c:23-50).  All arrays are HOST
 * arrays of n_streams entries holding device pointers / sizes; stream i uses exactly the arguments
 * pim_snappy_compress_device() would take for it (its own scratch, output and offset buffers) and gets exactly
 * the same bytes.  Up to 32 streams share ONE compress (decompress) kernel launch whose CTAs pull blocks of all
 * streams from one queue, so the GPU does not drain and refill between streams; larger batches take one launch
 * per 32 streams.  d_status: device u32[n_streams], one word per stream (zeroed by the call, OR-ed with 1 if a
 * block of that stream is malformed).  Asynchronous on `stream`. */
snappy_status pim_snappy_compress_batch_device(int n_streams, const uint8_t *const *d_in, const uint64_t *input_len, uint32_t block_size,
                                               uint8_t *const *d_scratch, const uint64_t *scratch_bytes, uint8_t *const *d_out,
                                               uint64_t *const *d_block_off, void *stream);
snappy_status pim_snappy_decompress_batch_device(int n_streams, const uint8_t *const *d_comp, const uint64_t *const *d_block_off,
                                                 const uint64_t *nblocks, uint32_t block_size, const uint64_t *out_len,
                                                 uint8_t *const *d_out, uint32_t *d_status, void *stream);

/* One contiguous block stream on ONE GPU from shards that were compressed on several GPUs
 * (the device-side form of the host stitch, snappy_compress.c:697-703).
 * Shard g, in block order, lives on device devices[g]: d_shard[g] = its block stream of bytes[g]
 * bytes, d_shard_off[g] = its u64 offsets [nblocks[g] + 1] relative to the shard (both as written by
 * pim_snappy_compress_device).  The shards are copied peer to peer (NVLink / NVSwitch where available)
 * to d_dst on dst_device, back to back, and d_dst_off[0 .. sum(nblocks)] receives the offsets rebased
 * to d_dst.  Asynchronous on `stream`, a stream of dst_device; the caller orders it after the shards'
 * producers (events) and passes the shard sizes it read back.  No collective library is involved. */
snappy_status pim_snappy_gather_shards(int n_shards, const int *devices, const uint8_t *const *d_shard, const uint64_t *bytes,
                                       const uint64_t *const *d_shard_off, const uint64_t *nblocks, int dst_device,
                                       uint8_t *d_dst, uint64_t *d_dst_off, void *stream);

/* Optional per-kernel timing.  When enabled, the two device entry points bracket their
 * kernels with CUDA events on the caller's stream; pim_snappy_profile_read() synchronises
 * those events and returns the accumulated milliseconds and launch counts since the last
 * read: ms[0]/n[0] compress kernel, [1] size scan, [2] compaction, [3] decompress kernel. */
void pim_snappy_profile(int enable);
int pim_snappy_profile_read(double ms[4], uint64_t n[4]);

/* Number of kernels of this library launched by the calling process so far
 * (bench.py reports the difference over the timed region as gpu_launches). */
uint64_t pim_snappy_launch_count(void);

const char *pim_snappy_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PIM_SNAPPY_H */
