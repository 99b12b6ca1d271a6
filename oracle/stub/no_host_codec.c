/* TEST INFRASTRUCTURE: lets the reference's unmodified main() (snappy/dpu_snappy.c) link against libpim_snappy.so.
 * Its non-"-d" branches call the reference's HOST codec, which the product deliberately does not contain (no CPU
 * fallback); these two symbols only satisfy the linker and fail loudly if that branch is taken. */
#include <stdio.h>
#include <stdint.h>
#include "dpu_snappy.h"

snappy_status snappy_compress_host(struct host_buffer_context *input, struct host_buffer_context *output, uint32_t block_size)
{
	(void)input; (void)output; (void)block_size;
	fprintf(stderr, "this build has no CPU codec: run with -d\n");
	return SNAPPY_INVALID_INPUT;
}

snappy_status snappy_decompress_host(struct host_buffer_context *input, struct host_buffer_context *output)
{
	(void)input; (void)output;
	fprintf(stderr, "this build has no CPU codec: run with -d\n");
	return SNAPPY_INVALID_INPUT;
}
