/*
 * dpu_snappy -- command-line front end, GPU build.
 *
 * Same options, defaults, exit codes and stdout lines as the reference CLI
 * (snappy/dpu_snappy.c:12,70-82,93-237) so that the reference's scripts, which parse this
 * output (snappy/scripts/asplos21/parse_output_file.py), keep working:
 *
 *   dpu_snappy [-d] [-c] [-b <block_size>] -i <input_file> [-o <output_file>]
 *
 * Differences, all deliberate:
 *   - there is no host codec in this build: the device (now the B200 GPUs) is always used;
 *     -d is accepted and ignored.  (The CPU oracle lives in oracle/ as test infrastructure.)
 *   - the number of devices is a run-time choice (-g N or $PIM_SNAPPY_NR_GPUS) instead of
 *     the compile-time NR_DPUS (snappy/Makefile:10-12).
 *   - the compressed file is written here from output.buffer; the reference's DPU path
 *     fwrite()s it inside snappy_compress_dpu (snappy_compress.c:626-627,697-703).
 *   - "Host time" prints the kernel time (the reference leaves it uninitialised with -d).
 */
#include <getopt.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pim_snappy.h"

static const char opts[] = "dcb:i:o:g:";

static void usage(const char *exe)
{
	fprintf(stderr, "Compress or decompress a file with Snappy (multi-block format) on NVIDIA B200 GPUs\n");
	fprintf(stderr, "usage: %s [-d] [-c] [-b <block_size>] [-g <gpus>] -i <input_file> [-o <output_file>]\n", exe);
	fprintf(stderr, "d: use the device (always on in this build; accepted for compatibility)\n");
	fprintf(stderr, "c: perform compression, by default performs decompression\n");
	fprintf(stderr, "b: block size used for compression, default is 32KB, ignored for decompression\n");
	fprintf(stderr, "g: number of GPUs to shard over, default all visible\n");
	fprintf(stderr, "i: input file\n");
	fprintf(stderr, "o: output file\n");
}

/* whole file into page-locked memory (what read_input_host, dpu_snappy.c:23-50, does with malloc) */
static int slurp(const char *path, struct host_buffer_context *in)
{
	FILE *f = fopen(path, "rb");
	if (!f) {
		fprintf(stderr, "Invalid input file: %s\n", path);
		return 1;
	}
	fseek(f, 0, SEEK_END);
	long sz = ftell(f);
	fseek(f, 0, SEEK_SET);
	if (sz < 0 || (unsigned long)sz > in->max) {
		fprintf(stderr, "input_size is too big (%ld > %ld)\n", sz, in->max);
		fclose(f);
		return 1;
	}
	in->length = (unsigned long)sz;
	/* page-locked memory pays for files the library streams (copies beside the kernels); locking a small buffer costs more than its copy */
	in->buffer = sz > (64l << 20) ? pim_snappy_host_alloc(((size_t)sz + 7) & ~(size_t)7) : malloc((((size_t)sz + 7) & ~(size_t)7) + 8);
	if (!in->buffer) {
		fprintf(stderr, "cannot allocate the input buffer (no CUDA device?)\n");
		fclose(f);
		return 1;
	}
	in->curr = in->buffer;
	size_t got = fread(in->buffer, 1, in->length, f);
	fclose(f);
	return got != in->length;
}

int main(int argc, char **argv)
{
	int opt, compress = 0, ngpu = 0;
	int block_size = 32 * 1024;
	char *in_name = NULL, *out_name = NULL;
	struct host_buffer_context input = {0}, output = {0};
	input.max = ULONG_MAX;
	output.max = ULONG_MAX;

	while ((opt = getopt(argc, argv, opts)) != -1) {
		switch (opt) {
		case 'd': break;
		case 'c': compress = 1; break;
		case 'b': block_size = atoi(optarg); break;
		case 'g': ngpu = atoi(optarg); break;
		case 'i': in_name = optarg; break;
		case 'o': out_name = optarg; break;
		default: usage(argv[0]); return -2;
		}
	}
	if (!in_name) {
		usage(argv[0]);
		return -1;
	}
	input.file_name = in_name;
	printf("Using input file %s\n", in_name);
	if (!out_name)
		out_name = "output.txt";
	output.file_name = out_name;
	printf("Using output file %s\n", out_name);

	if (ngpu > 0) {
		int ids[64];
		for (int i = 0; i < ngpu && i < 64; i++)
			ids[i] = i;
		pim_snappy_set_devices(ids, ngpu < 64 ? ngpu : 64);
	}
	if (pim_snappy_device_count() <= 0) {
		fprintf(stderr, "No usable sm_100 GPU found (this build has no CPU codec)\n");
		return -1;
	}
	if (slurp(in_name, &input))
		return -1;

	struct program_runtime rt;
	memset(&rt, 0, sizeof(rt));
	snappy_status st;
	if (compress) {
		setup_compression(&input, &output, &rt);
		st = snappy_compress_dpu(&input, &output, (uint32_t)block_size, &rt);
	} else {
		if (setup_decompression(&input, &output, &rt))
			return -1;
		st = snappy_decompress_dpu(&input, &output, &rt);
	}
	if (st != SNAPPY_OK) {
		fprintf(stderr, "Encountered Snappy error %u\n", st);
		return -1;
	}

	FILE *fo = fopen(out_name, "wb");
	if (!fo) {
		fprintf(stderr, "Cannot open output file: %s\n", out_name);
		return -1;
	}
	fwrite(output.buffer, 1, output.length, fo);
	fclose(fo);

	if (compress) {
		printf("Compressed %ld bytes to: %s\n", output.length, out_name);
		printf("Compression ratio: %f\n", 1 - (double)output.length / (double)input.length);
	} else {
		printf("Decompressed %ld bytes to: %s\n", output.length, out_name);
		printf("Compression ratio: %f\n", 1 - (double)input.length / (double)output.length);
	}
	printf("Pre-processing time: %f\n", rt.pre);
	printf("Alloc time: %f\n", rt.d_alloc);
	printf("Load time: %f\n", rt.load);
	printf("Copy in time: %f\n", rt.copy_in);
	printf("Host time: %f\n", rt.run);
	printf("Copy out time: %f\n", rt.copy_out);
	printf("Free time: %f\n", rt.d_free);
	return 0;
}
