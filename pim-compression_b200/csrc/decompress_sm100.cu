// Block-parallel Snappy decompression for sm_100a.
//
// Replaces the reference's DPU decoder (snappy/dpu-decompress/dpu_decompress.c:224-299,
// dpu_task.c:21-95) and must reproduce what its host twin snappy_decompress_host
// (snappy/snappy_decompress.c:218-289) produces on every valid stream.
//
// ONE WARP PER COMPRESSED BLOCK (decode_block_warp, decompress_warp.cuh: tag chain by speculative
// segment walks, 32 elements decoded per step, bytes moved by a gather per output word), PIM_DEC_WARPS
// independent warps per CTA, persistent over the grid: every warp pulls the next block of the batch from
// a counter in global memory, so a warp that drew slow blocks does not set the tail (the reference hands
// each tasklet a fixed contiguous run, dpu-decompress/dpu_task.c:42-77).
//
// Per-warp shared memory: a ring of recent output and a ring of staged compressed bytes (DW_RB bytes
// each), the queue of tag positions and the ring of element records.
//
// Element formats: snappy/README.md:19-33, snappy_decompress.c:244-283, dpu_snappy.h:10-13.

#include <cstring>
#include <mutex>

#include "common.cuh"
#include "decompress_warp.cuh"

namespace pimsnappy {
namespace {

#ifndef PIM_DEC_WARPS
#define PIM_DEC_WARPS 4
#endif
#ifndef PIM_DEC_MINB
#define PIM_DEC_MINB 5
#endif
constexpr int WARPS = PIM_DEC_WARPS; // independent warps (= blocks in flight) per CTA
constexpr uint32_t WS_BYTES = 2 * DW_RB + DW_SMALL_BYTES;
constexpr uint32_t SMEM_BYTES = WARPS * WS_BYTES + 256;

// ---------------------------------------------------------------------------
// Kernel: one launch decodes up to DEC_MAX_BATCH streams (pim_snappy_decompress_batch_device); the counter
// runs over the concatenation of the streams' blocks.
// ---------------------------------------------------------------------------
constexpr uint32_t DEC_MAX_BATCH = 32;
struct DecStream {
	const uint8_t *comp;
	const uint64_t *block_off;
	uint8_t *out;
	uint64_t out_len;
	uint64_t first_blk; // index of the stream's first block in the batch
};
struct DecArgs {
	DecStream s[DEC_MAX_BATCH];
	uint32_t n_streams;
	uint32_t block_size;
	uint64_t nblocks;   // of the whole batch
	uint32_t *status;   // one word per stream
	unsigned long long *counter; // next block of the batch (zero at launch)
};

__global__ void __launch_bounds__(32 * WARPS, PIM_DEC_MINB) snappy_decompress_kernel(const __grid_constant__ DecArgs a)
{
	extern __shared__ __align__(1024) uint8_t sm[];
	const int lane = lane_id();
	const int warp = threadIdx.x >> 5;

	DSm s;
	s.sm = sm;
	s.ring = warp * 2 * DW_RB;
	s.stg = s.ring + DW_RB;
	s.q = WARPS * 2 * DW_RB + warp * DW_SMALL_BYTES;
	s.rec = s.q + 2 * DW_QCAP;
	s.lut = WARPS * WS_BYTES;
	for (uint32_t t = threadIdx.x; t < 256; t += 32 * WARPS)
		sm[s.lut + t] = (uint8_t)dw_lut_foot(t);
	__syncthreads();

	for (;;) {
		unsigned long long gblk = 0;
		if (lane == 0)
			gblk = atomicAdd(a.counter, 1ull);
		gblk = __shfl_sync(FULL, gblk, 0);
		if (gblk >= a.nblocks)
			break;
		uint32_t si = 0;
		while (si + 1 < a.n_streams && gblk >= a.s[si + 1].first_blk)
			si++;
		const DecStream &ds = a.s[si];
		const uint64_t blk = gblk - ds.first_blk;
		const uint64_t off0 = ds.block_off[blk], off1 = ds.block_off[blk + 1];
		const uint64_t out_off = blk * (uint64_t)a.block_size;
		DBlk b;
		b.out_size = (uint32_t)min((uint64_t)a.block_size, ds.out_len - out_off);
		b.gdst = ds.out + out_off;
		b.gcomp = ds.comp + off0 + 4;
		b.csize = 0;
		bool ok = off1 >= off0 + 5; // 4-byte size + at least one element byte
		if (ok) {
			const uint8_t *h = ds.comp + off0;
			b.csize = (uint32_t)__ldg(h) | ((uint32_t)__ldg(h + 1) << 8) | ((uint32_t)__ldg(h + 2) << 16) | ((uint32_t)__ldg(h + 3) << 24);
			ok = (uint64_t)b.csize + 4 == off1 - off0;
		}
		if (ok)
			ok = decode_block_warp(s, b);
		if (!ok && lane == 0)
			atomicOr(a.status + si, 1u);
		__syncwarp();
	}
}

// Block counters: one 8-byte slot per launch, taken round robin from a per-device pool (a slot is zeroed
// on the launch's stream right before the kernel; POOL launches would have to be in flight for a clash).
constexpr int POOL = 4096;
struct CounterPool {
	unsigned long long *d = nullptr;
	unsigned next = 0;
	int max_ctas = 0;
};
std::mutex g_pool_mu;
CounterPool g_pool[64];

} // namespace
} // namespace pimsnappy

using namespace pimsnappy;

extern "C" snappy_status pim_snappy_decompress_batch_device(int n_streams, const uint8_t *const *d_comp, const uint64_t *const *d_block_off,
                                                            const uint64_t *nblocks, uint32_t block_size, const uint64_t *out_len,
                                                            uint8_t *const *d_out, uint32_t *d_status, void *stream)
{
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	if (block_size == 0 || block_size > PIM_SNAPPY_MAX_BLOCK_SIZE)
		return SNAPPY_BUFFER_TOO_SMALL;
	if (n_streams < 0)
		return SNAPPY_INVALID_INPUT;
	for (int i = 0; i < n_streams; i++)
		if (nblocks[i] != pim_snappy_num_blocks(out_len[i], block_size) || nblocks[i] > 0x7FFFFFFFull)
			return SNAPPY_INVALID_INPUT;
	if (n_streams == 0)
		return SNAPPY_OK;
	PIM_CUDA_TRY(cudaMemsetAsync(d_status, 0, sizeof(uint32_t) * (size_t)n_streams, st));
	int dev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&dev));
	if (dev >= 64)
		return SNAPPY_INVALID_INPUT;
	unsigned long long *counters = nullptr;
	int max_ctas = 0;
	{
		std::lock_guard<std::mutex> lk(g_pool_mu);
		CounterPool &cp = g_pool[dev];
		if (!cp.d) {
			PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
			PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
			int per_sm = 0, sms = 0;
			PIM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, snappy_decompress_kernel, 32 * WARPS, SMEM_BYTES));
			PIM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
			cp.max_ctas = per_sm * sms > 0 ? per_sm * sms : 1;
			PIM_CUDA_TRY(cudaMalloc(&cp.d, sizeof(unsigned long long) * POOL));
		}
		const int launches = (n_streams + (int)DEC_MAX_BATCH - 1) / (int)DEC_MAX_BATCH;
		if (cp.next + launches > (unsigned)POOL)
			cp.next = 0;
		counters = cp.d + cp.next;
		cp.next += launches;
		max_ctas = cp.max_ctas;
	}
	int launch_i = 0;
	for (int at = 0; at < n_streams; at += (int)DEC_MAX_BATCH) { // one launch per DEC_MAX_BATCH streams
		DecArgs da;
		memset(&da, 0, sizeof(da));
		da.block_size = block_size;
		da.status = d_status + at;
		uint64_t total = 0;
		const int k = n_streams - at < (int)DEC_MAX_BATCH ? n_streams - at : (int)DEC_MAX_BATCH;
		for (int i = 0; i < k; i++) { // (streams without blocks keep their slot: status words are indexed by stream)
			DecStream &ds = da.s[da.n_streams++];
			ds.comp = d_comp[at + i];
			ds.block_off = d_block_off[at + i];
			ds.out = d_out[at + i];
			ds.out_len = out_len[at + i];
			ds.first_blk = total;
			total += nblocks[at + i];
		}
		if (total == 0)
			continue;
		da.nblocks = total;
		da.counter = counters + launch_i++;
		PIM_CUDA_TRY(cudaMemsetAsync(da.counter, 0, sizeof(unsigned long long), st));
		const uint64_t want = (total + WARPS - 1) / WARPS;
		const unsigned grid = (unsigned)(want < (uint64_t)max_ctas ? want : (uint64_t)max_ctas);
		prof_begin(3, st);
		snappy_decompress_kernel<<<grid, 32 * WARPS, SMEM_BYTES, st>>>(da);
		prof_end(3, st);
		note_launch();
	}
	PIM_CUDA_TRY(cudaGetLastError());
	return SNAPPY_OK;
}

extern "C" snappy_status pim_snappy_decompress_device(const uint8_t *d_comp, const uint64_t *d_block_off, uint64_t nblocks,
                                                      uint32_t block_size, uint64_t out_len, uint8_t *d_out, uint32_t *d_status,
                                                      void *stream)
{
	return pim_snappy_decompress_batch_device(1, &d_comp, &d_block_off, &nblocks, block_size, &out_len, &d_out, d_status, stream);
}
