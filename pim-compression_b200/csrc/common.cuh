// Shared device helpers for the sm_100a Snappy kernels.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "../../include/pim_snappy.h"

namespace pimsnappy {

constexpr unsigned FULL = 0xFFFFFFFFu;

// ---- launch accounting (pim_snappy_launch_count) ---------------------------
void note_launch(unsigned n = 1);

// ---- largest block size the COMPRESSOR accepts: PIM_SNAPPY_MAX_BLOCK_SIZE, or PIM_SNAPPY_MAX_EXTENDED_BLOCK_SIZE after
// pim_snappy_set_extended_blocks(1) (the decoder never goes beyond the former)
uint32_t max_compress_block_size();

// ---- optional per-kernel event timing (pim_snappy_profile) ------------------
// which: 0 compress, 1 scan, 2 compact, 3 decompress.  No-ops unless profiling is enabled.
void prof_begin(int which, cudaStream_t st);
void prof_end(int which, cudaStream_t st);

// ---- compression with a placement hint (compress_sm100.cu) -----------------
// alone_on_gpu: no other compress kernel runs next to this one (one-piece calls): a stream with fewer blocks
// than the GPU has pair slots is spread over all SMs, fewer pairs per SM (each block then runs faster).
// The streamed host path passes false: several group kernels share the SMs there.
snappy_status compress_device_placed(const uint8_t *d_in, uint64_t input_len, uint32_t block_size, uint8_t *d_scratch, uint64_t scratch_bytes,
                                     uint8_t *d_out, uint64_t *d_block_off, cudaStream_t st, bool alone_on_gpu);
// the same for up to 32 streams in one compress launch (pim_snappy_compress_batch_device)
snappy_status compress_batch_placed(int n_streams, const uint8_t *const *d_in, const uint64_t *input_len, uint32_t block_size,
                                    uint8_t *const *d_scratch, const uint64_t *scratch_bytes, uint8_t *const *d_out,
                                    uint64_t *const *d_block_off, cudaStream_t st, bool alone_on_gpu);

// ---- shared-memory byte-addressed access -----------------------------------
// All kernels address their dynamic shared memory as one byte array `sm` and
// keep positions as byte offsets into it.

__device__ __forceinline__ uint32_t lds32(const uint8_t *sm, uint32_t aligned_off)
{
	return *reinterpret_cast<const uint32_t *>(sm + aligned_off);
}

// four bytes at an arbitrary byte offset (two aligned words + funnel shift)
__device__ __forceinline__ uint32_t lds32_unaligned(const uint8_t *sm, uint32_t off)
{
	const uint32_t a = off & ~3u;
	const uint32_t lo = lds32(sm, a);
	const uint32_t hi = lds32(sm, a + 4);
	return __funnelshift_r(lo, hi, (off & 3u) * 8u);
}

// ---- cp.async (LDGSTS) 16-byte global -> shared -----------------------------
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
	const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
	asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// Copy n bytes global -> shared where (smem_dst - gsrc) is a multiple of 16:
// byte loads for the ragged head/tail, cp.async for the aligned middle.
// Never touches a global byte outside [gsrc, gsrc+n).  Caller commits/waits.
__device__ __forceinline__ void stage_bytes_async(uint8_t *smem_dst, const uint8_t *gsrc, uint32_t n, int tid, int nthreads)
{
	const uint32_t mis = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(gsrc) & 15u);
	uint32_t head = (16u - mis) & 15u;
	if (head > n)
		head = n;
	const uint32_t mid = (n - head) & ~15u;
	const uint32_t tail = n - head - mid;
	for (uint32_t i = tid; i < head; i += nthreads)
		smem_dst[i] = __ldg(gsrc + i);
	for (uint32_t i = tid * 16u; i < mid; i += nthreads * 16u)
		cp_async16(smem_dst + head + i, gsrc + head + i);
	for (uint32_t i = tid; i < tail; i += nthreads)
		smem_dst[head + mid + i] = __ldg(gsrc + head + mid + i);
}

// ---- named barriers (producer / consumer between warps of one CTA) ----------
__device__ __forceinline__ void bar_sync(int id, int nthreads)
{
	asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int nthreads)
{
	asm volatile("bar.arrive %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

} // namespace pimsnappy

#define PIM_CUDA_TRY(expr)                                                                                  \
	do {                                                                                                    \
		cudaError_t e__ = (expr);                                                                           \
		if (e__ != cudaSuccess) {                                                                           \
			fprintf(stderr, "pim_snappy: %s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(e__), __FILE__, \
			        __LINE__);                                                                              \
			return SNAPPY_INVALID_INPUT;                                                                    \
		}                                                                                                   \
	} while (0)
