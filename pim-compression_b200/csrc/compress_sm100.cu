// Block-parallel Snappy compression for sm_100a.
//
// Replaces the reference's DPU compressor (snappy/dpu-compress/dpu_compress.c:338-504,
// dpu_task.c:22-94) but reproduces, bit for bit, the output of the reference's HOST compressor
// snappy_compress_host / compress_block (snappy/snappy_compress.c:284-413,455-485): same
// multiplicative hash (:161-166), same per-block table sizing (:139-146), same skip heuristic
// (:333-348), same order of table updates (:391-397), same element encodings (:202-272).
//
// ONE persistent CTA per SM made of (parser, emitter) warp PAIRS; every pair compresses one input
// block at a time and pulls the next from a global counter.  A block's parse is a serial dependency
// chain that leaves the SM mostly idle, so throughput is the number of blocks in flight per SM, and
// that is bounded by where the 16-bit hash tables live (32 KB for blocks of 16 KB and more):
//   * as many pairs as shared memory holds tables for (seven at 32 KB -- one CTA owning all of the
//     SM's shared memory fits one more table than six independent CTAs did) keep theirs there;
//   * as many pairs again keep their table in global memory (L2-resident, ld/st.cg): 0.6x the speed of
//     a shared-memory pair, and more than one of them per shared-memory pair saturates L2.
// The input is NOT staged: the parser prefetches it into L2 a few KB ahead of itself and reads it from there through L1.  The parser warp
// (compress_warp.cuh) fuses the reference's step 1 / step 3 probes into 32-lane passes that resolve
// several (literal, copy) pairs per shared-memory round trip, the emitter warp turns its per-pass
// records into bytes.
// Each block writes to a fixed worst-case slot; block sizes then go through a device
// exclusive scan, and a compaction kernel moves every block to its final offset, yielding
// the contiguous ( u32 size | elements )* stream of the file format.

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include <mutex>

#include "common.cuh"
#ifdef PIM_CW_PROFILE
namespace pimsnappy { __device__ unsigned long long g_cw_prof[16]; }
#endif
#include "compress_warp.cuh"

namespace pimsnappy {
namespace {

__host__ __device__ inline uint32_t table_entries(uint32_t n) { return cw_table_entries(n); }

__host__ __device__ inline uint32_t slot_stride(uint32_t block_size)
{
	return (4u + 32u + block_size + block_size / 6u + 15u) & ~15u;
}

// Persistent CTAs (one per SM): at most this many, each with at most MAX_GT pairs whose table is in
// global memory (the tables in scratch are sized by these two).
constexpr uint32_t MAX_PERSIST_CTAS = 160;
constexpr uint32_t MAX_GT = 8;
constexpr uint32_t MAX_PAIRS = 15;   // named barriers 1..15, one per pair
constexpr uint32_t PAIR_CTL_BYTES = (uint32_t)sizeof(EmitRing); // record ring + mail word

// Pairs of one CTA for this block size: n_sm with the table in shared memory, n_gt in global memory.
inline void plan_pairs(uint32_t block_size, uint32_t smem_budget, uint32_t *n_sm, uint32_t *n_gt)
{
	const uint32_t tbytes = 2u * table_entries(block_size);
	uint32_t ns = (smem_budget - 64u) / (tbytes + PAIR_CTL_BYTES);
	ns = std::max(1u, std::min(ns, MAX_PAIRS));
	const uint32_t ctl_room = (smem_budget - 64u - ns * tbytes) / PAIR_CTL_BYTES; // pairs the rest has control blocks for
	uint32_t ng = std::min(std::min(MAX_PAIRS - ns, MAX_GT), std::min(ns + 1u, ctl_room > ns ? ctl_room - ns : 0u));
	static const int env_nsm = [] { const char *e = getenv("PIM_SNAPPY_NSM"); return e ? atoi(e) : -1; }(); // development aids, read once
	static const int env_ngt = [] { const char *e = getenv("PIM_SNAPPY_NGT"); return e ? atoi(e) : -1; }();
	if (env_nsm >= 0)
		ns = std::max(1u, std::min(ns, (uint32_t)env_nsm));
	if (env_ngt >= 0)
		ng = std::min(ng, (uint32_t)env_ngt);
	*n_sm = ns;
	*n_gt = ng;
}

// The parse reads up to TAIL_SLACK bytes past a block's end.  Inside the stream that is the next
// block; the stream's last block is read from a copy in scratch that has the slack behind it.
constexpr uint32_t TAIL_SLACK = 32;
__host__ __device__ inline uint64_t tail_copy_bytes(uint32_t block_size) { return ((uint64_t)block_size + TAIL_SLACK + 16u + 255u) & ~255ull; }
__host__ __device__ inline uint64_t gtable_bytes(uint32_t block_size) { return (uint64_t)MAX_PERSIST_CTAS * MAX_GT * 2u * table_entries(block_size); }

#ifndef PIM_PAIR_INLINE
#define PIM_PAIR_INLINE __forceinline__
#endif

// One launch compresses up to MAX_BATCH streams (pim_snappy_compress_batch_device): the pairs pull block indices of the
// concatenation of the streams' blocks from one counter, so no SM idles between streams.
constexpr uint32_t MAX_BATCH = 32;
struct StreamArgs {
	const uint8_t *in;
	uint64_t in_len;
	const uint8_t *tail_copy; // the stream's last block with readable slack behind it
	uint8_t *slots;
	uint32_t *sizes;
	uint32_t first_blk;       // index of the stream's first block in the batch
	uint32_t nblocks;
};
struct CompressArgs {
	StreamArgs s[MAX_BATCH];
	uint32_t n_streams;
	uint32_t block_size, nblocks; // nblocks: of the whole batch
	uint32_t stride;
	uint32_t *counter;
};

// One (parser, emitter) warp pair: pulls blocks from the global counter until none are left.
// GT: the pair's hash table is in global memory (tbytes), else in shared memory.
// bar_id: the pair's named barrier (run-time ids make ptxas reserve all 16 barriers: fine for one CTA per SM).
template <bool GT>
__device__ PIM_PAIR_INLINE void compress_pair(const CompressArgs &a, uint8_t *tbytes, EmitRing *ring, volatile uint32_t *mail, int bar_id, uint32_t role)
{
	const int lane = lane_id();
	const bool parser = role == 0; // else emitter
	const int ptid = (int)role * 32 + lane;
	uint32_t seq = 0; // records this warp has published (parser) / consumed (emitter) so far
	if (ptid == 0)
		ring_init(ring);
	for (;;) {
		if (ptid == 0)
			*mail = atomicAdd(a.counter, 1u);
		bar_sync(bar_id, 64);
		const uint32_t gblk = *mail;
		if (gblk >= a.nblocks)
			break;
		uint32_t si = 0; // the stream this block belongs to (streams are in block order)
		while (si + 1 < a.n_streams && gblk >= a.s[si + 1].first_blk)
			si++;
		const StreamArgs &sa = a.s[si];
		const uint32_t blk = gblk - sa.first_blk;
		const uint64_t in_off = blk * (uint64_t)a.block_size;
		const uint32_t n = (uint32_t)min((uint64_t)a.block_size, sa.in_len - in_off);
		const uint8_t *gsrc = blk + 1 == sa.nblocks ? sa.tail_copy : sa.in + in_off;
		const uint32_t ts = table_entries(n);

		CBlock c;
		c.mis = (uint32_t)(reinterpret_cast<uintptr_t>(gsrc) & 3u);
		c.inw = reinterpret_cast<const uint32_t *>(gsrc - c.mis);
		c.table = reinterpret_cast<uint16_t *>(tbytes);
		c.shift = 32 - (31 - __clz(ts)); // snappy_compress.c:288
		c.out = sa.slots + blk * (uint64_t)a.stride;
		c.n = n;

		if (!PIM_PF_AHEAD)
			for (uint32_t i = ptid * 128u; i < n; i += 64u * 128u) // HBM -> L2 for the whole block, ahead of the parse
				asm volatile("prefetch.global.L2 [%0];" ::"l"(gsrc + i));
		for (uint32_t i = ptid * 16u; i < 2u * ts; i += 1024u) // get_hash_table, snappy_compress.c:139-146
			*reinterpret_cast<uint4 *>(tbytes + i) = make_uint4(0, 0, 0, 0);
		bar_sync(bar_id, 64);

		if (parser) { // the block's serial chain
			parse_block_warp<GT>(c, ring, seq);
		} else { // output bytes only
			const uint32_t o = emit_block_warp(c, ring, seq);
			if (lane == 0)
				sa.sizes[blk] = o;
		}
		bar_sync(bar_id, 64); // table, ring and mail are reused
	}
}

// The persistent CTA: pairs [0, n_sm) keep their table in shared memory, pairs [n_sm, n_sm + n_gt) in gtables.
__global__ void __launch_bounds__(64 * (MAX_PAIRS + 1), 1) snappy_compress_kernel(const __grid_constant__ CompressArgs a, uint8_t *__restrict__ gtables, uint32_t n_sm, uint32_t n_gt)
{
	extern __shared__ __align__(16) uint8_t sm[];
	// Warp w runs on scheduler w % 4.  The parsers are the busy warps (an emitter mostly waits), so the
	// roles are laid out in groups of four warps = two pairs: (parser, parser, emitter, emitter) in even
	// groups and (emitter, emitter, parser, parser) in odd ones -- every scheduler gets its share of parsers.
	const uint32_t w = threadIdx.x >> 5, g = w >> 2, k = w & 3, npairs = n_sm + n_gt;
	const uint32_t kk = (g & 1) ? (k ^ 2u) : k;
	const uint32_t pair = 2 * g + (kk & 1u), role = kk >> 1; // role 0: parser, 1: emitter
	if (pair >= npairs) // the CTA is launched with an even number of pair slots
		return;
	const uint32_t tbytes = 2u * table_entries(a.block_size);
	uint8_t *ctl = sm + n_sm * tbytes;
	EmitRing *ring = reinterpret_cast<EmitRing *>(ctl) + pair;
	volatile uint32_t *mail = reinterpret_cast<volatile uint32_t *>(&ring->rec[0].pad[0]); // (never written by the ring protocol)
	if (pair < n_sm)
		compress_pair<false>(a, sm + pair * tbytes, ring, mail, (int)pair + 1, role);
	else
		compress_pair<true>(a, gtables + ((uint64_t)blockIdx.x * n_gt + (pair - n_sm)) * tbytes, ring, mail, (int)pair + 1, role);
}

// Exclusive scan of the per-block sizes into 64-bit offsets (one CTA; nblocks is at most a few
// million).  off[i] = sum_{j<i} size[j], off[n] = total.
__global__ void __launch_bounds__(1024) scan_sizes_kernel(const uint32_t *__restrict__ sizes, uint64_t n, uint64_t *__restrict__ off)
{
	__shared__ uint64_t part[1024];
	const uint64_t per = (n + 1023) / 1024;
	const uint64_t lo = min(n, per * threadIdx.x), hi = min(n, lo + per);
	uint64_t s = 0;
	for (uint64_t i = lo; i < hi; i++)
		s += sizes[i];
	part[threadIdx.x] = s;
	__syncthreads();
	for (int d = 1; d < 1024; d <<= 1) {
		uint64_t v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
		__syncthreads();
		part[threadIdx.x] += v;
		__syncthreads();
	}
	uint64_t run = part[threadIdx.x] - s;
	for (uint64_t i = lo; i < hi; i++) {
		off[i] = run;
		run += sizes[i];
	}
	if (threadIdx.x == 1023)
		off[n] = part[1023];
}

// Move every block from its slot (16-byte aligned) to its final, arbitrarily aligned offset.
__global__ void __launch_bounds__(128) compact_kernel(const uint8_t *__restrict__ slots, uint32_t stride, const uint32_t *__restrict__ sizes,
                                                      const uint64_t *__restrict__ off, uint8_t *__restrict__ out)
{
	const uint64_t blk = blockIdx.x;
	const uint8_t *src = slots + blk * (uint64_t)stride;
	uint8_t *dst = out + off[blk];
	const uint32_t n = sizes[blk];
	const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 3u);
	uint32_t head = (4u - mis) & 3u;
	if (head > n)
		head = n;
	for (uint32_t i = threadIdx.x; i < head; i += 128)
		dst[i] = src[i];
	// destination words are aligned; the source is read as two aligned words and funnel-shifted
	const uint32_t words = (n - head) / 4;
	const uint32_t sh = (head & 3u) * 8u; // src + head is misaligned by head (src is 16-byte aligned)
	const uint32_t *s32 = reinterpret_cast<const uint32_t *>(src + (head & ~3u));
	uint32_t *d32 = reinterpret_cast<uint32_t *>(dst + head);
	for (uint32_t i = threadIdx.x; i < words; i += 128) {
		const uint32_t lo = __ldg(s32 + i);
		const uint32_t hi = sh ? __ldg(s32 + i + 1) : 0u;
		d32[i] = __funnelshift_r(lo, hi, sh);
	}
	for (uint32_t i = head + words * 4 + threadIdx.x; i < n; i += 128)
		dst[i] = src[i];
}

} // namespace
} // namespace pimsnappy

#ifdef PIM_CW_PROFILE // development aid: see compress_warp.cuh
extern "C" int pim_snappy_debug_cw_prof(unsigned long long out[16], int reset)
{
	if (cudaMemcpyFromSymbol(out, pimsnappy::g_cw_prof, sizeof(unsigned long long) * 16) != cudaSuccess)
		return 1;
	if (reset) {
		unsigned long long z[16] = {0};
		if (cudaMemcpyToSymbol(pimsnappy::g_cw_prof, z, sizeof(z)) != cudaSuccess)
			return 1;
	}
	return 0;
}
#endif

using namespace pimsnappy;

extern "C" uint64_t pim_snappy_num_blocks(uint64_t input_len, uint32_t block_size)
{
	return block_size ? (input_len + block_size - 1) / block_size : 0;
}

extern "C" uint64_t pim_snappy_compress_bound(uint64_t input_len, uint32_t block_size)
{
	return pim_snappy_num_blocks(input_len, block_size) * 40 + input_len + input_len / 6 + 64;
}

extern "C" uint64_t pim_snappy_compress_scratch_bytes(uint64_t input_len, uint32_t block_size)
{
	const uint64_t nb = pim_snappy_num_blocks(input_len, block_size);
	return ((nb * 4 + 255) & ~255ull) + 256 + tail_copy_bytes(block_size) + gtable_bytes(block_size) + nb * (uint64_t)slot_stride(block_size) + 256;
}

extern "C" snappy_status pim_snappy_compress_device(const uint8_t *d_in, uint64_t input_len, uint32_t block_size, uint8_t *d_scratch,
                                                    uint64_t scratch_bytes, uint8_t *d_out, uint64_t *d_block_off, void *stream)
{
	return compress_device_placed(d_in, input_len, block_size, d_scratch, scratch_bytes, d_out, d_block_off, static_cast<cudaStream_t>(stream), true);
}

snappy_status pimsnappy::compress_device_placed(const uint8_t *d_in, uint64_t input_len, uint32_t block_size, uint8_t *d_scratch,
                                                uint64_t scratch_bytes, uint8_t *d_out, uint64_t *d_block_off, cudaStream_t st, bool alone_on_gpu)
{
	return compress_batch_placed(1, &d_in, &input_len, block_size, &d_scratch, &scratch_bytes, &d_out, &d_block_off, st, alone_on_gpu);
}

extern "C" snappy_status pim_snappy_compress_batch_device(int n_streams, const uint8_t *const *d_in, const uint64_t *input_len, uint32_t block_size,
                                                          uint8_t *const *d_scratch, const uint64_t *scratch_bytes, uint8_t *const *d_out,
                                                          uint64_t *const *d_block_off, void *stream)
{
	if (n_streams < 0)
		return SNAPPY_INVALID_INPUT;
	for (int at = 0; at < n_streams; at += (int)MAX_BATCH) { // one launch per MAX_BATCH streams
		const int k = std::min<int>((int)MAX_BATCH, n_streams - at);
		const snappy_status r = compress_batch_placed(k, d_in + at, input_len + at, block_size, d_scratch + at, scratch_bytes + at, d_out + at,
		                                              d_block_off + at, static_cast<cudaStream_t>(stream), true);
		if (r != SNAPPY_OK)
			return r;
	}
	return SNAPPY_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// pim_snappy_compress_device_noalloc: pim_snappy_compress_device() with its scratch in a file-scope device array, for
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No
// counterpart in the reference.
//
// Every GPU that loads the library holds the array (PIM_SNAPPY_NOALLOC_SCRATCH_BYTES) until exit.  In host code the
// array's name is only a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
// device with that device current.  All calls on a device share the array: calls that overlap on the GPU race.  What
// must start at zero is cleared by every call (the block counter: cudaMemsetAsync in compress_batch_placed; the hash
// tables: zeroed per block by the kernel, get_hash_table snappy_compress.c:139-146); sizes and slots are written before
// they are read.  Calls no cudaMalloc / cudaFree / allocating routine.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) uint8_t g_noalloc_scratch[PIM_SNAPPY_NOALLOC_SCRATCH_BYTES];

extern "C" snappy_status pim_snappy_compress_device_noalloc(const uint8_t *d_in, uint64_t input_len, uint32_t block_size, uint8_t *d_out,
                                                            uint64_t *d_block_off, void *stream)
{
	if (block_size == 0 || block_size > max_compress_block_size())
		return SNAPPY_BUFFER_TOO_SMALL;
	const uint64_t need = pim_snappy_compress_scratch_bytes(input_len, block_size);
	if (need > (uint64_t)PIM_SNAPPY_NOALLOC_SCRATCH_BYTES) {
		fprintf(stderr, "pim_snappy_compress_device_noalloc: %llu input bytes at block size %u need %llu bytes of scratch, the built-in array holds %llu; "
		                "use pim_snappy_compress_device with scratch of your own\n",
		        (unsigned long long)input_len, block_size, (unsigned long long)need, (unsigned long long)PIM_SNAPPY_NOALLOC_SCRATCH_BYTES);
		return SNAPPY_BUFFER_TOO_SMALL;
	}
	static std::mutex mu;
	static uint8_t *scratch_addr[64] = {nullptr}; // per device: the array's device address
	int dev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&dev));
	if (dev < 0 || dev >= 64)
		return SNAPPY_INVALID_INPUT;
	uint8_t *scratch = nullptr;
	{
		std::lock_guard<std::mutex> lk(mu);
		if (!scratch_addr[dev]) {
			void *p = nullptr;
			PIM_CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch)); // (the current device's copy of the array)
			scratch_addr[dev] = static_cast<uint8_t *>(p);
		}
		scratch = scratch_addr[dev];
	}
	return compress_device_placed(d_in, input_len, block_size, scratch, PIM_SNAPPY_NOALLOC_SCRATCH_BYTES, d_out, d_block_off,
	                              static_cast<cudaStream_t>(stream), true);
}

// Streams [0, n_streams) (n_streams <= MAX_BATCH) in ONE compress launch, then one size scan + compaction per stream.
snappy_status pimsnappy::compress_batch_placed(int n_streams, const uint8_t *const *d_in, const uint64_t *input_len, uint32_t block_size,
                                               uint8_t *const *d_scratch, const uint64_t *scratch_bytes, uint8_t *const *d_out,
                                               uint64_t *const *d_block_off, cudaStream_t st, bool alone_on_gpu)
{
	if (block_size == 0 || block_size > max_compress_block_size())
		return SNAPPY_BUFFER_TOO_SMALL;
	if (n_streams < 0 || n_streams > (int)MAX_BATCH)
		return SNAPPY_INVALID_INPUT;
	const uint32_t stride = slot_stride(block_size);
	CompressArgs ca;
	memset(&ca, 0, sizeof(ca));
	ca.block_size = block_size;
	ca.stride = stride;
	uint8_t *gtables = nullptr;
	uint64_t total = 0;
	int idx[MAX_BATCH]; // streams with blocks, in order
	for (int i = 0; i < n_streams; i++) {
		const uint64_t nb = pim_snappy_num_blocks(input_len[i], block_size);
		if (nb > 0x7FFFFFFFull || total + nb > 0x7FFFFFFFull || scratch_bytes[i] < pim_snappy_compress_scratch_bytes(input_len[i], block_size))
			return SNAPPY_BUFFER_TOO_SMALL;
		if (nb == 0) {
			PIM_CUDA_TRY(cudaMemsetAsync(d_block_off[i], 0, sizeof(uint64_t), st));
			continue;
		}
		uint32_t *sizes = reinterpret_cast<uint32_t *>(d_scratch[i]);
		uint32_t *counter = reinterpret_cast<uint32_t *>(d_scratch[i] + ((nb * 4 + 255) & ~255ull));
		uint8_t *tail_copy = reinterpret_cast<uint8_t *>(counter) + 256;
		uint8_t *gt = tail_copy + tail_copy_bytes(block_size);
		uint8_t *slots = gt + gtable_bytes(block_size);
		if (!gtables) { // the batch's counter and global-memory hash tables are those of its first stream
			gtables = gt;
			ca.counter = counter;
			PIM_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
		}
		// the stream's last block, with readable slack behind it (same 4-byte phase as in the stream)
		const uint64_t last_off = (nb - 1) * (uint64_t)block_size;
		tail_copy += reinterpret_cast<uintptr_t>(d_in[i] + last_off) & 15u;
		PIM_CUDA_TRY(cudaMemcpyAsync(tail_copy, d_in[i] + last_off, input_len[i] - last_off, cudaMemcpyDeviceToDevice, st));
		StreamArgs &sa = ca.s[ca.n_streams];
		sa.in = d_in[i];
		sa.in_len = input_len[i];
		sa.tail_copy = tail_copy;
		sa.slots = slots;
		sa.sizes = sizes;
		sa.first_blk = (uint32_t)total;
		sa.nblocks = (uint32_t)nb;
		idx[ca.n_streams++] = i;
		total += nb;
	}
	if (total == 0)
		return SNAPPY_OK;
	ca.nblocks = (uint32_t)total;
	const uint64_t nb = total;

	static int smem_optin[64] = {0}; // per device: opt-in shared memory per CTA
	static int sm_count[64] = {0};
	int dev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&dev));
	if (dev >= 64)
		return SNAPPY_INVALID_INPUT;
	if (!smem_optin[dev]) {
		int optin = 0;
		PIM_CUDA_TRY(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_compress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_compress_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
		PIM_CUDA_TRY(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
		smem_optin[dev] = optin;
	}
	uint32_t n_sm = 0, n_gt = 0;
	plan_pairs(block_size, (uint32_t)smem_optin[dev], &n_sm, &n_gt);
	const uint64_t ctas_max = std::min<uint64_t>(MAX_PERSIST_CTAS, (uint64_t)sm_count[dev]);
	static const bool no_spread = getenv("PIM_SNAPPY_NO_SPREAD") != nullptr; // development aid
	if (alone_on_gpu && !no_spread && nb < ctas_max * (n_sm + n_gt)) { // fewer blocks than pair slots: all SMs, fewer pairs each, shared-memory pairs first
		const uint32_t need = (uint32_t)((nb + ctas_max - 1) / ctas_max);
		n_gt = need > n_sm ? std::min(n_gt, need - n_sm) : 0u;
		n_sm = std::min(n_sm, need);
	}
	const uint32_t npairs = n_sm + n_gt;
	const uint32_t smem = n_sm * 2u * table_entries(block_size) + npairs * PAIR_CTL_BYTES + 16u;
	const unsigned grid = (unsigned)std::min<uint64_t>(ctas_max, (nb + npairs - 1) / npairs);
	prof_begin(0, st);
	snappy_compress_kernel<<<grid, 64 * ((npairs + 1u) & ~1u), smem, st>>>(ca, gtables, n_sm, n_gt);
	prof_end(0, st);
	for (uint32_t k = 0; k < ca.n_streams; k++) {
		const StreamArgs &sa = ca.s[k];
		prof_begin(1, st);
		scan_sizes_kernel<<<1, 1024, 0, st>>>(sa.sizes, sa.nblocks, d_block_off[idx[k]]);
		prof_end(1, st);
		prof_begin(2, st);
		compact_kernel<<<sa.nblocks, 128, 0, st>>>(sa.slots, stride, sa.sizes, d_block_off[idx[k]], d_out[idx[k]]);
		prof_end(2, st);
	}
	note_launch(1 + 2 * ca.n_streams);
	PIM_CUDA_TRY(cudaGetLastError());
	return SNAPPY_OK;
}
