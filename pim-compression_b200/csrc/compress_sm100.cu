// Block-parallel Snappy compression for sm_100a.
//
// Replaces the reference's DPU compressor (snappy/dpu-compress/dpu_compress.c:338-504,
// dpu_task.c:22-94) but reproduces, bit for bit, the output of the reference's HOST compressor
// snappy_compress_host / compress_block (snappy/snappy_compress.c:284-413,455-485): same
// multiplicative hash (:161-166), same per-block table sizing (:139-146), same skip heuristic
// (:333-348), same order of table updates (:391-397), same element encodings (:202-272).
//
// One CTA (one warp) per input block.
//   * the block is staged into shared memory with cp.async; the 16-bit hash table lives next
//     to it in shared memory;
//   * step 1 (probing) runs 32 probes per warp pass -- lane k takes the k-th next probe
//     position of the reference's schedule (its stride grows by one every 32 probes, so one
//     pass is exactly one stride group).  Table hits that an earlier lane of the same pass
//     would have overwritten are forwarded with __match_any_sync, and only the last writer of
//     a slot stores, so the table evolves exactly as in the serial loop;
//   * step 3 (match extension) compares 32 single bytes first, then 16 bytes per lane per
//     pass, reduced with __ballot_sync/__ffs;
//   * literals are copied by the whole warp, copy tags of long matches are written one tag
//     per lane.
// Each block writes to a fixed worst-case slot; block sizes then go through a device
// exclusive scan, and a compaction kernel moves every block to its final offset, yielding
// the contiguous ( u32 size | elements )* stream of the file format.

#include "common.cuh"

namespace pimsnappy {
namespace {

constexpr uint32_t MAX_TABLE = 16384; // snappy_compress.c:16-17
constexpr uint32_t MARGIN = 15;       // snappy_compress.c:299

__host__ __device__ inline uint32_t table_entries(uint32_t n) // snappy_compress.c:139-146
{
	uint32_t ts = 256;
	while (ts < MAX_TABLE && ts < n)
		ts <<= 1;
	return ts;
}

__host__ __device__ inline uint32_t slot_stride(uint32_t block_size)
{
	return (4u + 32u + block_size + block_size / 6u + 15u) & ~15u;
}

__host__ __device__ inline uint32_t compress_smem_bytes(uint32_t block_size)
{
	return 2u * table_entries(block_size) + 16u + block_size + 48u;
}

struct BlockCtx {
	uint8_t *sm;       // shared memory base
	uint32_t in0;      // smem byte offset of input byte 0
	uint16_t *table;
	int shift;
	uint8_t *gout;     // this block's output slot
	int lane;
};

__device__ __forceinline__ uint32_t load32(const BlockCtx &c, uint32_t pos) { return lds32_unaligned(c.sm, c.in0 + pos); }
__device__ __forceinline__ uint32_t hash32(const BlockCtx &c, uint32_t bytes) { return (bytes * 0x1e35a7bdu) >> c.shift; }

// snappy_compress.c:202-225 -- returns the new output cursor
__device__ __forceinline__ uint32_t emit_literal(const BlockCtx &c, uint32_t o, uint32_t from, uint32_t len)
{
	const uint32_t n = len - 1;
	uint32_t hdr;
	if (n < 60) {
		hdr = 1;
		if (c.lane == 0)
			c.gout[o] = (uint8_t)(n << 2);
	} else {
		const uint32_t extra = n < 256 ? 1 : (n < 65536 ? 2 : 3); // n <= 65535 in the valid domain
		hdr = 1 + extra;
		if (c.lane == 0)
			c.gout[o] = (uint8_t)((59 + extra) << 2);
		if (c.lane < (int)extra)
			c.gout[o + 1 + c.lane] = (uint8_t)(n >> (8 * c.lane));
	}
	uint8_t *dst = c.gout + o + hdr;
	const uint8_t *src = c.sm + c.in0 + from;
	for (uint32_t i = c.lane; i < len; i += 32)
		dst[i] = src[i];
	return o + hdr + len;
}

// snappy_compress.c:234-245
__device__ __forceinline__ uint32_t emit_copy_upto64(uint8_t *p, uint32_t off, uint32_t len)
{
	if (len < 12 && off < 2048) {
		p[0] = (uint8_t)(1 + ((len - 4) << 2) + ((off >> 8) << 5));
		p[1] = (uint8_t)off;
		return 2;
	}
	p[0] = (uint8_t)(2 + ((len - 1) << 2));
	p[1] = (uint8_t)off;
	p[2] = (uint8_t)(off >> 8);
	return 3;
}

// snappy_compress.c:254-272.  `while (len >= 68) emit 64` runs (len-4)/64 times.
__device__ __forceinline__ uint32_t emit_copy(const BlockCtx &c, uint32_t o, uint32_t off, uint32_t len)
{
	const uint32_t n64 = len >= 68 ? (len - 4) / 64 : 0;
	for (uint32_t t = c.lane; t < n64; t += 32) {
		uint8_t *p = c.gout + o + 3 * t;
		p[0] = (uint8_t)(2 + (63 << 2));
		p[1] = (uint8_t)off;
		p[2] = (uint8_t)(off >> 8);
	}
	o += 3 * n64;
	len -= 64 * n64;
	const bool two = len > 64;
	const uint32_t l1 = two ? 60 : len;
	const uint32_t s1 = (l1 < 12 && off < 2048) ? 2 : 3;
	uint32_t s2 = 0;
	if (two) {
		const uint32_t l2 = len - 60;
		s2 = (l2 < 12 && off < 2048) ? 2 : 3;
	}
	if (c.lane == 0) {
		emit_copy_upto64(c.gout + o, off, l1);
		if (two)
			emit_copy_upto64(c.gout + o + s1, off, len - 60);
	}
	return o + s1 + s2;
}

// sixteen bytes at an arbitrary shared-memory byte offset
__device__ __forceinline__ void load128_unaligned(const uint8_t *sm, uint32_t off, uint32_t v[4])
{
	const uint32_t a = off & ~3u;
	const uint32_t sh = (off & 3u) * 8u;
	const uint32_t w0 = lds32(sm, a), w1 = lds32(sm, a + 4), w2 = lds32(sm, a + 8), w3 = lds32(sm, a + 12), w4 = lds32(sm, a + 16);
	v[0] = __funnelshift_r(w0, w1, sh);
	v[1] = __funnelshift_r(w1, w2, sh);
	v[2] = __funnelshift_r(w2, w3, sh);
	v[3] = __funnelshift_r(w3, w4, sh);
}

// snappy_compress.c:176-193: common prefix of in[a..] and in[b..], at most maxn bytes (b > a).
__device__ __forceinline__ uint32_t match_length(const BlockCtx &c, uint32_t a, uint32_t b, uint32_t maxn)
{
	// pass 0: one byte per lane
	{
		const uint32_t j = c.lane;
		const bool eq = j < maxn && c.sm[c.in0 + a + j] == c.sm[c.in0 + b + j];
		const uint32_t ne = __ballot_sync(FULL, !eq);
		if (ne)
			return __ffs(ne) - 1;
	}
	// long match: sixteen bytes per lane per pass
	for (uint32_t base = 32;; base += 512) {
		const uint32_t j = base + 16u * c.lane;
		uint32_t cnt = 0; // equal bytes at the front of this lane's 16
		if (j < maxn) {
			uint32_t x[4], y[4];
			load128_unaligned(c.sm, c.in0 + a + j, x);
			load128_unaligned(c.sm, c.in0 + b + j, y);
			cnt = 16;
#pragma unroll
			for (int w = 3; w >= 0; w--) {
				const uint32_t d = x[w] ^ y[w];
				if (d)
					cnt = 4 * w + ((__ffs(d) - 1) >> 3);
			}
			cnt = min(cnt, maxn - j);
		}
		const uint32_t stop = __ballot_sync(FULL, cnt < 16);
		if (stop) {
			const int l = __ffs(stop) - 1;
			return base + 16u * l + __shfl_sync(FULL, cnt, l);
		}
	}
}

__global__ void __launch_bounds__(32) snappy_compress_kernel(const uint8_t *__restrict__ in, uint64_t in_len, uint32_t block_size,
                                                             uint8_t *__restrict__ slots, uint32_t stride, uint32_t *__restrict__ sizes)
{
	extern __shared__ __align__(16) uint8_t sm[];
	const int lane = lane_id();
	const uint64_t blk = blockIdx.x;
	const uint64_t in_off = blk * (uint64_t)block_size;
	const uint32_t n = (uint32_t)min((uint64_t)block_size, in_len - in_off);
	const uint8_t *gsrc = in + in_off;

	const uint32_t ts_max = table_entries(block_size);
	const uint32_t ts = table_entries(n);

	BlockCtx c;
	c.sm = sm;
	c.in0 = 2u * ts_max + (uint32_t)(reinterpret_cast<uintptr_t>(gsrc) & 15u);
	c.table = reinterpret_cast<uint16_t *>(sm);
	c.shift = 32 - (31 - __clz(ts)); // snappy_compress.c:288
	c.gout = slots + blk * (uint64_t)stride;
	c.lane = lane;

	stage_bytes_async(sm + c.in0, gsrc, n, lane, 32);
	cp_async_commit();
	for (uint32_t i = lane * 16u; i < 2u * ts; i += 512u)
		*reinterpret_cast<uint4 *>(sm + i) = make_uint4(0, 0, 0, 0);
	cp_async_wait<0>();
	__syncwarp();

	uint32_t o = 4;    // output cursor (4 bytes reserved for the size, :291)
	uint32_t emit = 0; // first input byte not yet emitted

	if (n >= MARGIN) {
		const uint32_t limit = n - MARGIN;
		uint32_t ip = 1;
		for (;;) {
			// ---- step 1: probe, 32 positions per pass --------------------------------
			uint32_t bstart = ip, cand = 0;
			bool out_of_input = false;
			for (uint32_t s = 1;; s++) {
				const uint32_t p = bstart + lane * s;
				const bool valid = p + s <= limit;
				uint32_t h = 0xFFFF0000u | lane, cnd = 0, cur = 0;
				if (valid) {
					cur = load32(c, p);
					h = hash32(c, cur);
					cnd = c.table[h];
				}
				const uint32_t same = __match_any_sync(FULL, h);
				const uint32_t lower = same & ((1u << lane) - 1u);
				if (lower)
					cnd = bstart + (31 - __clz(lower)) * s; // an earlier lane of this pass owns the slot
				const bool hit = valid && load32(c, cnd) == cur;
				const uint32_t hits = __ballot_sync(FULL, hit);
				const uint32_t inval = __ballot_sync(FULL, !valid);
				const int f = hits ? __ffs(hits) - 1 : 32;
				const int iv = inval ? __ffs(inval) - 1 : 32;
				if (f < iv) {
					const uint32_t later = same & ~((2u << lane) - 1u) & ((2u << f) - 1u);
					if (lane <= f && !later)
						c.table[h] = (uint16_t)p;
					ip = bstart + f * s;
					cand = __shfl_sync(FULL, cnd, f);
					__syncwarp();
					break;
				}
				if (iv < 32) {
					out_of_input = true;
					break;
				}
				if (!(same & ~((2u << lane) - 1u)))
					c.table[h] = (uint16_t)p;
				__syncwarp();
				bstart += 32 * s;
			}
			if (out_of_input)
				break;

			// ---- step 2: pending literal ---------------------------------------------
			o = emit_literal(c, o, emit, ip - emit);

			// ---- step 3: copy, then try to chain another copy ------------------------
			bool done = false;
			uint32_t cand_bytes;
			do {
				const uint32_t start = ip;
				const uint32_t m = 4 + match_length(c, cand + 4, ip + 4, n - ip - 4);
				ip += m;
				o = emit_copy(c, o, start - cand, m);
				emit = ip;
				if (ip >= limit) {
					done = true;
					break;
				}
				c.table[hash32(c, load32(c, ip - 1))] = (uint16_t)(ip - 1);
				__syncwarp();
				const uint32_t h = hash32(c, load32(c, ip));
				cand = c.table[h];
				cand_bytes = load32(c, cand);
				__syncwarp();
				c.table[h] = (uint16_t)ip;
				__syncwarp();
			} while (load32(c, ip) == cand_bytes);
			if (done)
				break;
			ip++;
		}
	}
	if (emit < n)
		o = emit_literal(c, o, emit, n - emit);
	if (lane < 4)
		c.gout[lane] = (uint8_t)((o - 4) >> (8 * lane));
	if (lane == 0)
		sizes[blk] = o;
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
	return ((nb * 4 + 255) & ~255ull) + nb * (uint64_t)slot_stride(block_size) + 256;
}

extern "C" snappy_status pim_snappy_compress_device(const uint8_t *d_in, uint64_t input_len, uint32_t block_size, uint8_t *d_scratch,
                                                    uint64_t scratch_bytes, uint8_t *d_out, uint64_t *d_block_off, void *stream)
{
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	if (block_size == 0 || block_size > PIM_SNAPPY_MAX_BLOCK_SIZE)
		return SNAPPY_BUFFER_TOO_SMALL;
	const uint64_t nb = pim_snappy_num_blocks(input_len, block_size);
	if (nb > 0x7FFFFFFFull || scratch_bytes < pim_snappy_compress_scratch_bytes(input_len, block_size))
		return SNAPPY_BUFFER_TOO_SMALL;
	if (nb == 0) {
		PIM_CUDA_TRY(cudaMemsetAsync(d_block_off, 0, sizeof(uint64_t), st));
		return SNAPPY_OK;
	}
	uint32_t *sizes = reinterpret_cast<uint32_t *>(d_scratch);
	uint8_t *slots = d_scratch + ((nb * 4 + 255) & ~255ull);
	const uint32_t stride = slot_stride(block_size);
	const uint32_t smem = compress_smem_bytes(block_size);

	static int configured_for[64] = {0};
	int dev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&dev));
	if (dev < 64 && !configured_for[dev]) {
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_compress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                  (int)compress_smem_bytes(PIM_SNAPPY_MAX_BLOCK_SIZE)));
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_compress_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
		configured_for[dev] = 1;
	}
	prof_begin(0, st);
	snappy_compress_kernel<<<(unsigned)nb, 32, smem, st>>>(d_in, input_len, block_size, slots, stride, sizes);
	prof_end(0, st);
	prof_begin(1, st);
	scan_sizes_kernel<<<1, 1024, 0, st>>>(sizes, nb, d_block_off);
	prof_end(1, st);
	prof_begin(2, st);
	compact_kernel<<<(unsigned)nb, 128, 0, st>>>(slots, stride, sizes, d_block_off, d_out);
	prof_end(2, st);
	note_launch(3);
	PIM_CUDA_TRY(cudaGetLastError());
	return SNAPPY_OK;
}
