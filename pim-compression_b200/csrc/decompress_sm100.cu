// Block-parallel Snappy decompression for sm_100a.
//
// Replaces the reference's DPU decoder (snappy/dpu-decompress/dpu_decompress.c:224-299,
// dpu_task.c:21-95) and must reproduce what its host twin snappy_decompress_host
// (snappy/snappy_decompress.c:218-289) produces on every valid stream.
//
// One CTA (two warps) per compressed block:
//   * the PARSER warp stages the block's compressed bytes through shared memory in
//     1 KB chunks (cp.async), walks the element tags and emits one 8-byte record per
//     element into a double-buffered record queue;
//   * the COPIER warp consumes 32 records per step, one element per lane, writing into a
//     shared-memory tile that holds the block's whole output (so back-references are
//     shared-memory reads).  Elements whose source overlaps output of the same step are
//     resolved in-warp in dependency rounds; self-overlapping matches (offset < length)
//     are replicated cooperatively by the whole warp;
//   * both warps then flush the tile to HBM with 16-byte stores.
// The warps hand batches over with named barriers (bar.sync / bar.arrive).
//
// Element formats: snappy/README.md:19-33, snappy_decompress.c:244-283, dpu_snappy.h:10-13.

#include "common.cuh"

namespace pimsnappy {
namespace {

constexpr int CHUNK = 1024;           // compressed bytes parsed per batch
constexpr int TAIL = 64;              // staged look-ahead past the chunk (tag bytes + short literals)
constexpr int STAGE_BYTES = CHUNK + TAIL + 16;
constexpr int MAX_REC = CHUNK / 2;    // an element occupies >= 2 compressed bytes
constexpr int SHORT_LEN = 16;         // elements up to this long take the per-lane register path

// shared-memory map (byte offsets)
constexpr uint32_t SM_NREC = 0;       // u32[2]
constexpr uint32_t SM_LAST = 8;       // u32[2]
constexpr uint32_t SM_ERR = 16;       // u32
constexpr uint32_t SM_SEGBASE = 64;   // u32[2][32]
constexpr uint32_t SM_LUT = 320;      // u8[256]: compressed footprint of an element by tag byte (0 = long literal)
constexpr uint32_t SM_EX = 576;       // u16[32][32]: parser's exit table, [offset in segment][segment]
constexpr uint32_t SM_STAGE = SM_EX + 2048; // u8[2][STAGE_BYTES]
constexpr uint32_t SM_REC = SM_STAGE + 2 * STAGE_BYTES; // uint2[2][MAX_REC]
constexpr uint32_t SM_TILE = SM_REC + 2 * MAX_REC * 8;
static_assert(SM_REC % 16 == 0 && SM_TILE % 16 == 0, "alignment");

constexpr int BAR_FULL = 1;  // +b
constexpr int BAR_EMPTY = 3; // +b

// record word 1: type(2) | segment(5) | src(24)
constexpr uint32_t T_LIT_SMEM = 0; // literal, payload staged in shared memory (src = smem byte offset)
constexpr uint32_t T_COPY = 1;     // back-reference (src = offset)
constexpr uint32_t T_LIT_GMEM = 2; // literal read from HBM (src = offset from the block's first element byte)

__host__ __device__ inline uint32_t decompress_smem_bytes(uint32_t block_size)
{
	return SM_TILE + 16 + block_size + 32;
}

// ---------------------------------------------------------------------------
// Element decode (snappy_decompress.c:242-283, :64-133).  `so` = shared-memory offset of the
// tag byte, `cp` = its position in the block's compressed bytes.  Returns false if malformed.
// ---------------------------------------------------------------------------
struct Element {
	uint32_t len;  // bytes produced
	uint32_t foot; // compressed bytes occupied
	uint32_t type; // T_*
	uint32_t src;  // see T_*
};

__device__ __forceinline__ bool decode_element(const uint8_t *sm, uint32_t so, uint32_t cp, uint32_t csize, uint32_t staged_end, Element &e)
{
	const uint32_t v = lds32_unaligned(sm, so); // tag + next three bytes
	const uint32_t tag = v & 0xFF;
	const uint32_t kind = tag & 3;
	if (kind == 0) {
		const uint32_t n6 = tag >> 2;
		uint32_t hdr = 1;
		if (n6 < 60) {
			e.len = n6 + 1;
		} else {
			const uint32_t nb = n6 - 59;
			hdr = 1 + nb;
			if (cp + hdr > csize)
				return false;
			const uint32_t lm1 = (nb == 4) ? ((v >> 8) | (uint32_t(sm[so + 4]) << 24)) : ((v >> 8) & ((1u << (8 * nb)) - 1));
			if (lm1 >= 65536u)
				return false;
			e.len = lm1 + 1;
		}
		if (e.len > csize - cp - hdr)
			return false;
		e.foot = hdr + e.len;
		if (e.len <= SHORT_LEN && cp + e.foot <= staged_end) {
			e.type = T_LIT_SMEM;
			e.src = so + hdr;
		} else {
			e.type = T_LIT_GMEM;
			e.src = cp + hdr;
		}
		return true;
	}
	e.type = T_COPY;
	if (kind == 1) {
		e.foot = 2;
		e.len = ((tag >> 2) & 7) + 4;
		e.src = ((tag >> 5) << 8) | ((v >> 8) & 0xFF);
	} else if (kind == 2) {
		e.foot = 3;
		e.len = (tag >> 2) + 1;
		e.src = (v >> 8) & 0xFFFF;
	} else {
		e.foot = 5;
		e.len = (tag >> 2) + 1;
		e.src = (v >> 8) | (uint32_t(sm[so + 4]) << 24);
		if (e.src > 65535u)
			return false;
	}
	return cp + e.foot <= csize;
}

// footprint of a long literal (tag>>2 >= 60); BIG_FOOT if malformed.  Must agree with decode_element.
constexpr uint32_t BIG_FOOT = 0x100000u;
__device__ __forceinline__ uint32_t long_literal_foot(const uint8_t *sm, uint32_t so, uint32_t cp, uint32_t csize)
{
	const uint32_t v = lds32_unaligned(sm, so);
	const uint32_t nb = ((v & 0xFF) >> 2) - 59;
	if (cp + 1 + nb > csize)
		return BIG_FOOT;
	const uint32_t lm1 = (nb == 4) ? ((v >> 8) | (uint32_t(sm[so + 4]) << 24)) : ((v >> 8) & ((1u << (8 * nb)) - 1));
	return lm1 >= 65536u ? BIG_FOOT : 2 + nb + lm1;
}

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, int lane)
{
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t t = __shfl_up_sync(FULL, v, d);
		if (lane >= d)
			v += t;
	}
	return v;
}

// ---------------------------------------------------------------------------
// Parser: the tag chain of a 1 KB chunk is inherently serial (an element's length says where
// the next tag is), so the warp resolves it in three data-parallel phases:
//   1. lane s owns compressed bytes [32s, 32s+32) of the chunk.  Walking its bytes backwards
//      it fills an "exit table": for EVERY byte offset j, where a chain entering the segment
//      at j leaves it, and how many elements it passes (a dynamic program over
//      next(j) = j + footprint(j), with footprints from a 256-entry table by tag byte);
//   2. the true chain is then 32 table look-ups, one per segment, starting from the known
//      entry position of the chunk;
//   3. every lane decodes the elements of its own segment from its true entry and writes
//      their records; prefix sums over the lanes give record slots and output positions.
// Returns the compressed position of the first element of the next chunk (0xFFFFFFFF on a
// malformed element).
// ---------------------------------------------------------------------------
constexpr uint32_t EX_FAR = 0x7FFu; // exit "beyond anything this chunk cares about"

__device__ __forceinline__ uint32_t parse_chunk_parallel(uint8_t *sm, int b, uint32_t stage_off, uint32_t cs, uint32_t cp, uint32_t csize,
                                                         uint32_t out_base, uint32_t *out_total, int lane)
{
	uint16_t *ex = reinterpret_cast<uint16_t *>(sm + SM_EX);
	const uint8_t *lut = sm + SM_LUT;
	const uint32_t seg_so = stage_off + 32u * lane; // smem offset of my segment
	const uint32_t seg_cp = cs + 32u * lane;        // compressed position of my segment

	// ---- phase 1: exit table -------------------------------------------------------------
#pragma unroll 4
	for (int j = 31; j >= 0; j--) {
		uint32_t v = EX_FAR; // positions past the end of the block: no element, exit far
		if (seg_cp + j < csize) {
			uint32_t f = lut[sm[seg_so + j]];
			if (f == 0)
				f = long_literal_foot(sm, seg_so + j, seg_cp + j, csize);
			const uint32_t n = j + f;
			if (n < 32)
				v = ex[n * 32 + lane] + (1u << 11); // same exit, one more element
			else
				v = min(n, EX_FAR) | (1u << 11);
		}
		ex[j * 32 + lane] = (uint16_t)v;
	}
	__syncwarp();

	// ---- phase 2: follow the true chain across the segments ----------------------------------
	uint32_t x = cp - cs; // entry, relative to the chunk
	const uint32_t csz = csize - cs;
	uint32_t my_x = 0, my_cnt = 0;
	for (int s = 0; s < 32; s++) {
		if (x >= 1024u || x >= csz)
			break;
		if (x < 32u * s + 32u) {
			const uint32_t w = ex[(x - 32u * s) * 32 + s];
			if (lane == s) {
				my_x = x;
				my_cnt = w >> 11;
			}
			const uint32_t e = w & 0x7FFu;
			x = e == EX_FAR ? 0xFFFFFFu : 32u * s + e;
		}
	}

	// ---- phase 3: decode my segment's elements into records ----------------------------------
	const uint32_t cnt_incl = warp_inclusive_scan(my_cnt, lane);
	const uint32_t nrec = __shfl_sync(FULL, cnt_incl, 31);
	uint2 *rec = reinterpret_cast<uint2 *>(sm + SM_REC) + b * MAX_REC + (cnt_incl - my_cnt);
	const uint32_t staged_end = cs + CHUNK + TAIL;
	uint32_t p = cs + my_x, local = 0;
	bool bad = false;
	const uint32_t max_cnt = __reduce_max_sync(FULL, my_cnt);
	for (uint32_t k = 0; k < max_cnt; k++) {
		if (k < my_cnt && !bad) {
			Element e;
			if (!decode_element(sm, stage_off + (p - cs), p, csize, staged_end, e) || local > 0xFFFFu) {
				bad = true;
				e.len = 1; e.foot = 1; e.type = T_LIT_SMEM; e.src = 0;
			}
			rec[k] = make_uint2((local & 0xFFFFu) | ((e.len - 1) << 16), (e.type << 30) | ((uint32_t)lane << 24) | e.src);
			local += e.len;
			p += e.foot;
		}
	}
	const uint32_t out_incl = warp_inclusive_scan(local, lane);
	reinterpret_cast<uint32_t *>(sm + SM_SEGBASE)[b * 32 + lane] = out_base + out_incl - local;
	*out_total = out_base + __shfl_sync(FULL, out_incl, 31);
	// A malformed element leaves the rest of its lane's record slots unwritten (stale): the
	// batch is then dropped as a whole -- the block is reported invalid anyway.
	const bool any_bad = __any_sync(FULL, bad);
	if (lane == 0)
		reinterpret_cast<uint32_t *>(sm + SM_NREC)[b] = any_bad ? 0u : nrec;
	if (any_bad && lane == 0)
		*reinterpret_cast<volatile uint32_t *>(sm + SM_ERR) = 1;
	const uint32_t next_cp = __reduce_max_sync(FULL, my_cnt ? p : 0u);
	return any_bad ? 0xFFFFFFFFu : next_cp;
}

// ---------------------------------------------------------------------------
// Copier: whole warp copies one long / self-overlapping element.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void coop_copy(uint8_t *sm, const uint8_t *gcomp, uint32_t type, uint32_t dst, uint32_t src, uint32_t len, int lane)
{
	if (type == T_LIT_GMEM) {
		for (uint32_t k = lane; k < len; k += 32)
			sm[dst + k] = __ldg(gcomp + src + k);
	} else if (type == T_LIT_SMEM) {
		for (uint32_t k = lane; k < len; k += 32)
			sm[dst + k] = sm[src + k];
	} else {
		const uint32_t off = src;
		const uint32_t s = dst - off;
		if (off >= 32) {
			for (uint32_t k0 = 0; k0 < len; k0 += 32) {
				const uint32_t k = k0 + lane;
				if (k < len)
					sm[dst + k] = sm[s + k];
				if (off < len)
					__syncwarp(); // the next 32 bytes may read what this pass wrote
			}
		} else {
			// pattern of period `off` (< 32) that already exists in [dst-off, dst)
			for (uint32_t k = lane; k < len; k += 32)
				sm[dst + k] = sm[s + (k % off)];
		}
	}
}

__device__ __forceinline__ void copier_batch(uint8_t *sm, const uint8_t *gcomp, int b, uint32_t tile0, uint32_t out_size, int lane)
{
	const uint32_t n = reinterpret_cast<const uint32_t *>(sm + SM_NREC)[b];
	const uint2 *rec = reinterpret_cast<const uint2 *>(sm + SM_REC) + b * MAX_REC;
	const uint32_t *segbase = reinterpret_cast<const uint32_t *>(sm + SM_SEGBASE) + b * 32;

	for (uint32_t base = 0; base < n; base += 32) {
		const uint32_t i = base + lane;
		bool act = i < n;
		uint32_t len = 1, type = T_LIT_SMEM, src = 0, out_pos = 0;
		if (act) {
			const uint2 r = rec[i];
			len = (r.x >> 16) + 1;
			type = r.y >> 30;
			src = r.y & 0xFFFFFFu;
			out_pos = segbase[(r.y >> 24) & 31] + (r.x & 0xFFFF);
		}
		const uint32_t dst = tile0 + out_pos;
		const bool is_copy = type == T_COPY;
		if (act && (out_pos + len > out_size || (is_copy && (src == 0 || src > out_pos)))) {
			*reinterpret_cast<volatile uint32_t *>(sm + SM_ERR) = 1;
			act = false; // malformed element: never touch memory for it
		}
		// sa: shared-memory source address for per-lane copies
		const uint32_t sa = is_copy ? dst - src : src;
		const bool small = len <= SHORT_LEN && type != T_LIT_GMEM && !(is_copy && src < len);
		// first output address this element needs to be complete before it can run
		const uint32_t need = is_copy ? min(sa + len, dst) : 0u;

		uint32_t undone = __ballot_sync(FULL, act);
		while (undone) {
			const int u = __ffs(undone) - 1;
			const uint32_t frontier = __shfl_sync(FULL, dst, u); // everything below is written
			const bool ready = ((undone >> lane) & 1u) && need <= frontier;
			if (ready && small) {
				const uint32_t a = sa & ~3u;
				const uint32_t sh = (sa & 3u) * 8u;
				uint32_t w0 = lds32(sm, a), w1 = lds32(sm, a + 4), w2 = lds32(sm, a + 8), w3 = lds32(sm, a + 12), w4 = lds32(sm, a + 16);
				uint32_t v[4];
				v[0] = __funnelshift_r(w0, w1, sh);
				v[1] = __funnelshift_r(w1, w2, sh);
				v[2] = __funnelshift_r(w2, w3, sh);
				v[3] = __funnelshift_r(w3, w4, sh);
#pragma unroll
				for (int k = SHORT_LEN - 1; k >= 0; k--) {
					if (k < (int)len)
						sm[dst + k] = (uint8_t)(v[k >> 2] >> (8 * (k & 3)));
				}
			}
			uint32_t big = __ballot_sync(FULL, ready && !small);
			while (big) {
				const int l = __ffs(big) - 1;
				big &= big - 1;
				const uint32_t btype = __shfl_sync(FULL, type, l);
				const uint32_t bdst = __shfl_sync(FULL, dst, l);
				const uint32_t bsrc = __shfl_sync(FULL, src, l);
				const uint32_t blen = __shfl_sync(FULL, len, l);
				coop_copy(sm, gcomp, btype, bdst, bsrc, blen, lane);
			}
			__syncwarp();
			undone &= ~__ballot_sync(FULL, ready);
		}
	}
}

// ---------------------------------------------------------------------------
// Kernel: grid = number of blocks, 64 threads.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(64, 8) snappy_decompress_kernel(const uint8_t *__restrict__ comp, const uint64_t *__restrict__ block_off,
                                                               uint32_t block_size, uint64_t out_len, uint8_t *__restrict__ out,
                                                               uint32_t *__restrict__ status)
{
	extern __shared__ __align__(16) uint8_t sm[];
	const int lane = lane_id();
	const int warp = threadIdx.x >> 5;
	const uint64_t blk = blockIdx.x;

	const uint64_t off0 = block_off[blk], off1 = block_off[blk + 1];
	const uint64_t out_off = blk * (uint64_t)block_size;
	const uint32_t out_size = (uint32_t)min((uint64_t)block_size, out_len - out_off);
	uint8_t *gdst = out + out_off;
	const uint32_t tile0 = SM_TILE + (uint32_t)(reinterpret_cast<uintptr_t>(gdst) & 15u);

	if (threadIdx.x == 0)
		*reinterpret_cast<uint32_t *>(sm + SM_ERR) = 0;
	for (uint32_t t = threadIdx.x; t < 256; t += 64) {
		const uint32_t kind = t & 3, n6 = t >> 2;
		sm[SM_LUT + t] = (uint8_t)(kind == 0 ? (n6 < 60 ? n6 + 2 : 0) : (kind == 1 ? 2 : (kind == 2 ? 3 : 5)));
	}
	__syncthreads();

	bool header_ok = off1 >= off0 + 4;
	uint32_t csize = 0;
	const uint8_t *gcomp = comp + off0 + 4;
	if (header_ok) {
		const uint8_t *h = comp + off0;
		csize = (uint32_t)__ldg(h) | ((uint32_t)__ldg(h + 1) << 8) | ((uint32_t)__ldg(h + 2) << 16) | ((uint32_t)__ldg(h + 3) << 24);
		header_ok = (uint64_t)csize + 4 == off1 - off0;
	}

	if (!header_ok) {
		if (threadIdx.x == 0)
			*reinterpret_cast<uint32_t *>(sm + SM_ERR) = 1;
	} else if (warp == 0) {
		// ------------------------------ parser ------------------------------
		const uint32_t amis = (uint32_t)(reinterpret_cast<uintptr_t>(gcomp) & 15u);
		uint32_t cp = 0, out_base = 0;
		uint32_t k = 0;
		for (;; k++) {
			const int b = k & 1;
			if (k >= 2)
				bar_sync(BAR_EMPTY + b, 64);
			const uint32_t cs = (cp / CHUNK) * CHUNK;
			const uint32_t stage_off = SM_STAGE + b * STAGE_BYTES + amis;
			const uint32_t nstage = min((uint32_t)(CHUNK + TAIL), csize - cs);
			stage_bytes_async(sm + stage_off, gcomp + cs, nstage, lane, 32);
			cp_async_commit();
			cp_async_wait<0>();
			__syncwarp();
			uint32_t total = 0;
			cp = parse_chunk_parallel(sm, b, stage_off, cs, cp, csize, out_base, &total, lane);
			out_base = total;
			if (out_base > out_size) { // decodes to more than the block holds: stop before positions can wrap
				if (lane == 0)
					*reinterpret_cast<volatile uint32_t *>(sm + SM_ERR) = 1;
				cp = 0xFFFFFFFFu;
			}
			bool last = cp >= csize; // also true for the error marker
			if (last && lane == 0) {
				if (cp != csize || out_base != out_size)
					*reinterpret_cast<volatile uint32_t *>(sm + SM_ERR) = 1;
			}
			if (lane == 0)
				reinterpret_cast<uint32_t *>(sm + SM_LAST)[b] = last;
			__threadfence_block();
			bar_arrive(BAR_FULL + b, 64);
			if (last)
				break;
		}
		// consume the copier's outstanding "buffer free" arrivals
		const uint32_t nb = k + 1;
		for (uint32_t j = (nb >= 2 ? nb - 2 : 0); j < nb; j++)
			bar_sync(BAR_EMPTY + (j & 1), 64);
	} else {
		// ------------------------------ copier ------------------------------
		for (uint32_t k = 0;; k++) {
			const int b = k & 1;
			bar_sync(BAR_FULL + b, 64);
			const uint32_t last = reinterpret_cast<volatile uint32_t *>(sm + SM_LAST)[b];
			copier_batch(sm, gcomp, b, tile0, out_size, lane);
			__threadfence_block();
			bar_arrive(BAR_EMPTY + b, 64);
			if (last)
				break;
		}
	}
	__syncthreads();

	const uint32_t err = *reinterpret_cast<volatile uint32_t *>(sm + SM_ERR);
	if (err) {
		if (threadIdx.x == 0)
			atomicOr(status, 1u);
		return;
	}
	// flush the tile: tile0 and gdst are congruent mod 16
	{
		const uint32_t mis = tile0 & 15u;
		uint32_t head = (16u - mis) & 15u;
		if (head > out_size)
			head = out_size;
		const uint32_t mid = (out_size - head) & ~15u;
		for (uint32_t i = threadIdx.x; i < head; i += 64)
			gdst[i] = sm[tile0 + i];
		const uint4 *s4 = reinterpret_cast<const uint4 *>(sm + tile0 + head);
		uint4 *g4 = reinterpret_cast<uint4 *>(gdst + head);
		for (uint32_t i = threadIdx.x; i < mid / 16; i += 64)
			g4[i] = s4[i];
		for (uint32_t i = head + mid + threadIdx.x; i < out_size; i += 64)
			gdst[i] = sm[tile0 + i];
	}
}

} // namespace
} // namespace pimsnappy

using namespace pimsnappy;

extern "C" snappy_status pim_snappy_decompress_device(const uint8_t *d_comp, const uint64_t *d_block_off, uint64_t nblocks,
                                                      uint32_t block_size, uint64_t out_len, uint8_t *d_out, uint32_t *d_status,
                                                      void *stream)
{
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	if (block_size == 0 || block_size > PIM_SNAPPY_MAX_BLOCK_SIZE)
		return SNAPPY_BUFFER_TOO_SMALL;
	if (nblocks != pim_snappy_num_blocks(out_len, block_size) || nblocks > 0x7FFFFFFFull)
		return SNAPPY_INVALID_INPUT;
	PIM_CUDA_TRY(cudaMemsetAsync(d_status, 0, sizeof(uint32_t), st));
	if (nblocks == 0)
		return SNAPPY_OK;
	const uint32_t smem = decompress_smem_bytes(block_size);
	static int configured_for[64] = {0}; // per device: largest smem opted in
	int dev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&dev));
	if (dev < 64 && configured_for[dev] < (int)smem) {
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                  (int)decompress_smem_bytes(PIM_SNAPPY_MAX_BLOCK_SIZE)));
		PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
		configured_for[dev] = (int)decompress_smem_bytes(PIM_SNAPPY_MAX_BLOCK_SIZE);
	}
	prof_begin(3, st);
	snappy_decompress_kernel<<<(unsigned)nblocks, 64, smem, st>>>(d_comp, d_block_off, block_size, out_len, d_out, d_status);
	prof_end(3, st);
	note_launch();
	PIM_CUDA_TRY(cudaGetLastError());
	return SNAPPY_OK;
}
