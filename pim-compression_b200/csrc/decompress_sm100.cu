// Block-parallel Snappy decompression for sm_100a.
//
// Replaces the reference's DPU decoder (snappy/dpu-decompress/dpu_decompress.c:224-299,
// dpu_task.c:21-95) and must reproduce what its host twin snappy_decompress_host
// (snappy/snappy_decompress.c:218-289) produces on every valid stream.
//
// One CTA (three warps) per compressed block:
//   * two PARSER warps take the block's compressed bytes in 512-byte chunks, alternately.
//     A chunk is staged into shared memory with cp.async; its serial tag chain (an element's
//     length says where the next tag is) is resolved data-parallel: lane s owns bytes
//     [16s, 16s+16) and fills an "exit table" (for EVERY entry offset: where a chain entering
//     there leaves the segment and how many elements it passes), then the 32 lanes iterate
//     x[s+1] = exit_s(x[s]) to the fixed point, which is the true chain.  The chunk's entry
//     position is the only thing handed from one parser to the other, so the exit tables of
//     consecutive chunks are built concurrently.  Each lane then decodes the elements of its own
//     segment into 8-byte records;
//   * the COPIER warp executes the records against a shared-memory tile that holds the block's
//     whole output, so back-references (write_copy_host, snappy_decompress.c:164-184) are
//     shared-memory reads.  It keeps a sliding window of 32 in-flight elements, one per lane: an
//     element runs as soon as every output byte its source needs has been written, finished
//     lanes are refilled from the record queue, so a copy that depends on its neighbours delays
//     only itself.  Self-overlapping matches (offset < length) are widened in-lane by doubling;
//   * all three warps flush the tile to HBM with 16-byte stores.
// Hand-over: named barriers (bar.sync / bar.arrive) for "records ready" and "entry known",
// a shared counter for "slot free".
//
// Element formats: snappy/README.md:19-33, snappy_decompress.c:244-283, dpu_snappy.h:10-13.

#include "common.cuh"

namespace pimsnappy {
namespace {

constexpr int SEG = 16;                 // compressed bytes per lane
constexpr int CHUNK = 32 * SEG;         // compressed bytes per parse step
constexpr int TAIL = 64;                // staged look-ahead past the chunk (tag bytes + short literal payloads)
constexpr int STAGE_BYTES = CHUNK + TAIL + 16;
constexpr int MAX_REC = CHUNK / 2;      // an element occupies >= 2 compressed bytes
constexpr int NS = 5;                   // staging slots
constexpr int NR = 3;                   // record slots
constexpr int NPARSE = 2;
constexpr int NTHREADS = 32 * (NPARSE + 1);
constexpr uint32_t SHORT_LIT = 60;      // literals up to this long are served from the staged bytes

// shared-memory map (byte offsets)
constexpr uint32_t SM_ERR = 0;          // u32
constexpr uint32_t SM_CONSUMED = 4;     // u32: chunks the copier has finished
constexpr uint32_t SM_ENTRY = 8;        // u32[2][2]: {cp, done} handed to the parser of the next chunk
constexpr uint32_t SM_META = 32;        // u32[NR][4]: {nrec, out bytes, last, -}
constexpr uint32_t SM_SEGBASE = 128;    // u32[NR][32]
constexpr uint32_t SM_LUT = SM_SEGBASE + NR * 128; // u8[256]: compressed footprint by tag byte (0 = long literal)
constexpr uint32_t SM_EX = SM_LUT + 256;           // u16[NPARSE][SEG][32]: exit tables
constexpr uint32_t SM_STAGE = SM_EX + NPARSE * SEG * 64;
constexpr uint32_t SM_REC = (SM_STAGE + NS * STAGE_BYTES + 15) & ~15u; // uint2[NR][MAX_REC]
constexpr uint32_t SM_TILE = SM_REC + NR * MAX_REC * 8;
static_assert(SM_STAGE % 16 == 0 && SM_REC % 16 == 0 && SM_TILE % 16 == 0, "alignment");

constexpr int BAR_FULL = 1;   // +record slot
constexpr int BAR_ENTRY = 4;  // +(chunk sequence number & 1)

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

__device__ __forceinline__ bool decode_element(const uint8_t *sm, uint32_t so, uint32_t cp, uint32_t csize, Element &e)
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
		if (e.len <= SHORT_LIT) { // hdr + len <= 5 + 60: inside the staged look-ahead
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

__device__ __forceinline__ void set_error(uint8_t *sm) { *reinterpret_cast<volatile uint32_t *>(sm + SM_ERR) = 1; }
__device__ __forceinline__ uint32_t vload(const uint8_t *sm, uint32_t off) { return *reinterpret_cast<const volatile uint32_t *>(sm + off); }
__device__ __forceinline__ void vstore(uint8_t *sm, uint32_t off, uint32_t v) { *reinterpret_cast<volatile uint32_t *>(sm + off) = v; }

// ---------------------------------------------------------------------------
// Parser, phase 1: exit table of one chunk.  Lane s owns compressed bytes [16s, 16s+16) of the
// chunk.  Walking its bytes backwards it fills, for EVERY byte offset j, where a chain entering
// the segment at j leaves it (relative to the segment start; EX_FAR = "a long literal, far
// away") and how many elements it passes: a dynamic program over next(j) = j + footprint(j),
// with footprints from a 256-entry table by tag byte.  Entry: exit | count << 8.
// ---------------------------------------------------------------------------
constexpr uint32_t EX_FAR = 0xFFu;

__device__ __forceinline__ void build_exit_table(uint8_t *sm, uint16_t *ex, uint32_t stage_off, uint32_t cs, uint32_t csize, int lane)
{
	const uint8_t *lut = sm + SM_LUT;
	const uint32_t seg_so = stage_off + SEG * lane;
	const uint32_t seg_cp = cs + SEG * lane;
	// footprints of all 16 positions first (independent loads), then the dependent chain
	uint32_t foot[SEG];
#pragma unroll
	for (int j = 0; j < SEG; j++)
		foot[j] = lut[sm[seg_so + j]];
#pragma unroll
	for (int j = SEG - 1; j >= 0; j--) {
		uint32_t v = j; // positions past the end of the block: no element, the chain stops here
		if (seg_cp + j < csize) {
			uint32_t f = foot[j];
			if (f == 0)
				f = long_literal_foot(sm, seg_so + j, seg_cp + j, csize);
			const uint32_t n = j + f;
			if (n < SEG)
				v = ex[n * 32 + lane] + 0x100u; // same exit, one more element
			else
				v = min(n, EX_FAR) | 0x100u;
		}
		ex[j * 32 + lane] = (uint16_t)v;
	}
	__syncwarp();
}

// ---------------------------------------------------------------------------
// Parser, phase 2: the true chain.  x[s] = first element start at or after segment s (relative to
// the chunk); x[0] is known, x[s+1] = exit_s(x[s]) if the chain touches segment s, else x[s].
// All lanes iterate this map from a guess until nothing changes; lane s is right after at most
// s rounds, usually after a few because chains entering a segment at different offsets merge.
// Returns this lane's entry (my_x), its element count, and the position after the chunk's last
// element (relative to the chunk; 0xFFFFFF if a long literal ends the chain).
// ---------------------------------------------------------------------------
constexpr uint32_t X_FAR = 0xFFFFFFu;

__device__ __forceinline__ void resolve_chain(const uint16_t *ex, uint32_t x0, uint32_t csz, int lane, uint32_t *my_x, uint32_t *my_cnt, bool *my_far,
                                              uint32_t *next_x)
{
	const uint32_t base = SEG * lane;
	uint32_t x = lane == 0 ? x0 : base;
	uint32_t f = x, cnt = 0;
	bool far = false;
	for (int it = 0; it < 33; it++) {
		f = x;
		cnt = 0;
		far = false;
		if (x < base + SEG && x < csz) {
			const uint32_t e = ex[(x - base) * 32 + lane];
			cnt = e >> 8;
			far = (e & 0xFFu) == EX_FAR;
			f = far ? X_FAR : base + (e & 0xFFu);
		}
		uint32_t nx = __shfl_up_sync(FULL, f, 1);
		if (lane == 0)
			nx = x0;
		const bool changed = nx != x;
		x = nx;
		if (!__any_sync(FULL, changed))
			break;
	}
	*my_x = x;
	*my_cnt = cnt;
	*my_far = far;
	*next_x = __shfl_sync(FULL, f, 31);
}

// ---------------------------------------------------------------------------
// Parser, phase 3: every lane decodes the elements of its own segment from its true entry and
// writes their records; prefix sums over the lanes give record slots and output positions
// (relative to the chunk -- the copier keeps the running base).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void emit_records(uint8_t *sm, int rslot, uint32_t stage_off, uint32_t cs, uint32_t csize, uint32_t my_x, uint32_t my_cnt,
                                             bool last, int lane)
{
	const uint32_t cnt_incl = warp_inclusive_scan(my_cnt, lane);
	const uint32_t nrec = __shfl_sync(FULL, cnt_incl, 31);
	uint2 *rec = reinterpret_cast<uint2 *>(sm + SM_REC) + rslot * MAX_REC + (cnt_incl - my_cnt);
	uint32_t p = cs + my_x, local = 0;
	bool bad = false;
	const uint32_t max_cnt = __reduce_max_sync(FULL, my_cnt);
	for (uint32_t k = 0; k < max_cnt; k++) {
		if (k < my_cnt && !bad) {
			Element e;
			if (!decode_element(sm, stage_off + (p - cs), p, csize, e)) {
				bad = true;
				e.len = 1; e.foot = 1; e.type = T_LIT_SMEM; e.src = SM_LUT;
			}
			rec[k] = make_uint2((local & 0xFFFFu) | ((e.len - 1) << 16), (e.type << 30) | ((uint32_t)lane << 24) | e.src);
			local += e.len;
			p += e.foot;
		}
	}
	const uint32_t out_incl = warp_inclusive_scan(local, lane);
	reinterpret_cast<uint32_t *>(sm + SM_SEGBASE)[rslot * 32 + lane] = out_incl - local;
	// A malformed element leaves the rest of its lane's record slots unwritten (stale): the batch
	// is then dropped as a whole -- the block is reported invalid anyway.
	const bool any_bad = __any_sync(FULL, bad);
	if (lane == 31) {
		uint32_t *meta = reinterpret_cast<uint32_t *>(sm + SM_META) + rslot * 4;
		meta[0] = any_bad ? 0u : nrec;
		meta[1] = out_incl;
		meta[2] = last ? 1u : 0u; // (a bad chunk is not the last one: the chain runs on to the block's end)
	}
	if (any_bad && lane == 0)
		set_error(sm);
}

// One parser warp.  Chunks are numbered in the order they are parsed (q); parser w takes
// q = w, w+2, ...  Chunk q normally covers the 512 bytes after chunk q-1's, which is what the
// parser stages and tabulates BEFORE chunk q-1's parser has told it the entry position; after a
// long literal the guess is wrong and the right chunk is staged again.
__device__ __forceinline__ void parser_warp(uint8_t *sm, const uint8_t *gcomp, uint32_t csize, int w, int lane)
{
	uint16_t *ex = reinterpret_cast<uint16_t *>(sm + SM_EX) + w * SEG * 32;
	const uint32_t amis = (uint32_t)(reinterpret_cast<uintptr_t>(gcomp) & 15u);
	uint32_t guess = w; // chunk index (position / CHUNK) expected for my next q
	for (uint32_t q = w;; q += NPARSE) {
		const int sslot = q % NS, rslot = q % NR;
		// both slots must have been released by the copier
		if (q >= NR) {
			while ((int)(vload(sm, SM_CONSUMED) + NR - 1 - q) < 0)
				__nanosleep(40);
		}
		const uint32_t stage_off = SM_STAGE + sslot * STAGE_BYTES + amis;
		uint32_t cs = guess * CHUNK;
		auto stage = [&]() {
			const uint32_t nstage = cs < csize ? min((uint32_t)(CHUNK + TAIL), csize - cs) : 0u;
			stage_bytes_async(sm + stage_off, gcomp + cs, nstage, lane, 32);
			cp_async_commit();
			cp_async_wait<0>();
			__syncwarp();
			build_exit_table(sm, ex, stage_off, cs, csize, lane);
		};
		stage();

		uint32_t cp = 0, done = 0;
		if (q > 0) {
			bar_sync(BAR_ENTRY + (q & 1), 64);
			cp = vload(sm, SM_ENTRY + (q & 1) * 8);
			done = vload(sm, SM_ENTRY + (q & 1) * 8 + 4);
		}
		if (done)
			break;
		if (cp / CHUNK != guess) {
			guess = cp / CHUNK;
			cs = guess * CHUNK;
			stage();
		}

		uint32_t my_x, my_cnt, next_x;
		bool my_far;
		resolve_chain(ex, cp - cs, csize - cs, lane, &my_x, &my_cnt, &my_far, &next_x);
		uint32_t next_cp = cs + next_x;
		if (__any_sync(FULL, my_far)) {
			// the chain leaves the chunk through a long literal: find where it ends exactly
			uint32_t far_end = 0;
			if (my_far) {
				uint32_t p = cs + my_x;
				for (uint32_t k = 0; k + 1 < my_cnt; k++) {
					uint32_t f = sm[SM_LUT + sm[stage_off + (p - cs)]];
					p += f; // only the last element of a segment can be a long literal
				}
				far_end = p + long_literal_foot(sm, stage_off + (p - cs), p, csize);
			}
			next_cp = __reduce_max_sync(FULL, far_end);
		}
		bool last = next_cp >= csize;
		if ((last && next_cp != csize) || next_cp <= cp) {
			if (lane == 0)
				set_error(sm); // the last element overruns the block (or the chain does not advance)
			last = true;
		}
		if (lane == 0) {
			vstore(sm, SM_ENTRY + ((q + 1) & 1) * 8, next_cp);
			vstore(sm, SM_ENTRY + ((q + 1) & 1) * 8 + 4, last ? 1u : 0u);
		}
		__threadfence_block();
		bar_arrive(BAR_ENTRY + ((q + 1) & 1), 64);

		emit_records(sm, rslot, stage_off, cs, csize, my_x, my_cnt, last, lane);
		__threadfence_block();
		bar_arrive(BAR_FULL + rslot, 64);
		if (last)
			break;
		guess += NPARSE;
	}
}

// ---------------------------------------------------------------------------
// Copier.
// ---------------------------------------------------------------------------

// Whole warp copies one long literal (> 60 bytes) from HBM into the tile: the destination is
// written in aligned 16-byte shared-memory stores, the source is read as aligned 32-bit words
// and funnel-shifted.  Every word read contains at least one byte of this block's compressed
// data (a long literal's payload is preceded by >= 6 bytes of headers), so no access can fault.
__device__ __forceinline__ void long_literal_copy(uint8_t *sm, uint32_t dst, const uint8_t *g, uint32_t len, int lane)
{
	uint32_t head = (16u - (dst & 15u)) & 15u;
	if (head > len)
		head = len;
	for (uint32_t k = lane; k < head; k += 32)
		sm[dst + k] = __ldg(g + k);
	const uint32_t body = (len - head) & ~15u;
	const uint8_t *s = g + head;                                         // source of the first full 16 bytes
	const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(s) & 3u); // constant over the body
	const uint32_t sh = mis * 8u;
	const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s - mis);
	uint4 *d4 = reinterpret_cast<uint4 *>(sm + dst + head);
	const uint32_t nchunk = body / 16;
#pragma unroll 4
	for (uint32_t c = lane; c < nchunk; c += 32) {
		const uint32_t w0 = __ldg(s32 + 4 * c), w1 = __ldg(s32 + 4 * c + 1), w2 = __ldg(s32 + 4 * c + 2), w3 = __ldg(s32 + 4 * c + 3);
		const uint32_t w4 = mis ? __ldg(s32 + 4 * c + 4) : 0u;
		d4[c] = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
	}
	for (uint32_t k = head + body + lane; k < len; k += 32)
		sm[dst + k] = __ldg(g + k);
	__syncwarp();
}

// Up to 16 bytes from shared-memory byte offset sa to dst (both arbitrary), n <= 16.
__device__ __forceinline__ void copy_upto16(uint8_t *sm, uint32_t dst, uint32_t sa, uint32_t n)
{
	const uint32_t a = sa & ~3u;
	const uint32_t sh = (sa & 3u) * 8u;
	const uint32_t w0 = lds32(sm, a), w1 = lds32(sm, a + 4), w2 = lds32(sm, a + 8), w3 = lds32(sm, a + 12), w4 = lds32(sm, a + 16);
	uint32_t v[4];
	v[0] = __funnelshift_r(w0, w1, sh);
	v[1] = __funnelshift_r(w1, w2, sh);
	v[2] = __funnelshift_r(w2, w3, sh);
	v[3] = __funnelshift_r(w3, w4, sh);
#pragma unroll
	for (int k = 0; k < 16; k++) {
		if (k < (int)n)
			sm[dst + k] = (uint8_t)(v[k >> 2] >> (8 * (k & 3)));
	}
}

struct Window { // one in-flight element per lane
	bool pend;
	uint32_t dst;  // tile byte offset of the element's first output byte
	uint32_t len;
	uint32_t off;  // copy: back-reference distance; literal: 0
	uint32_t src;  // literal: shared-memory offset of the payload
	uint32_t need; // runs once every tile byte below this has been written
};

__device__ __forceinline__ void copier_warp(uint8_t *sm, const uint8_t *gcomp, uint32_t tile0, uint32_t out_size, int lane)
{
	Window w;
	w.pend = false;
	w.dst = w.len = w.off = w.src = w.need = 0;
	const uint32_t lt_mask = (1u << lane) - 1u;
	uint32_t chunk_base = 0;            // output bytes of the chunks before this one
	uint32_t fetched_end = tile0;       // tile offset just past the last fetched element

	for (uint32_t q = 0;; q++) {
		const int rslot = q % NR;
		bar_sync(BAR_FULL + rslot, 64);
		const uint32_t n = vload(sm, SM_META + rslot * 16), chunk_out = vload(sm, SM_META + rslot * 16 + 4), last = vload(sm, SM_META + rslot * 16 + 8);
		const uint2 *rec = reinterpret_cast<const uint2 *>(sm + SM_REC) + rslot * MAX_REC;
		const uint32_t *segbase = reinterpret_cast<const uint32_t *>(sm + SM_SEGBASE) + rslot * 32;
		uint32_t next = 0;
		bool drained = false;

		while (!drained) {
			// ---- refill free lanes with the next records ------------------------------------
			uint32_t pm = __ballot_sync(FULL, w.pend);
			uint32_t type = T_COPY;
			bool took = false;
			if (next < n && pm != FULL) {
				const uint32_t freem = ~pm;
				const uint32_t idx = next + __popc(freem & lt_mask);
				took = !w.pend && idx < n;
				if (took) {
					const uint2 r = rec[idx];
					w.len = (r.x >> 16) + 1;
					type = r.y >> 30;
					const uint32_t s = r.y & 0xFFFFFFu;
					const uint32_t out_pos = chunk_base + segbase[(r.y >> 24) & 31] + (r.x & 0xFFFF);
					w.dst = tile0 + out_pos;
					const bool is_copy = type == T_COPY;
					w.off = is_copy ? s : 0u;
					w.src = s;
					// copy: source [dst-off, dst-off+len) must be complete up to dst
					w.need = is_copy ? min(w.dst - s + w.len, w.dst) : 0u;
					w.pend = type != T_LIT_GMEM;
					if (out_pos + w.len > out_size || (is_copy && (s == 0 || s > out_pos))) {
						set_error(sm); // malformed element: never touch memory for it
						w.pend = false;
						type = T_COPY;
					}
				}
				next = min(n, next + (uint32_t)__popc(freem));
				fetched_end = max(fetched_end, __reduce_max_sync(FULL, took ? w.dst + w.len : 0u));
			}
			// ---- long literals: straight from HBM, whole warp, no dependencies -------------
			uint32_t big = __ballot_sync(FULL, took && type == T_LIT_GMEM);
			while (big) {
				const int l = __ffs(big) - 1;
				big &= big - 1;
				long_literal_copy(sm, __shfl_sync(FULL, w.dst, l), gcomp + __shfl_sync(FULL, w.src, l), __shfl_sync(FULL, w.len, l), lane);
			}
			// ---- run every element whose source is complete --------------------------------
			pm = __ballot_sync(FULL, w.pend);
			if (pm) {
				const uint32_t frontier = min(__reduce_min_sync(FULL, w.pend ? w.dst : 0xFFFFFFFFu), fetched_end);
				const bool ready = w.pend && w.need <= frontier;
				uint32_t pos = 0, o = w.off ? w.off : 0x10000u;
				while (__any_sync(FULL, ready && pos < w.len)) {
					if (ready && pos < w.len) {
						const uint32_t nb = min(min(16u, w.len - pos), o);
						const uint32_t sa = w.off ? w.dst + pos - o : w.src + pos;
						copy_upto16(sm, w.dst + pos, sa, nb);
						pos += nb;
						if (o < 16u)
							o <<= 1; // [dst-o, dst+pos) is periodic: twice the distance is as good
					}
				}
				if (ready)
					w.pend = false;
				__syncwarp();
			}
			drained = next >= n && (!last || !__any_sync(FULL, w.pend));
		}
		chunk_base += chunk_out;
		__threadfence_block();
		if (lane == 0)
			vstore(sm, SM_CONSUMED, q + 1);
		if (last)
			break;
	}
	if (chunk_base != out_size && lane == 0)
		set_error(sm);
}

// ---------------------------------------------------------------------------
// Kernel: grid = number of blocks, 96 threads (2 parsers + 1 copier).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(NTHREADS, 5) snappy_decompress_kernel(const uint8_t *__restrict__ comp, const uint64_t *__restrict__ block_off,
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

	if (threadIdx.x < 8)
		reinterpret_cast<uint32_t *>(sm)[threadIdx.x] = 0; // error flag, consumed counter, entry mailboxes
	for (uint32_t t = threadIdx.x; t < 256; t += NTHREADS) {
		const uint32_t kind = t & 3, n6 = t >> 2;
		sm[SM_LUT + t] = (uint8_t)(kind == 0 ? (n6 < 60 ? n6 + 2 : 0) : (kind == 1 ? 2 : (kind == 2 ? 3 : 5)));
	}
	__syncthreads();

	bool header_ok = off1 >= off0 + 5; // 4-byte size + at least one element byte
	uint32_t csize = 0;
	const uint8_t *gcomp = comp + off0 + 4;
	if (header_ok) {
		const uint8_t *h = comp + off0;
		csize = (uint32_t)__ldg(h) | ((uint32_t)__ldg(h + 1) << 8) | ((uint32_t)__ldg(h + 2) << 16) | ((uint32_t)__ldg(h + 3) << 24);
		header_ok = (uint64_t)csize + 4 == off1 - off0;
	}

	if (!header_ok) {
		if (threadIdx.x == 0)
			set_error(sm);
	} else if (warp < NPARSE) {
		parser_warp(sm, gcomp, csize, warp, lane);
	} else {
		copier_warp(sm, gcomp, tile0, out_size, lane);
	}
	__syncthreads();

	if (vload(sm, SM_ERR)) {
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
		for (uint32_t i = threadIdx.x; i < head; i += NTHREADS)
			gdst[i] = sm[tile0 + i];
		const uint4 *s4 = reinterpret_cast<const uint4 *>(sm + tile0 + head);
		uint4 *g4 = reinterpret_cast<uint4 *>(gdst + head);
		for (uint32_t i = threadIdx.x; i < mid / 16; i += NTHREADS)
			g4[i] = s4[i];
		for (uint32_t i = head + mid + threadIdx.x; i < out_size; i += NTHREADS)
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
	snappy_decompress_kernel<<<(unsigned)nblocks, NTHREADS, smem, st>>>(d_comp, d_block_off, block_size, out_len, d_out, d_status);
	prof_end(3, st);
	note_launch();
	PIM_CUDA_TRY(cudaGetLastError());
	return SNAPPY_OK;
}
