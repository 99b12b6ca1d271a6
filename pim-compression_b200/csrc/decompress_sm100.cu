// Block-parallel Snappy decompression for sm_100a.
//
// Replaces the reference's DPU decoder (snappy/dpu-decompress/dpu_decompress.c:224-299,
// dpu_task.c:21-95) and must reproduce what its host twin snappy_decompress_host
// (snappy/snappy_decompress.c:218-289) produces on every valid stream.
//
// ONE WARP PER COMPRESSED BLOCK, four independent warps per CTA, persistent over the grid.
// A block's decode is a chain of a few hundred dependent rounds (a copy can only run once its
// source bytes exist), so throughput comes from having many blocks in flight per SM, not from
// many warps on one block.  Per-warp shared memory is therefore kept small:
//   * a 4 KB RING holds the most recent output.  Elements write into it bytewise; every round
//     the completed 16-byte words below the write frontier are flushed to HBM with coalesced
//     16-byte stores.  Back-references (write_copy_host, snappy_decompress.c:164-184) that reach
//     back further than the ring are read from HBM/L2 (ld.global.cg), which this warp wrote
//     itself a while ago;
//   * the compressed bytes are staged in 1 KB chunks (cp.async, double-buffered).  The serial tag
//     chain of a chunk (an element's length says where the next tag is) is found by 32 lanes
//     WALKING 32-byte segments speculatively from their segment starts, then iterating
//     x[s+1] = exit_s(x[s]) to the fixed point: a lane whose true entry lies on the chain it
//     already walked is done, otherwise it re-walks until it merges with that chain;
//   * elements are executed through a sliding window of 32 in-flight elements, one per lane:
//     an element runs as soon as every output byte below its source end is complete, finished
//     lanes are refilled, so a copy that depends on its neighbours delays only itself.
//     Self-overlapping matches (offset < length) are widened in-lane by doubling;
//   * literals longer than 60 bytes go HBM -> HBM directly.
//
// Element formats: snappy/README.md:19-33, snappy_decompress.c:244-283, dpu_snappy.h:10-13.

#include <cstring>
#include <mutex>

#include "common.cuh"

namespace pimsnappy {
namespace {

#ifndef PIM_DEC_WARPS
#define PIM_DEC_WARPS 4
#endif
#ifndef PIM_DEC_RING
#define PIM_DEC_RING 4096
#endif
#ifndef PIM_DEC_MINB
#define PIM_DEC_MINB 6
#endif
constexpr int WARPS = PIM_DEC_WARPS;     // independent warps (= blocks in flight) per CTA
constexpr uint32_t RING = PIM_DEC_RING;  // bytes of recent output kept in shared memory
constexpr uint32_t RMASK = RING - 1;
constexpr uint32_t SEGB = 32;            // compressed bytes per lane
constexpr uint32_t CHUNK = 32 * SEGB;    // compressed bytes per parse step
constexpr uint32_t TAIL = 80;            // staged look-ahead past the chunk (tag + length bytes + short literal payload)
constexpr uint32_t STAGE_BYTES = CHUNK + TAIL + 16;
constexpr uint32_t MAX_REC = CHUNK / 2;  // an element occupies >= 2 compressed bytes
constexpr uint32_t SHORT_LIT = 60;       // literals up to this long are served from the staged bytes
#ifndef PIM_DEC_RUNAHEAD
#define PIM_DEC_RUNAHEAD 256 // (a short run-ahead leaves more history in the ring: fewer back-references go to L2.  Measured on B200, decode GB/s
                             // corpus / text / mix: 1920: 128.5 / 203.2 / 322.8, 1024: 128.8 / 203.1 / 323.2, 512: 134.1 / 204.5 / 324.8, 256: 141.4 / 202.6 / 321.3,
                             // 128: 141.8 / 197.7 / 316.2, 64: 138.7 / 191.9 / 306.2)
#endif
constexpr uint32_t RUNAHEAD = PIM_DEC_RUNAHEAD; // the window stops fetching when this far ahead of the frontier
constexpr uint32_t KEEP = 80;            // bytes of a long literal's end that are also put into the ring
// One refill fetches up to 32 elements of at most 64 bytes on top of the run-ahead: everything between the
// write frontier and the last fetched byte must fit the ring (measured on B200: a 2 KB ring with 7 or 8 CTAs
// per SM is only ~3 % faster on SYN-MIX and breaks this bound, so the ring stays at 4 KB).
static_assert((RING & (RING - 1)) == 0 && RING >= RUNAHEAD + 32 * 64 + 64, "ring too small for the element window");

// per-warp shared memory (byte offsets from the warp's base)
constexpr uint32_t WS_RING = 0;                         // RING + 32 (over-read slack)
constexpr uint32_t WS_STAGE = RING + 32;                // 2 x STAGE_BYTES
constexpr uint32_t WS_REC = WS_STAGE + 2 * STAGE_BYTES; // u32[MAX_REC]: cpos | local_out << 16
constexpr uint32_t WS_SEGBASE = WS_REC + 4 * MAX_REC;   // u32[32]
constexpr uint32_t WS_BYTES = WS_SEGBASE + 128;
static_assert(WS_STAGE % 16 == 0 && WS_REC % 16 == 0 && WS_BYTES % 16 == 0, "alignment");
// CTA-shared
constexpr uint32_t SM_LUT_FOOT = 0;   // u8[256]: compressed footprint by tag byte (0 = long literal)
constexpr uint32_t SM_LUT_LEN = 256;  // u8[256]: output length by tag byte (0 = long literal)
constexpr uint32_t SM_WARP0 = 512;
constexpr uint32_t SMEM_BYTES = SM_WARP0 + WARPS * WS_BYTES;

constexpr uint32_t BIG_FOOT = 0x100000u;

__device__ __forceinline__ uint32_t ldcg32(const uint32_t *p)
{
	uint32_t v;
	asm volatile("ld.global.cg.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ uint32_t ldcg8(const uint8_t *p)
{
	uint32_t v;
	asm volatile("ld.global.cg.u8 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}

// footprint / output length of a long literal (tag>>2 >= 60) whose tag is at shared-memory
// offset so = compressed position cp.  BIG_FOOT / 0 if malformed.
__device__ __forceinline__ uint32_t long_literal(const uint8_t *sm, uint32_t so, uint32_t cp, uint32_t csize, uint32_t *len)
{
	const uint32_t v = lds32_unaligned(sm, so);
	const uint32_t nb = ((v & 0xFF) >> 2) - 59;
	*len = 0;
	if (cp + 1 + nb > csize)
		return BIG_FOOT;
	const uint32_t lm1 = (nb == 4) ? ((v >> 8) | (uint32_t(sm[so + 4]) << 24)) : ((v >> 8) & ((1u << (8 * nb)) - 1));
	if (lm1 >= 65536u)
		return BIG_FOOT;
	*len = lm1 + 1;
	return 2 + nb + lm1;
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
// Per-block decoder state (one warp).  Output positions are offsets in the block's output.
// ---------------------------------------------------------------------------
struct Dec {
	uint8_t *sm;          // CTA shared memory
	uint8_t *ws;          // this warp's shared memory
	const uint8_t *gcomp; // first element byte of the block
	uint32_t csize;       // compressed bytes
	uint8_t *gdst;        // block's output in HBM
	uint32_t out_size;
	uint32_t g15;         // gdst & 15: ring offset of output byte p is (p + g15) & RMASK
	int lane;
	bool err;
	// write-back state
	uint32_t fw;          // 16-byte words (of gdst - g15) flushed so far
	uint32_t ring_lo;     // output bytes below this are not in the ring (they went to HBM directly)
	bool runs;            // (warp-uniform) some pending element continues a run of copies
};

__device__ __forceinline__ uint32_t ring_off(const Dec &d, uint32_t p) { return (p + d.g15) & RMASK; }
__device__ __forceinline__ uint32_t flushed_pos(const Dec &d) { return d.fw * 16u > d.g15 ? d.fw * 16u - d.g15 : 0u; }

// Flush the complete 16-byte words below output position `upto` from the ring to HBM.
__device__ __forceinline__ void flush_words(Dec &d, uint32_t upto)
{
	const uint32_t target = (upto + d.g15) >> 4;
	uint8_t *gal = d.gdst - d.g15;
	for (uint32_t w = d.fw + d.lane; w < target; w += 32) {
		const uint4 v = *reinterpret_cast<const uint4 *>(d.ws + WS_RING + ((w << 4) & RMASK));
		if (w == 0 && d.g15) { // the block starts inside this word
			const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
			for (uint32_t k = d.g15; k < 16; k++)
				gal[k] = (uint8_t)(vv[k >> 2] >> (8 * (k & 3)));
		} else {
			*reinterpret_cast<uint4 *>(gal + 16ull * w) = v;
		}
	}
	if (target > d.fw)
		d.fw = target;
}

// ---------------------------------------------------------------------------
// Parse one staged chunk: find the element starts.  Positions are relative to the chunk.
//   stage : shared-memory offset (in d.ws) of the chunk's first byte
//   cs    : compressed position of the chunk's first byte;  x0: entry (first element start)
// Returns the number of elements; fills WS_REC / WS_SEGBASE; *chunk_out = bytes they produce,
// *next_cp = compressed position of the first element after the chunk.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t foot_at(const Dec &d, uint32_t stage, uint32_t p, uint32_t cs)
{
	uint32_t f = d.sm[SM_LUT_FOOT + d.ws[stage + p]];
	if (f == 0) {
		uint32_t len;
		f = long_literal(d.ws, stage + p, cs + p, d.csize, &len);
	}
	return f;
}

__device__ __forceinline__ uint32_t parse_chunk(Dec &d, uint32_t stage, uint32_t cs, uint32_t x0, uint32_t *chunk_out, uint32_t *next_cp)
{
	const int lane = d.lane;
	const uint32_t csz = d.csize - cs;             // bytes from the chunk start to the block's end
	const uint32_t base = SEGB * lane;
	const uint32_t lim = min(base + SEGB, csz);    // my segment is [base, lim)

	// ---- speculative walk from the segment start ------------------------------------------
	uint32_t V = 0;      // bit b: an element of my chain starts at base + b
	uint32_t exitA = base;
	{
		uint32_t p = base;
		while (p < lim) {
			V |= 1u << (p - base);
			p += foot_at(d, stage, p, cs);
		}
		exitA = p;
	}
	// ---- fixed point over the lanes --------------------------------------------------------
	uint32_t x = lane == 0 ? x0 : base; // first element start at or after my segment
	uint32_t eval_x = base, r = 0;
	uint32_t f_out = 0;
	for (int it = 0; it < 34; it++) {
		const bool touched = x < lim;  // (x >= base always)
		if (touched && x != eval_x) {
			eval_x = x;
			r = x - base;
			if (!((V >> r) & 1u)) { // not on the chain I know: walk until it merges with it
				uint32_t p = x, nb = 0;
				while (p < lim && !((V >> (p - base)) & 1u)) {
					nb |= 1u << (p - base);
					p += foot_at(d, stage, p, cs);
				}
				if (p < lim) {
					V = nb | (V & ~((1u << (p - base)) - 1u));
				} else {
					V = nb;
					exitA = p;
				}
			}
		}
		f_out = touched ? exitA : x;
		uint32_t nx = __shfl_up_sync(FULL, f_out, 1);
		if (lane == 0)
			nx = x0;
		const bool changed = nx != x;
		x = nx;
		if (!__any_sync(FULL, changed))
			break;
	}
	const bool touched = x < lim;
	uint32_t bits = touched ? (V & ~((1u << r) - 1u)) : 0u;
	const uint32_t nx_all = __shfl_sync(FULL, f_out, 31);
	*next_cp = nx_all >= BIG_FOOT ? 0xFFFFFFFFu : cs + nx_all;

	// ---- records: compressed position + output offset inside my segment ---------------------
	const uint32_t cnt = __popc(bits);
	const uint32_t cnt_incl = warp_inclusive_scan(cnt, lane);
	uint32_t *rec = reinterpret_cast<uint32_t *>(d.ws + WS_REC) + (cnt_incl - cnt);
	uint32_t local = 0;
	while (bits) {
		const uint32_t b = __ffs(bits) - 1;
		bits &= bits - 1;
		const uint32_t p = base + b;
		uint32_t len = d.sm[SM_LUT_LEN + d.ws[stage + p]];
		*rec++ = p | (local << 16);
		if (len == 0) { // long literal: always the last element of a segment
			long_literal(d.ws, stage + p, cs + p, d.csize, &len);
			if (len == 0)
				d.err = true;
		}
		local += len;
	}
	const uint32_t out_incl = warp_inclusive_scan(local, lane);
	reinterpret_cast<uint32_t *>(d.ws + WS_SEGBASE)[lane] = out_incl - local;
	*chunk_out = __shfl_sync(FULL, out_incl, 31);
	__syncwarp();
	return __shfl_sync(FULL, cnt_incl, 31);
}

// ---------------------------------------------------------------------------
// Copy engine.
// ---------------------------------------------------------------------------
constexpr uint32_t K_LIT = 0;  // literal, payload in the staged chunk (src = shared-memory offset in d.ws)
constexpr uint32_t K_COPY = 1; // back-reference (off = distance)
constexpr uint32_t K_LONG = 2; // literal longer than SHORT_LIT (src = compressed position of the payload)

struct Window { // one in-flight element per lane
	bool pend;
	uint32_t kind;
	uint32_t dst; // output position of the element's first byte
	uint32_t len;
	uint32_t off; // K_COPY: distance
	uint32_t src; // K_LIT / K_LONG
	uint32_t base; // K_COPY continuing a run of copies with the same distance: output position where the run started, else NO_RUN
};
constexpr uint32_t NO_RUN = 0xFFFFFFFFu;

// sixteen bytes at an arbitrary shared-memory byte offset (reads 20 bytes from off & ~3)
__device__ __forceinline__ void lds128_unaligned(const uint8_t *sm, uint32_t off, uint32_t v[4])
{
	const uint32_t a = off & ~3u;
	const uint32_t sh = (off & 3u) * 8u;
	const uint32_t w0 = lds32(sm, a), w1 = lds32(sm, a + 4), w2 = lds32(sm, a + 8), w3 = lds32(sm, a + 12), w4 = lds32(sm, a + 16);
	v[0] = __funnelshift_r(w0, w1, sh);
	v[1] = __funnelshift_r(w1, w2, sh);
	v[2] = __funnelshift_r(w2, w3, sh);
	v[3] = __funnelshift_r(w3, w4, sh);
}

// sixteen bytes at an arbitrary global address, through L2 (the warp wrote them earlier)
__device__ __forceinline__ void ldg128_unaligned_cg(const uint8_t *g, uint32_t v[4])
{
	const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(g) & 3u);
	const uint32_t sh = mis * 8u;
	const uint32_t *a = reinterpret_cast<const uint32_t *>(g - mis);
	const uint32_t w0 = ldcg32(a), w1 = ldcg32(a + 1), w2 = ldcg32(a + 2), w3 = ldcg32(a + 3);
	const uint32_t w4 = mis ? ldcg32(a + 4) : 0u;
	v[0] = __funnelshift_r(w0, w1, sh);
	v[1] = __funnelshift_r(w1, w2, sh);
	v[2] = __funnelshift_r(w2, w3, sh);
	v[3] = __funnelshift_r(w3, w4, sh);
}

// n <= 16 bytes of v to ring offset ro (no wrap: ro + n <= RING)
__device__ __forceinline__ void ring_store(uint8_t *ring, uint32_t ro, const uint32_t v[4], uint32_t n)
{
#pragma unroll
	for (int k = 0; k < 16; k++) {
		if (k < (int)n)
			ring[ro + k] = (uint8_t)(v[k >> 2] >> (8 * (k & 3)));
	}
}

// Byte-serial copy of one element: any source (ring or HBM per byte), wrapping ring addresses,
// overlap handled by the order.  Needs every output byte below the element's source end complete.
__device__ __forceinline__ void slow_copy(const Dec &d, const Window &w, uint32_t ring_floor)
{
	uint8_t *ring = d.ws + WS_RING;
	for (uint32_t k = 0; k < w.len; k++) {
		uint32_t b;
		if (w.kind == K_LIT) {
			b = d.ws[w.src + k];
		} else {
			const uint32_t p = w.dst + k - w.off;
			b = p >= ring_floor ? ring[ring_off(d, p)] : ldcg8(d.gdst + p);
		}
		ring[ring_off(d, w.dst + k)] = (uint8_t)b;
	}
}

// A copy that continues a run of back-to-back copies with the same distance o (zero pages, repeated records:
// every such element would otherwise wait for the one before it).  From the run's start on the output is
// periodic with period o, so the element is read from the o bytes before the run's start, whatever its own
// position: out[dst + k] = out[base - o + ((dst - base + k) mod o)].  Needs only the output below `base`.
__device__ __forceinline__ void run_copy(const Dec &d, const Window &w)
{
	uint8_t *ring = d.ws + WS_RING;
	const uint32_t pat = w.base - w.off;
	uint32_t idx = (w.dst - w.base) % w.off;
	for (uint32_t k = 0; k < w.len; k++) {
		ring[ring_off(d, w.dst + k)] = ring[ring_off(d, pat + idx)];
		idx = idx + 1 == w.off ? 0u : idx + 1;
	}
}

// Long literal (K_LONG) at output position p: everything before it is complete and flushed up to
// the last full word.  The first bytes (to the next 16-byte boundary of HBM) and the last KEEP+
// bytes go through the ring, the aligned middle goes HBM -> HBM.
__device__ __forceinline__ void long_literal_copy(Dec &d, uint32_t p, const uint8_t *g, uint32_t len)
{
	const int lane = d.lane;
	uint8_t *ring = d.ws + WS_RING;
	uint32_t head = (16u - ((p + d.g15) & 15u)) & 15u;
	if (head > len)
		head = len;
	uint32_t tail = len - head;                       // bytes after the head
	uint32_t body = 0;
	if (tail > KEEP + 16) {
		body = (tail - KEEP) & ~15u;                  // whole words that bypass the ring
		tail -= body;
	}
	for (uint32_t k = lane; k < head; k += 32)
		ring[ring_off(d, p + k)] = __ldg(g + k);
	__syncwarp();
	flush_words(d, p + head);                         // the word the literal started in is complete now
	if (body) {
		const uint8_t *s = g + head;
		const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(s) & 3u);
		const uint32_t sh = mis * 8u;
		const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s - mis);
		uint4 *d4 = reinterpret_cast<uint4 *>(d.gdst + p + head);
		const uint32_t nchunk = body / 16;
#pragma unroll 4
		for (uint32_t c = lane; c < nchunk; c += 32) {
			const uint32_t w0 = __ldg(s32 + 4 * c), w1 = __ldg(s32 + 4 * c + 1), w2 = __ldg(s32 + 4 * c + 2), w3 = __ldg(s32 + 4 * c + 3);
			const uint32_t w4 = mis ? __ldg(s32 + 4 * c + 4) : 0u;
			d4[c] = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
		}
		d.fw += nchunk;
		d.ring_lo = p + head + body;
	}
	for (uint32_t k = lane; k < tail; k += 32)
		ring[ring_off(d, p + head + body + k)] = __ldg(g + head + body + k);
	__syncwarp();
}

// Decode the element whose tag is at chunk position p (snappy_decompress.c:242-283, :64-133).
// Returns false if malformed.  w.src for literals: shared-memory offset (K_LIT) / compressed
// position (K_LONG) of the payload.
__device__ __forceinline__ bool decode_at(const Dec &d, uint32_t stage, uint32_t cs, uint32_t p, Window &w)
{
	const uint32_t csz = d.csize - cs;
	const uint32_t v = lds32_unaligned(d.ws, stage + p); // tag + next three bytes
	const uint32_t tag = v & 0xFF;
	const uint32_t kind = tag & 3;
	if (kind == 0) {
		const uint32_t n6 = tag >> 2;
		uint32_t hdr = 1;
		if (n6 < 60) {
			w.len = n6 + 1;
		} else {
			const uint32_t nb = n6 - 59;
			hdr = 1 + nb;
			if (p + hdr > csz)
				return false;
			const uint32_t lm1 = (nb == 4) ? ((v >> 8) | (uint32_t(d.ws[stage + p + 4]) << 24)) : ((v >> 8) & ((1u << (8 * nb)) - 1));
			if (lm1 >= 65536u)
				return false;
			w.len = lm1 + 1;
		}
		if (w.len > csz - p - hdr)
			return false;
		w.off = 0;
		if (w.len <= SHORT_LIT) { // p < CHUNK, hdr + len <= 65: inside the staged look-ahead
			w.kind = K_LIT;
			w.src = stage + p + hdr;
		} else {
			w.kind = K_LONG;
			w.src = cs + p + hdr;
		}
		return true;
	}
	w.kind = K_COPY;
	w.src = 0;
	uint32_t foot;
	if (kind == 1) {
		foot = 2;
		w.len = ((tag >> 2) & 7) + 4;
		w.off = ((tag >> 5) << 8) | ((v >> 8) & 0xFF);
	} else if (kind == 2) {
		foot = 3;
		w.len = (tag >> 2) + 1;
		w.off = (v >> 8) & 0xFFFF;
	} else {
		foot = 5;
		w.len = (tag >> 2) + 1;
		if (p + 5 > csz)
			return false;
		w.off = (v >> 8) | (uint32_t(d.ws[stage + p + 4]) << 24);
		if (w.off > 65535u)
			return false;
	}
	return p + foot <= csz;
}

// One window round: run every pending element whose source is complete.  fetched_end = output
// position just past the last fetched element.
__device__ __forceinline__ void window_round(Dec &d, Window &w, uint32_t fetched_end)
{
	uint8_t *ring = d.ws + WS_RING;
	const uint32_t F = min(__reduce_min_sync(FULL, w.pend ? w.dst : 0xFFFFFFFFu), fetched_end); // everything below is complete
	flush_words(d, F);
	__syncwarp();
	const uint32_t fl = flushed_pos(d);
	const uint32_t ring_floor = max(d.ring_lo, fetched_end > RING ? fetched_end - RING : 0u); // the ring holds [ring_floor, F)
	const bool is_copy = w.kind == K_COPY;
	const uint32_t src_a = w.dst - w.off;
	const uint32_t src_e = is_copy ? min(src_a + w.len, w.dst) : 0u;
	// a run member is served from the pattern before its run as soon as that is complete and in the ring
	if (d.runs) {
		const bool by_run = w.pend && is_copy && w.base != NO_RUN && w.base <= F && w.base - w.off >= ring_floor && !(src_e <= F);
		if (by_run) {
			run_copy(d, w);
			w.pend = false;
		}
		d.runs = __any_sync(FULL, w.pend && w.base != NO_RUN);
	}
	const bool ready = w.pend && src_e <= F;
	// which path
	const uint32_t dro = ring_off(d, w.dst);
	const bool dst_flat = dro + w.len <= RING;
	bool fast = dst_flat, far = false;
	uint32_t sro = 0;
	if (is_copy) {
		if (src_a >= ring_floor) {
			sro = ring_off(d, src_a);
			fast = fast && sro + w.len <= RING; // (for an overlapping copy the bytes read stay inside [sro, dro + len))
		} else {
			far = true;
			fast = fast && w.off >= w.len && src_a + w.len <= fl;
		}
	}
	if (ready && !fast)
		slow_copy(d, w, ring_floor);
	uint32_t pos = 0, o = is_copy ? w.off : 0x10000u;
	while (__any_sync(FULL, ready && fast && pos < w.len)) {
		if (ready && fast && pos < w.len) {
			const uint32_t nb = min(min(16u, w.len - pos), o);
			uint32_t v[4];
			if (!is_copy)
				lds128_unaligned(d.ws, w.src + pos, v);
			else if (!far)
				lds128_unaligned(ring, sro + pos + w.off - o, v); // = ring offset of (dst + pos - o)
			else
				ldg128_unaligned_cg(d.gdst + src_a + pos, v);
			ring_store(ring, dro + pos, v, nb);
			pos += nb;
			if (o < 16u)
				o <<= 1; // [dst-o, dst+pos) is periodic: twice the distance is as good
		}
	}
	if (ready)
		w.pend = false;
	__syncwarp();
}

// Decode one block (one warp).
__device__ __forceinline__ void decode_block(Dec &d)
{
	const int lane = d.lane;
	const uint32_t lt_mask = (1u << lane) - 1u;
	const uint32_t amis = (uint32_t)(reinterpret_cast<uintptr_t>(d.gcomp) & 15u);
	Window w;
	w.pend = false;
	w.kind = K_LIT;
	w.dst = w.len = w.off = w.src = 0;
	w.base = NO_RUN;
	uint32_t fetched_end = 0; // output position just past the last fetched element
	uint32_t chunk_base = 0;  // output bytes of the chunks before this one
	uint32_t cp = 0;
	int buf = 0;
	d.fw = 0;
	d.ring_lo = 0;
	d.runs = false;

	auto stage_chunk = [&](int b, uint32_t cs) {
		const uint32_t nstage = cs < d.csize ? min(CHUNK + TAIL, d.csize - cs) : 0u;
		stage_bytes_async(d.ws + WS_STAGE + b * STAGE_BYTES + amis, d.gcomp + cs, nstage, lane, 32);
		cp_async_commit();
	};
	uint32_t staged_cs = 0; // chunk in flight / present in `buf`
	stage_chunk(0, 0);

	for (;;) {
		const uint32_t cs = (cp / CHUNK) * CHUNK;
		if (cs != staged_cs) { // a long literal jumped over the prefetched chunk
			cp_async_wait<0>();
			__syncwarp();
			stage_chunk(buf, cs);
			staged_cs = cs;
		}
		cp_async_wait<0>();
		__syncwarp();
		const uint32_t stage = WS_STAGE + buf * STAGE_BYTES + amis;
		uint32_t chunk_out, next_cp;
		const uint32_t n = parse_chunk(d, stage, cs, cp - cs, &chunk_out, &next_cp);
		// prefetch the next chunk into the other buffer (its previous content is no longer needed:
		// literals of the chunk before this one were executed when they were fetched)
		stage_chunk(buf ^ 1, cs + CHUNK);
		staged_cs = cs + CHUNK;

		const uint32_t *rec = reinterpret_cast<const uint32_t *>(d.ws + WS_REC);
		const uint32_t *segbase = reinterpret_cast<const uint32_t *>(d.ws + WS_SEGBASE);
		uint32_t next = 0;
		while (next < n) {
			// ---- refill free lanes with the next records ------------------------------------
			const uint32_t pm = __ballot_sync(FULL, w.pend);
			const uint32_t F0 = min(__reduce_min_sync(FULL, w.pend ? w.dst : 0xFFFFFFFFu), fetched_end);
			bool took = false;
			if (pm != FULL && fetched_end - F0 < RUNAHEAD) {
				const uint32_t freem = ~pm;
				const uint32_t idx = next + __popc(freem & lt_mask);
				took = !w.pend && idx < n;
				if (took) {
					const uint32_t r = rec[idx];
					const uint32_t p = r & 0xFFFFu;
					w.dst = chunk_base + segbase[p / SEGB] + (r >> 16);
					bool ok = decode_at(d, stage, cs, p, w);
					ok = ok && w.dst + w.len <= d.out_size && (w.kind != K_COPY || (w.off != 0 && w.off <= w.dst));
					if (!ok) {
						d.err = true; // malformed element: never touch memory for it
						w.kind = K_LIT;
						w.len = 0;
					}
					w.pend = ok && w.kind != K_LONG;
				}
				// runs: records of one refill are consecutive elements in lane order, so an element continues a run
				// iff the closest lower fetching lane holds a copy with the same distance
				if (took)
					w.base = NO_RUN;
				if (__any_sync(FULL, took && w.pend && w.kind == K_COPY && w.off < w.len)) { // (only runs of self-overlapping copies are looked for)
					const uint32_t tookm = __ballot_sync(FULL, took);
					const bool cp = took && w.pend && w.kind == K_COPY;
					const uint32_t lowt = tookm & lt_mask;
					const int prev = lowt ? 31 - __clz(lowt) : lane;
					const uint32_t key = cp ? w.off : 0u; // 0 is never a valid distance
					const uint32_t pkey = __shfl_sync(FULL, key, prev);
					const bool chained = cp && lowt != 0u && pkey == key;
					const uint32_t heads = __ballot_sync(FULL, took && !chained) & (lt_mask | (1u << lane));
					const uint32_t hdst = __shfl_sync(FULL, w.dst, heads ? 31 - __clz(heads) : lane);
					if (took)
						w.base = chained ? hdst : NO_RUN;
					d.runs = d.runs || __any_sync(FULL, chained);
				}
				next = min(n, next + (uint32_t)__popc(freem));
				fetched_end = max(fetched_end, __reduce_max_sync(FULL, took ? w.dst + w.len : 0u));
			}
			// ---- long literals: drain the window, then HBM -> HBM -----------------------------
			uint32_t big = __ballot_sync(FULL, took && w.kind == K_LONG);
			if (big) {
				const int l = __ffs(big) - 1; // (at most one per refill would be typical; handle the first, refetch the rest)
				const uint32_t lp = __shfl_sync(FULL, w.dst, l), ll = __shfl_sync(FULL, w.len, l), ls = __shfl_sync(FULL, w.src, l);
				// elements fetched after it must wait for it: un-fetch them, run everything
				// fetched before the literal to completion
				const bool later = took && w.dst > lp;
				const uint32_t nlater = __popc(__ballot_sync(FULL, later));
				if (later)
					w.pend = false;
				next -= nlater; // they are refetched after the literal
				fetched_end = lp;
				while (__any_sync(FULL, w.pend))
					window_round(d, w, fetched_end);
				long_literal_copy(d, lp, d.gcomp + ls, ll);
				fetched_end = lp + ll;
				continue;
			}
			window_round(d, w, fetched_end);
		}
		chunk_base += chunk_out;
		const bool last = next_cp >= d.csize;
		if ((last && next_cp != d.csize) || next_cp <= cp)
			d.err = true; // the last element overruns the block, or the chain does not advance
		if (last || next_cp <= cp)
			break;
		cp = next_cp;
		buf ^= 1;
	}
	cp_async_wait<0>();
	while (__any_sync(FULL, w.pend))
		window_round(d, w, fetched_end);
	if (chunk_base != d.out_size || fetched_end != d.out_size)
		d.err = true;
	if (__any_sync(FULL, d.err))
		return;
	// final flush: whole words, then the last partial word bytewise
	flush_words(d, d.out_size);
	for (uint32_t p = flushed_pos(d) + lane; p < d.out_size; p += 32)
		d.gdst[p] = d.ws[WS_RING + ring_off(d, p)];
}

// ---------------------------------------------------------------------------
// Kernel: persistent CTAs of WARPS independent warps over the concatenation of the batch's streams' blocks (one launch
// decodes up to DEC_MAX_BATCH streams, pim_snappy_decompress_batch_device, so that no SM idles between streams).
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

#ifndef PIM_DEC_QUEUE
#define PIM_DEC_QUEUE 0 // block order of the kernel below: 0 fixed stride, 2 stride + counter for the tail
#endif
#if PIM_DEC_QUEUE != 0
// the next index of the block counter, for the whole warp
__device__ __noinline__ unsigned long long pull_block(unsigned long long *counter)
{
	unsigned long long g = 0;
	if (lane_id() == 0)
		g = atomicAdd(counter, 1ull);
	return __shfl_sync(FULL, g, 0);
}
#endif

__global__ void __launch_bounds__(32 * WARPS, PIM_DEC_MINB) snappy_decompress_kernel(const __grid_constant__ DecArgs a)
{
	extern __shared__ __align__(16) uint8_t sm[];
	const int lane = lane_id();
	const int warp = threadIdx.x >> 5;

	for (uint32_t t = threadIdx.x; t < 256; t += 32 * WARPS) {
		const uint32_t kind = t & 3, n6 = t >> 2;
		sm[SM_LUT_FOOT + t] = (uint8_t)(kind == 0 ? (n6 < 60 ? n6 + 2 : 0) : (kind == 1 ? 2 : (kind == 2 ? 3 : 5)));
		sm[SM_LUT_LEN + t] = (uint8_t)(kind == 0 ? (n6 < 60 ? n6 + 1 : 0) : (kind == 1 ? (n6 & 7) + 4 : n6 + 1));
	}
	__syncthreads();

	Dec d;
	d.sm = sm;
	d.ws = sm + SM_WARP0 + warp * WS_BYTES;
	d.lane = lane;
	// Block order.  Default (PIM_DEC_QUEUE 0): warp g of the grid decodes blocks g, g + warps, ... of the batch.
	// PIM_DEC_QUEUE 2: all but the last full round that way, the rest pulled from a counter in global memory, so that a
	// warp that drew slow blocks does not set the tail (the reference hands each tasklet a fixed contiguous run,
	// dpu-decompress/dpu_task.c:42-77).  Measured on B200 in round 2 (profiles/r2_decoder_experiments.md): every loop
	// shape that contains the counter makes the compiler lay the decoder out differently (12,100 instead of 9,900 SASS
	// lines, twice the local-memory accesses) and costs 5-15 % on uniform data (SYN-MIX 320-357 vs 375 GB/s, text 191 vs
	// 213) for 3 % on the tiled fixtures (152 vs 147), so the plain stride stays the default.
	const uint64_t total_warps = (uint64_t)gridDim.x * WARPS;
#if PIM_DEC_QUEUE == 0
	for (uint64_t gblk = (uint64_t)blockIdx.x * WARPS + warp; gblk < a.nblocks; gblk += total_warps) {
#else
	const uint64_t rounds = a.nblocks / total_warps;
	const uint64_t static_rounds = rounds > 1 ? rounds - 1 : 1;
	uint64_t done = 0; // rounds this warp has finished
	for (uint64_t gblk = (uint64_t)blockIdx.x * WARPS + warp; gblk < a.nblocks;
	     gblk = ++done < static_rounds ? gblk + total_warps : static_rounds * total_warps + pull_block(a.counter)) {
#endif
		uint32_t si = 0;
		while (si + 1 < a.n_streams && gblk >= a.s[si + 1].first_blk)
			si++;
		const DecStream &ds = a.s[si];
		const uint64_t blk = gblk - ds.first_blk;
		const uint64_t off0 = ds.block_off[blk], off1 = ds.block_off[blk + 1];
		const uint64_t out_off = blk * (uint64_t)a.block_size;
		d.out_size = (uint32_t)min((uint64_t)a.block_size, ds.out_len - out_off);
		d.gdst = ds.out + out_off;
		d.g15 = (uint32_t)(reinterpret_cast<uintptr_t>(d.gdst) & 15u);
		d.gcomp = ds.comp + off0 + 4;
		d.err = false;
		bool header_ok = off1 >= off0 + 5; // 4-byte size + at least one element byte
		d.csize = 0;
		if (header_ok) {
			const uint8_t *h = ds.comp + off0;
			d.csize = (uint32_t)__ldg(h) | ((uint32_t)__ldg(h + 1) << 8) | ((uint32_t)__ldg(h + 2) << 16) | ((uint32_t)__ldg(h + 3) << 24);
			header_ok = (uint64_t)d.csize + 4 == off1 - off0;
		}
		if (header_ok)
			decode_block(d);
		else
			d.err = true;
		if (__any_sync(FULL, d.err) && lane == 0)
			atomicOr(a.status + si, 1u);
		__syncwarp();
	}
}

// Block counters: one 8-byte slot per launch, taken round robin from a per-device pool and zeroed on the launch's
// stream right before the kernel (POOL launches would have to be in flight for a clash).
constexpr int POOL = 4096;
struct DecDevice {
	unsigned long long *counters = nullptr;
	unsigned next = 0;
	int max_ctas = 0;
};
std::mutex g_dec_mu;
DecDevice g_dec[64];

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
	const int launches = (n_streams + (int)DEC_MAX_BATCH - 1) / (int)DEC_MAX_BATCH;
	if (dev >= 64 || launches > POOL)
		return SNAPPY_INVALID_INPUT;
	unsigned long long *counters = nullptr;
	int max_ctas = 0;
	{
		std::lock_guard<std::mutex> lk(g_dec_mu);
		DecDevice &dd = g_dec[dev];
		if (!dd.max_ctas) {
			PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
			PIM_CUDA_TRY(cudaFuncSetAttribute(snappy_decompress_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
			int per_sm = 0, sms = 0;
			PIM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, snappy_decompress_kernel, 32 * WARPS, SMEM_BYTES));
			PIM_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
			dd.max_ctas = per_sm * sms > 0 ? per_sm * sms : 1;
			if (PIM_DEC_QUEUE != 0)
				PIM_CUDA_TRY(cudaMalloc(&dd.counters, sizeof(unsigned long long) * POOL));
		}
		if (dd.next + launches > (unsigned)POOL)
			dd.next = 0;
		counters = dd.counters ? dd.counters + dd.next : nullptr;
		dd.next += launches;
		max_ctas = dd.max_ctas;
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
		if (counters) {
			da.counter = counters + launch_i++;
			PIM_CUDA_TRY(cudaMemsetAsync(da.counter, 0, sizeof(unsigned long long), st));
		}
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
