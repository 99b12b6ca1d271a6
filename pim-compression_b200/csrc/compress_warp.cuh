// Per-warp Snappy block compressor: one warp compresses one input block and reproduces, bit for
// bit, the reference's HOST compressor compress_block (snappy/snappy_compress.c:284-413): same
// hash (:161-166), same skip schedule (:333-348), same order of table updates (:391-397), same
// element encodings (:202-272).
//
// The reference loop alternates   step 1: probe positions until a 4-byte match is found
//                                 step 2: emit the pending literal
//                                 step 3: emit the copy, insert ip-1, probe ip; repeat while it matches.
// All three are driven here by ONE kind of pass over 32 lanes:
//
//   * "ic" layout (after a match that ended at q):  lane 0 = insert-only position q-1 (:388-389),
//     lane 1 = the step-3 continuation probe at q (:391-394), lane 2+i = step-1 probe i at q+1+i
//     (the first 32 probes of step 1 have stride 1, :333-340);
//   * "probe" layout (block start, or a step 1 that needs more than the ic pass covers):
//     lane i = step-1 probe k0+i at its position in the reference's schedule (stride (32+k)>>5).
//
// Every lane loads the 16 bytes at its position, hashes the first four, reads its table slot
// and compares 16 bytes at the candidate, so a lane knows whether it hits AND its match length
// up to 16 without another shared-memory round trip.  Table updates that earlier lanes of the
// same pass would have made are forwarded with __match_any_sync (the candidate of a lane is the
// closest lower lane with the same hash that is really inserted, else the table's entry).
//
// In a dense pass (all lanes at consecutive positions) the warp then resolves SEVERAL
// (literal, copy) pairs without touching shared memory again: first hit lane f, match length m
// -> the next pair's insert-only lane is f+m-1, its continuation probe f+m, its probes follow;
// lanes inside the match are skipped (never inserted).  One table update (last writer per slot
// wins) closes the pass.  Matches of 16 bytes or more are extended by the whole warp (16 bytes
// per lane per step).  All literals and copy tags found in a pass are written by one batched
// emission (prefix sums over ballot masks) -- by a SECOND warp, fed through a ring of per-pass
// records in shared memory, so that emission is off the parser's dependency chain.
//
// This header is written against warp_prims.cuh so that the identical source also runs on the
// host lock-step emulator (tests/emu) against the oracle.
#pragma once

#include "warp_prims.cuh"

#ifdef PIM_WARP_EMU_STATS // emulator-only counters (tests/emu): passes, pairs, ... per run
namespace pimsnappy { extern unsigned long cw_stats[8]; }
#define CW_STAT(i) do { if (wp::lane() == 0) pimsnappy::cw_stats[i]++; } while (0)
#else
#define CW_STAT(i) do { } while (0)
#endif

// Development aid (-DPIM_CW_PROFILE, device only): cycles of the parser's pass, by segment, summed over all passes of all
// shared-memory-table pairs into g_cw_prof (read with pim_snappy_debug_cw_prof, compress_sm100.cu).
#if defined(PIM_CW_PROFILE) && !defined(PIM_WARP_EMU)
// (g_cw_prof is defined by the including .cu file before this header)
#define CW_T_DECL long long cw_tprev = clock64(); unsigned long long cw_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; unsigned long long cw_passes = 0
#define CW_T(i) do { const long long t__ = clock64(); cw_acc[i] += (unsigned long long)(t__ - cw_tprev); cw_tprev = t__; } while (0)
#define CW_T_USE(v) do { if (wp::ballot((v) == 0x9E3779B9u) == 0xFFFFFFFFu) cw_passes += 1000000; } while (0) /* waits for v */
#define CW_T_PASS do { cw_passes++; } while (0)
#define CW_T_FLUSH do { if (!GT && wp::lane() == 0) { for (int i__ = 0; i__ < 8; i__++) atomicAdd(&pimsnappy::g_cw_prof[i__], cw_acc[i__]); atomicAdd(&pimsnappy::g_cw_prof[8], cw_passes); } } while (0)
#else
#define CW_T_DECL do { } while (0)
#define CW_T(i) do { } while (0)
#define CW_T_USE(v) do { } while (0)
#define CW_T_PASS do { } while (0)
#define CW_T_FLUSH do { } while (0)
#endif

namespace pimsnappy {

constexpr uint32_t CW_MAX_TABLE = 16384; // snappy_compress.c:16-17
constexpr uint32_t CW_MARGIN = 15;       // snappy_compress.c:299
#ifndef PIM_CW_CHAIN_MAX
#define PIM_CW_CHAIN_MAX 30 // (30 = as long as a continuation lane exists; measured on B200, text: 20: 34.4, 24: 37.6, 28: 40.2, 30: 41.2 GB/s)
#endif
constexpr int CW_CHAIN_MAX = PIM_CW_CHAIN_MAX; // a new pair is resolved in the same pass if its insert lane is <= this
#ifndef PIM_PF_AHEAD
#define PIM_PF_AHEAD 1024 // the parser prefetches the input into L2, 4 KB at a time, when it gets this close to the end of what it has prefetched
                          // (0: whole block up front, compress_sm100.cu -- with 2,220 blocks in flight that is 72 MB of input parked in L2 next
                          // to tables and output slots, and part of it was evicted before use: DRAM reads per 256 MiB 529 -> 445 MB, speed +-1 %)
#endif
#ifndef PIM_GT_MATCH
#define PIM_GT_MATCH 1 // global-memory tables: slot sharing inside a pass found with match.any (0: store + read-back, as for shared memory)
#endif

#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint32_t cw_table_entries(uint32_t n) // snappy_compress.c:139-146
{
	uint32_t ts = 256;
	while (ts < CW_MAX_TABLE && ts < n)
		ts <<= 1;
	return ts;
}

struct CBlock {
	const uint32_t *inw; // 4-byte aligned; input byte i is byte (i + mis) from here (global memory, L2-resident, on the device)
	uint32_t mis;        // 0..15; 24 readable bytes follow the input
	uint16_t *table;     // zeroed, cw_table_entries(n) entries (shared memory, or global memory for parse_block_warp<true>)
	int shift;           // 32 - log2(entries), snappy_compress.c:288
	uint8_t *out;        // output slot: 4-byte size field + elements
	uint32_t n;          // input bytes
};

namespace cw {

// Input access goes through word indices only (no pointer <-> integer casts), so that the
// compiler keeps the shared-memory address space and emits LDS with 32-bit addresses.
PW_FN uint32_t in_byte(const CBlock &c, uint32_t pos)
{
	const uint32_t b = pos + c.mis;
	return (c.inw[b >> 2] >> ((b & 3u) * 8u)) & 0xFFu;
}
PW_FN uint32_t in_load32(const CBlock &c, uint32_t pos)
{
	const uint32_t b = pos + c.mis;
	const uint32_t *w = c.inw + (b >> 2);
	return wp::funnel_r(w[0], w[1], (b & 3u) * 8u);
}
PW_FN void in_load128(const CBlock &c, uint32_t pos, uint32_t v[4])
{
	const uint32_t b = pos + c.mis;
	const uint32_t *w = c.inw + (b >> 2);
	const uint32_t sh = (b & 3u) * 8u;
	const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
	v[0] = wp::funnel_r(w0, w1, sh);
	v[1] = wp::funnel_r(w1, w2, sh);
	v[2] = wp::funnel_r(w2, w3, sh);
	v[3] = wp::funnel_r(w3, w4, sh);
}

// number of equal leading bytes (0..16) of in[cand..] and the sixteen bytes `cur`; no FLO/BREV
// (those cost ~40 cycles on the pass's critical path)
PW_FN uint32_t prefix16(const uint32_t cwd[4], const uint32_t cur[4])
{
	const uint32_t d0 = cur[0] ^ cwd[0], d1 = cur[1] ^ cwd[1], d2 = cur[2] ^ cwd[2], d3 = cur[3] ^ cwd[3];
	const uint32_t d = d0 ? d0 : (d1 ? d1 : (d2 ? d2 : d3));
	const uint32_t base = d0 ? 0u : (d1 ? 4u : (d2 ? 8u : 12u));
	const uint32_t inword = ((d & 0xFFu) == 0u) + ((d & 0xFFFFu) == 0u) + ((d & 0xFFFFFFu) == 0u);
	return d ? base + inword : 16u;
}
PW_FN uint32_t common_prefix16(const CBlock &c, uint32_t cand, const uint32_t cur[4])
{
	uint32_t cwd[4];
	in_load128(c, cand, cwd);
	return prefix16(cwd, cur);
}

// lanes a..b inclusive (empty if b < a); 0 <= a, b <= 31
PW_FN uint32_t lanes(int a, int b)
{
	return b < a ? 0u : (((2u << b) - 1u) & ~((1u << a) - 1u));
}

// snappy_compress.c:202-225 -- returns the new output cursor
PW_FN uint32_t emit_literal(const CBlock &c, uint32_t o, uint32_t from, uint32_t len, int lane)
{
	const uint32_t n1 = len - 1;
	uint32_t hdr;
	if (n1 < 60) {
		hdr = 1;
		if (lane == 0)
			c.out[o] = (uint8_t)(n1 << 2);
	} else {
		const uint32_t extra = n1 < 256 ? 1 : (n1 < 65536 ? 2 : 3); // n1 <= 65535 in the valid domain
		hdr = 1 + extra;
		if (lane == 0)
			c.out[o] = (uint8_t)((59 + extra) << 2);
		if (lane < (int)extra)
			c.out[o + 1 + lane] = (uint8_t)(n1 >> (8 * lane));
	}
	uint8_t *dst = c.out + o + hdr;
	uint32_t head = (4u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u;
	if (head > len)
		head = len;
	if ((uint32_t)lane < head)
		dst[lane] = (uint8_t)in_byte(c, from + lane);
	const uint32_t words = (len - head) >> 2;
	uint32_t *d32 = reinterpret_cast<uint32_t *>(dst + head);
	for (uint32_t i = lane; i < words; i += 32)
		d32[i] = in_load32(c, from + head + 4u * i);
	const uint32_t done = head + 4u * words;
	if (done + lane < len)
		dst[done + lane] = (uint8_t)in_byte(c, from + done + lane);
	return o + hdr + len;
}

// snappy_compress.c:234-245
PW_FN uint32_t emit_copy_upto64(uint8_t *p, uint32_t off, uint32_t len)
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
PW_FN uint32_t emit_copy(const CBlock &c, uint32_t o, uint32_t off, uint32_t len, int lane)
{
	const uint32_t n64 = len >= 68 ? (len - 4) / 64 : 0;
	for (uint32_t t = lane; t < n64; t += 32) {
		uint8_t *p = c.out + o + 3 * t;
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
	if (lane == 0) {
		emit_copy_upto64(c.out + o, off, l1);
		if (two)
			emit_copy_upto64(c.out + o + s1, off, len - 60);
	}
	return o + s1 + s2;
}

// snappy_compress.c:176-193: common prefix of in[a..] and in[b..], at most maxn bytes (b > a),
// sixteen bytes per lane per step.
PW_FN uint32_t match_length(const CBlock &c, uint32_t a, uint32_t b, uint32_t maxn, int lane)
{
	for (uint32_t base = 0;; base += 512) {
		const uint32_t j = base + 16u * lane;
		uint32_t cnt = 0; // equal bytes at the front of this lane's 16
		if (j < maxn) {
			uint32_t x[4];
			in_load128(c, b + j, x);
			cnt = wp::umin(common_prefix16(c, a + j, x), maxn - j);
		}
		const uint32_t stop = wp::ballot(cnt < 16);
		if (stop) {
			const int l = wp::ffs32(stop) - 1;
			return base + 16u * l + wp::shfl(cnt, l);
		}
	}
}

constexpr uint32_t CW_SAT = 0xFFFFFFFFu; // "next insert lane" of a hit lane whose match is >= 16 bytes

// What the pair walk of one pass decided (uniform across the warp).
struct Walk {
	uint32_t V;        // insert lanes: lane 0 in the ic layout, then lane f+mlen-1 of every match the walk continued after
	uint32_t H;        // hit lanes
	uint32_t firstbit; // first hit lane (0: none)
	uint32_t lastbit;  // last hit lane (0: none)
	uint32_t satbit;   // hit lane whose match is >= 16 bytes (always the last one), else 0
	uint32_t insbit;   // insert lane of the pair the walk stopped in (0: none)
	bool first;        // the walk stopped in the pass's first pair
	bool redo;         // outcome 0 without a hit: the pair the walk stopped in gets a pass of its own
	int outcome;       // 0: next pass in ic layout   1: keep probing   2: remainder   3: long match at satbit
};

// Fast walk: every lane's candidate is fixed (table entry, or the closest lower lane with the same
// hash).  The reference's loop (:370-398) goes hit -> end of its match -> next hit -> ...; here every lane works out, in
// parallel and once per pass, where the chain would go after a hit at ITS position (the next hit lane at or above its
// continuation lane, its insert lane, whether the pass may continue there), and the chain is then followed with one
// shuffle and one branch per pair.  (Round 1 fetched the hit lane's data with a warp reduction and re-derived the rest
// per pair: 27 % of the kernel's instructions.)
PW_FN void walk_fast(Walk &w, bool has_ic, bool dense, uint32_t hits, uint32_t inval, uint32_t mynins, uint32_t mybit)
{
	(void)mybit;
	const bool sat = mynins == CW_SAT;
	const uint32_t nv = sat ? 0u : mynins; // my match's insert lane (:388-389), 0 if I am no hit or it lies beyond the window
	// continue in this pass after my match if the next pair's insert lane and continuation lane are usable
	const bool cont = dense && nv != 0u && nv <= (1u << CW_CHAIN_MAX) && ((nv << 1) & ~inval);
	const uint32_t after = hits & (0u - (nv << 1)); // hit lanes at or above my continuation lane
	const uint32_t nx = after ? (uint32_t)wp::ffs32(after) - 1u : 32u;
	// bits 0-7: next hit lane (32: none), 8-12: my insert lane, 16: the pass may continue after me, 17: match of >= 16 bytes,
	// 18: ... and does continue (there is a next hit): the only test inside the loop
	const uint32_t word = nx | ((nv ? 31u - (uint32_t)wp::clz32(nv) : 0u) << 8) | (cont ? 1u << 16 : 0u) | (sat ? 1u << 17 : 0u) |
	                      (cont && nx != 32u ? 1u << 18 : 0u);
	w.V = has_ic ? 1u : 0u;
	w.H = w.firstbit = w.lastbit = w.satbit = 0;
	w.insbit = w.V;
	w.first = true;
	w.redo = false;
	const uint32_t ab = hits & (has_ic ? ~1u : ~0u);
	if (!ab) { // inval: :341 / :386, the rest of the block is one literal; otherwise keep probing
		w.outcome = inval ? 2 : 1;
		return;
	}
	uint32_t cur = (uint32_t)wp::ffs32(ab) - 1u;
	w.firstbit = 1u << cur;
	uint32_t x;
	for (;;) {
		CW_STAT(2);
		x = wp::shfl(word, (int)cur);
		w.H |= 1u << cur;
		if (!(x & (1u << 18)))
			break;
		const uint32_t nvb = 1u << ((x >> 8) & 31u);
		w.V |= nvb;
		w.insbit = nvb;
		cur = x & 0xFFu;
	}
	w.lastbit = 1u << cur;
	w.first = (w.V & ~1u) == 0u; // no pair was continued after (an insert lane of a match is lane 3 or higher)
	if (x & (1u << 17)) { // a match of 16 bytes or more ends the walk (the whole warp extends it)
		CW_STAT(4);
		w.satbit = w.lastbit;
		w.outcome = 3;
	} else if (!(x & (1u << 16))) {
		w.outcome = 0;
	} else { // the pass continues after the last hit but the window holds no further hit: that pair gets a pass of its own
		const uint32_t nvb = 1u << ((x >> 8) & 31u);
		w.V |= nvb;
		w.insbit = nvb;
		w.first = false;
		w.outcome = inval ? 2 : 0;
		w.redo = !inval;
	}
}

// Serial, always-exact walk: used when a lane's forwarded candidate (a lower lane of the same pass
// with the same hash) turned out to lie inside a match.  cand / mlen of a lane are re-evaluated
// whenever the set of skipped lanes changes its closest inserted lane.  Dense passes only.
PW_FN void walk_serial(const CBlock &c, Walk &w, bool has_ic, bool valid, uint32_t inval, uint32_t p, uint32_t tc, const uint32_t cur[4],
                       uint32_t lowmask, uint32_t &cand, uint32_t &mlen, int lane)
{
	w.V = has_ic ? 1u : 0u;
	w.H = w.firstbit = w.lastbit = w.satbit = 0;
	w.first = true;
	w.redo = false;
	uint32_t SK = 0; // lanes inside a match (never inserted)
	int ins = has_ic ? 0 : -1;
	int loaded = -2; // force an evaluation in the first round
	for (;;) {
		const int cont = ins + 1;
		const uint32_t eff = lowmask & ~SK;
		const int want = eff ? 31 - wp::clz32(eff) : -1;
		if (want != loaded && valid) {
			loaded = want;
			cand = want < 0 ? tc : (p - lane + want) & 0xFFFFu; // (what the table would hold: entries are 16 bits wide, :139)
			mlen = common_prefix16(c, cand, cur);
		}
		const uint32_t hits = wp::ballot(valid && lane >= cont && mlen >= 4);
		w.insbit = ins < 0 ? 0u : 1u << ins;
		if (!hits) {
			w.outcome = inval ? 2 : (w.first ? 1 : 0);
			w.redo = !inval && !w.first;
			return;
		}
		const int f = wp::ffs32(hits) - 1;
		const uint32_t mf = wp::shfl(mlen, f);
		w.H |= 1u << f;
		w.lastbit = 1u << f;
		if (w.first)
			w.firstbit = 1u << f;
		if (mf == 16) {
			w.satbit = 1u << f;
			w.outcome = 3;
			return;
		}
		const int ins2 = f + (int)mf - 1;
		const bool go = ins2 <= CW_CHAIN_MAX && !((inval >> (ins2 + 1)) & 1u);
		if (!go) {
			w.outcome = 0;
			return;
		}
		SK |= lanes(f + 1, ins2 - 1);
		w.V |= 1u << ins2;
		ins = ins2;
		w.first = false;
	}
}

} // namespace cw

// ---------------------------------------------------------------------------------------------
// Two warps per block.  The PARSER warp runs the passes (probe, walk, table update): that is the
// block's serial dependency chain.  Everything that only produces output bytes is handed to the
// EMITTER warp through a small ring of per-pass records in shared memory, so it leaves the chain.
// ---------------------------------------------------------------------------------------------
#ifndef PIM_CW_RING
#define PIM_CW_RING 1
#endif
constexpr uint32_t CW_RING = PIM_CW_RING; // records in flight between the two warps

struct EmitRec {
	uint32_t lanew[32]; // per lane: first byte at its position | (match length - 4) << 8 (4 bits) | copy offset << 12 (20 bits: blocks up to 1 MiB)
	uint32_t L;         // in-window literal bytes
	uint32_t H;         // hit lanes
	uint32_t V;         // insert lanes
	uint32_t satbit;    // hit lane whose copy is not written from lanew (long match), else 0
	uint32_t lit_from, lit_len;   // literal emitted before the window's elements (len 0: none)
	uint32_t copy_off, copy_len;  // copy emitted after them (len 0: none)
	uint32_t done;      // last record of the block
	uint32_t pad[7];    // (192-byte records: measurably faster than packed ones; pad[0] of record 0 carries the pair's block index)
};

// Hand-off protocol.  On the device every slot has two mbarriers: `full` (the 32 parser lanes arrive after writing the
// record, the emitter lanes block in mbarrier.try_wait until the phase completes) and `empty` (the other way round), so
// a waiting warp sleeps in hardware instead of polling shared memory -- in round 1 the emitters' poll loop was a quarter
// of all issued instructions.  The record counters run over the whole kernel (phase parity = (count / CW_RING) & 1).
// On the emulator the two counters live in the ring and the warps poll them.
struct EmitRing {
	EmitRec rec[CW_RING];
#ifdef PIM_WARP_EMU
	volatile uint32_t head; // records published by the parser
	volatile uint32_t tail; // records consumed by the emitter
#else
	unsigned long long full[CW_RING], empty[CW_RING];
#endif
};

#ifndef PIM_WARP_EMU
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity)
{
	asm volatile("{\n"
	             ".reg .pred p;\n"
	             "LAB_WAIT:\n"
	             "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n" // (with a suspend-time hint: without one the wait comes back every few hundred cycles)
	             "@p bra LAB_DONE;\n"
	             "bra LAB_WAIT;\n"
	             "LAB_DONE:\n"
	             "}" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity), "r"(0x989680u) : "memory");
}
PW_FN void ring_init(EmitRing *ring) // by one thread, before the pair's first barrier
{
	for (uint32_t i = 0; i < CW_RING; i++) {
		mbar_init(&ring->full[i], 32);
		mbar_init(&ring->empty[i], 32);
	}
}
#endif

PW_FN void ring_push(EmitRing *ring, uint32_t &head, uint32_t lanew, uint32_t L, uint32_t H, uint32_t V, uint32_t satbit, uint32_t lit_from,
                     uint32_t lit_len, uint32_t copy_off, uint32_t copy_len, uint32_t done, int lane)
{
#ifdef PIM_WARP_EMU
	while (head - ring->tail >= CW_RING)
		wp::spin_pause();
#else
	mbar_wait(&ring->empty[head % CW_RING], ((head / CW_RING) & 1u) ^ 1u); // (passes at once the first time round)
#endif
	EmitRec &r = ring->rec[head % CW_RING];
	r.lanew[lane] = lanew;
	if (lane == 0) {
		r.L = L;
		r.H = H;
		r.V = V;
		r.satbit = satbit;
		r.lit_from = lit_from;
		r.lit_len = lit_len;
		r.copy_off = copy_off;
		r.copy_len = copy_len;
		r.done = done;
	}
#ifdef PIM_WARP_EMU
	wp::fence_block();
	wp::syncwarp();
	head++;
	if (lane == 0)
		ring->head = head;
#else
	mbar_arrive(&ring->full[head % CW_RING]); // release: my stores above are visible to whoever sees the phase complete
	head++;
#endif
}

// Emitter warp: turns the parser's records into output bytes.  Returns the bytes written (4 + elements).
// `tail`: records consumed so far by this pair (runs over the whole kernel on the device).
PW_FN uint32_t emit_block_warp(const CBlock &c, EmitRing *ring, uint32_t &tail)
{
	using namespace cw;
	const int lane = wp::lane();
	const uint32_t mybit = 1u << lane;
	const uint32_t lt = mybit - 1u;
	const uint32_t le = lt | mybit;
	uint32_t o = 4; // output cursor (4 bytes reserved for the size, :291)
	for (;; tail++) {
#ifdef PIM_WARP_EMU
		while (ring->head == tail)
			wp::spin_pause();
		wp::fence_block();
#else
		mbar_wait(&ring->full[tail % CW_RING], (tail / CW_RING) & 1u);
#endif
		const EmitRec &r = ring->rec[tail % CW_RING];
		const uint32_t lanew = r.lanew[lane];
		const uint32_t L = r.L, H = r.H, V = r.V, satbit = r.satbit, lit_from = r.lit_from, lit_len = r.lit_len, copy_off = r.copy_off,
		               copy_len = r.copy_len, done = r.done;
#ifdef PIM_WARP_EMU
		wp::syncwarp();
		if (lane == 0)
			ring->tail = tail + 1; // the slot may be reused
#else
		mbar_arrive(&ring->empty[tail % CW_RING]); // (my loads above are done: their values are in registers)
#endif
		if (lit_len) // :355 / :405-409
			o = emit_literal(c, o, lit_from, lit_len, lane);
		const uint32_t LS = L & (V << 1); // first byte of each in-window literal
		const uint32_t Hb = H & ~satbit;  // hit lanes whose tag is written here
		if (L | Hb) {
			const uint32_t mlen = 4u + ((lanew >> 8) & 15u), off = lanew >> 12;
			const bool inHb = (Hb >> lane) & 1u;
			const bool three = inHb && !(mlen < 12 && off < 2048);
			const uint32_t H3 = wp::ballot(three);
			const uint32_t pre = wp::popc32((L | Hb) & lt) + wp::popc32((LS & le) | (Hb & lt)) + wp::popc32(H3 & lt);
			uint8_t *dst = c.out + o + pre;
			if ((L >> lane) & 1u) {
				dst[0] = (uint8_t)lanew;
			} else if ((H >> lane) & 1u) {
				const uint32_t vlow = V & lt;
				if (vlow) { // literal header: the literal runs from behind the pair's insert lane to me
					const uint32_t litlen = (uint32_t)lane - 1u - (uint32_t)(31 - wp::clz32(vlow));
					if (litlen)
						dst[-(int)litlen - 1] = (uint8_t)((litlen - 1) << 2);
				}
				if (inHb) {
					if (!three) { // :238-240
						dst[0] = (uint8_t)(1 + ((mlen - 4) << 2) + ((off >> 8) << 5));
						dst[1] = (uint8_t)off;
					} else { // :242-244
						dst[0] = (uint8_t)(2 + ((mlen - 1) << 2));
						dst[1] = (uint8_t)off;
						dst[2] = (uint8_t)(off >> 8);
					}
				}
			}
			o += wp::popc32(L | Hb) + wp::popc32(LS | Hb) + wp::popc32(H3);
		}
		if (copy_len)
			o = emit_copy(c, o, copy_off, copy_len, lane);
		if (done) {
			tail++;
			break;
		}
	}
	if (lane < 4) // :411
		c.out[lane] = (uint8_t)((o - 4) >> (8 * lane));
	return o;
}

// Parser warp: runs the passes of one block and publishes one record per pass.
// GT: the hash table lives in global memory (L2) instead of shared memory.
template <bool GT>
// `head`: records published so far by this pair (runs over the whole kernel on the device).
PW_FN void parse_block_warp(const CBlock &c, EmitRing *ring, uint32_t &head)
{
	using namespace cw;
	const int lane = wp::lane();
	const uint32_t mybit = 1u << lane;
	const uint32_t lt = mybit - 1u;
	const uint32_t le = lt | mybit;
	const uint32_t n = c.n;
	uint32_t emit = 0; // first input byte not yet emitted (next_emit)
	uint32_t pf = 0;   // input below this position has been prefetched into L2

	CW_T_DECL;
	if (n >= CW_MARGIN) {
		const uint32_t limit = n - CW_MARGIN;
		uint32_t q = 1;      // ic layout: where the last match ended; probe layout: position of probe 0
		bool has_ic = false; // layout of the next pass
		uint32_t k0 = 0;     // probe layout: step-1 probes already done for this pair
		for (;;) {
			if (has_ic && q >= limit) // :386
				break;
			CW_STAT(has_ic ? 0 : 1);
			CW_T(7); // (loop control, between passes)
			CW_T_PASS;
			if (PIM_PF_AHEAD && pf < n && pf < q + PIM_PF_AHEAD) { // HBM -> L2 a few KB ahead of the parse (32 lines at a time)
				const uint32_t at = pf + 128u * lane;
				if (at < n + 24u)
					wp::prefetch_l2(reinterpret_cast<const uint8_t *>(c.inw) + c.mis + at);
				pf += 4096u;
			}
			// ---- lane layout -----------------------------------------------------------------
			uint32_t p, stride = 1;
			if (has_ic) {
				p = q - 1 + lane;
			} else {
				const uint32_t K = k0 + lane + 32, s = K >> 5, r = K & 31;
				p = q + 16u * s * (s - 1) + r * s;
				stride = s;
			}
			const bool dense = has_ic || k0 == 0;
			bool valid = p + stride <= limit; // :341 (probe), :386 (continuation)
			if (has_ic && lane == 0)
				valid = true;
			// ---- per-lane probe --------------------------------------------------------------
			uint32_t cur[4] = {0, 0, 0, 0};
			uint32_t h = 0, tc = 0, cand = 0, mlen = 0;
			CW_T(7);
#if defined(PIM_CW_PROFILE) && !defined(PIM_WARP_EMU) // (the same probe with stamps between its steps; all lanes reach the votes)
			if (valid) {
				in_load128(c, p, cur);
				h = (cur[0] * 0x1e35a7bdu) >> c.shift;
			}
			CW_T_USE(h);
			CW_T(6); // own bytes arrived
			if (valid)
				tc = wp::tbl_ld<GT>(c.table, h);
			CW_T_USE(tc);
			CW_T(0); // table entry arrived
#else
			if (valid) {
				in_load128(c, p, cur);
				h = (cur[0] * 0x1e35a7bdu) >> c.shift; // :161-166
				tc = wp::tbl_ld<GT>(c.table, h);
			}
#endif
			wp::syncwarp(); // every lane has read its slot before anyone writes
			// Speculative insert + read-back: a lane that does not read its own position back shares
			// its slot with another lane of this pass.
			uint32_t w1 = p & 0xFFFFu;
			uint32_t same = mybit, lowmask = 0, dup1;
			bool involved;
			if (GT && PIM_GT_MATCH) {
				// Table in global memory: a read-back is one more L2 round trip on the chain.  The lanes that share a
				// slot are found with match.any instead, which executes while the candidate bytes are on their way,
				// and nothing is stored speculatively: the inserted lanes write their slots once the pass is resolved.
				if (valid)
					cand = tc;
				uint32_t cwd[4];
				in_load128(c, tc, cwd); // (tc = 0 on lanes that do not probe)
				same = wp::match_any(valid ? h : 0xFFFFFFFFu);
				mlen = valid ? prefix16(cwd, cur) : 0u;
				involved = valid && same != mybit;
				dup1 = wp::ballot(involved);
			} else {
				if (valid) {
					wp::tbl_st<GT>(c.table, h, p);
					cand = tc;
					mlen = common_prefix16(c, tc, cur);
				}
				wp::syncwarp();
				if (valid)
					w1 = wp::tbl_ld<GT>(c.table, h);
				dup1 = wp::ballot(w1 != (p & 0xFFFFu)); // (table entries are 16 bits wide, snappy_compress.c:139: above 64 KB the position comes back truncated)
				involved = false;
			}
			const uint32_t inval = wp::ballot(!valid);
			CW_T_USE(mlen);
			CW_T(1); // speculative insert + read-back, candidate bytes arrived, match lengths
			if (dup1) { // some lanes share a slot: find the groups (a second round tells the winners)
				CW_STAT(6);
				if (!(GT && PIM_GT_MATCH)) {
					uint32_t w2 = p & 0xFFFFu;
					if (valid && w1 != (p & 0xFFFFu))
						wp::tbl_st<GT>(c.table, h, p);
					wp::syncwarp();
					if (valid)
						w2 = wp::tbl_ld<GT>(c.table, h);
					involved = valid && (w1 != (p & 0xFFFFu) || w2 != (p & 0xFFFFu));
					same = wp::match_any(involved ? h : 0xFFFFFFFFu);
				}
				if (!involved)
					same = mybit;
				lowmask = same & lt;
				const int src = lowmask ? 31 - wp::clz32(lowmask) : lane;
				const uint32_t pfwd = wp::shfl(p, src);
				if (lowmask) { // the closest lower lane with my hash is my candidate (if it is really inserted)
					cand = pfwd & 0xFFFFu; // (the value the reference would read back from its 16-bit table)
					mlen = common_prefix16(c, cand, cur);
				}
			}

			// ---- resolve the pairs of this pass ------------------------------------------------
			CW_T(2); // shared slots resolved
			Walk w;
			bool inI;       // my position is inserted into the table
			uint32_t Imask;
			for (bool serial = false;; serial = true) {
				if (!serial) {
					const bool hit = valid && mlen >= 4 && !(has_ic && lane == 0);
					const uint32_t mynins = hit ? (mlen == 16 ? CW_SAT : mybit << (mlen - 1)) : 0u;
					walk_fast(w, has_ic, dense, wp::ballot(hit), inval, mynins, mybit);
				} else {
					CW_STAT(5);
					walk_serial(c, w, has_ic, valid, inval, p, tc, cur, lowmask, cand, mlen, lane);
				}
				// a lane was probed iff the closest marker at or below it is an insert lane (not a hit lane)
				const uint32_t Vle = w.V & le, Hlt = w.H & lt;
				const bool probed = Vle > Hlt || Hlt == 0u;
				inI = w.outcome == 1 || (probed && mybit <= w.lastbit);
				Imask = wp::ballot(inI);
				if (!dup1 || serial)
					break;
				// was every forwarded candidate really inserted (or probed, in an unfinished last pair)?
				const bool evaluated = inI || (w.outcome == 2 && valid && mybit > w.insbit);
				const uint32_t E = wp::ballot(evaluated);
				const bool wrong = evaluated && lowmask && !((E >> (31 - wp::clz32(lowmask))) & 1u);
				if (!wp::ballot(wrong))
					break;
			}
			if (w.redo)
				CW_STAT(3);
			CW_T(3); // the pair walk, forwarded candidates checked

			// ---- table: undo the speculative inserts of lanes that were not inserted -----------
			if (GT && PIM_GT_MATCH) { // (nothing was stored yet: the last inserted lane of every slot writes it)
				if (w.outcome != 2 && valid && inI && !(same & Imask & ~le))
					wp::tbl_st<GT>(c.table, h, p);
			} else if (w.outcome != 2) {
				if (valid && !inI)
					wp::tbl_st<GT>(c.table, h, tc);
				if (dup1) { // shared slots: the last inserted lane of the group owns the slot
					wp::syncwarp();
					if (inI && !(same & Imask & ~le))
						wp::tbl_st<GT>(c.table, h, p);
				}
			}

			CW_T(4); // table updates
			// ---- where the pass ended ----------------------------------------------------------
			const bool by_ins = w.redo || (w.outcome == 2 && !w.first); // just past the last complete match
			const bool have_pos = by_ins || w.outcome == 0 || w.outcome == 3;
			uint32_t pos_last = 0;
			if (have_pos) {
				const uint32_t sel = by_ins ? w.insbit : w.lastbit;
				const uint32_t add = by_ins ? 1u : (w.outcome == 3 ? 0u : mlen);
				pos_last = wp::reduce_max(mybit == sel ? p + add : 0u);
			}

			CW_T(4); // position of the pass's end
			// ---- hand the pass's output to the emitter warp -------------------------------------
			uint32_t lit_from = 0, lit_len = 0, copy_off = 0, copy_len = 0;
			if (!has_ic && w.firstbit) { // the first pair's literal started before this window
				const uint32_t lit_to = wp::reduce_max(mybit == w.firstbit ? p : 0u);
				lit_from = emit;
				lit_len = lit_to - emit;
			}
			const uint32_t before_last = w.lastbit ? w.lastbit - 1u : 0u;
			const uint32_t first_region = has_ic ? 0u : w.firstbit - 1u; // (firstbit == 0: nothing is emitted anyway)
			const uint32_t L = Imask & ~w.H & ~w.V & ~first_region & before_last; // in-window literal bytes
			int outcome = w.outcome;
			if (outcome == 3) { // match of >= 16 bytes: extend with the whole warp
				const uint32_t cf = wp::reduce_max(mybit == w.satbit ? cand : 0u);
				const uint32_t total = 16u + match_length(c, cf + 16, pos_last + 16, n - pos_last - 16, lane);
				copy_off = pos_last - cf;
				copy_len = total;
				q = pos_last + total;
				emit = q;
				outcome = 0;
			} else if (have_pos) {
				q = pos_last;
				emit = q;
			}
			if (L | w.H | lit_len)
				ring_push(ring, head, (cur[0] & 0xFFu) | (((mlen - 4u) & 15u) << 8) | ((p - cand) << 12), L, w.H, w.V, w.satbit, lit_from, lit_len, copy_off,
				          copy_len, 0u, lane);
			wp::syncwarp(); // table writes visible to the next pass
			CW_T(5); // literal bounds, long matches, record to the emitter

			if (outcome == 2)
				break;
			if (outcome == 0) {
				has_ic = true;
				k0 = 0;
			} else if (has_ic) { // 30 probes done behind the continuation; probe 0 sat at q+1
				has_ic = false;
				q = q + 1;
				k0 = 30;
			} else {
				k0 += 32;
			}
		}
	}
	CW_T_FLUSH;
	// the rest of the block is one literal (:405-409); this record also ends the block
	ring_push(ring, head, 0u, 0u, 0u, 0u, 0u, emit, n - emit, 0u, 0u, 1u, lane);
}

} // namespace pimsnappy
