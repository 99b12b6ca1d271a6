// Per-warp Snappy block decoder: one warp decodes one compressed block of the reference's
// multi-block format (snappy/README.md:19-33) and reproduces what snappy_decompress_host
// (snappy/snappy_decompress.c:218-289) and the DPU twin dpu_uncompress
// (snappy/dpu-decompress/dpu_decompress.c:224-299) produce on every valid stream.
//
// The reference decodes one element at a time (tag switch, :242-283) and copies byte by byte
// (writer_append_host :142-154, write_copy_host :164-184).  Here the three things that are
// serial in that loop are each made 32 lanes wide:
//
//   1. TAG CHAIN (where does the next element start?).  The compressed bytes are staged in
//      chunks of DW_CH bytes.  Every lane walks its DW_SEG-byte segment of the chunk from the
//      segment's first byte (a guess), remembers which positions it visited, and the lanes then
//      iterate  entry[s+1] = exit_s(entry[s])  to the fixed point: a lane whose true entry lies
//      on the chain it walked is done, otherwise it re-walks until it merges.  The confirmed
//      tag positions go into a queue.
//   2. ELEMENT DECODE, 32 elements at a time, one per lane: tag -> (length, distance) as in
//      snappy_decompress.c:244-283 / dpu_snappy.h:10-13, output positions by a warp prefix sum,
//      and one 16-byte RECORD per element {end position, source delta, source region | flags}.
//      Back-to-back copies with one distance (zero pages, repeated records) and self-overlapping
//      copies (distance < length, write_copy_host's byte order :176-183) are marked PERIODIC with
//      the start of their run: their bytes are read from the period before the run, so they do
//      not wait for each other.
//   3. BYTE MOVEMENT is a GATHER BY OUTPUT WORD, not a copy by element: in one pass lane j owns
//      the aligned DW_BPL-byte word j of the next 32 * DW_BPL output bytes, finds the element
//      covering its first byte (binary search over the records' end positions with shuffles),
//      walks the records across its bytes, reads every byte from where it comes from -- the
//      staged literal bytes, or the ring of recent output at (position - distance) -- and issues
//      ONE word store to the ring and ONE word store to HBM.  All lanes do the same work whatever
//      the element lengths are.  Bytes whose source lies inside the pass's own window wait for
//      the (rare) fix-up rounds after the store; distances beyond the ring are read back from
//      HBM/L2, which this warp wrote itself.
//
// Literals with explicit length bytes (> 60 bytes; incompressible blocks are one such literal)
// bypass all of this: the warp copies them HBM -> HBM 16 bytes per lane.
//
// Coordinates:  a = compressed position + (address of the first element byte & 15)   ("a")
//               y = output position     + (address of the block's output & 15)       ("y")
// so that 16-byte pieces of both rings line up with 16-byte pieces of global memory.
//
// Written against warp_prims.cuh: the identical source runs on the host lock-step emulator
// (tests/emu) against the oracle.
#pragma once

#include "warp_prims.cuh"

#ifndef PIM_DW_RB
#define PIM_DW_RB 4096
#endif
#ifndef PIM_DW_BPL
#define PIM_DW_BPL 8
#endif
#ifndef PIM_DW_WSEG
#define PIM_DW_WSEG 128
#endif
#ifndef PIM_DW_NREC
#define PIM_DW_NREC 128
#endif

#ifdef PIM_WARP_EMU_STATS
namespace pimsnappy { extern unsigned long dw_stats[16]; }
#define DW_STAT(i, n) do { if (wp::lane() == 0) pimsnappy::dw_stats[i] += (n); } while (0)
#else
#define DW_STAT(i, n) do { } while (0)
#endif

namespace pimsnappy {

constexpr uint32_t DW_RB = PIM_DW_RB;                 // bytes of the output ring = bytes of the stage ring
constexpr uint32_t DW_RMASK = DW_RB - 1;
constexpr uint32_t DW_CH = DW_RB / 4;                 // compressed bytes per parse chunk (the stage ring holds 4)
constexpr uint32_t DW_SEG = DW_CH / 32;               // compressed positions per lane when the starts are queued
constexpr uint32_t DW_WSEG = PIM_DW_WSEG;             // compressed bytes per walking lane of a parse
constexpr uint32_t DW_NWK = DW_CH / DW_WSEG;          // walking lanes
constexpr uint32_t DW_NV = DW_WSEG / 32;              // registers of a walker's position set
constexpr uint32_t DW_BPL = PIM_DW_BPL;               // output bytes per lane of a gather pass
constexpr uint32_t DW_WB = 32 * DW_BPL;               // output bytes per gather pass
constexpr uint32_t DW_NREC = PIM_DW_NREC;             // decoded elements kept (ring of 16-byte records)
constexpr uint32_t DW_QCAP = 512;                     // queue of tag positions (u16)
constexpr uint32_t DW_NEAR_MAX = DW_RB - DW_WB - 8;   // copies with a larger distance read their source from HBM/L2
constexpr uint32_t DW_BIG = 0x100000u;                // "footprint" of a malformed element
constexpr uint32_t DW_F_FAR = 1, DW_F_NEAR = 2, DW_F_PERIODIC = 4, DW_F_ALL = 7;

static_assert((DW_RB & (DW_RB - 1)) == 0 && DW_SEG >= 16 && DW_SEG <= 32, "ring size: 2048 or 4096");
static_assert(DW_WSEG % 32 == 0 && DW_WSEG >= 32 && DW_WSEG <= 128 && DW_CH % DW_WSEG == 0, "walker segment: 32, 64 or 128");
static_assert(DW_BPL == 4 || DW_BPL == 8, "4 or 8 output bytes per lane");
static_assert((DW_NREC & (DW_NREC - 1)) == 0 && DW_NREC >= 64 && DW_NREC <= 256, "record ring");
static_assert(DW_QCAP >= DW_NREC + 32 + 64, "queue must hold the record window, a leftover batch and at least some lanes of a chunk");

// shared memory of one warp (byte offsets into the CTA's array `sm`); the two rings sit at multiples of DW_RB
struct DSm {
	uint8_t *sm;
	uint32_t ring; // DW_RB bytes: output byte y at ring + (y & DW_RMASK)
	uint32_t stg;  // DW_RB bytes: compressed byte a at stg + (a & DW_RMASK)
	uint32_t q;    // u16[DW_QCAP]: tag positions (a & 0xFFFF), element g at index g & (DW_QCAP - 1)
	uint32_t rec;  // 16 bytes x DW_NREC, element g at index g & (DW_NREC - 1)
	uint32_t lut;  // u8[256]: compressed footprint by tag byte, 0 = literal with explicit length bytes (CTA-shared)
};
constexpr uint32_t DW_SMALL_BYTES = 2 * DW_QCAP + 16 * DW_NREC; // q + rec per warp

struct DBlk {
	const uint8_t *gcomp; // first element byte
	uint32_t csize;       // bytes of elements
	uint8_t *gdst;
	uint32_t out_size;    // bytes the block must decode to
};

struct alignas(16) DRec {
	uint32_t pn;  // y just past the element
	uint32_t dl;  // source of byte y: region + ((y + dl) & DW_RMASK)
	uint32_t w2;  // region (ring or stg, multiple of DW_RB) | DW_F_*
	uint32_t w3;  // periodic: distance | run base y << 15
};

// the compressed footprint by tag byte that DSm::lut holds (0 = literal with explicit length bytes)
#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint32_t dw_lut_foot(uint32_t tag)
{
	const uint32_t kind = tag & 3, n6 = tag >> 2;
	return kind == 0 ? (n6 < 60 ? n6 + 2 : 0) : (kind == 1 ? 2 : (kind == 2 ? 3 : 5));
}

namespace dw {

struct State {
	// block constants
	uint32_t amis, a_end, ybeg, yend;
	const uint8_t *gcomp_al; // a-indexed
	uint8_t *gdst_al;        // y-indexed
	// parser
	uint32_t pa;        // a of the next tag to find
	uint32_t st_issued; // chunks below this index have been requested into the stage ring
	// element counters: pushed (qt), decoded (e_hi), first record still needed (e_lo)
	uint32_t qt, e_hi, e_lo;
	uint32_t cons_a; // a just past the last decoded element
	uint32_t dec_y;  // y just past the last decoded element
	uint32_t gy;     // output complete below this y
	bool lo_dead;    // the bytes of gy's word below gy cannot be recomputed (block start, after a bulk literal)
	uint32_t c_off, c_base; // previous element: copy distance (0 = not a copy) and the start of its run
	bool bulk;              // a literal with explicit length bytes waits for everything before it
	uint32_t b_src, b_len;  // its payload (a) and length
	bool err;
};

struct alignas(8) U2 {
	uint32_t x, y;
};
struct alignas(16) U4 {
	uint32_t x, y, z, w;
};

PW_FN uint32_t lds8(const DSm &s, uint32_t off) { return s.sm[off]; }
PW_FN uint32_t lds32(const DSm &s, uint32_t off) { return *reinterpret_cast<const uint32_t *>(s.sm + off); }

PW_FN uint32_t scan_incl(uint32_t v, int lane)
{
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t t = wp::shfl(v, lane - d);
		if (lane >= d)
			v += t;
	}
	return v;
}

// ---- staging ---------------------------------------------------------------------------------
// Request chunk j (a in [j * DW_CH, (j + 1) * DW_CH)) into its slot of the stage ring.  Only bytes of
// the block are read.
PW_FN void stage_chunk(const DSm &s, const State &st, uint32_t j, int lane)
{
	for (uint32_t piece = lane; piece < DW_CH / 16; piece += 32) {
		const uint32_t a0 = j * DW_CH + 16 * piece;
		const uint32_t lo = wp::umax(a0, st.amis), hi = wp::umin(a0 + 16, st.a_end);
		if (lo >= hi)
			continue;
		if (hi - lo == 16) {
			wp::cp16(s.sm + s.stg + (a0 & DW_RMASK), st.gcomp_al + a0);
		} else {
			for (uint32_t p = lo; p < hi; p++)
				s.sm[s.stg + (p & DW_RMASK)] = (uint8_t)wp::ldg8(st.gcomp_al + p);
		}
	}
	wp::cp_commit();
}

// ---- 1. tag chain ----------------------------------------------------------------------------
// compressed footprint of the element whose tag is at a = p (p < a_end); DW_BIG if malformed
PW_FN uint32_t foot_at(const DSm &s, const State &st, uint32_t p)
{
	const uint32_t tag = lds8(s, s.stg + (p & DW_RMASK));
	uint32_t f = lds8(s, s.lut + tag);
	if (f == 0) { // literal with 1..4 length bytes (snappy_decompress.c:64-74)
		const uint32_t nb = (tag >> 2) - 59;
		if (p + 1 + nb > st.a_end)
			return DW_BIG;
		uint32_t lm1 = 0;
		for (uint32_t i = 0; i < nb; i++)
			lm1 |= lds8(s, s.stg + ((p + 1 + i) & DW_RMASK)) << (8 * i);
		if (lm1 >= 65536u)
			return DW_BIG;
		f = 2 + nb + lm1;
	}
	return f;
}

// bit sets of DW_WSEG positions (one walker's segment) in DW_NV registers; indices are compile-time after unrolling
PW_FN bool v_test(const uint32_t V[DW_NV], uint32_t r)
{
	uint32_t w = 0;
#pragma unroll
	for (uint32_t i = 0; i < DW_NV; i++)
		if ((r >> 5) == i)
			w = V[i];
	return (w >> (r & 31u)) & 1u;
}
PW_FN void v_set(uint32_t V[DW_NV], uint32_t r)
{
#pragma unroll
	for (uint32_t i = 0; i < DW_NV; i++)
		if ((r >> 5) == i)
			V[i] |= 1u << (r & 31u);
}
// V = nb | (V without the bits below r)
PW_FN void v_merge(uint32_t V[DW_NV], const uint32_t nb[DW_NV], uint32_t r)
{
#pragma unroll
	for (uint32_t i = 0; i < DW_NV; i++) {
		const uint32_t keep = i < (r >> 5) ? 0u : (i == (r >> 5) ? ~((1u << (r & 31u)) - 1u) : 0xFFFFFFFFu);
		V[i] = nb[i] | (V[i] & keep);
	}
}

// Find the element starts of the chunk that holds st.pa, from st.pa on; push them; advance st.pa.
// The first DW_NWK lanes walk DW_WSEG bytes each (segments long enough that a walk from a guessed start
// has merged with the true chain before it leaves the segment); every lane then takes the DW_SEG positions
// of its own and queues the starts among them.
// A literal with explicit length bytes ends the parse: it is reported in st.bulk / b_src / b_len and
// st.pa continues after it.
PW_FN void parse_chunk(const DSm &s, State &st)
{
	const int lane = wp::lane();
	const uint32_t kc = st.pa / DW_CH, cbase = kc * DW_CH;
	const uint32_t x0 = st.pa - cbase;
	const uint32_t cend = wp::umin(DW_CH, st.a_end - cbase); // chunk-relative end of the block's data
	const bool walker = (uint32_t)lane < DW_NWK;
	const uint32_t wbase = walker ? DW_WSEG * lane : DW_CH;
	uint32_t lim = wp::umin(wbase + DW_WSEG, cend); // (<= wbase past the block's end: such a lane is never entered)
	if (wbase + DW_WSEG <= x0 || !walker)
		lim = wbase; // nothing of mine is at or after the entry
	const int lastl = (int)((cend - 1) / DW_WSEG); // last walker with data
	DW_STAT(0, 1);

	// speculative walk from the segment start
	uint32_t V[DW_NV]; // bit b: an element of the chain I know starts at wbase + b
#pragma unroll
	for (uint32_t i = 0; i < DW_NV; i++)
		V[i] = 0;
	uint32_t exitA = wbase;
	{
		uint32_t p = wbase;
		while (p < lim) {
			v_set(V, p - wbase);
			p += foot_at(s, st, cbase + p);
			DW_STAT(11, 1);
		}
		exitA = p;
	}
	// fixed point over the walkers: x = first element start at or after my segment's first byte
	uint32_t x = wp::umax(wbase, x0); // (always >= wbase: lanes before the entry pass the entry along)
	uint32_t eval_x = wbase, f_out = 0;
	for (int it = 0; it < 34; it++) {
		const bool touched = x < lim;
		if (touched && x != eval_x) {
			eval_x = x;
			if (!v_test(V, x - wbase)) { // not on the chain I know: walk until it merges with it
				uint32_t p = x, nb[DW_NV];
#pragma unroll
				for (uint32_t i = 0; i < DW_NV; i++)
					nb[i] = 0;
				while (p < lim && !v_test(V, p - wbase)) {
					v_set(nb, p - wbase);
					p += foot_at(s, st, cbase + p);
					DW_STAT(12, 1);
				}
				if (p < lim) {
					v_merge(V, nb, p - wbase);
				} else {
#pragma unroll
					for (uint32_t i = 0; i < DW_NV; i++)
						V[i] = nb[i];
					exitA = p;
				}
			}
		}
		f_out = touched ? exitA : x;
		uint32_t nx = wp::shfl(f_out, lane - 1);
		if (lane == 0)
			nx = x0;
		nx = wp::umax(nx, wbase);
		const bool changed = nx != x && lane <= lastl;
		x = nx;
		DW_STAT(1, 1);
		if (!wp::ballot(changed))
			break;
	}
	{ // drop what lies before my true entry
		const bool touched = x < lim;
		uint32_t none[DW_NV];
#pragma unroll
		for (uint32_t i = 0; i < DW_NV; i++)
			none[i] = 0;
		if (touched)
			v_merge(V, none, x - wbase);
		else {
#pragma unroll
			for (uint32_t i = 0; i < DW_NV; i++)
				V[i] = 0;
		}
	}
	uint32_t next = wp::shfl(f_out, lastl); // chunk-relative start of the first element after the chunk
	// every lane takes its DW_SEG positions from the walker that owns them
	const uint32_t base = DW_SEG * lane;
	uint32_t bits = 0;
	{
		const int owner = (int)(base / DW_WSEG);
		const uint32_t r = base % DW_WSEG;
#pragma unroll
		for (uint32_t i = 0; i < DW_NV; i++) {
			const uint32_t t = wp::shfl(V[i], owner);
			if ((r >> 5) == i)
				bits = t;
		}
		bits >>= (r & 31u);
		if (DW_SEG < 32)
			bits &= (1u << DW_SEG) - 1u;
	}

	// a literal with explicit length bytes: the last element of its segment (its footprint is > 62 bytes)
	uint32_t b_a = 0, b_nb = 0, b_len = 0;
	bool isbulk = false;
	const uint32_t first_pos = bits ? base + (uint32_t)wp::ffs32(bits) - 1u : 0xFFFFFFFFu; // my first element start
	if (bits) {
		const uint32_t hb = 31u - (uint32_t)wp::clz32(bits);
		const uint32_t p = cbase + base + hb;
		const uint32_t tag = lds8(s, s.stg + (p & DW_RMASK));
		if (lds8(s, s.lut + tag) == 0) {
			isbulk = true;
			bits &= ~(1u << hb);
			b_a = p;
			b_nb = (tag >> 2) - 59;
			const uint32_t f = foot_at(s, st, p);
			b_len = f >= DW_BIG ? 0u : f - 1 - b_nb; // 0: malformed
		}
	}
	const uint32_t bm = wp::ballot(isbulk);
	int lb = 32;
	if (bm) {
		lb = wp::ffs32(bm) - 1;
		if (lane > lb)
			bits = 0;
	}
	// queue room: entries from e_lo on stay readable (the stage guard looks at the oldest one)
	const uint32_t cnt = (uint32_t)wp::popc32(bits);
	const uint32_t incl = scan_incl(cnt, lane);
	const uint32_t room = DW_QCAP - (st.qt - st.e_lo);
	const uint32_t fitm = wp::ballot(incl <= room);
	uint32_t total;
	if (fitm != 0xFFFFFFFFu) { // (never when the caller keeps the queue short; kept for any stream)
		const int m = wp::popc32(fitm); // lanes 0..m-1 fit (incl is monotonic)
		if (lane >= m)
			bits = 0;
		total = wp::shfl(incl, m ? m - 1 : 0);
		if (!m)
			total = 0;
		const uint32_t fn = wp::reduce_min(lane >= m ? first_pos : 0xFFFFFFFFu); // first start that was not queued
		if (fn != 0xFFFFFFFFu)
			next = fn;
		if (lb >= m)
			lb = 32;
	} else {
		total = wp::shfl(incl, 31);
	}
	if (lb < 32) {
		const uint32_t ba = wp::shfl(b_a, lb), bnb = wp::shfl(b_nb, lb), bl = wp::shfl(b_len, lb);
		st.bulk = true;
		st.b_src = ba + 1 + bnb;
		st.b_len = bl;
		if (bl == 0)
			st.err = true;
		next = st.b_src + bl - cbase;
	}
	uint32_t idx = st.qt + incl - cnt;
	while (bits) {
		const uint32_t b = (uint32_t)wp::ffs32(bits) - 1u;
		bits &= bits - 1u;
		*reinterpret_cast<uint16_t *>(s.sm + s.q + 2u * (idx & (DW_QCAP - 1))) = (uint16_t)(cbase + base + b);
		idx++;
	}
	st.qt += total;
	if (next >= DW_BIG || cbase + next <= st.pa)
		st.err = true; // malformed element on the chain, or no progress
	else
		st.pa = cbase + next;
	wp::syncwarp();
}

// ---- 2. element decode -----------------------------------------------------------------------
PW_FN void decode_batch(const DSm &s, State &st)
{
	const int lane = wp::lane();
	const uint32_t n = wp::umin(32u, st.qt - st.e_hi);
	const bool valid = (uint32_t)lane < n;
	DW_STAT(2, 1);
	DW_STAT(3, n);
	uint32_t a = st.cons_a;
	if (valid) {
		const uint32_t qa = *reinterpret_cast<const uint16_t *>(s.sm + s.q + 2u * ((st.e_hi + lane) & (DW_QCAP - 1)));
		a = st.cons_a + ((qa - st.cons_a) & 0xFFFFu);
	}
	// tag and the three bytes after it (snappy_decompress.c:242-283)
	const uint32_t al = a & ~3u;
	const uint32_t w0 = lds32(s, s.stg + (al & DW_RMASK)), w1 = lds32(s, s.stg + ((al + 4) & DW_RMASK));
	const uint32_t v = wp::funnel_r(w0, w1, (a & 3u) * 8u);
	const uint32_t kind = v & 3u, n6 = (v >> 2) & 63u;
	uint32_t len, off, foot;
	bool bad = false;
	if (kind == 0) {            // literal, length in the tag (explicit length bytes never get here)
		len = n6 + 1;
		off = 0;
		foot = len + 1;
	} else if (kind == 1) {     // dpu_snappy.h:10-11
		len = 4 + (n6 & 7u);
		off = ((v >> 5) & 7u) << 8 | ((v >> 8) & 0xFFu);
		foot = 2;
	} else if (kind == 2) {     // dpu_snappy.h:12
		len = n6 + 1;
		off = (v >> 8) & 0xFFFFu;
		foot = 3;
	} else {                    // snappy_decompress.c:118-133
		len = n6 + 1;
		foot = 5;
		const uint32_t b4 = lds8(s, s.stg + ((a + 4) & DW_RMASK));
		off = (v >> 8) | (b4 << 24);
		if (off > 65535u) {
			bad = true;
			off = 1;
		}
	}
	if (!valid) {
		len = 0;
		foot = 0;
		bad = false;
	}
	bad = bad || (valid && a + foot > st.a_end);
	const uint32_t incl = scan_incl(len, lane);
	const uint32_t pn = st.dec_y + incl, p = pn - len;
	const bool is_copy = valid && kind != 0;
	bad = bad || (valid && pn > st.yend) || (is_copy && (off == 0 || off > p - st.ybeg));
	if (wp::ballot(bad)) {
		st.err = true;
		return;
	}
	// runs of copies with one distance
	const uint32_t key = is_copy ? off : 0u;
	uint32_t prev = wp::shfl(key, lane - 1);
	if (lane == 0)
		prev = st.c_off;
	const bool chained = is_copy && off == prev && off < 32768u;
	const uint32_t heads = wp::ballot(!chained) & ((2u << lane) - 1u);
	const uint32_t hp = wp::shfl(p, heads ? 31 - wp::clz32(heads) : lane);
	const uint32_t rbase = chained ? (heads ? hp : st.c_base) : p;
	const bool periodic = is_copy && pn - rbase > off;
	DRec rc;
	rc.pn = pn;
	rc.dl = is_copy ? 0u - off : a + 1u - p;
	rc.w2 = is_copy ? (s.ring | (periodic ? DW_F_PERIODIC : (off > DW_NEAR_MAX ? DW_F_FAR : (off < DW_WB ? DW_F_NEAR : 0u)))) : s.stg;
	rc.w3 = periodic ? (off | (rbase << 15)) : 0u;
	if (!valid) { // sentinel after the last record: never ends, reads the ring (lanes past the decoded range walk into it)
		rc.pn = 0xFFFFFFFFu;
		rc.dl = 0;
		rc.w2 = s.ring;
		rc.w3 = 0;
	}
	if ((uint32_t)lane <= n)
		*reinterpret_cast<DRec *>(s.sm + s.rec + 16u * ((st.e_hi + lane) & (DW_NREC - 1))) = rc;
	if (n == 32 && lane == 0) {
		DRec sn;
		sn.pn = 0xFFFFFFFFu;
		sn.dl = 0;
		sn.w2 = s.ring;
		sn.w3 = 0;
		*reinterpret_cast<DRec *>(s.sm + s.rec + 16u * ((st.e_hi + 32) & (DW_NREC - 1))) = sn;
	}
	const int last = (int)n - 1;
	st.c_off = wp::shfl(key, last);
	st.c_base = wp::shfl(rbase, last);
	st.cons_a = wp::shfl(a + foot, last);
	st.dec_y = wp::shfl(pn, last);
	st.e_hi += n;
	wp::syncwarp();
}

// ---- 3. gather by output word ----------------------------------------------------------------
PW_FN uint32_t word_get(const uint32_t w[DW_BPL / 4], uint32_t k) { return (w[k >> 2] >> (8u * (k & 3u))) & 0xFFu; }
PW_FN void word_put(uint32_t w[DW_BPL / 4], uint32_t k, uint32_t v)
{
	const uint32_t sh = 8u * (k & 3u);
	w[k >> 2] = (w[k >> 2] & ~(0xFFu << sh)) | (v << sh);
}

// One pass: output bytes [lo, hi) of the window [wst, wst + DW_WB), wst = lo rounded down to DW_BPL.
// Bytes of lo's word below lo keep their ring content; bytes at and above hi are not valid afterwards
// (the word that holds hi is computed again by the next pass).
PW_FN void gather_pass(const DSm &s, State &st, uint32_t wst, uint32_t lo, uint32_t hi)
{
	const int lane = wp::lane();
	const uint32_t x0 = wst + DW_BPL * (uint32_t)lane;
	DW_STAT(4, 1);
	DW_STAT(5, hi - lo);

	// covering element of x0: number of records that end at or before x0
	uint32_t e0 = st.e_hi;
	{
		bool found = x0 >= st.dec_y; // nothing decoded covers me: every record ends at or before x0
		for (uint32_t eb = st.e_lo;; eb += 32) {
			const uint32_t idx = eb + (uint32_t)lane;
			const uint32_t cand = idx < st.e_hi ? lds32(s, s.rec + 16u * (idx & (DW_NREC - 1))) : 0xFFFFFFFFu;
			uint32_t c = 0;
#pragma unroll
			for (int step = 16; step >= 1; step >>= 1) {
				const uint32_t t = wp::shfl(cand, (int)c + step - 1);
				if (t <= x0)
					c += step;
			}
			const uint32_t lastc = wp::shfl(cand, 31);
			if (!found && lastc > x0) {
				e0 = eb + c;
				found = true;
			}
			DW_STAT(6, 1);
			if (!wp::ballot(!found))
				break;
		}
	}
	uint32_t e = e0; // (a lane past the decoded range reads the sentinel record: its bytes are not alive)
	DRec rc = *reinterpret_cast<const DRec *>(s.sm + s.rec + 16u * (e & (DW_NREC - 1)));
	uint32_t w[DW_BPL / 4];
	uint32_t dlk[DW_BPL], w2k[DW_BPL];
	uint32_t adv = 0, slow = 0;
#pragma unroll
	for (int i = 0; i < (int)DW_BPL / 4; i++)
		w[i] = 0;
#pragma unroll
	for (uint32_t k = 0; k < DW_BPL; k++) {
		const uint32_t x = x0 + k;
		if (x >= rc.pn) { // (an element is at least one byte long: one step per byte is enough)
			e++;
			adv |= 1u << k;
			rc = *reinterpret_cast<const DRec *>(s.sm + s.rec + 16u * (e & (DW_NREC - 1)));
		}
		dlk[k] = rc.dl;
		w2k[k] = rc.w2;
		slow |= rc.w2;
		const uint32_t addr = ((x + rc.dl) & DW_RMASK) | (rc.w2 & ~DW_RMASK);
		w[k >> 2] |= lds8(s, addr) << (8u * (k & 3u));
	}
	// which of my bytes are computed by this pass
	uint32_t alive = 0;
#pragma unroll
	for (uint32_t k = 0; k < DW_BPL; k++)
		alive |= (x0 + k >= lo && x0 + k < hi) ? 1u << k : 0u;

	// far sources, sources inside the window, periodic elements
	uint32_t pend = 0;
	if (wp::ballot((slow & DW_F_ALL) != 0 && alive != 0)) {
		DW_STAT(7, 1);
#pragma unroll
		for (uint32_t k = 0; k < DW_BPL; k++) {
			const uint32_t fl = w2k[k] & DW_F_ALL;
			if (fl && ((alive >> k) & 1u)) {
				const uint32_t x = x0 + k;
				uint32_t src;
				if (fl & DW_F_PERIODIC) {
					const uint32_t ek = e0 + (uint32_t)wp::popc32(adv & ((2u << k) - 1u));
					const uint32_t w3 = lds32(s, s.rec + 16u * (ek & (DW_NREC - 1)) + 12u);
					const uint32_t poff = w3 & 0x7FFFu, pbase = w3 >> 15;
					src = pbase - poff + (x - pbase) % poff;
					dlk[k] = src - x;
				} else {
					src = x + dlk[k];
				}
				if (src >= wst) {
					pend |= 1u << k;
				} else if (x - src > DW_NEAR_MAX) {
					word_put(w, k, wp::ld8_l2(st.gdst_al + src));
				} else if (fl & DW_F_PERIODIC) {
					word_put(w, k, lds8(s, s.ring + (src & DW_RMASK)));
				}
			}
		}
	}
	// keep the bytes below lo (lane 0 of a pass that starts inside a word)
	const uint32_t ro = s.ring + (x0 & DW_RMASK);
	if (lo > wst && x0 < lo) {
#pragma unroll
		for (uint32_t k = 0; k < DW_BPL; k++)
			if (x0 + k < lo)
				word_put(w, k, lds8(s, ro + k));
	}
	if (DW_BPL == 8) {
		U2 t;
		t.x = w[0];
		t.y = w[DW_BPL / 4 - 1];
		*reinterpret_cast<U2 *>(s.sm + ro) = t;
	} else {
		*reinterpret_cast<uint32_t *>(s.sm + ro) = w[0];
	}
	wp::syncwarp();

	// fix-up rounds: a byte whose source lies inside the window waits until everything below the first
	// waiting byte is final
	const bool had_pend = pend != 0;
	for (;;) {
		const uint32_t first = wp::reduce_min(pend ? x0 + (uint32_t)wp::ffs32(pend) - 1u : 0xFFFFFFFFu);
		if (first == 0xFFFFFFFFu)
			break;
		DW_STAT(8, 1);
#pragma unroll
		for (uint32_t k = 0; k < DW_BPL; k++) {
			if ((pend >> k) & 1u) {
				const uint32_t src = x0 + k + dlk[k];
				if (src < first) {
					s.sm[ro + k] = (uint8_t)lds8(s, s.ring + (src & DW_RMASK));
					pend &= ~(1u << k);
				}
			}
		}
		wp::syncwarp();
	}
	if (had_pend) {
		w[0] = lds32(s, ro);
		if (DW_BPL == 8)
			w[DW_BPL / 4 - 1] = lds32(s, ro + 4);
	}
	// one word store to HBM; the words at the two ends of [lo, hi) go bytewise
	uint8_t *g = st.gdst_al + x0;
	if (alive == (1u << DW_BPL) - 1u) {
		if (DW_BPL == 8) {
			U2 t;
			t.x = w[0];
			t.y = w[DW_BPL / 4 - 1];
			*reinterpret_cast<U2 *>(g) = t;
		} else {
			*reinterpret_cast<uint32_t *>(g) = w[0];
		}
	} else if (alive) {
#pragma unroll
		for (uint32_t k = 0; k < DW_BPL; k++)
			if ((alive >> k) & 1u)
				wp::st_global_u8(g + k, word_get(w, k));
	}
	// first record the next pass needs: the one covering the word that holds hi
	const uint32_t nw = (hi - wst) / DW_BPL; // lane that owns hi's word, 32 = the pass was full
	const uint32_t e_next = nw < 32 ? wp::shfl(e0, (int)nw) : wp::shfl(e, 31);
	st.e_lo = wp::umin(e_next, st.e_hi);
	st.gy = hi;
	st.lo_dead = false;
	wp::syncwarp();
}

// ---- literals with explicit length bytes: HBM -> HBM -------------------------------------------
// The last DW_RB - DW_WB bytes also go into the ring (later copies may reach back into them).
PW_FN void bulk_literal(const DSm &s, State &st)
{
	const int lane = wp::lane();
	const uint32_t y0 = st.gy, len = st.b_len;
	DW_STAT(9, 1);
	if (y0 + len > st.yend || st.b_src + len > st.a_end) {
		st.err = true;
		return;
	}
	const uint8_t *src = st.gcomp_al + st.b_src;
	const uint32_t yend = y0 + len;
	const uint32_t keep_from = len > DW_RB - DW_WB ? yend - (DW_RB - DW_WB) : y0;
	uint32_t head = (16u - (y0 & 15u)) & 15u;
	if (head > len)
		head = len;
	const uint32_t body = (len - head) & ~15u;
	for (uint32_t k = lane; k < head; k += 32) {
		const uint32_t b = wp::ldg8(src + k);
		wp::st_global_u8(st.gdst_al + y0 + k, b);
		if (y0 + k >= keep_from)
			s.sm[s.ring + ((y0 + k) & DW_RMASK)] = (uint8_t)b;
	}
	if (body) {
		const uint8_t *sb = src + head;
		const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(sb) & 3u);
		const uint32_t sh = mis * 8u;
		const uint32_t *s32 = reinterpret_cast<const uint32_t *>(sb - mis);
		const uint32_t yb = y0 + head;
		for (uint32_t c = lane; c < body / 16; c += 32) {
			const uint32_t q0 = wp::ldg32(s32 + 4 * c), q1 = wp::ldg32(s32 + 4 * c + 1), q2 = wp::ldg32(s32 + 4 * c + 2), q3 = wp::ldg32(s32 + 4 * c + 3);
			const uint32_t q4 = mis ? wp::ldg32(s32 + 4 * c + 4) : 0u;
			U4 t;
			t.x = wp::funnel_r(q0, q1, sh);
			t.y = wp::funnel_r(q1, q2, sh);
			t.z = wp::funnel_r(q2, q3, sh);
			t.w = wp::funnel_r(q3, q4, sh);
			const uint32_t y = yb + 16 * c;
			*reinterpret_cast<U4 *>(st.gdst_al + y) = t;
			if (y + 16 > keep_from) // (a 16-byte piece that straddles keep_from is kept whole)
				*reinterpret_cast<U4 *>(s.sm + s.ring + (y & DW_RMASK)) = t;
		}
	}
	for (uint32_t k = head + body + lane; k < len; k += 32) {
		const uint32_t b = wp::ldg8(src + k);
		wp::st_global_u8(st.gdst_al + y0 + k, b);
		if (y0 + k >= keep_from)
			s.sm[s.ring + ((y0 + k) & DW_RMASK)] = (uint8_t)b;
	}
	st.gy = st.dec_y = yend;
	st.cons_a = st.pa;
	st.e_lo = st.e_hi;
	st.lo_dead = true;
	st.c_off = 0;
	st.bulk = false;
	wp::syncwarp();
}

} // namespace dw

// Decode one block with the calling warp.  Returns false if the block is malformed (the output is then
// unspecified but nothing outside [gdst, gdst + out_size) was written).
PW_FN bool decode_block_warp(const DSm &s, const DBlk &b)
{
	using namespace dw;
	const int lane = wp::lane();
	State st;
	st.amis = (uint32_t)(reinterpret_cast<uintptr_t>(b.gcomp) & 15u);
	st.a_end = st.amis + b.csize;
	st.ybeg = (uint32_t)(reinterpret_cast<uintptr_t>(b.gdst) & 15u);
	st.yend = st.ybeg + b.out_size;
	st.gcomp_al = b.gcomp - st.amis;
	st.gdst_al = b.gdst - st.ybeg;
	st.pa = st.amis;
	st.st_issued = 0;
	st.qt = st.e_hi = st.e_lo = 0;
	st.cons_a = st.amis;
	st.dec_y = st.gy = st.ybeg;
	st.lo_dead = true;
	st.c_off = st.c_base = 0;
	st.bulk = false;
	st.b_src = st.b_len = 0;
	st.err = false;
	const uint32_t nchunks = (st.a_end + DW_CH - 1) / DW_CH;

	// request chunks below `upto` unless a slot still holds bytes that queued / undecoded elements need
	auto stage_upto = [&](uint32_t upto) -> bool {
		if (st.st_issued >= upto)
			return true;
		uint32_t oldest = st.pa;
		if (st.e_lo < st.qt) {
			const uint32_t qa = *reinterpret_cast<const uint16_t *>(s.sm + s.q + 2u * (st.e_lo & (DW_QCAP - 1)));
			oldest = st.e_lo < st.e_hi ? st.cons_a - ((st.cons_a - qa) & 0xFFFFu)  // decoded: before cons_a
			                           : st.cons_a + ((qa - st.cons_a) & 0xFFFFu); // still queued: at or after it
		}
		while (st.st_issued < upto) {
			if (st.st_issued >= 3 && oldest < (st.st_issued - 3) * DW_CH)
				return false;
			stage_chunk(s, st, st.st_issued, lane);
			st.st_issued++;
		}
		return true;
	};

	for (;;) {
		if (st.err)
			return false;
		const bool more = !st.bulk && st.pa < st.a_end;
		bool force = false;
		if (more && st.qt - st.e_hi < 32) {
			const uint32_t kc = st.pa / DW_CH;
			if (st.st_issued < kc)
				st.st_issued = kc; // chunks a bulk literal jumped over are never needed
			if (stage_upto(wp::umin(kc + 2, nchunks))) {
				wp::cp_wait_all();
				wp::syncwarp();
				parse_chunk(s, st);
				stage_upto(wp::umin(kc + 3, nchunks)); // prefetch
				continue;
			}
			force = true; // the stage ring is full of bytes still needed: consume first
			DW_STAT(10, 1);
		}
		const uint32_t qn = st.qt - st.e_hi;
		if (qn > 0 && (st.e_hi - st.e_lo) + 33 <= DW_NREC && (qn >= 32 || !more || force)) {
			decode_batch(s, st);
			if (st.err)
				return false;
		}
		const bool final = st.qt == st.e_hi && !more;
		while (st.gy < st.dec_y) {
			const uint32_t wst = st.gy & ~(DW_BPL - 1);
			const bool full = st.dec_y >= wst + DW_WB;
			const bool recfull = (st.e_hi - st.e_lo) + 33 > DW_NREC;
			if (!(full || final || force || recfull))
				break;
			gather_pass(s, st, wst, st.lo_dead ? st.gy : wst, wp::umin(st.dec_y, wst + DW_WB));
			if (st.qt - st.e_hi >= 32 && (st.e_hi - st.e_lo) + 33 <= DW_NREC)
				break; // room for another batch: keep the record window full
		}
		if (final && st.gy == st.dec_y) {
			if (!st.bulk)
				break;
			bulk_literal(s, st);
		}
	}
	wp::cp_wait_all();
	return !st.err && st.pa == st.a_end && st.gy == st.yend;
}

} // namespace pimsnappy
