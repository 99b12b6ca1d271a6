// Lock-step warp emulator for the per-warp compressor (test infrastructure, CPU only).
//
// Compiles pim-compression_b200/csrc/compress_warp.cuh -- the very source the CUDA kernel
// includes -- with PIM_WARP_EMU, and runs its 32 lanes as 32 cooperative contexts that meet
// at every warp primitive (ballot / shfl / match_any / syncwarp).  emu_compress() builds the
// whole file image (header varints + blocks, snappy/README.md:19-33 of the reference) so a
// test can compare it with the oracle byte for byte without a GPU.
//
// Build: g++ -O1 -shared -fPIC -DPIM_WARP_EMU tests/emu/warp_emu.cpp -o tests/emu/_build/libwarp_emu.so

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ucontext.h>
#include <vector>

#define PIM_WARP_EMU_STATS
#include "../../pim-compression_b200/csrc/compress_warp.cuh"

namespace pimsnappy {
unsigned long cw_stats[8]; // 0 ic passes, 1 probe passes, 2 pairs, 3 redone pairs, 4 long matches, 5 serial re-resolves, 6 passes with shared slots
}
extern "C" void emu_set_order(int order);
extern "C" void emu_stats(unsigned long out[8], int reset)
{
	for (int i = 0; i < 8; i++) {
		out[i] = pimsnappy::cw_stats[i];
		if (reset)
			pimsnappy::cw_stats[i] = 0;
	}
}

namespace {

constexpr int NL = 32;  // lanes per warp
constexpr int NW = 2;   // warps: 0 parser, 1 emitter
constexpr int NC = NL * NW;
constexpr size_t STACK = 256 * 1024;

ucontext_t g_main, g_ctx[NC];
int g_cur = 0;
bool g_done[NC];
uint32_t g_slot[NC];
int g_op[NC];          // which primitive each lane is in (divergence check)
unsigned g_opseq[NC];  // how many primitives each lane has executed
int g_global_table = 0; // 1: run parse_block_warp<true>
int g_order = 0; // 0: lanes run 0..31 between primitives (lane 31's store wins a conflict), 1: 31..0, 2: shuffled
void (*g_body)(void *) = nullptr;
void *g_arg = nullptr;

void yield() { swapcontext(&g_ctx[g_cur], &g_main); }

// Publish v, wait until every lane of the warp has published for the same primitive, combine.
// Lanes may arrive rounds apart (a wait loop on another warp's flag lets them leave at different
// times, as on the GPU, where they reconverge at the next primitive).  Slots are double-buffered
// by the parity of the primitive's sequence number: a lane can only overwrite a slot two
// primitives later, which needs every lane to have published the next one, i.e. to be done
// reading this one.
uint32_t g_slot2[2][NC];
int g_op2[2][NC];

template <class F>
uint32_t exchange(int op, uint32_t v, F &&combine)
{
	const unsigned seq = ++g_opseq[g_cur];
	const int par = seq & 1;
	g_slot2[par][g_cur] = v;
	g_op2[par][g_cur] = op;
	const int w0 = (g_cur / NL) * NL;
	for (;;) {
		bool all = true;
		for (int l = w0; l < w0 + NL; l++) {
			if (g_opseq[l] < seq) {
				if (g_done[l]) {
					fprintf(stderr, "warp_emu: lane %d finished while lane %d waits at a warp primitive\n", l, g_cur);
					abort();
				}
				all = false;
			}
		}
		if (all)
			break;
		yield();
	}
	for (int l = w0; l < w0 + NL; l++) {
		if (g_op2[par][l] != op) {
			fprintf(stderr, "warp_emu: lanes diverged at a warp primitive (lane %d op %d vs lane %d op %d, seq %u)\n", g_cur, op, l, g_op2[par][l], seq);
			abort();
		}
	}
	for (int l = 0; l < NL; l++)
		g_slot[w0 + l] = g_slot2[par][w0 + l]; // combine() reads g_slot of the own warp
	return combine();
}

void trampoline()
{
	g_body(g_arg);
	g_done[g_cur] = true;
	swapcontext(&g_ctx[g_cur], &g_main);
}

void run_warp(void (*body)(void *), void *arg)
{
	static std::vector<char> stacks(NC * STACK);
	g_body = body;
	g_arg = arg;
	for (int l = 0; l < NC; l++) {
		getcontext(&g_ctx[l]);
		g_ctx[l].uc_stack.ss_sp = stacks.data() + l * STACK;
		g_ctx[l].uc_stack.ss_size = STACK;
		g_ctx[l].uc_link = &g_main;
		makecontext(&g_ctx[l], trampoline, 0);
		g_done[l] = false;
		g_op[l] = -1;
		g_opseq[l] = 0;
	}
	for (;;) {
		bool any = false;
		static unsigned rnd = 12345;
		rnd = rnd * 1664525u + 1013904223u;
		for (int i = 0; i < NC; i++) {
			const int l = g_order == 0 ? i : (g_order == 1 ? NC - 1 - i : (int)((i * 13u + (rnd >> 16)) & (NC - 1)));
			if (g_done[l])
				continue;
			any = true;
			g_cur = l;
			swapcontext(&g_main, &g_ctx[l]);
		}
		if (!any)
			break;
	}
}

} // namespace

namespace pimsnappy {
namespace wp {
int lane() { return g_cur % NL; }
static int wbase() { return (g_cur / NL) * NL; }
uint32_t ballot(bool pred)
{
	return exchange(1, pred ? 1u : 0u, [] {
		uint32_t m = 0;
		for (int l = 0; l < NL; l++)
			m |= (g_slot[wbase() + l] & 1u) << l;
		return m;
	});
}
uint32_t shfl(uint32_t v, int src_lane)
{
	const int s = src_lane & 31;
	return exchange(2, v, [s] { return g_slot[wbase() + s]; });
}
uint32_t match_any(uint32_t v)
{
	return exchange(3, v, [v] {
		uint32_t m = 0;
		for (int l = 0; l < NL; l++)
			m |= (g_slot[wbase() + l] == v ? 1u : 0u) << l;
		return m;
	});
}
uint32_t reduce_max(uint32_t v)
{
	return exchange(5, v, [] {
		uint32_t m = 0;
		for (int l = 0; l < NL; l++)
			m = g_slot[wbase() + l] > m ? g_slot[wbase() + l] : m;
		return m;
	});
}
void syncwarp()
{
	exchange(4, 0, [] { return 0u; });
}
void spin_pause() { yield(); }
} // namespace wp
} // namespace pimsnappy

namespace {

struct Job {
	pimsnappy::CBlock blk;
	pimsnappy::EmitRing ring;
	uint32_t result;
};

void body(void *arg)
{
	Job *j = static_cast<Job *>(arg);
	uint32_t seq = 0; // records of this block (the emulator's ring counters restart with every block)
	if (g_cur < NL) {
		if (g_global_table)
			pimsnappy::parse_block_warp<true>(j->blk, &j->ring, seq); // the code path of the pairs whose table lives in global memory
		else
			pimsnappy::parse_block_warp<false>(j->blk, &j->ring, seq);
	} else {
		const uint32_t r = pimsnappy::emit_block_warp(j->blk, &j->ring, seq);
		if (pimsnappy::wp::lane() == 0)
			j->result = r;
	}
}

size_t put_varint(uint8_t *p, uint32_t v) // snappy_compress.c:69-98
{
	size_t k = 0;
	while (v >= 128) {
		p[k++] = (uint8_t)(v | 128);
		v >>= 7;
	}
	p[k++] = (uint8_t)v;
	return k;
}

} // namespace

extern "C" void emu_set_order(int order) { g_order = order; }
extern "C" void emu_set_global_table(int on) { g_global_table = on; }

// Compress in[0..len) into out (capacity >= 32 + len + len/6 + 40 * blocks); returns the file size.
extern "C" uint64_t emu_compress(const uint8_t *in, uint64_t len, uint32_t block_size, uint8_t *out)
{
	using namespace pimsnappy;
	uint64_t o = put_varint(out, (uint32_t)len);
	o += put_varint(out + o, block_size);
	std::vector<uint32_t> tablemem(CW_MAX_TABLE / 2 + 16);
	std::vector<uint32_t> inmem((block_size + 128) / 4 + 16);
	std::vector<uint32_t> slotmem((block_size + block_size / 6 + 128) / 4 + 16);
	for (uint64_t at = 0; at < len; at += block_size) {
		const uint32_t n = (uint32_t)((len - at < block_size) ? len - at : block_size);
		// vary the alignment of input and output like the device does (input keeps the global address' low 4 bits)
		uint8_t *inb = reinterpret_cast<uint8_t *>(inmem.data() + 8) + (at & 15);
		memset(inmem.data(), 0xA5, inmem.size() * 4);
		memcpy(inb, in + at, n);
		const uint32_t ts = cw_table_entries(n);
		memset(tablemem.data(), 0, 2 * ts);
		Job j;
		j.blk.inw = inmem.data() + 8;
		j.blk.mis = (uint32_t)(at & 15);
		j.blk.table = reinterpret_cast<uint16_t *>(tablemem.data());
		j.blk.shift = 32 - (31 - __builtin_clz(ts));
		j.blk.out = reinterpret_cast<uint8_t *>(slotmem.data());
		j.blk.n = n;
		j.result = 0;
		j.ring.head = j.ring.tail = 0;
		run_warp(body, &j);
		memcpy(out + o, j.blk.out, j.result);
		o += j.result;
	}
	return o;
}
