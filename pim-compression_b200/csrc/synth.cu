// Seeded, integer-only workload generators -- bit-identical on host and device.
//
// The corpus files BASELINE.json names (dickens, mozilla, spamfile) are absent from the reference
// checkout (.MISSING_LARGE_BLOBS), so benchmarks and tests run on the stand-ins of SURVEY.md
// section 8(d).  Definition (also in DESIGN.md):
//
//   PRNG      xorshift64*:  x ^= x>>12; x ^= x<<25; x ^= x>>27; return x * 0x2545F4914F6CDD1D
//   segments  the stream is cut into 64 KiB generator segments; segment g of a stream with seed S
//             starts from x = S ^ ((g+1) * 0x9E3779B97F4A7C15) (0 replaced by a constant), two
//             warm-up draws.  Segments are independent, so any sub-range can be produced alone.
//   random    little-endian bytes of successive draws.
//   text      vocabulary of 2048 words (seed 0x5EED0000; word length 2 + ((r & 15) % 11), letters
//             'a' + ((draw >> 33) % 26)); each emitted word index comes from one draw r with
//             a,b,c = its three low 16-bit fields: idx = ((((a*b)>>16)*c)>>16) * 2048 >> 16
//             (heavily skewed to small indices); words are separated by ' ', or by '\n' once the
//             line is longer than 72 characters.
//   zero      0x00.   counter: u32le(base + 4*i), base = one draw.   record: one 37-byte random
//             record repeated.
//   kinds     TEXT / RANDOM: every segment alike.  MIX: 1 MiB (16 segments) text, 1 MiB random,
//             alternating.  BIN: segment g is {random, zero, text, counter, record}[g % 5].

#include <thread>
#include <vector>
#include <mutex>
#include <cstring>

#include <cstdint>
#include <algorithm>
#include <cuda_runtime.h>

#include "../../include/pim_synth.h"

namespace pimsnappy {
namespace {

constexpr uint32_t SEG = 65536;
constexpr uint32_t VOCAB = 2048;
constexpr uint32_t WORD_MAX = 12;

struct Rng {
	uint64_t x;
	__host__ __device__ uint64_t next()
	{
		x ^= x >> 12;
		x ^= x << 25;
		x ^= x >> 27;
		return x * 0x2545F4914F6CDD1Dull;
	}
};

__host__ __device__ inline Rng seg_rng(uint64_t seed, uint64_t g)
{
	Rng r;
	r.x = seed ^ ((g + 1) * 0x9E3779B97F4A7C15ull);
	if (r.x == 0)
		r.x = 0x106689D45497FDB5ull;
	r.next();
	r.next();
	return r;
}

__host__ __device__ inline void build_vocab(uint8_t *words /* VOCAB*WORD_MAX */, uint8_t *lens /* VOCAB */)
{
	Rng r;
	r.x = 0x5EED0000ull;
	for (uint32_t k = 0; k < VOCAB; k++) {
		const uint32_t len = 2 + (uint32_t)((r.next() & 15) % 11);
		lens[k] = (uint8_t)len;
		for (uint32_t i = 0; i < WORD_MAX; i++)
			words[k * WORD_MAX + i] = i < len ? (uint8_t)('a' + (r.next() >> 33) % 26) : 0;
	}
}

// packs bytes into 8-byte stores (segment starts are 8-byte aligned relative to `out`)
struct Emit {
	uint8_t *out;
	uint32_t pos, end;
	uint64_t acc;
	__host__ __device__ void put(uint8_t b)
	{
		acc |= (uint64_t)b << (8 * (pos & 7));
		pos++;
		if ((pos & 7) == 0) {
			flush8();
		}
	}
	__host__ __device__ void flush8()
	{
#ifdef __CUDA_ARCH__
		*reinterpret_cast<uint64_t *>(out + pos - 8) = acc;
#else
		memcpy(out + pos - 8, &acc, 8);
#endif
		acc = 0;
	}
	__host__ __device__ void finish()
	{
		const uint32_t r = pos & 7;
		for (uint32_t i = 0; i < r; i++)
			out[pos - r + i] = (uint8_t)(acc >> (8 * i));
	}
};

enum { K_TEXT = 0, K_RANDOM = 1, K_ZERO = 2, K_COUNTER = 3, K_RECORD = 4 };

__host__ __device__ inline int seg_kind(int kind, uint64_t g)
{
	switch (kind) {
	case PIM_SYNTH_TEXT: return K_TEXT;
	case PIM_SYNTH_RANDOM: return K_RANDOM;
	case PIM_SYNTH_MIX: return ((g >> 4) & 1) ? K_RANDOM : K_TEXT;
	default: {
		const int t[5] = {K_RANDOM, K_ZERO, K_TEXT, K_COUNTER, K_RECORD};
		return t[g % 5];
	}
	}
}

__host__ __device__ inline void gen_segment(int kind, uint64_t seed, uint64_t g, uint8_t *seg_out, uint32_t n, const uint8_t *words,
                                            const uint8_t *lens)
{
	Rng r = seg_rng(seed, g);
	Emit e{seg_out, 0, n, 0};
	switch (seg_kind(kind, g)) {
	case K_RANDOM:
		while (e.pos < n) {
			uint64_t v = r.next();
			for (int i = 0; i < 8 && e.pos < n; i++, v >>= 8)
				e.put((uint8_t)v);
		}
		break;
	case K_ZERO:
		while (e.pos < n)
			e.put(0);
		break;
	case K_COUNTER: {
		const uint32_t base = (uint32_t)r.next();
		for (uint32_t i = 0; e.pos < n; i++) {
			uint32_t v = base + 4 * i;
			for (int b = 0; b < 4 && e.pos < n; b++, v >>= 8)
				e.put((uint8_t)v);
		}
		break;
	}
	case K_RECORD: {
		uint8_t rec[40];
		for (int i = 0; i < 40; i += 8) {
			uint64_t v = r.next();
			for (int b = 0; b < 8; b++, v >>= 8)
				rec[i + b] = (uint8_t)v;
		}
		for (uint32_t i = 0; e.pos < n; i = (i + 1 == 37 ? 0 : i + 1))
			e.put(rec[i]);
		break;
	}
	default: { // text
		uint32_t col = 0;
		while (e.pos < n) {
			const uint64_t d = r.next();
			const uint32_t a = (uint32_t)(d & 0xFFFF), b = (uint32_t)((d >> 16) & 0xFFFF), c = (uint32_t)((d >> 32) & 0xFFFF);
			const uint32_t idx = (((((a * b) >> 16) * c) >> 16) * VOCAB) >> 16;
			const uint32_t len = lens[idx];
			for (uint32_t i = 0; i < len && e.pos < n; i++)
				e.put(words[idx * WORD_MAX + i]);
			col += len;
			if (e.pos < n) {
				if (col > 72) {
					e.put('\n');
					col = 0;
				} else {
					e.put(' ');
					col++;
				}
			}
		}
		break;
	}
	}
	e.finish();
}

__device__ uint8_t g_words[VOCAB * WORD_MAX];
__device__ uint8_t g_lens[VOCAB];

__global__ void synth_vocab_kernel() { build_vocab(g_words, g_lens); }

__global__ void __launch_bounds__(128) synth_kernel(int kind, uint64_t seed, uint8_t *out, uint64_t n)
{
	const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint64_t lo = g * SEG;
	if (lo >= n)
		return;
	const uint32_t len = (uint32_t)min((uint64_t)SEG, n - lo);
	gen_segment(kind, seed, g, out + lo, len, g_words, g_lens);
}

} // namespace
} // namespace pimsnappy

using namespace pimsnappy;

extern "C" int pim_snappy_synth_host(int kind, uint64_t seed, uint8_t *out, uint64_t n)
{
	if (kind < 0 || kind > PIM_SYNTH_BIN)
		return 1;
	static std::once_flag once;
	static std::vector<uint8_t> words(VOCAB * WORD_MAX), lens(VOCAB);
	std::call_once(once, [] { build_vocab(words.data(), lens.data()); });
	const uint64_t nseg = (n + SEG - 1) / SEG;
	unsigned nt = std::thread::hardware_concurrency();
	nt = nt == 0 ? 1 : (nt > 32 ? 32 : nt);
	if (nseg < nt)
		nt = (unsigned)(nseg ? nseg : 1);
	auto work = [&](unsigned t) {
		for (uint64_t g = t; g < nseg; g += nt) {
			const uint64_t lo = g * SEG;
			gen_segment(kind, seed, g, out + lo, (uint32_t)std::min<uint64_t>(SEG, n - lo), words.data(), lens.data());
		}
	};
	if (nt == 1) {
		work(0);
	} else {
		std::vector<std::thread> th;
		for (unsigned t = 0; t < nt; t++)
			th.emplace_back(work, t);
		for (auto &t : th)
			t.join();
	}
	return 0;
}

extern "C" int pim_snappy_synth_device(int kind, uint64_t seed, uint8_t *d_out, uint64_t n, void *stream)
{
	if (kind < 0 || kind > PIM_SYNTH_BIN)
		return 1;
	if ((reinterpret_cast<uintptr_t>(d_out) & 7) != 0)
		return 1;
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	static bool vocab_ready[64] = {false};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess)
		return 1;
	if (dev >= 64 || !vocab_ready[dev]) {
		synth_vocab_kernel<<<1, 1, 0, st>>>();
		if (dev < 64)
			vocab_ready[dev] = true;
	}
	const uint64_t nseg = (n + SEG - 1) / SEG;
	if (nseg)
		synth_kernel<<<(unsigned)((nseg + 127) / 128), 128, 0, st>>>(kind, seed, d_out, n);
	return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
