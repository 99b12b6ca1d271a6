// C-ABI shim: the reference's DPU entry points, served by B200 GPUs.
//
//   snappy_compress_dpu    replaces snappy/snappy_compress.c:487-714
//   snappy_decompress_dpu  replaces snappy/snappy_decompress.c:292-493
//   setup_compression      replaces snappy/snappy_compress.c:418-453
//   setup_decompression    replaces snappy/snappy_decompress.c:187-215
//
// Like the reference, the work is split into contiguous block ranges, one per device
// (snappy_compress.c:494-520, snappy_decompress.c:306-340) -- here one host thread and one
// CUDA stream per GPU, no inter-GPU traffic, and a host-side stitch of the compressed shards
// (snappy_compress.c:697-703).  There is no CPU codec in this library: if no usable GPU is
// present the entry points fail.

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>
#include <vector>
#include <atomic>

#include "common.cuh"

namespace pimsnappy {

static std::atomic<uint64_t> g_launches{0};
static std::atomic<int> g_extended{-1}; // -1: not set yet ($PIM_SNAPPY_EXTENDED_BLOCKS), 0 / 1
uint32_t max_compress_block_size()
{
	int v = g_extended.load(std::memory_order_relaxed);
	if (v < 0) {
		const char *e = getenv("PIM_SNAPPY_EXTENDED_BLOCKS");
		v = e && atoi(e) != 0 ? 1 : 0;
		g_extended.store(v);
	}
	return v ? PIM_SNAPPY_MAX_EXTENDED_BLOCK_SIZE : PIM_SNAPPY_MAX_BLOCK_SIZE;
}

void note_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- per-kernel event timing --------------------------------------------------
static std::atomic<int> g_prof{0};
static std::mutex g_prof_mu;
struct ProfSpan { int which; cudaEvent_t a, b; };
static std::vector<ProfSpan> g_spans;       // completed begin/end pairs awaiting a read
struct ProfOpen { int which; cudaStream_t st; cudaEvent_t a; };
static std::vector<ProfOpen> g_open;        // begun, not yet ended: keyed by (kernel kind, stream)

void prof_begin(int which, cudaStream_t st)
{
	if (!g_prof.load(std::memory_order_relaxed))
		return;
	std::lock_guard<std::mutex> lk(g_prof_mu);
	cudaEvent_t e;
	if (cudaEventCreate(&e) != cudaSuccess)
		return;
	cudaEventRecord(e, st);
	g_open.push_back({which, st, e});
}

void prof_end(int which, cudaStream_t st)
{
	if (!g_prof.load(std::memory_order_relaxed))
		return;
	std::lock_guard<std::mutex> lk(g_prof_mu);
	for (size_t i = g_open.size(); i-- > 0;) {
		if (g_open[i].which != which || g_open[i].st != st)
			continue;
		cudaEvent_t e;
		if (cudaEventCreate(&e) != cudaSuccess)
			return;
		cudaEventRecord(e, st);
		g_spans.push_back({which, g_open[i].a, e});
		g_open.erase(g_open.begin() + (long)i);
		return;
	}
}

namespace {

double now_s()
{
	using namespace std::chrono;
	return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// ---- pinned host allocations made through pim_snappy_host_alloc -------------
std::mutex g_alloc_mu;
struct HostAlloc {
	size_t bytes;
	bool pinned;
};
std::map<void *, HostAlloc> g_allocs;

size_t owned_capacity(const void *p) // capacity from p to the end of our allocation, 0 if not ours
{
	std::lock_guard<std::mutex> lk(g_alloc_mu);
	auto it = g_allocs.upper_bound(const_cast<void *>(p));
	if (it == g_allocs.begin())
		return 0;
	--it;
	const uint8_t *base = static_cast<const uint8_t *>(it->first);
	const uint8_t *q = static_cast<const uint8_t *>(p);
	if (q >= base + it->second.bytes)
		return 0;
	return it->second.bytes - (size_t)(q - base);
}

// ---- devices ----------------------------------------------------------------
struct DeviceCtx {
	int dev = -1;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
	// grow-only device buffers, reused across calls
	uint8_t *d_a = nullptr; size_t cap_a = 0;       // uncompressed side
	uint8_t *d_b = nullptr; size_t cap_b = 0;       // compressed side
	uint8_t *d_scratch = nullptr; size_t cap_s = 0; // compress slots + sizes
	uint64_t *d_off = nullptr; size_t cap_off = 0;  // block offsets (entries)
	uint32_t *d_status = nullptr;
	uint64_t *h_small = nullptr;                    // pinned: [0] total, [1] status
	// chunked pipeline (H2D | kernels | D2H on three streams, two buffers each)
	cudaStream_t s_in = nullptr, s_out = nullptr, stream2 = nullptr;
	cudaStream_t cs_more[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; // further compute streams of the compress pipeline
	std::vector<cudaEvent_t> ev_pool;               // per-group events of the streamed paths (untimed)
	cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_k[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
	cudaEvent_t ev_span[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; // first/last of each stream
	uint64_t *h_pin = nullptr; size_t cap_pin = 0;  // pinned: per-chunk totals / block offsets (entries)
	uint32_t *d_statv = nullptr; size_t cap_statv = 0;
};

std::mutex g_api_mu; // the host entry points run one at a time (they share the per-device streams and grow-only buffers)
std::atomic<int> g_write_file{-1}; // -1: not set yet ($PIM_SNAPPY_WRITE_FILE), 0 / 1
std::mutex g_dev_mu;
std::vector<DeviceCtx> g_devs;
bool g_devs_ready = false;

bool device_usable(int id)
{
	cudaDeviceProp p;
	if (cudaGetDeviceProperties(&p, id) != cudaSuccess)
		return false;
	return p.major >= 10; // the kernels are compiled for sm_100a only
}

void release_device(DeviceCtx &d)
{
	if (d.dev < 0)
		return;
	cudaSetDevice(d.dev);
	if (d.stream) cudaStreamDestroy(d.stream);
	for (auto &e : d.ev) if (e) cudaEventDestroy(e);
	cudaFree(d.d_a); cudaFree(d.d_b); cudaFree(d.d_scratch); cudaFree(d.d_off); cudaFree(d.d_status); cudaFree(d.d_statv);
	if (d.h_small) cudaFreeHost(d.h_small);
	if (d.h_pin) cudaFreeHost(d.h_pin);
	if (d.s_in) cudaStreamDestroy(d.s_in);
	if (d.s_out) cudaStreamDestroy(d.s_out);
	if (d.stream2) cudaStreamDestroy(d.stream2);
	for (auto &c : d.cs_more) if (c) cudaStreamDestroy(c);
	for (auto &e : d.ev_pool) if (e) cudaEventDestroy(e);
	for (auto &e : d.ev_in) if (e) cudaEventDestroy(e);
	for (auto &e : d.ev_k) if (e) cudaEventDestroy(e);
	for (auto &e : d.ev_out) if (e) cudaEventDestroy(e);
	for (auto &e : d.ev_span) if (e) cudaEventDestroy(e);
	cudaGetLastError();
}

int init_devices_locked(const int *ids, int n)
{
	for (auto &d : g_devs)
		release_device(d);
	g_devs.clear();
	std::vector<int> want;
	if (ids && n > 0) {
		want.assign(ids, ids + n);
	} else {
		int count = 0;
		if (cudaGetDeviceCount(&count) != cudaSuccess)
			count = 0;
		if (const char *e = getenv("PIM_SNAPPY_NR_GPUS")) {
			int lim = atoi(e);
			if (lim > 0 && lim < count)
				count = lim;
		}
		for (int i = 0; i < count; i++)
			want.push_back(i);
	}
	for (int id : want) {
		if (!device_usable(id)) {
			fprintf(stderr, "pim_snappy: CUDA device %d is not an sm_100-class GPU; skipped\n", id);
			continue;
		}
		DeviceCtx d;
		d.dev = id;
		if (cudaSetDevice(id) != cudaSuccess) continue;
		if (cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess) continue;
		bool all = true; // every resource, or the device is not used (what was created is released)
		for (auto &e : d.ev) all = all && cudaEventCreate(&e) == cudaSuccess;
		all = all && cudaStreamCreateWithFlags(&d.s_in, cudaStreamNonBlocking) == cudaSuccess;
		all = all && cudaStreamCreateWithFlags(&d.s_out, cudaStreamNonBlocking) == cudaSuccess;
		all = all && cudaStreamCreateWithFlags(&d.stream2, cudaStreamNonBlocking) == cudaSuccess;
		for (auto &c : d.cs_more) all = all && cudaStreamCreateWithFlags(&c, cudaStreamNonBlocking) == cudaSuccess;
		for (auto &e : d.ev_in) all = all && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
		for (auto &e : d.ev_k) all = all && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
		for (auto &e : d.ev_out) all = all && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
		for (auto &e : d.ev_span) all = all && cudaEventCreate(&e) == cudaSuccess;
		all = all && cudaMalloc(&d.d_status, 256) == cudaSuccess;
		all = all && cudaHostAlloc(&d.h_small, 256, cudaHostAllocDefault) == cudaSuccess;
		if (!all) {
			fprintf(stderr, "pim_snappy: cannot set up CUDA device %d (%s); skipped\n", id, cudaGetErrorString(cudaGetLastError()));
			release_device(d);
			continue;
		}
		g_devs.push_back(d);
	}
	g_devs_ready = true;
	return (int)g_devs.size();
}

int ensure_devices()
{
	std::lock_guard<std::mutex> lk(g_dev_mu);
	if (!g_devs_ready)
		init_devices_locked(nullptr, 0);
	return (int)g_devs.size();
}

template <typename T>
bool grow(T *&p, size_t &cap, size_t need)
{
	if (need <= cap)
		return true;
	if (p)
		cudaFree(p);
	p = nullptr;
	cap = 0;
	need = (need + (need >> 3) + 4095) & ~size_t(4095);
	if (cudaMalloc(&p, need * sizeof(T)) != cudaSuccess) {
		fprintf(stderr, "pim_snappy: cudaMalloc(%zu) failed\n", need * sizeof(T));
		return false;
	}
	cap = need;
	return true;
}

bool grow_pinned(uint64_t *&p, size_t &cap, size_t need)
{
	if (need <= cap)
		return true;
	if (p)
		cudaFreeHost(p);
	p = nullptr;
	cap = 0;
	need = (need + (need >> 2) + 511) & ~size_t(511);
	if (cudaHostAlloc(&p, need * sizeof(uint64_t), cudaHostAllocMapped) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	cap = need;
	return true;
}

// Shards larger than two chunks are streamed: chunk i+1 is copied in while chunk i is coded and
// chunk i-1 is copied out.  $PIM_SNAPPY_CHUNK_MIB overrides the chunk size (0 disables streaming).
std::atomic<uint64_t> g_chunk_bytes{~uint64_t(0)}; // ~0: not set yet
uint64_t chunk_bytes()
{
	uint64_t v = g_chunk_bytes.load(std::memory_order_relaxed);
	if (v == ~uint64_t(0)) {
		const char *e = getenv("PIM_SNAPPY_CHUNK_MIB");
		const long mib = e ? atol(e) : 32;
		v = mib <= 0 ? uint64_t(0) : uint64_t(mib) << 20;
		g_chunk_bytes.store(v, std::memory_order_relaxed);
	}
	return v;
}

// Compute streams the streamed compress path spreads its groups over ($PIM_SNAPPY_STREAMS, 1..8).
int compress_streams()
{
	static const int ns = [] {
		const char *e = getenv("PIM_SNAPPY_STREAMS");
		const int v = e ? atoi(e) : 8;
		return v < 1 ? 1 : (v > 8 ? 8 : v);
	}();
	return ns;
}

size_t put_varint32(uint8_t *p, uint32_t v) // snappy_compress.c:69-98
{
	size_t k = 0;
	while (v >= 0x80) {
		p[k++] = (uint8_t)(v | 0x80);
		v >>= 7;
	}
	p[k++] = (uint8_t)v;
	return k;
}

bool get_varint32(const uint8_t *&p, const uint8_t *end, uint32_t *v) // snappy_decompress.c:23-37
{
	uint32_t acc = 0;
	for (int k = 0; k < 5 && p < end; k++) {
		const uint8_t c = *p++;
		acc |= (uint32_t)(c & 0x7F) << (7 * k);
		if (!(c & 0x80)) {
			*v = acc;
			return true;
		}
	}
	return false;
}

// First failed CUDA call of a streamed path, by name (the streams are only synchronised at the end).
struct CudaCheck {
	const char *what = nullptr;
	cudaError_t err = cudaSuccess;
	void operator()(cudaError_t e, const char *w)
	{
		if (e != cudaSuccess && !what) {
			what = w;
			err = e;
		}
	}
	bool failed(int dev) const
	{
		if (what)
			fprintf(stderr, "pim_snappy: %s failed on GPU %d: %s\n", what, dev, cudaGetErrorString(err));
		return what != nullptr;
	}
};
#define CK(call) ck((call), #call)

struct ShardTimes {
	float copy_in = 0, run = 0, copy_out = 0; // ms
	double alloc = 0;                         // s
};

// contiguous block range of shard g out of ng (snappy_compress.c:495)
void shard_range(uint64_t nblocks, int ng, int g, uint64_t *lo, uint64_t *hi)
{
	const uint64_t per = (nblocks + ng - 1) / ng;
	*lo = std::min(nblocks, per * (uint64_t)g);
	*hi = std::min(nblocks, *lo + per);
}

template <typename F>
void for_each_shard(int ng, F &&fn)
{
	if (ng == 1) {
		fn(0);
		return;
	}
	std::vector<std::thread> th;
	for (int g = 0; g < ng; g++)
		th.emplace_back([&, g] { fn(g); });
	for (auto &t : th)
		t.join();
}

inline uint64_t align256(uint64_t v) { return (v + 255) & ~uint64_t(255); }

// ---- streamed shards ---------------------------------------------------------------------------
// A shard larger than two chunks is processed as a sequence of block GROUPS with the whole shard
// resident on the device (streams are < 4 GiB by format): the H2D copies of all groups are queued
// up front and run ahead, each group's kernels start as soon as its bytes have arrived, and its
// result goes back to the host as soon as its kernels are done.  Group sizes ramp up from
// chunk/8 (short pipeline fill) to the chunk size and, for compression, back down (short drain).

cudaEvent_t pool_event(DeviceCtx &d, size_t i)
{
	while (d.ev_pool.size() <= i) {
		cudaEvent_t e = nullptr;
		cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
		d.ev_pool.push_back(e);
	}
	return d.ev_pool[i];
}

// Group sizes (in units) for `total` units with groups of at most `chunk` units.
std::vector<uint64_t> plan_groups(uint64_t total, uint64_t chunk, bool ramp_down)
{
	std::vector<uint64_t> g;
	const uint64_t small = std::max<uint64_t>(1, chunk / 8);
	uint64_t left = total, next = small;
	const uint64_t tail = ramp_down ? chunk : 0; // units kept for the ramp down
	while (left > tail) {
		const uint64_t take = std::min(next, left - tail);
		g.push_back(take);
		left -= take;
		next = std::min(chunk, next * 2);
	}
	while (left) { // halves, down to chunk/8
		const uint64_t take = left <= small ? left : std::max(small, left / 2);
		g.push_back(take);
		left -= take;
	}
	return g;
}

// Compress one shard (a whole number of blocks).  host_out != nullptr: the compressed groups land back to
// back at host_out as soon as each is done.  host_out == nullptr (several GPUs: a shard's place in the file is
// only known once the shards before it are done): the groups stay on the device, `kept` lists them (device
// offset in d_b, bytes) for copy_kept_groups().  Returns the shard's compressed bytes in *total.
int compress_shard_streamed(DeviceCtx &d, const uint8_t *host_in, uint64_t in_bytes, uint32_t block_size, uint8_t *host_out,
                            uint64_t *total, ShardTimes &tm, std::vector<std::pair<uint64_t, uint64_t>> *kept = nullptr)
{
	const uint64_t nblocks = (in_bytes + block_size - 1) / block_size;
	const uint64_t cblocks = std::max<uint64_t>(1, chunk_bytes() / block_size);
	const std::vector<uint64_t> groups = plan_groups(nblocks, cblocks, true);
	const uint64_t ng = groups.size();
	const uint64_t gmax = *std::max_element(groups.begin(), groups.end());
	const uint64_t scratch_stride = align256(pim_snappy_compress_scratch_bytes(gmax * block_size, block_size));
	// The groups' kernels go round robin over NS compute streams.  A block takes about a millisecond
	// and the GPU holds ~1800 of them at a time, so a group's kernel is short of blocks at its start
	// and its end: several groups have to be in flight for the SMs to stay full.
	const int NS = compress_streams();
	cudaStream_t cs[8] = {d.stream, d.stream2, d.cs_more[0], d.cs_more[1], d.cs_more[2], d.cs_more[3], d.cs_more[4], d.cs_more[5]};
	// group g: blocks [blo[g], blo[g+1]), compacted output at d_b + out_at[g], offsets at d_off + off_at[g]
	std::vector<uint64_t> blo(ng + 1, 0), out_at(ng + 1, 0), off_at(ng + 1, 0);
	for (uint64_t g = 0; g < ng; g++) {
		blo[g + 1] = blo[g] + groups[g];
		out_at[g + 1] = out_at[g] + align256(pim_snappy_compress_bound(groups[g] * block_size, block_size));
		off_at[g + 1] = off_at[g] + groups[g] + 1;
	}
	const double a0 = now_s();
	const bool ok = grow(d.d_a, d.cap_a, in_bytes + 64) && grow(d.d_b, d.cap_b, out_at[ng]) && grow(d.d_scratch, d.cap_s, NS * scratch_stride) &&
	                grow(d.d_off, d.cap_off, off_at[ng]) && grow_pinned(d.h_pin, d.cap_pin, ng);
	tm.alloc = now_s() - a0;
	if (!ok)
		return SNAPPY_BUFFER_TOO_SMALL;
	CudaCheck ck;
	// all H2D copies, queued now; event g: group g has arrived
	CK(cudaEventRecord(d.ev_span[0], d.s_in));
	for (uint64_t g = 0; g < ng; g++) {
		const uint64_t lo = blo[g] * block_size, bytes = std::min(in_bytes, blo[g + 1] * block_size) - lo;
		CK(cudaMemcpyAsync(d.d_a + lo, host_in + lo, bytes, cudaMemcpyHostToDevice, d.s_in));
		CK(cudaEventRecord(pool_event(d, g), d.s_in));
	}
	CK(cudaEventRecord(d.ev_span[1], d.s_in));
	int rc = SNAPPY_OK;
	uint64_t at = 0; // host offset of the next compressed group
	bool out_started = false;
	auto finish = [&](uint64_t g) { // group g's size is known once its kernels are done: queue its D2H
		if (cudaEventSynchronize(pool_event(d, ng + g)) != cudaSuccess) {
			ck(cudaGetLastError(), "cudaEventSynchronize(group done)");
			rc = SNAPPY_INVALID_INPUT;
			return;
		}
		const uint64_t t = d.h_pin[g];
		if (host_out) {
			if (!out_started)
				CK(cudaEventRecord(d.ev_span[4], d.s_out));
			out_started = true;
			CK(cudaMemcpyAsync(host_out + at, d.d_b + out_at[g], t, cudaMemcpyDeviceToHost, d.s_out));
		} else if (kept) {
			kept->emplace_back(out_at[g], t);
		}
		at += t;
	};
	CK(cudaEventRecord(d.ev_span[2], cs[0]));
	uint64_t finished = 0; // groups whose D2H is queued (in order: the groups land back to back)
	for (uint64_t g = 0; g < ng && rc == SNAPPY_OK && !ck.what; g++) {
		cudaStream_t st = cs[g % NS];
		const uint64_t lo = blo[g] * block_size, bytes = std::min(in_bytes, blo[g + 1] * block_size) - lo;
		CK(cudaStreamWaitEvent(st, pool_event(d, g), 0));
		rc = compress_device_placed(d.d_a + lo, bytes, block_size, d.d_scratch + (g % NS) * scratch_stride, scratch_stride, d.d_b + out_at[g],
		                            d.d_off + off_at[g], st, false);
		CK(cudaMemcpyAsync(d.h_pin + g, d.d_off + off_at[g] + groups[g], sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		CK(cudaEventRecord(pool_event(d, ng + g), st));
		while (rc == SNAPPY_OK && finished + (uint64_t)(NS - 1) <= g) // keep NS - 1 groups queued behind the running one
			finish(finished++);
	}
	while (rc == SNAPPY_OK && !ck.what && finished < ng)
		finish(finished++);
	if (out_started)
		CK(cudaEventRecord(d.ev_span[5], d.s_out));
	// span 3 closes the kernels' interval: after every compute stream
	for (uint64_t g = ng > (uint64_t)NS ? ng - NS : 0; g < ng; g++)
		CK(cudaStreamWaitEvent(cs[0], pool_event(d, ng + g), 0));
	CK(cudaEventRecord(d.ev_span[3], cs[0]));
	CK(cudaStreamSynchronize(d.s_in));
	for (int k = 0; k < NS; k++)
		CK(cudaStreamSynchronize(cs[k]));
	CK(cudaStreamSynchronize(d.s_out));
	if (ck.failed(d.dev))
		return SNAPPY_INVALID_INPUT;
	if (rc != SNAPPY_OK)
		return rc;
	cudaEventElapsedTime(&tm.copy_in, d.ev_span[0], d.ev_span[1]);
	cudaEventElapsedTime(&tm.run, d.ev_span[2], d.ev_span[3]);
	if (out_started)
		cudaEventElapsedTime(&tm.copy_out, d.ev_span[4], d.ev_span[5]);
	*total = at;
	return SNAPPY_OK;
}

// The compressed groups a shard kept on its device, back to back to host_out.
int copy_kept_groups(DeviceCtx &d, const std::vector<std::pair<uint64_t, uint64_t>> &kept, uint8_t *host_out, ShardTimes &tm)
{
	CudaCheck ck;
	CK(cudaEventRecord(d.ev_span[4], d.s_out));
	uint64_t at = 0;
	for (const auto &k : kept) {
		CK(cudaMemcpyAsync(host_out + at, d.d_b + k.first, k.second, cudaMemcpyDeviceToHost, d.s_out));
		at += k.second;
	}
	CK(cudaEventRecord(d.ev_span[5], d.s_out));
	CK(cudaStreamSynchronize(d.s_out));
	if (ck.failed(d.dev))
		return SNAPPY_INVALID_INPUT;
	cudaEventElapsedTime(&tm.copy_out, d.ev_span[4], d.ev_span[5]);
	return SNAPPY_OK;
}

// Block offsets of a group travel through mapped pinned memory: one small kernel on the compute
// stream instead of a copy that would queue behind the bulk H2D transfers.
__global__ void fetch_offsets_kernel(const uint64_t *__restrict__ host_mapped, uint64_t *__restrict__ d_off, uint64_t n)
{
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
		d_off[i] = host_mapped[i];
}

// Decompress a whole stream body (nblocks blocks, out_bytes output bytes).  The serial walk over
// the 4-byte block headers (snappy_decompress.c:317-340) runs on the host WHILE the compressed
// bytes are uploaded: a group of blocks is launched as soon as the upload has passed its end.
int decompress_stream_resident(DeviceCtx &d, const uint8_t *body, uint64_t body_len, uint64_t nblocks, uint32_t block_size,
                               uint64_t out_bytes, uint8_t *host_out, ShardTimes &tm)
{
	if (body_len < 5 * nblocks || nblocks == 0)
		return SNAPPY_INVALID_INPUT; // every block needs its 4-byte size and at least one element byte (a truncated file)
	const std::vector<uint64_t> pieces = plan_groups(body_len, std::max<uint64_t>(chunk_bytes(), 8), false); // compressed bytes
	const uint64_t np = pieces.size();
	const double a0 = now_s();
	const bool ok = grow(d.d_a, d.cap_a, out_bytes + 64) && grow(d.d_b, d.cap_b, body_len + 64) && grow(d.d_off, d.cap_off, nblocks + 1) &&
	                grow_pinned(d.h_pin, d.cap_pin, nblocks + 1 + np) && grow(d.d_statv, d.cap_statv, np);
	tm.alloc = now_s() - a0;
	if (!ok)
		return SNAPPY_BUFFER_TOO_SMALL;
	uint64_t *h_off = d.h_pin;
	uint32_t *h_stat = reinterpret_cast<uint32_t *>(d.h_pin + nblocks + 1);
	uint64_t *h_off_dev = nullptr;
	if (cudaHostGetDevicePointer(&h_off_dev, h_off, 0) != cudaSuccess)
		return SNAPPY_INVALID_INPUT;
	CudaCheck ck;
	// all uploads, queued now; event j: bytes [0, pend[j]) have arrived
	std::vector<uint64_t> pend(np);
	CK(cudaEventRecord(d.ev_span[0], d.s_in));
	for (uint64_t j = 0, lo = 0; j < np; j++) {
		CK(cudaMemcpyAsync(d.d_b + lo, body + lo, pieces[j], cudaMemcpyHostToDevice, d.s_in));
		CK(cudaEventRecord(pool_event(d, j), d.s_in));
		lo += pieces[j];
		pend[j] = lo;
	}
	CK(cudaEventRecord(d.ev_span[1], d.s_in));
	CK(cudaMemsetAsync(d.d_statv, 0, np * sizeof(uint32_t), d.stream));
	CK(cudaEventRecord(d.ev_span[2], d.stream));

	int rc = SNAPPY_OK;
	uint64_t launched = 0; // blocks [0, launched) are on their way
	bool first_out = true;
	int64_t last_piece = -1; // last piece a launch was recorded for
	auto launch = [&](uint64_t bhi, uint64_t piece) { // blocks [launched, bhi) end inside pieces 0..piece
		if (bhi == launched || rc != SNAPPY_OK)
			return;
		const uint64_t blo = launched, nb = bhi - blo;
		const uint64_t o_lo = blo * (uint64_t)block_size, o_bytes = std::min(out_bytes, bhi * (uint64_t)block_size) - o_lo;
		CK(cudaStreamWaitEvent(d.stream, pool_event(d, piece), 0));
		fetch_offsets_kernel<<<(unsigned)std::min<uint64_t>(64, (nb + 256) / 256), 256, 0, d.stream>>>(h_off_dev + blo, d.d_off + blo, nb + 1);
		note_launch(1);
		rc = pim_snappy_decompress_device(d.d_b, d.d_off + blo, nb, block_size, o_bytes, d.d_a + o_lo, d.d_statv + piece, d.stream);
		CK(cudaEventRecord(pool_event(d, np + piece), d.stream));
		CK(cudaStreamWaitEvent(d.s_out, pool_event(d, np + piece), 0));
		if (first_out)
			CK(cudaEventRecord(d.ev_span[4], d.s_out));
		first_out = false;
		CK(cudaMemcpyAsync(host_out + o_lo, d.d_a + o_lo, o_bytes, cudaMemcpyDeviceToHost, d.s_out));
		launched = bhi;
		last_piece = (int64_t)piece;
	};
	uint64_t pos = 0, piece = 0;
	for (uint64_t i = 0; i < nblocks && rc == SNAPPY_OK; i++) {
		if (pos + 4 > body_len) {
			rc = SNAPPY_INVALID_INPUT;
			break;
		}
		h_off[i] = pos;
		const uint32_t cs = (uint32_t)body[pos] | ((uint32_t)body[pos + 1] << 8) | ((uint32_t)body[pos + 2] << 16) | ((uint32_t)body[pos + 3] << 24);
		pos += 4 + (uint64_t)cs;
		if (pos > body_len) {
			rc = SNAPPY_INVALID_INPUT;
			break;
		}
		if (pos > pend[piece]) { // block i ends in a later piece: everything before it goes now
			launch(i, piece);
			while (pos > pend[piece])
				piece++;
		}
	}
	if (rc == SNAPPY_OK && pos != body_len)
		rc = SNAPPY_INVALID_INPUT;
	if (rc == SNAPPY_OK) {
		h_off[nblocks] = pos;
		launch(nblocks, np - 1);
	}
	CK(cudaEventRecord(d.ev_span[3], d.stream));
	if (last_piece >= 0) { // (nothing was launched for a stream that turned out malformed in its first piece)
		CK(cudaStreamWaitEvent(d.s_out, pool_event(d, np + (uint64_t)last_piece), 0));
		CK(cudaMemcpyAsync(h_stat, d.d_statv, np * sizeof(uint32_t), cudaMemcpyDeviceToHost, d.s_out));
	}
	CK(cudaEventRecord(d.ev_span[5], d.s_out));
	CK(cudaStreamSynchronize(d.s_in));
	CK(cudaStreamSynchronize(d.stream));
	CK(cudaStreamSynchronize(d.s_out));
	if (ck.failed(d.dev))
		return SNAPPY_INVALID_INPUT;
	if (rc != SNAPPY_OK)
		return rc;
	for (uint64_t j = 0; j < np; j++)
		if (h_stat[j] != 0)
			return SNAPPY_INVALID_INPUT;
	cudaEventElapsedTime(&tm.copy_in, d.ev_span[0], d.ev_span[1]);
	cudaEventElapsedTime(&tm.run, d.ev_span[2], d.ev_span[3]);
	if (!first_out)
		cudaEventElapsedTime(&tm.copy_out, d.ev_span[4], d.ev_span[5]);
	return SNAPPY_OK;
}

} // namespace
} // namespace pimsnappy

using namespace pimsnappy;

// ------------------------------------------------------------------------ misc

extern "C" const char *pim_snappy_version(void) { return "pim-compression_b200 0.1 (sm_100a)"; }
extern "C" uint64_t pim_snappy_launch_count(void) { return g_launches.load(); }

extern "C" void pim_snappy_profile(int enable) { g_prof.store(enable ? 1 : 0); }

extern "C" int pim_snappy_profile_read(double ms[4], uint64_t n[4])
{
	std::lock_guard<std::mutex> lk(g_prof_mu);
	for (int i = 0; i < 4; i++) {
		ms[i] = 0;
		n[i] = 0;
	}
	int rc = 0;
	for (auto &s : g_spans) {
		float t = 0;
		if (cudaEventSynchronize(s.b) == cudaSuccess && cudaEventElapsedTime(&t, s.a, s.b) == cudaSuccess) {
			ms[s.which] += t;
			n[s.which] += 1;
		} else {
			rc = 1;
		}
		cudaEventDestroy(s.a);
		cudaEventDestroy(s.b);
	}
	g_spans.clear();
	return rc;
}

extern "C" double get_runtime(struct timeval *start, struct timeval *end) // dpu_snappy.c:87-91
{
	return (double)(end->tv_sec - start->tv_sec) + (double)(end->tv_usec - start->tv_usec) / 1e6;
}

extern "C" void *pim_snappy_host_alloc(size_t bytes)
{
	void *p = nullptr;
	if (bytes == 0)
		bytes = 1;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
		cudaGetLastError();
		return nullptr; // no pageable substitute: the caller sees the failure
	}
	std::lock_guard<std::mutex> lk(g_alloc_mu);
	g_allocs[p] = {bytes, true};
	return p;
}

extern "C" void pim_snappy_host_free(void *p)
{
	if (!p)
		return;
	bool pinned = true;
	{
		std::lock_guard<std::mutex> lk(g_alloc_mu);
		auto it = g_allocs.find(p);
		if (it != g_allocs.end()) {
			pinned = it->second.pinned;
			g_allocs.erase(it);
		}
	}
	if (pinned)
		cudaFreeHost(p);
	else
		free(p);
}

namespace pimsnappy {
namespace {
// The buffers setup_compression / setup_decompression hand out (what the reference mallocs, snappy_compress.c:446-447,
// snappy_decompress.c:206-207).  Page-locking costs about half a millisecond per MB (lock + unlock) and pays only where
// the copies must run beside the kernels, i.e. for shards that are streamed (> 2 chunks): a 10 MB file spent 12 of its
// 17 ms in cudaHostAlloc / cudaFreeHost.  Smaller buffers are plain malloc; pim_snappy_host_free() knows which is which.
void *setup_alloc(size_t bytes)
{
	if (bytes == 0)
		bytes = 1;
	const uint64_t chunk = chunk_bytes();
	if (chunk && bytes > 2 * chunk)
		return pim_snappy_host_alloc(bytes);
	void *p = malloc(bytes);
	if (!p)
		return nullptr;
	std::lock_guard<std::mutex> lk(g_alloc_mu);
	g_allocs[p] = {bytes, false};
	return p;
}
} // namespace
} // namespace pimsnappy

extern "C" int pim_snappy_set_devices(const int *device_ids, int n)
{
	std::lock_guard<std::mutex> api(g_api_mu); // never while an entry point is using the devices
	std::lock_guard<std::mutex> lk(g_dev_mu);
	return init_devices_locked(device_ids, n);
}

extern "C" int pim_snappy_device_count(void) { return ensure_devices(); }
extern "C" void pim_snappy_shard_range(uint64_t nblocks, int n_shards, int shard, uint64_t *block_lo, uint64_t *block_hi)
{
	if (n_shards <= 0 || shard < 0 || shard >= n_shards) {
		*block_lo = *block_hi = 0;
		return;
	}
	shard_range(nblocks, n_shards, shard, block_lo, block_hi);
}
namespace pimsnappy {
namespace {
__global__ void rebase_offsets_kernel(uint64_t *off, uint64_t n, uint64_t base)
{
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
		off[i] += base;
}
} // namespace
} // namespace pimsnappy

extern "C" snappy_status pim_snappy_gather_shards(int n_shards, const int *devices, const uint8_t *const *d_shard, const uint64_t *bytes,
                                                  const uint64_t *const *d_shard_off, const uint64_t *nblocks, int dst_device, uint8_t *d_dst,
                                                  uint64_t *d_dst_off, void *stream)
{
	if (n_shards <= 0 || !devices || !d_shard || !bytes || !d_shard_off || !nblocks || !d_dst || !d_dst_off)
		return SNAPPY_INVALID_INPUT;
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	int prev = 0;
	PIM_CUDA_TRY(cudaGetDevice(&prev));
	PIM_CUDA_TRY(cudaSetDevice(dst_device));
	snappy_status rc = SNAPPY_OK;
	uint64_t base = 0, blk = 0;
	for (int g = 0; g < n_shards && rc == SNAPPY_OK; g++) {
		if (devices[g] != dst_device) { // direct peer access if the link allows it; the copies below work either way
			int can = 0;
			if (cudaDeviceCanAccessPeer(&can, dst_device, devices[g]) == cudaSuccess && can && cudaDeviceEnablePeerAccess(devices[g], 0) != cudaSuccess)
				cudaGetLastError(); // already enabled
		}
		const bool last = g + 1 == n_shards;
		const uint64_t nent = nblocks[g] + (last ? 1 : 0); // the closing entry of a shard is the first of the next
		if (cudaMemcpyPeerAsync(d_dst + base, dst_device, d_shard[g], devices[g], bytes[g], st) != cudaSuccess ||
		    (nent && cudaMemcpyPeerAsync(d_dst_off + blk, dst_device, d_shard_off[g], devices[g], nent * sizeof(uint64_t), st) != cudaSuccess)) {
			rc = SNAPPY_INVALID_INPUT;
			break;
		}
		if (nent && base) {
			rebase_offsets_kernel<<<(unsigned)std::min<uint64_t>(256, (nent + 255) / 256), 256, 0, st>>>(d_dst_off + blk, nent, base);
			note_launch(1);
		}
		base += bytes[g];
		blk += nblocks[g];
	}
	if (rc == SNAPPY_OK && cudaGetLastError() != cudaSuccess)
		rc = SNAPPY_INVALID_INPUT;
	cudaSetDevice(prev);
	return rc;
}

extern "C" void pim_snappy_set_chunk_bytes(uint64_t bytes) { g_chunk_bytes.store(bytes == ~uint64_t(0) ? bytes - 1 : bytes); }

// ----------------------------------------------------------------------- setup

extern "C" void setup_compression(struct host_buffer_context *input, struct host_buffer_context *output, struct program_runtime *runtime)
{
	const double t0 = now_s();
	// The reference allocates 32 + n + n/6 (snappy_compress.c:55-60,446-447), which forgets the
	// 4-byte block headers; size for the smallest block size this library accepts being 1 would
	// be 5n, so the bound is taken for block_size >= 64 and re-checked in snappy_compress_dpu.
	const size_t n = input->length;
	const size_t cap = 128 + n + n / 6 + 40 * (n / 64 + 1); // (room for the header beside the streamed path's bound)
	output->buffer = static_cast<uint8_t *>(setup_alloc(cap));
	output->curr = output->buffer;
	output->length = 0;
	runtime->pre = now_s() - t0;
}

extern "C" snappy_status setup_decompression(struct host_buffer_context *input, struct host_buffer_context *output,
                                             struct program_runtime *runtime)
{
	const double t0 = now_s();
	uint32_t dlength = 0;
	const uint8_t *p = input->curr;
	if (!get_varint32(p, input->buffer + input->length, &dlength)) {
		fprintf(stderr, "Failed to read decompressed length\n");
		return SNAPPY_INVALID_INPUT;
	}
	input->curr = const_cast<uint8_t *>(p);
	if (dlength > output->max) {
		fprintf(stderr, "Output length is to big: max=%ld len=%d\n", output->max, dlength);
		return SNAPPY_BUFFER_TOO_SMALL;
	}
	output->buffer = static_cast<uint8_t *>(setup_alloc(((size_t)dlength + 7) | 0x7FF));
	if (!output->buffer)
		return SNAPPY_INVALID_INPUT;
	output->curr = output->buffer;
	output->length = dlength;
	runtime->pre = now_s() - t0;
	return SNAPPY_OK;
}

// -------------------------------------------------------------------- compress

extern "C" snappy_status snappy_compress_gpu(struct host_buffer_context *input, struct host_buffer_context *output, uint32_t block_size,
                                             struct program_runtime *runtime)
{
	std::lock_guard<std::mutex> api(g_api_mu);
	double t0 = now_s();
	runtime->d_alloc = runtime->load = runtime->copy_in = runtime->run = runtime->copy_out = runtime->d_free = 0;
	if (block_size == 0 || block_size > max_compress_block_size()) {
		fprintf(stderr, "pim_snappy: block size %u is outside the format's valid domain (1..%u)\n", block_size, PIM_SNAPPY_MAX_BLOCK_SIZE);
		return SNAPPY_BUFFER_TOO_SMALL;
	}
	if (input->length > 0xFFFFFFFFul)
		return SNAPPY_BUFFER_TOO_SMALL; // lengths are varint32 (snappy_compress.c:461)
	if (!output->buffer)
		return SNAPPY_INVALID_INPUT;

	const uint64_t len = input->length;
	const uint64_t nblocks = pim_snappy_num_blocks(len, block_size);
	// capacity we may assume: ours if we allocated it, else the reference's own bound
	size_t cap = owned_capacity(output->buffer);
	if (cap == 0)
		cap = len ? 32 + len + len / 6 : 16;

	uint8_t hdr[16];
	size_t hlen = put_varint32(hdr, (uint32_t)len); // snappy_compress.c:523-525
	hlen += put_varint32(hdr + hlen, block_size);
	if (cap < hlen)
		return SNAPPY_BUFFER_TOO_SMALL;
	memcpy(output->buffer, hdr, hlen);
	output->curr = output->buffer + hlen;
	output->length = hlen;
	runtime->pre += now_s() - t0;

	t0 = now_s();
	int ng = ensure_devices();
	runtime->load = now_s() - t0;
	if (ng <= 0) {
		fprintf(stderr, "pim_snappy: no usable sm_100 GPU; there is no CPU fallback\n");
		return SNAPPY_INVALID_INPUT;
	}
	if (nblocks == 0)
		return SNAPPY_OK;
	if ((uint64_t)ng > nblocks)
		ng = (int)nblocks;

	std::vector<ShardTimes> tm(ng);
	std::vector<uint64_t> total(ng, 0);
	std::vector<int> rc(ng, SNAPPY_OK);

	// One device, a shard of several chunks and room for the worst case: stream the chunks
	// (copies overlap the kernels, compressed chunks go straight to their final place).
	if (ng == 1 && chunk_bytes() && len > 2 * chunk_bytes() && hlen + pim_snappy_compress_bound(len, block_size) <= cap) {
		DeviceCtx &d = g_devs[0];
		cudaSetDevice(d.dev);
		const int r = compress_shard_streamed(d, input->buffer, len, block_size, output->buffer + hlen, &total[0], tm[0]);
		if (r != SNAPPY_OK)
			return (snappy_status)r;
		runtime->d_alloc = tm[0].alloc;
		runtime->copy_in = tm[0].copy_in * 1e-3;
		runtime->run = tm[0].run * 1e-3;
		runtime->copy_out = tm[0].copy_out * 1e-3;
		output->length = hlen + total[0];
		output->curr = output->buffer + output->length;
		return SNAPPY_OK;
	}

	// Several GPUs, large shards: every GPU streams its shard (uploads overlap the kernels) and keeps the
	// compressed groups on the device until the shards before it are done and its place in the file is known.
	uint64_t max_shard = 0;
	for (int g = 0; g < ng; g++) {
		uint64_t lo, hi;
		shard_range(nblocks, ng, g, &lo, &hi);
		max_shard = std::max(max_shard, (hi - lo) * (uint64_t)block_size);
	}
	const bool stream_shards = chunk_bytes() && max_shard > 2 * chunk_bytes();
	std::vector<std::vector<std::pair<uint64_t, uint64_t>>> kept(ng);

	// phase 1: H2D + kernels + read back each shard's compressed size
	for_each_shard(ng, [&](int g) {
		DeviceCtx &d = g_devs[g];
		uint64_t lo, hi;
		shard_range(nblocks, ng, g, &lo, &hi);
		const uint64_t in_lo = lo * (uint64_t)block_size;
		const uint64_t in_bytes = std::min(len, hi * (uint64_t)block_size) - in_lo;
		const uint64_t nb = hi - lo;
		cudaSetDevice(d.dev);
		if (stream_shards) {
			rc[g] = nb ? compress_shard_streamed(d, input->buffer + in_lo, in_bytes, block_size, nullptr, &total[g], tm[g], &kept[g]) : SNAPPY_OK;
			return;
		}
		const double a0 = now_s();
		bool ok = grow(d.d_a, d.cap_a, in_bytes + 64) && grow(d.d_b, d.cap_b, pim_snappy_compress_bound(in_bytes, block_size)) &&
		          grow(d.d_scratch, d.cap_s, pim_snappy_compress_scratch_bytes(in_bytes, block_size)) && grow(d.d_off, d.cap_off, nb + 1);
		tm[g].alloc = now_s() - a0;
		if (!ok) { rc[g] = SNAPPY_BUFFER_TOO_SMALL; return; }
		CudaCheck ck;
		CK(cudaEventRecord(d.ev[0], d.stream));
		CK(cudaMemcpyAsync(d.d_a, input->buffer + in_lo, in_bytes, cudaMemcpyHostToDevice, d.stream));
		CK(cudaEventRecord(d.ev[1], d.stream));
		rc[g] = pim_snappy_compress_device(d.d_a, in_bytes, block_size, d.d_scratch, d.cap_s, d.d_b, d.d_off, d.stream);
		CK(cudaEventRecord(d.ev[2], d.stream));
		CK(cudaMemcpyAsync(d.h_small, d.d_off + nb, sizeof(uint64_t), cudaMemcpyDeviceToHost, d.stream));
		CK(cudaStreamSynchronize(d.stream));
		if (ck.failed(d.dev)) {
			rc[g] = SNAPPY_INVALID_INPUT;
			return;
		}
		total[g] = d.h_small[0];
		kept[g].assign(1, {0, total[g]});
		cudaEventElapsedTime(&tm[g].copy_in, d.ev[0], d.ev[1]);
		cudaEventElapsedTime(&tm[g].run, d.ev[1], d.ev[2]);
	});
	for (int g = 0; g < ng; g++)
		if (rc[g] != SNAPPY_OK)
			return (snappy_status)rc[g];

	// host-side stitch: shard g lands after the header and the shards before it
	std::vector<uint64_t> at(ng + 1, hlen);
	for (int g = 0; g < ng; g++)
		at[g + 1] = at[g] + total[g];
	if (at[ng] > cap) {
		// Only tiny block sizes get here (4 header bytes per block dominate).  A buffer that
		// setup_compression() allocated is ours to replace; a foreign one is not.
		if (owned_capacity(output->buffer) == 0 || at[ng] > output->max) {
			fprintf(stderr, "pim_snappy: compressed size %llu exceeds the output buffer (%zu)\n", (unsigned long long)at[ng], cap);
			return SNAPPY_BUFFER_TOO_SMALL;
		}
		uint8_t *bigger = static_cast<uint8_t *>(pim_snappy_host_alloc(at[ng] + 64));
		if (!bigger)
			return SNAPPY_BUFFER_TOO_SMALL;
		memcpy(bigger, output->buffer, hlen);
		pim_snappy_host_free(output->buffer);
		output->buffer = bigger;
	}

	// phase 2: D2H straight into the final position
	for_each_shard(ng, [&](int g) {
		DeviceCtx &d = g_devs[g];
		cudaSetDevice(d.dev);
		rc[g] = copy_kept_groups(d, kept[g], output->buffer + at[g], tm[g]);
	});
	for (int g = 0; g < ng; g++) {
		if (rc[g] != SNAPPY_OK)
			return (snappy_status)rc[g];
		runtime->d_alloc = std::max(runtime->d_alloc, tm[g].alloc);
		runtime->copy_in = std::max(runtime->copy_in, tm[g].copy_in * 1e-3);
		runtime->run = std::max(runtime->run, tm[g].run * 1e-3);
		runtime->copy_out = std::max(runtime->copy_out, tm[g].copy_out * 1e-3);
	}
	output->length = at[ng];
	output->curr = output->buffer + at[ng];
	return SNAPPY_OK;
}

// ------------------------------------------------------------------ decompress

extern "C" snappy_status snappy_decompress_gpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                               struct program_runtime *runtime)
{
	std::lock_guard<std::mutex> api(g_api_mu);
	double t0 = now_s();
	runtime->d_alloc = runtime->load = runtime->copy_in = runtime->run = runtime->copy_out = runtime->d_free = 0;
	const uint8_t *end = input->buffer + input->length;
	const uint8_t *p = input->curr;
	uint32_t block_size = 0;
	if (!get_varint32(p, end, &block_size)) { // snappy_decompress.c:299-303
		fprintf(stderr, "Failed to read decompressed block size\n");
		return SNAPPY_INVALID_INPUT;
	}
	const uint64_t out_len = output->length;
	if (out_len == 0)
		return p == end ? SNAPPY_OK : SNAPPY_INVALID_INPUT;
	if (block_size == 0)
		return SNAPPY_INVALID_INPUT;
	if (block_size > PIM_SNAPPY_MAX_BLOCK_SIZE) {
		fprintf(stderr, "pim_snappy: block size %u is outside the format's valid domain (1..%u)\n", block_size, PIM_SNAPPY_MAX_BLOCK_SIZE);
		return SNAPPY_BUFFER_TOO_SMALL;
	}
	if (!output->buffer)
		return SNAPPY_INVALID_INPUT;
	const uint64_t nblocks = pim_snappy_num_blocks(out_len, block_size);

	const uint8_t *body = p;
	const uint64_t body_len = (uint64_t)(end - body);
	runtime->pre += now_s() - t0;

	t0 = now_s();
	int ng = ensure_devices();
	runtime->load = now_s() - t0;
	if (ng <= 0) {
		fprintf(stderr, "pim_snappy: no usable sm_100 GPU; there is no CPU fallback\n");
		return SNAPPY_INVALID_INPUT;
	}
	if ((uint64_t)ng > nblocks)
		ng = (int)nblocks;

	// One device and a large stream: upload, header walk, kernels and download all overlap.
	if (ng == 1 && chunk_bytes() && out_len > 2 * chunk_bytes()) {
		DeviceCtx &d = g_devs[0];
		cudaSetDevice(d.dev);
		ShardTimes t1;
		const int r = decompress_stream_resident(d, body, body_len, nblocks, block_size, out_len, output->buffer, t1);
		if (r != SNAPPY_OK)
			return (snappy_status)r;
		runtime->d_alloc = t1.alloc;
		runtime->copy_in = t1.copy_in * 1e-3;
		runtime->run = t1.run * 1e-3;
		runtime->copy_out = t1.copy_out * 1e-3;
		input->curr = const_cast<uint8_t *>(end);
		output->curr = output->buffer + out_len;
		return SNAPPY_OK;
	}

	// serial walk over the 4-byte block headers (snappy_decompress.c:317-340)
	t0 = now_s();
	std::vector<uint64_t> off(nblocks + 1);
	uint64_t pos = 0;
	for (uint64_t i = 0; i < nblocks; i++) {
		if (pos + 4 > body_len)
			return SNAPPY_INVALID_INPUT;
		off[i] = pos;
		const uint32_t cs = (uint32_t)body[pos] | ((uint32_t)body[pos + 1] << 8) | ((uint32_t)body[pos + 2] << 16) | ((uint32_t)body[pos + 3] << 24);
		pos += 4 + (uint64_t)cs;
		if (pos > body_len)
			return SNAPPY_INVALID_INPUT;
	}
	off[nblocks] = pos;
	if (pos != body_len)
		return SNAPPY_INVALID_INPUT;
	runtime->pre += now_s() - t0;

	std::vector<ShardTimes> tm(ng);
	std::vector<int> rc(ng, SNAPPY_OK);
	for_each_shard(ng, [&](int g) {
		DeviceCtx &d = g_devs[g];
		uint64_t lo, hi;
		shard_range(nblocks, ng, g, &lo, &hi);
		const uint64_t nb = hi - lo;
		const uint64_t c_lo = off[lo], c_bytes = off[hi] - off[lo];
		const uint64_t o_lo = lo * (uint64_t)block_size;
		const uint64_t o_bytes = std::min(out_len, hi * (uint64_t)block_size) - o_lo;
		cudaSetDevice(d.dev);
		if (nb == 0)
			return;
		if (chunk_bytes() && o_bytes > 2 * chunk_bytes()) { // a large shard: upload, kernels and download overlap
			rc[g] = decompress_stream_resident(d, body + c_lo, c_bytes, nb, block_size, o_bytes, output->buffer + o_lo, tm[g]);
			return;
		}
		const double a0 = now_s();
		bool ok = grow(d.d_a, d.cap_a, o_bytes + 64) && grow(d.d_b, d.cap_b, c_bytes + 64) && grow(d.d_off, d.cap_off, nb + 1);
		tm[g].alloc = now_s() - a0;
		if (!ok) { rc[g] = SNAPPY_BUFFER_TOO_SMALL; return; }
		std::vector<uint64_t> rel(nb + 1);
		for (uint64_t i = 0; i <= nb; i++)
			rel[i] = off[lo + i] - c_lo;
		CudaCheck ck;
		CK(cudaEventRecord(d.ev[0], d.stream));
		CK(cudaMemcpyAsync(d.d_b, body + c_lo, c_bytes, cudaMemcpyHostToDevice, d.stream));
		CK(cudaMemcpyAsync(d.d_off, rel.data(), (nb + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, d.stream));
		CK(cudaEventRecord(d.ev[1], d.stream));
		rc[g] = pim_snappy_decompress_device(d.d_b, d.d_off, nb, block_size, o_bytes, d.d_a, d.d_status, d.stream);
		CK(cudaEventRecord(d.ev[2], d.stream));
		CK(cudaMemcpyAsync(output->buffer + o_lo, d.d_a, o_bytes, cudaMemcpyDeviceToHost, d.stream));
		CK(cudaMemcpyAsync(d.h_small + 1, d.d_status, sizeof(uint32_t), cudaMemcpyDeviceToHost, d.stream));
		CK(cudaEventRecord(d.ev[3], d.stream));
		CK(cudaStreamSynchronize(d.stream));
		if (ck.failed(d.dev)) {
			rc[g] = SNAPPY_INVALID_INPUT;
			return;
		}
		if (rc[g] == SNAPPY_OK && (uint32_t)d.h_small[1] != 0)
			rc[g] = SNAPPY_INVALID_INPUT;
		cudaEventElapsedTime(&tm[g].copy_in, d.ev[0], d.ev[1]);
		cudaEventElapsedTime(&tm[g].run, d.ev[1], d.ev[2]);
		cudaEventElapsedTime(&tm[g].copy_out, d.ev[2], d.ev[3]);
	});
	for (int g = 0; g < ng; g++) {
		if (rc[g] != SNAPPY_OK)
			return (snappy_status)rc[g];
		runtime->d_alloc = std::max(runtime->d_alloc, tm[g].alloc);
		runtime->copy_in = std::max(runtime->copy_in, tm[g].copy_in * 1e-3);
		runtime->run = std::max(runtime->run, tm[g].run * 1e-3);
		runtime->copy_out = std::max(runtime->copy_out, tm[g].copy_out * 1e-3);
	}
	input->curr = const_cast<uint8_t *>(end);
	output->curr = output->buffer + out_len;
	return SNAPPY_OK;
}

extern "C" void pim_snappy_set_write_file(int enable) { g_write_file.store(enable ? 1 : 0); }
extern "C" void pim_snappy_set_extended_blocks(int enable) { g_extended.store(enable ? 1 : 0); }

// The reference's names.  Its DPU compressor writes output->file_name itself (snappy_compress.c:626-627,697-703) and its
// main() then skips write_output_host (dpu_snappy.c:209); with pim_snappy_set_write_file(1) / $PIM_SNAPPY_WRITE_FILE=1
// this entry point does the same, so that the reference's UNMODIFIED main() produces the file.
extern "C" snappy_status snappy_compress_dpu(struct host_buffer_context *input, struct host_buffer_context *output, uint32_t block_size,
                                             struct program_runtime *runtime)
{
	const snappy_status rc = snappy_compress_gpu(input, output, block_size, runtime);
	int wf = g_write_file.load();
	if (wf < 0) {
		const char *e = getenv("PIM_SNAPPY_WRITE_FILE");
		wf = e && atoi(e) != 0 ? 1 : 0;
		g_write_file.store(wf);
	}
	if (rc == SNAPPY_OK && wf == 1 && output->file_name) {
		FILE *f = fopen(output->file_name, "w");
		if (!f || fwrite(output->buffer, 1, output->length, f) != output->length) {
			fprintf(stderr, "pim_snappy: cannot write %s\n", output->file_name);
			if (f)
				fclose(f);
			return SNAPPY_INVALID_INPUT;
		}
		fclose(f);
	}
	return rc;
}

extern "C" snappy_status snappy_decompress_dpu(struct host_buffer_context *input, struct host_buffer_context *output,
                                               struct program_runtime *runtime)
{
	return snappy_decompress_gpu(input, output, runtime);
}
