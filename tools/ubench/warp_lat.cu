// Dependent-chain latency of the warp primitives the codec kernels lean on (development aid).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N 512
template <int OP> __global__ void k(uint32_t *out, uint32_t seed, long long *cyc)
{
	__shared__ uint32_t sm[4096];
	for (int i = threadIdx.x; i < 4096; i += 32) sm[i] = (i * 2654435761u) >> 20;
	__syncwarp();
	uint32_t v = seed + threadIdx.x * (OP == 1 ? 1 : 7);
	long long t0 = clock64();
#pragma unroll 8
	for (int i = 0; i < N; i++) {
		if (OP == 0) v = __match_any_sync(0xffffffffu, v) + i;                 // distinct values
		if (OP == 1) v = __match_any_sync(0xffffffffu, v >> 3) + (v & ~7u) + 8; // groups of 8 equal
		if (OP == 2) v = __shfl_sync(0xffffffffu, v, (v + 1) & 31) + 1;
		if (OP == 3) v = __ballot_sync(0xffffffffu, v & 1) + v;
		if (OP == 4) v = __reduce_or_sync(0xffffffffu, v) + 1;
		if (OP == 5) v = sm[v & 4095];
		if (OP == 6) { sm[(v & 1023) * 4 + 0] = v; __syncwarp(); v = sm[((v + 1) & 1023) * 4] + 1; }
		if (OP == 7) v = v * 0x1e35a7bdu + 1;
		if (OP == 8) v = __reduce_min_sync(0xffffffffu, v) + 1;
		if (OP == 9) v = __popc(v) + v;
		if (OP == 10) v = __ffs(v) + v;
		if (OP == 11) v = (v >> 3) ^ (v + 5);
	}
	long long t1 = clock64();
	out[threadIdx.x] = v;
	if (threadIdx.x == 0) cyc[OP] = t1 - t0;
}
int main()
{
	uint32_t *o; long long *c;
	cudaMalloc(&o, 128); cudaMallocManaged(&c, 16 * 8);
	const char *names[] = {"match_any distinct", "match_any groups8", "shfl idx", "ballot", "redux or", "lds chain", "sts+syncwarp+lds", "imad", "redux min", "popc", "ffs", "shift-xor-add(2)"};
#define RUN(OP) k<OP><<<1, 32>>>(o, 12345, c); k<OP><<<1, 32>>>(o, 12345, c); cudaDeviceSynchronize(); printf("%-22s %.1f cycles/op\n", names[OP], (double)c[OP] / N);
	RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(10) RUN(11)
	printf("%s\n", cudaGetErrorString(cudaGetLastError()));
	return 0;
}
