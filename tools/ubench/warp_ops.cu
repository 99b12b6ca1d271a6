// Development aid: dependent-chain latency (cycles per op) of the warp collectives the parser uses, on one warp.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o warp_ops warp_ops.cu && ./warp_ops
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define N 256
template <int OP>
__global__ void k(uint32_t seed, long long *out, uint32_t *sink)
{
	uint32_t v = seed + threadIdx.x * 2654435761u;
	const int lane = threadIdx.x & 31;
	__shared__ uint32_t sm[64];
	sm[lane] = v;
	sm[lane + 32] = v * 3;
	__syncwarp();
	long long t0 = clock64();
#pragma unroll 16
	for (int i = 0; i < N; i++) {
		if (OP == 0) v = __ballot_sync(0xFFFFFFFFu, (v >> lane) & 1u) + lane;               // vote + dependent use
		if (OP == 1) v = __shfl_sync(0xFFFFFFFFu, v, (v + 1) & 31) + 1;                       // shfl, index depends on v
		if (OP == 2) v = __reduce_max_sync(0xFFFFFFFFu, v ^ lane) + lane;                     // redux.max
		if (OP == 3) v = __reduce_or_sync(0xFFFFFFFFu, v & (1u << lane)) + lane;              // redux.or
		if (OP == 4) v = __match_any_sync(0xFFFFFFFFu, v & 7u) + lane;                        // match.any, 8 groups
		if (OP == 5) { __syncwarp(); v = v * 3 + 1; }                                         // bar.warp.sync + 1 op
		if (OP == 6) v = sm[(v & 31)] + 1;                                                     // dependent LDS
		if (OP == 7) v = v * 0x1e35a7bdu + lane;                                               // dependent IMAD
		if (OP == 8) { sm[(v & 31) + 32] = v; __syncwarp(); v = sm[((v + 1) & 31) + 32] + 1; } // STS, syncwarp, LDS
		if (OP == 9) v = (__ffs(v | 0x80000000u) + (31 - __clz(v | 1u))) ^ v;                  // ffs + clz
		if (OP == 10) { if (__any_sync(0xFFFFFFFFu, v == 0x12345u)) v ^= 7; v += lane; }       // vote.any + branch
		if (OP == 11) v = __match_any_sync(0xFFFFFFFFu, v) + lane;                             // match.any, 32 distinct
	}
	long long t1 = clock64();
	if (threadIdx.x == 0)
		out[OP] = t1 - t0;
	sink[threadIdx.x] = v;
}

int main()
{
	long long *d_out, h[16] = {0};
	uint32_t *d_sink;
	cudaMalloc(&d_out, sizeof(h));
	cudaMalloc(&d_sink, 4 * 64);
	const char *names[] = {"ballot", "shfl (data-dependent index)", "redux.max", "redux.or", "match.any (8 groups)", "bar.warp.sync + imad", "dependent LDS", "dependent IMAD",
	                       "STS + syncwarp + LDS", "ffs + clz + 2 alu", "vote.any + uniform branch", "match.any (32 distinct)"};
	for (int rep = 0; rep < 2; rep++) {
		k<0><<<1, 32>>>(rep, d_out, d_sink); k<1><<<1, 32>>>(rep, d_out, d_sink); k<2><<<1, 32>>>(rep, d_out, d_sink); k<3><<<1, 32>>>(rep, d_out, d_sink);
		k<4><<<1, 32>>>(rep, d_out, d_sink); k<5><<<1, 32>>>(rep, d_out, d_sink); k<6><<<1, 32>>>(rep, d_out, d_sink); k<7><<<1, 32>>>(rep, d_out, d_sink);
		k<8><<<1, 32>>>(rep, d_out, d_sink); k<9><<<1, 32>>>(rep, d_out, d_sink); k<10><<<1, 32>>>(rep, d_out, d_sink); k<11><<<1, 32>>>(rep, d_out, d_sink);
		if (cudaDeviceSynchronize() != cudaSuccess) { printf("failed\n"); return 1; }
	}
	cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
	for (int i = 0; i < 12; i++)
		printf("%-32s %6.1f cycles per iteration\n", names[i], (double)h[i] / N);
	return 0;
}
