// Warp primitives used by the per-warp compressor (compress_warp.cuh).
//
// Device build: thin wrappers over the sm_100a intrinsics.
// Host build with PIM_WARP_EMU: the same names run on a lock-step emulator (32 cooperative
// contexts, tests/emu/warp_emu.cpp) so the warp algorithm can be checked against the oracle
// on a machine without a GPU.  The emulator is test infrastructure only; the product never
// compiles with PIM_WARP_EMU.
#pragma once

#include <cstdint>

#ifdef PIM_WARP_EMU

#define PW_FN inline
namespace pimsnappy {
namespace wp {
// implemented by the emulator
int lane();
uint32_t ballot(bool pred);
uint32_t shfl(uint32_t v, int src_lane);
uint32_t match_any(uint32_t v);
uint32_t reduce_max(uint32_t v);
void syncwarp();
void spin_pause();  // inside a wait loop on shared memory written by another warp
inline void fence_block() {}
inline void prefetch_l2(const void *) {}
template <bool GT> inline uint32_t tbl_ld(const uint16_t *t, uint32_t h) { return t[h]; }
template <bool GT> inline void tbl_st(uint16_t *t, uint32_t h, uint32_t v) { t[h] = (uint16_t)v; }
inline uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)
{
	sh &= 31u;
	return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
}
inline int ffs32(uint32_t v) { return __builtin_ffs((int)v); }
inline int clz32(uint32_t v) { return v ? __builtin_clz(v) : 32; }
inline int popc32(uint32_t v) { return __builtin_popcount(v); }
inline uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }
} // namespace wp
} // namespace pimsnappy

#else

#define PW_FN __device__ __forceinline__
namespace pimsnappy {
namespace wp {
__device__ __forceinline__ int lane() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t ballot(bool pred) { return __ballot_sync(0xFFFFFFFFu, pred); }
__device__ __forceinline__ uint32_t shfl(uint32_t v, int src_lane) { return __shfl_sync(0xFFFFFFFFu, v, src_lane); }
__device__ __forceinline__ uint32_t match_any(uint32_t v) { return __match_any_sync(0xFFFFFFFFu, v); }
__device__ __forceinline__ uint32_t reduce_max(uint32_t v) { return __reduce_max_sync(0xFFFFFFFFu, v); }
__device__ __forceinline__ void syncwarp() { __syncwarp(); }
#ifndef PIM_SPIN_NS
#define PIM_SPIN_NS 400
#endif
__device__ __forceinline__ void spin_pause() { __nanosleep(PIM_SPIN_NS); }
__device__ __forceinline__ void fence_block() { __threadfence_block(); }
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// hash-table access: shared memory (GT = false), or global memory served by L2 (GT = true; .cg so that
// the lanes of the warp always see each other's stores)
template <bool GT> __device__ __forceinline__ uint32_t tbl_ld(const uint16_t *t, uint32_t h)
{
	if (!GT)
		return t[h];
	uint16_t v;
	asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(v) : "l"(t + h) : "memory");
	return v;
}
template <bool GT> __device__ __forceinline__ void tbl_st(uint16_t *t, uint32_t h, uint32_t v)
{
	if (!GT)
		t[h] = (uint16_t)v;
	else
		asm volatile("st.global.cg.u16 [%0], %1;" ::"l"(t + h), "h"((uint16_t)v) : "memory");
}
__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) { return __funnelshift_r(lo, hi, sh); }
__device__ __forceinline__ int ffs32(uint32_t v) { return __ffs((int)v); }
__device__ __forceinline__ int clz32(uint32_t v) { return __clz((int)v); }
__device__ __forceinline__ int popc32(uint32_t v) { return __popc(v); }
__device__ __forceinline__ uint32_t umin(uint32_t a, uint32_t b) { return min(a, b); }
} // namespace wp
} // namespace pimsnappy

#endif
