// common.cuh -- shared device helpers for libpress_gpu (sm_100a).
//
// A batch = all reads concatenated; a tile = a fixed number of samples of ONE read.
// The svb encoder deals tiles round-robin to persistent CTAs and resolves the one
// loop-carried quantity of the reference's serial loop (the output byte position) with
// 64-bit status words (2 flag bits + 62 value bits, one relaxed store each); the other
// kernels hand whole reads (or segments of long reads) to CTAs by an atomic ticket.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pg {

constexpr int NT = 256;			// threads per CTA
constexpr int NW = NT / 32;		// warps per CTA

constexpr uint64_t LB_AGG = 1ull << 62;	// status: tile aggregate available
constexpr uint64_t LB_INC = 2ull << 62;	// status: inclusive prefix available
constexpr uint64_t LB_VAL = (1ull << 62) - 1;
constexpr uint64_t SENTINEL = ~0ull;	// "not written yet" for published offsets

enum : uint32_t { ERR_CAPACITY = 1, ERR_MALFORMED = 2, ERR_EXZD_CAP = 4 };

// Scratch header (first 256 bytes of d_scratch).
struct Hdr {
	uint32_t ticket;	// next tile to hand out (pass 1)
	uint32_t total_tiles;	// written by the plan kernel
	uint32_t error;		// ERR_* bits
	uint32_t ticket2;	// next tile to hand out (pass 2, two-pass methods)
	uint64_t fixed_total;	// sum of the per-read fixed bytes (kx[nreads])
	uint64_t aux[29];
};
static_assert(sizeof(Hdr) == 256, "Hdr is 256 bytes");

__device__ __forceinline__ uint64_t ld_relaxed(const uint64_t *p)
{
	uint64_t v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_relaxed(uint64_t *p, uint64_t v)
{
	asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// 16-byte asynchronous copy global -> shared (L2 only), and its group fences
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t) __cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
		     : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ uint64_t warp_sum64(uint64_t v)
{
#pragma unroll
	for (int o = 16; o; o >>= 1)
		v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Exclusive scan over the CTA in thread order (C must be 1; the template keeps the
// call sites' array form).  s_warp: NW words, 16-byte aligned.  ONE __syncthreads():
// every warp posts its total, then every thread adds up the totals of the warps
// before its own.  The caller must put a barrier between this call's reads of s_warp
// and the next write to the same array (every call site has one: the loop-end barrier).
template <int C>
__device__ __forceinline__ void block_scan(const uint32_t (&v)[C], uint32_t (&excl)[C],
					   uint32_t &total, uint32_t *s_warp)
{
	static_assert(C == 1 && NW == 8, "one value per thread, 8 warps");
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t x = v[0];
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
		if (lane >= o)
			x += y;
	}
	if (lane == 31)
		s_warp[warp] = x;
	__syncthreads();
	// the 8 warp totals scanned by shuffles in every warp (a third of the instructions of
	// adding them up under per-warp predicates)
	const uint32_t t = s_warp[lane & 7];
	uint32_t s = t;
#pragma unroll
	for (int o = 1; o < 8; o <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
		if ((lane & 7) >= o)
			s += y;
	}
	const uint32_t before = __shfl_sync(0xffffffffu, s - t, warp);
	const uint32_t all = __shfl_sync(0xffffffffu, s, 7);
	excl[0] = before + x - v[0];
	total = all;
}

// ---- SWAR on two 16-bit lanes per 32-bit word -----------------------------

// per-halfword a - b and a + b (mod 2^16): sm_100 has a native VIADD.16x2
__device__ __forceinline__ uint32_t sub16x2(uint32_t a, uint32_t b) { return __vsub2(a, b); }
__device__ __forceinline__ uint32_t add16x2(uint32_t a, uint32_t b) { return __vadd2(a, b); }
// per-halfword zigzag: (d << 1) ^ (d >> 15 arithmetic)      (trans.c:75-78)
// PRMT with the sign-replicate selector turns each halfword's top byte into 0x0000 / 0xffff.
__device__ __forceinline__ uint32_t zigzag16x2(uint32_t d)
{
	uint32_t sign;	// __byte_perm() masks the selector's replicate bit, prmt.b32 honours it
	asm("prmt.b32 %0, %1, %2, %3;" : "=r"(sign) : "r"(d), "r"(0u), "r"(0xbb99u));
	return __vadd2(d, d) ^ sign;
}
// per-halfword inverse: (z >> 1) ^ -(z & 1)                  (trans.c:80-83)
__device__ __forceinline__ uint32_t unzigzag16x2(uint32_t z)
{
	return ((z >> 1) & 0x7fff7fffu) ^ ((z & 0x00010001u) * 0xffffu);
}
// bit 0 / bit 1 = (low / high halfword >= 256)
__device__ __forceinline__ uint32_t ge256x2(uint32_t z)
{
	uint32_t t = (((z >> 8) & 0x00ff00ffu) + 0x00ff00ffu) & 0x01000100u;
	return ((t >> 8) | (t >> 23)) & 3u;
}

// 8 one-bit flags -> 8 two-bit codes (svb32 key layout for codes in {0,1})
__device__ __forceinline__ uint32_t spread8(uint32_t f)
{
	f = (f | (f << 4)) & 0x0f0fu;
	f = (f | (f << 2)) & 0x3333u;
	f = (f | (f << 1)) & 0x5555u;
	return f;
}
// inverse of spread8 on the low bit of every 2-bit code
__device__ __forceinline__ uint32_t squeeze16(uint32_t k)
{
	k &= 0x5555u;
	k = (k | (k >> 1)) & 0x3333u;
	k = (k | (k >> 2)) & 0x0f0fu;
	k = (k | (k >> 4)) & 0x00ffu;
	return k;
}

// ask L2 for the line holding p (no register waits for it): the per-read walks keep one step in
// registers / shared memory and a few more on their way from HBM
__device__ __forceinline__ void prefetch_l2(const void *p)
{
	asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

}  // namespace pg
