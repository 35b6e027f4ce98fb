// common.cuh -- shared device helpers for libpress_gpu (sm_100a).
//
// Work decomposition used by every kernel in this library:
//   * a batch = all reads concatenated; a TILE = 4096 samples of ONE read;
//   * persistent CTAs of 256 threads take tiles in ticket order (atomicAdd), so
//     tile g only ever waits on tiles < g, which are running or finished;
//   * every loop-carried quantity of the reference's serial loops (output byte
//     position, previous sample, running sum) becomes a prefix over tiles,
//     resolved in ONE pass with decoupled look-back on a 64-bit status word
//     (2 flag bits + 62 value bits, written with a single relaxed store).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pg {

constexpr int TILE = 4096;		// samples per tile
constexpr int NT = 256;			// threads per CTA
constexpr int NW = NT / 32;		// warps per CTA
constexpr int GROUPS = TILE / 8;	// 8-sample groups per tile (one 16-byte vector each)
constexpr int CHUNKS = GROUPS / NT;	// groups per thread

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

__device__ __forceinline__ uint64_t warp_sum64(uint64_t v)
{
#pragma unroll
	for (int o = 16; o; o >>= 1)
		v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Decoupled look-back, executed by ONE full warp.  Returns (in every lane) the
// sum of the aggregates of tiles [first, tile).  Tiles below `first` are
// treated as an inclusive prefix of 0 (segmented scans restart at a read's
// first tile).
__device__ __forceinline__ uint64_t lookback(const uint64_t *status, uint32_t tile, uint32_t first)
{
	const int lane = threadIdx.x & 31;
	uint64_t sum = 0;
	int64_t idx = (int64_t) tile - 1 - lane;
	for (;;) {
		uint64_t w = LB_INC;	// out of range == inclusive 0
		if (idx >= (int64_t) first) {
			do {
				w = ld_relaxed(status + idx);
			} while ((w >> 62) == 0);
		}
		unsigned inc = __ballot_sync(0xffffffffu, (w >> 62) == 2);
		uint64_t v = w & LB_VAL;
		if (inc) {
			int stop = __ffs(inc) - 1;
			if (lane > stop)
				v = 0;
			sum += warp_sum64(v);
			break;
		}
		sum += warp_sum64(v);
		idx -= 32;
	}
	return sum;
}

// Exclusive scan over the CTA in (chunk, thread) order: element (c, tid) comes
// before (c, tid + 1) and all of chunk c before chunk c + 1.
// s_warp needs C * NW + 1 words.  Contains two __syncthreads().
template <int C>
__device__ __forceinline__ void block_scan(const uint32_t (&v)[C], uint32_t (&excl)[C],
					   uint32_t &total, uint32_t *s_warp)
{
	static_assert(C * NW <= 32, "one warp scans the warp totals");
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t inc[C];
#pragma unroll
	for (int c = 0; c < C; c++) {
		uint32_t x = v[c];
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
			if (lane >= o)
				x += y;
		}
		inc[c] = x;
		if (lane == 31)
			s_warp[c * NW + warp] = x;
	}
	__syncthreads();
	if (warp == 0) {
		uint32_t x = lane < C * NW ? s_warp[lane] : 0, orig = x;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
			if (lane >= o)
				x += y;
		}
		if (lane < C * NW)
			s_warp[lane] = x - orig;
		if (lane == C * NW - 1)
			s_warp[C * NW] = x;
	}
	__syncthreads();
#pragma unroll
	for (int c = 0; c < C; c++)
		excl[c] = s_warp[c * NW + warp] + inc[c] - v[c];
	total = s_warp[C * NW];
}

// ---- SWAR on two 16-bit lanes per 32-bit word -----------------------------

// per-halfword a - b (mod 2^16)
__device__ __forceinline__ uint32_t sub16x2(uint32_t a, uint32_t b)
{
	return ((a | 0x80008000u) - (b & 0x7fff7fffu)) ^ ((a ^ ~b) & 0x80008000u);
}
// per-halfword a + b (mod 2^16)
__device__ __forceinline__ uint32_t add16x2(uint32_t a, uint32_t b)
{
	return ((a & 0x7fff7fffu) + (b & 0x7fff7fffu)) ^ ((a ^ b) & 0x80008000u);
}
// per-halfword zigzag: (d << 1) ^ (d >> 15 arithmetic)      (trans.c:75-78)
__device__ __forceinline__ uint32_t zigzag16x2(uint32_t d)
{
	return ((d << 1) & 0xfffefffeu) ^ (((d >> 15) & 0x00010001u) * 0xffffu);
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

}  // namespace pg
