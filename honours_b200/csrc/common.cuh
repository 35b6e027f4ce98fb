// common.cuh -- shared device helpers for libpress_gpu (sm_100a).
//
// Work decomposition used by every kernel in this library:
//   * a batch = all reads concatenated; a tile = a fixed number of samples of ONE read;
//   * a persistent, cooperatively launched grid (all CTAs resident) walks the
//     tiles round-robin, warp by warp, so tiles start in index order and tile g
//     only ever waits on tiles < g, which are running or finished;
//   * every loop-carried quantity of the reference's serial loops (output byte
//     position, previous sample, running sum) becomes a prefix over tiles,
//     resolved in ONE pass with decoupled look-back on a 64-bit status word
//     (2 flag bits + 62 value bits, written with a single relaxed store).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pg {

constexpr int NT = 256;			// threads per CTA
constexpr int NW = NT / 32;		// warps per CTA

// Warp tiles: the unit of work of the svb kernels.  One warp owns 1024
// samples (128 sixteen-byte groups, 4 per lane, lane-interleaved so that every
// load / store instruction of the warp covers 512 contiguous bytes).  Nothing
// in a tile's life needs a CTA barrier, so the 32+ resident warps of an SM
// hide each other's load, look-back and store latencies.
constexpr int WT = 1024;		// samples per warp tile
constexpr int WG = WT / 8;		// groups per warp tile
constexpr int WC = WG / 32;		// groups per lane

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

// Spin until the status word at p has a flag, return it.
__device__ __forceinline__ uint64_t lb_wait(const uint64_t *p)
{
	uint64_t w;
	do {
		w = ld_relaxed(p);
	} while ((w >> 62) == 0);
	return w;
}

// One window of a look-back: every lane holds one status word `w`, lane 0 the
// nearest predecessor.  Adds the values up to and including the first
// inclusive word to `sum`; returns true when one was found.
__device__ __forceinline__ bool lb_window(uint64_t w, uint64_t &sum)
{
	const int lane = threadIdx.x & 31;
	const unsigned inc = __ballot_sync(0xffffffffu, (w >> 62) == 2);
	uint64_t v = w & LB_VAL;
	if (inc && lane > __ffs(inc) - 1)
		v = 0;
	sum += warp_sum64(v);
	return inc != 0;
}

// Two-level decoupled look-back, executed by ONE full warp for tile g.
// Publishes the tile's aggregate `my`, returns (in every lane) the sum of the
// aggregates of tiles [first, g), then publishes the inclusive prefix.
//   tile[g]  : per-tile status word
//   block[b] : status of the 32 tiles [32 b, 32 b + 32), written by the block's
//              LAST tile: aggregate / inclusive prefix counted from
//              max(32 b, first) of that tile's segment.
// A tile first inspects its predecessors inside its own block (one window) and
// then whole blocks, 1024 tiles per window, so the walk stays 2-3 L2 round
// trips however many tiles are in flight.  `first` is the first tile of the
// segment (a read, or 0 for a batch-wide scan); every tile of a segment passes
// the same `first`.
__device__ __forceinline__ uint64_t scan_tile(uint64_t *tile, uint64_t *block, uint32_t g,
					      uint32_t first, uint64_t my)
{
	const int lane = threadIdx.x & 31;
	const uint32_t b = g >> 5;
	const bool closes = (g & 31) == 31;	// this tile publishes its block's status
	uint64_t D = 0;
	if (g > first) {
		if (lane == 0)
			st_relaxed(tile + g, LB_AGG | my);
		const int64_t lo = (int64_t) ((b << 5) > first ? (b << 5) : first);
		const int64_t idx = (int64_t) g - 1 - lane;
		// lanes before the window: nothing there (inclusive 0) when the segment starts
		// inside this block, else "look in the block words"
		uint64_t w = lo == (int64_t) first ? LB_INC : LB_AGG;
		if (idx >= lo)
			w = lb_wait(tile + idx);
		if (!lb_window(w, D)) {
			if (closes && lane == 0)
				st_relaxed(block + b, LB_AGG | (D + my));
			const int64_t bfirst = first >> 5;
			for (int64_t bidx = (int64_t) b - 1 - lane;; bidx -= 32) {
				w = LB_INC;	// below the segment's first block: inclusive 0
				if (bidx >= bfirst)
					w = lb_wait(block + bidx);
				if (lb_window(w, D))
					break;
			}
		}
	}
	if (lane == 0) {
		st_relaxed(tile + g, LB_INC | (D + my));
		if (closes)
			st_relaxed(block + b, LB_INC | (D + my));
	}
	return D;
}

// Row look-back for the persistent CTA-tile kernels (pipe.cuh): tiles are dealt
// round-robin, tile g = row * G + col with G = gridDim.x, so "row" = iteration
// and col = blockIdx.x.  Executed by ONE full warp for tile g:
//   publishes the tile's aggregate, returns (in every lane) the sum of the
//   aggregates of ALL tiles < g (a batch-wide exclusive prefix); the row's last
//   tile publishes its inclusive prefix in rowpref[row].
// All words of the own row and the previous row's prefix are requested at once
// (one L2 round trip in the common case); when rowpref[row - 1] is not there
// yet the walk continues over the aggregates of whole earlier rows instead of
// waiting for it, so no chain of dependent prefixes forms.
// Sum of the aggregates tile[base + j], j in [0, count): all loads are issued
// before the first flag is looked at (one L2 round trip when everything has
// been published), words that were still empty are polled afterwards.  One warp;
// returns this lane's partial sum.
__device__ __forceinline__ uint64_t lb_row_sum(const uint64_t *tile, uint64_t base, uint32_t count)
{
	const int lane = threadIdx.x & 31;
	uint64_t sum = 0;
	for (uint32_t j0 = 0; j0 < count; j0 += 256) {
		uint64_t w[8];
#pragma unroll
		for (int q = 0; q < 8; q++) {
			const uint32_t j = j0 + 32 * q + lane;
			w[q] = j < count ? ld_relaxed(tile + base + j) : LB_AGG;
		}
#pragma unroll
		for (int q = 0; q < 8; q++) {
			const uint32_t j = j0 + 32 * q + lane;
			while ((w[q] >> 62) == 0)
				w[q] = ld_relaxed(tile + base + j);
			sum += w[q] & LB_VAL;
		}
	}
	return sum;
}

__device__ __forceinline__ uint64_t row_scan(uint64_t *tile, uint64_t *rowpref, uint32_t g, uint32_t G,
					     uint32_t total, uint64_t my)
{
	const int lane = threadIdx.x & 31;
	const uint32_t row = g / G, col = g - row * G;
	if (lane == 0)
		st_relaxed(tile + g, LB_AGG | my);
	uint64_t wp = 0;
	if (row > 0 && lane == 0)	// in flight together with the own-row words
		wp = ld_relaxed(rowpref + row - 1);
	uint64_t D = lb_row_sum(tile, (uint64_t) row * G, col);	// own row, tiles before this one
	for (int64_t r = (int64_t) row - 1; r >= 0; r--) {
		const uint64_t w = __shfl_sync(0xffffffffu, wp, 0);
		if ((w >> 62) == 2) {
			if (lane == 0)
				D += w & LB_VAL;
			break;
		}
		if (r > 0 && lane == 0)
			wp = ld_relaxed(rowpref + r - 1);
		D += lb_row_sum(tile, (uint64_t) r * G, G);	// earlier rows are always complete rows
	}
	D = warp_sum64(D);
	if (lane == 0 && (col == G - 1 || g == total - 1))	// tile[] keeps aggregates only
		st_relaxed(rowpref + row, LB_INC | (D + my));
	return D;
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

}  // namespace pg
