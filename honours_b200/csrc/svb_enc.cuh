// svb_enc.cuh -- fused zigzag-delta + streamvbyte ENCODE: one persistent CTA per
// SM, warp-specialised dataflow over a shared-memory ring filled by TMA.
//
//   KEYBITS = 1 : ONT svb16, 1 key bit / value   (svb16/encode_scalar.hpp:14-62; press.c:1678-1687)
//   KEYBITS = 2 : classical streamvbyte over zd widened to u32; zd < 65536 so only
//                 codes 0/1 occur (streamvbyte_encode.c:11-59; press.c:1590-1598)
//   FRAME       : the [n u32 LE] prefix of the pre-zstd VBZ stream (press.c:1875-1879, 2035-2039)
//
// Stream of read r:  [frame][keys: ceil(n / (8 / KEYBITS)) bytes][data bytes].
//
// One pass over HBM.  A tile is 8192 samples of ONE read cut on 16-byte
// boundaries of the sample array; tiles are dealt round-robin to the CTAs.
// Every tile walks through the ring stage it was loaded into:
//
//   producer warp   tile metadata -> ONE cp.async.bulk (the tile, the group before
//                   it for the previous sample, the group after it for the last key)
//   16 worker warps in two teams (team = tiles of one parity), 1024 samples per
//                   warp and tile, two phases per tile:
//                   A  4 x LDS.128 per lane, delta + zigzag on 16-bit pairs
//                      (VIADD.16x2 / PRMT); no two-byte value in the block (the
//                      common case on nanopore signal): the 8 data bytes of each
//                      group stay packed in registers, the block's byte count
//                      (1024) goes to shared memory; else the exact count does
//                   B  EN_K - 1 team tiles later, when the tile's container
//                      position is known: aligned 64-bit stores of the packed
//                      bytes re-phased to the destination, or (blocks with
//                      two-byte values, first / last blocks of a read) per-sample
//                      placement through a staging line
//   publish warp    the tile's byte count -> global status word, as soon as it is known
//   prefix warp     batch-wide prefix of the byte counts (the key sections are data
//                   independent, plan.cuh) by a sliding window: P(g) = P(g - G) +
//                   the G aggregates in between, G = grid size; the window holds
//                   other CTAs' tiles only and is requested several tiles ahead;
//                   per-block positions, key pointer, capacity checks, frame
//
// Phase A of later tiles runs while the prefix of earlier ones resolves, so the
// look-back latency (a few L2 round trips) never stalls a worker.
#pragma once

#include "common.cuh"
#include "pipe.cuh"
#include "svb_zd.cuh"

namespace pg {

constexpr int EN_TS = 8192;			// samples per tile
constexpr int EN_TG = EN_TS / 8;		// 16-byte groups per tile
constexpr int EN_NB = 8;			// blocks per tile = worker warps per team
constexpr int EN_TEAMS = 2;
constexpr int EN_NWORK = EN_NB * EN_TEAMS;
constexpr int EN_K = 3;				// team tiles between phase A and phase B, + 1
constexpr int EN_EG = EN_TG / EN_NB;		// groups per block (128)
constexpr int EN_EC = EN_EG / 32;		// groups per lane (4)
constexpr int EN_W_SCAN = EN_NWORK;		// prefix warp; + 1: publish warp
constexpr int EN_W_PROD = EN_W_SCAN + 2;
constexpr int EN_THREADS = (EN_W_PROD + 1) * 32;
constexpr int EN_STAGES = 11;
constexpr int EN_STAGE_BYTES = (EN_TG + 2) * 16 + 96;	// slots 0 .. TG+1, padded to 128
constexpr int EN_LINE_BYTES = 2 * 8 * EN_EG + 16;	// per worker staging line (general path)
constexpr int EN_PD = 4;				// look-back windows in flight (prefix warp)
constexpr int EN_WW = 5;				// window words per lane: grid <= 32 * EN_WW + 1
constexpr int EN_MAX_GRID = 32 * EN_WW + 1;
static_assert(EN_STAGES >= EN_TEAMS * EN_K + 2, "ring too short for the A/B distance");

struct EnTileInfo {
	uint64_t off0, rend;	// sample range of the tile's read
	uint64_t kx0, kx1;	// fixed-byte prefix before / through the read (plan.cuh)
	uint64_t k0;		// first 16-byte group of the tile
	uint32_t g, r, t;	// tile id, read id, tile index inside the read
	uint32_t flags;		// bit b < 8: block b and the group after it lie inside the read, past its
				// first group; bit 31: last tile of the batch
};

struct EnSmem {
	alignas(128) uint8_t in[EN_STAGES][EN_STAGE_BYTES];
	alignas(16) uint8_t line[EN_NWORK][EN_LINE_BYTES];
	EnTileInfo info[EN_STAGES];
	uint64_t full[EN_STAGES], cnt_done[EN_STAGES], scan_done[EN_STAGES], empty[EN_STAGES];
	uint64_t blkP[EN_STAGES][EN_NB];	// container position of each block's data (prefix warp);
						// SENTINEL: the tile does not fit, write nothing
	uint64_t kbase[EN_STAGES];		// container position of the read's key section
	uint32_t blkL[EN_STAGES][EN_NB];	// data bytes of each 1024-sample block (phase A)
};

// word holding the sample before each of x's words' low halves
__device__ __forceinline__ void prev_words(const uint4 &x, uint32_t prev_half, uint32_t (&p)[4])
{
	p[0] = __byte_perm(prev_half, x.x, 0x5410);
	p[1] = __funnelshift_l(x.x, x.y, 16);
	p[2] = __funnelshift_l(x.y, x.z, 16);
	p[3] = __funnelshift_l(x.z, x.w, 16);
}

// valid-sample mask of group k of a read [off0, rend)
__device__ __forceinline__ uint32_t group_valid(uint64_t k, uint64_t off0, uint64_t rend)
{
	const int64_t lo64 = (int64_t) off0 - (int64_t) (8 * k);
	const int64_t left = (int64_t) rend - (int64_t) (8 * k);
	const int lo = lo64 > 0 ? (lo64 < 8 ? (int) lo64 : 8) : 0;
	const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
	return range_mask(lo, hi);
}

// ---------------------------------------------------------------------------
// worker warp: one 1024-sample block of a tile

// what a worker keeps in registers between phase A and phase B of a block
struct EnBlk {
	uint64_t B[EN_EC];	// the 8 data bytes of each of the lane's groups (fast blocks)
	uint32_t fast;		// warp-uniform
};

// zd words of the lane's EN_EC groups of block bw (+ of the group after the block)
__device__ __forceinline__ void en_block_zd(const EnTileInfo &ti, const uint8_t *st, int bw, int lane,
					    uint32_t (&z)[EN_EC][4], uint32_t (&zh)[4])
{
	const uint64_t A = ti.off0 >> 3;
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	const uint32_t g0 = bw * EN_EG;
	uint4 x[EN_EC];
#pragma unroll
	for (int c = 0; c < EN_EC; c++)
		x[c] = *reinterpret_cast<const uint4 *>(st + 16 * (g0 + 32 * c + lane + 1));
	const uint32_t prev0 = *reinterpret_cast<const uint16_t *>(st + 16 * g0 + 14);
	const uint4 xh = *reinterpret_cast<const uint4 *>(st + 16 * (g0 + EN_EG + 1));
	// only the read's first group needs the "previous sample is 0" rule (trans.c:221)
	const bool first = ti.t == 0 && bw == 0;
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		uint32_t prev = __shfl_up_sync(0xffffffffu, x[c].w >> 16, 1);
		if (c > 0) {
			const uint32_t wrap = __shfl_sync(0xffffffffu, x[c > 0 ? c - 1 : 0].w >> 16, 31);
			if (lane == 0)
				prev = wrap;
		} else if (lane == 0) {
			prev = prev0;
		}
		zd_group(x[c], prev, (first && c == 0 && lane == 0) ? (int) m : -1, z[c]);
	}
	zd_group(xh, __shfl_sync(0xffffffffu, x[EN_EC - 1].w >> 16, 31), -1, zh);
	(void) A;
}

// Phase A: byte count of the block -> blkL[bw]; fast blocks keep their bytes in R.
__device__ __forceinline__ void en_phase_a(const EnTileInfo &ti, const uint8_t *st, uint32_t *blkL, int bw,
					   int lane, EnBlk &R)
{
	uint32_t z[EN_EC][4], zh[4];
	en_block_zd(ti, st, bw, lane, z, zh);
	uint32_t high = zh[0] | zh[1] | zh[2] | zh[3];
#pragma unroll
	for (int c = 0; c < EN_EC; c++)
		high |= z[c][0] | z[c][1] | z[c][2] | z[c][3];
	const bool inside = (ti.flags >> bw) & 1u;
	const bool fast = inside && !__any_sync(0xffffffffu, (high & 0xff00ff00u) != 0);
	R.fast = fast;
	if (fast) {
#pragma unroll
		for (int c = 0; c < EN_EC; c++) {
			const uint32_t w0 = __byte_perm(z[c][0], z[c][1], 0x6420);
			const uint32_t w1 = __byte_perm(z[c][2], z[c][3], 0x6420);
			R.B[c] = ((uint64_t) w1 << 32) | w0;
		}
		if (lane == 0)
			blkL[bw] = 8 * EN_EG;
		return;
	}
	// exact: valid samples + two-byte values
	const uint64_t kw = ti.k0 + bw * EN_EG;
	uint32_t cnt = 0;
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		const uint32_t valid = group_valid(kw + 32 * c + lane, ti.off0, ti.rend);
		cnt += __popc(valid) + __popc(flags_group(z[c]) & valid);
	}
#pragma unroll
	for (int o = 16; o; o >>= 1)
		cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
	if (lane == 0)
		blkL[bw] = cnt;
}

// Phase B: the block's bytes to the container.  P = position of its data, L = their
// number, kbase = position of the read's key section.
template <int KEYBITS>
__device__ __forceinline__ void en_phase_b(const SvbEncArgs &a, const EnTileInfo &ti, const uint8_t *st,
					   uint64_t P, uint64_t kbase, uint32_t L, uint8_t *line, int bw,
					   int lane, const EnBlk &R)
{
	const uint64_t jw = (uint64_t) ti.t * EN_TG + bw * EN_EG;	// read-relative group of the block
	uint8_t *S = a.out + P;
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(S) & 7);
	if (R.fast) {
		// chunk i = 32 c + lane holds exactly 8 data bytes; aligned word i of the output =
		// top s bytes of chunk i - 1 | low 8 - s bytes of chunk i
		uint64_t *W = reinterpret_cast<uint64_t *>(S - s);
		if (s == 0) {
#pragma unroll
			for (int c = 0; c < EN_EC; c++)
				W[32 * c + lane] = R.B[c];
		} else {
			const uint32_t sh = 8 * s;
#pragma unroll
			for (int c = 0; c < EN_EC; c++) {
				uint64_t pb = __shfl_up_sync(0xffffffffu, R.B[c], 1);
				if (c > 0) {
					const uint64_t wrap = __shfl_sync(0xffffffffu, R.B[c > 0 ? c - 1 : 0], 31);
					if (lane == 0)
						pb = wrap;
				}
				const uint64_t v = (pb >> (64 - sh)) | (R.B[c] << sh);
				if (c > 0 || lane > 0)
					W[32 * c + lane] = v;
			}
			// partial words at both ends, one byte per lane: the low 8 - s bytes of
			// chunk 0 and the top s bytes of the last chunk
			const uint64_t b0 = __shfl_sync(0xffffffffu, R.B[0], 0);
			const uint64_t bl = __shfl_sync(0xffffffffu, R.B[EN_EC - 1], 31) >> (64 - sh);
			if ((uint32_t) lane < 8 - s)
				S[lane] = (uint8_t) (b0 >> (8 * lane));
			if ((uint32_t) lane < s)
				reinterpret_cast<uint8_t *>(W + EN_EG)[lane] = (uint8_t) (bl >> (8 * lane));
		}
		warp_zero(a.out + kbase + jw * KEYBITS, EN_EG * KEYBITS, lane);
		return;
	}

	// ---- general path: everything again from the staged tile ---------------------------
	const uint64_t kw = ti.k0 + bw * EN_EG;
	if (8 * kw >= ti.rend)
		return;			// block past the end of the read
	const uint64_t n = ti.rend - ti.off0;
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	uint32_t z[EN_EC][4], zh[4], valid[EN_EC], fl[EN_EC];
	en_block_zd(ti, st, bw, lane, z, zh);
	uint32_t run = 0;
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		valid[c] = group_valid(kw + 32 * c + lane, ti.off0, ti.rend);
		fl[c] = flags_group(z[c]) & valid[c];
		const uint32_t cnt = __popc(valid[c]) + __popc(fl[c]);
		const uint32_t inc = warp_inc_scan(cnt, lane);
		uint8_t *p = line + s + run + inc - cnt;
		run += __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
		for (int e = 0; e < 8; e++) {
			if ((valid[c] >> e) & 1) {
				const uint32_t v = (z[c][e >> 1] >> (16 * (e & 1))) & 0xffffu;
				*p++ = (uint8_t) v;
				if ((fl[c] >> e) & 1)
					*p++ = (uint8_t) (v >> 8);
			}
		}
	}
	const uint32_t fh = flags_group(zh) & group_valid(kw + EN_EG, ti.off0, ti.rend);
	__syncwarp();
	warp_copy_out(line, S, L, lane);
	__syncwarp();
	// keys: the read-relative 8-sample block of group k is k - A; key block j = bits m..7
	// of group j's flags + bits 0..m-1 of group j + 1's
	uint8_t *kout = a.out + kbase;
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		uint32_t fnext = __shfl_down_sync(0xffffffffu, fl[c], 1);
		const uint32_t wrap = c + 1 < EN_EC ? __shfl_sync(0xffffffffu, fl[c + 1 < EN_EC ? c + 1 : c], 0) : fh;
		if (lane == 31)
			fnext = wrap;
		const uint64_t j = jw + 32 * c + lane;
		const uint32_t f = ((fl[c] >> m) | (fnext << (8 - m))) & 0xffu;
		if (KEYBITS == 1) {
			if (8 * j < n)
				kout[j] = (uint8_t) f;
		} else {
			const uint32_t kk = spread8(f);
			if (8 * j < n)
				kout[2 * j] = (uint8_t) kk;
			if (8 * j + 4 < n)
				kout[2 * j + 1] = (uint8_t) (kk >> 8);
		}
	}
}

// number of window words of the CTA's i-th tile (tiles before it that are not its own)
__device__ __forceinline__ uint32_t g_window(uint32_t i, uint32_t G)
{
	return i == 0 ? blockIdx.x : G - 1;
}

// ----------------------------------------------------------------------------
// the kernel

// role timing (cycles summed over warps and CTAs into hdr->aux[], lane 0 only);
// compiled in with -DPG_ROLE_TIMING, see tools/role_timing.py
#ifdef PG_ROLE_TIMING
#define RT_DECL long long rt_t0 = 0, rt_acc[6] = {0, 0, 0, 0, 0, 0}
#define RT_BEGIN() (rt_t0 = clock64())
#define RT_END(k) (rt_acc[k] += clock64() - rt_t0)
#define RT_FLUSH(b)                                                                          \
	do {                                                                                 \
		if (lane == 0)                                                               \
			for (int k_ = 0; k_ < 6; k_++)                                       \
				atomicAdd((unsigned long long *) &a.hdr->aux[(b) + k_],      \
					  (unsigned long long) rt_acc[k_]);                  \
	} while (0)
#else
#define RT_DECL
#define RT_BEGIN()
#define RT_END(k)
#define RT_FLUSH(b)
#endif

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(EN_THREADS, 1) svb_enc_kernel(const __grid_constant__ SvbEncArgs a)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	EnSmem &sm = *reinterpret_cast<EnSmem *>(smem_raw);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t total = a.hdr->total_tiles;
	const uint32_t G = gridDim.x;
	// tiles of this CTA: blockIdx.x + i * G, i < nt
	const uint32_t nt = blockIdx.x < total ? (total - blockIdx.x + G - 1) / G : 0;

	if (threadIdx.x == 0) {
		for (int s = 0; s < EN_STAGES; s++) {
			mbar_init(&sm.full[s], 1);
			mbar_init(&sm.cnt_done[s], EN_NB);
			mbar_init(&sm.scan_done[s], 1);
			mbar_init(&sm.empty[s], EN_NB);
		}
		mbar_fence_init();
	}
	__syncthreads();
	if (nt == 0)
		return;

	if (warp == EN_W_PROD) {
		// ---- producer: metadata 32 tiles at a time, one bulk copy per tile ----------
		const uint64_t nfull = a.nsamples >> 3;
		RT_DECL;
		for (uint32_t i0 = 0; i0 < nt; i0 += 32) {
			const uint32_t i = i0 + lane;
			EnTileInfo me = {};
			if (i < nt) {
				me.g = blockIdx.x + i * G;
				me.r = __ldg(a.tile_read + me.g);
				me.t = me.g - __ldg(a.tile_start + me.r);
				me.off0 = __ldg(a.offsets + me.r);
				me.rend = __ldg(a.offsets + me.r + 1);
				me.kx0 = __ldg(a.kx + me.r);
				me.kx1 = __ldg(a.kx + me.r + 1);
				const uint64_t A = me.off0 >> 3;
				me.k0 = A + (uint64_t) me.t * EN_TG;
				uint32_t fl = me.g == total - 1 ? 0x80000000u : 0u;
				for (int b = 0; b < EN_NB; b++) {
					const uint64_t kw = me.k0 + (uint64_t) b * EN_EG;
					if (kw > A && 8 * (kw + EN_EG + 1) <= me.rend)
						fl |= 1u << b;
				}
				me.flags = fl;
			}
			const uint32_t cnt = nt - i0 < 32 ? nt - i0 : 32;
			for (uint32_t l = 0; l < cnt; l++) {
				const uint32_t it = i0 + l, s = it % EN_STAGES;
				// the whole warp waits: a lane running ahead of its turn would see the
				// parity of a phase that has not started yet as "complete"
				RT_BEGIN();
				if (it >= EN_STAGES)
					mbar_wait(&sm.empty[s], ((it / EN_STAGES) & 1) ^ 1);
				RT_END(0);
				if (lane == (int) l) {
					sm.info[s] = me;
					// groups [kf, ke) of the sample array -> slots kf - (k0 - 1) ...
					const uint64_t kf = me.k0 > 0 ? me.k0 - 1 : 0;
					uint64_t kr = (me.rend + 7) >> 3;	// groups holding samples of the read
					uint64_t ke = me.k0 + EN_TG + 1;
					ke = ke < kr ? ke : kr;
					if (ke > nfull) {	// the array's last, partial group: by hand, zero filled
						const uint4 v = load_group(a.samples, a.nsamples, nfull);
						*reinterpret_cast<uint4 *>(sm.in[s] + 16 * (nfull + 1 - me.k0)) = v;
						ke = nfull;
					}
					if (ke > kf) {
						const uint32_t bytes = (uint32_t) (ke - kf) * 16;
						mbar_arrive_expect_tx(&sm.full[s], bytes);
						tma_load_1d(sm.in[s] + 16 * (kf + 1 - me.k0), a.samples + kf, bytes,
							    &sm.full[s]);
					} else {
						mbar_arrive(&sm.full[s]);
					}
				}
				__syncwarp();
			}
		}
		RT_FLUSH(0);
		return;
	}

	if (warp == EN_W_SCAN + 1) {
		// ---- publish warp: the tile's byte count goes out the moment it is known ---------
		for (uint32_t i = 0; i < nt; i++) {
			const uint32_t s = i % EN_STAGES;
			mbar_wait(&sm.cnt_done[s], (i / EN_STAGES) & 1);
			uint32_t L = lane < EN_NB ? sm.blkL[s][lane] : 0u;
#pragma unroll
			for (int o = EN_NB / 2; o; o >>= 1)
				L += __shfl_xor_sync(0xffffffffu, L, o);
			if (lane == 0)
				st_relaxed(a.status + blockIdx.x + (uint64_t) i * G, LB_AGG | L);
		}
		return;
	}

	if (warp == EN_W_SCAN) {
		// ---- prefix warp: P(g) = P(g - G) + L(g - G) + sum of the G - 1 aggregates in between.
		// The window holds other CTAs' tiles only, so it is requested EN_PD tiles ahead;
		// nothing here waits for another CTA's prefix, only for its counts.
		RT_DECL;
		uint64_t w[EN_PD][EN_WW];
		auto issue = [&](uint32_t i, uint64_t(&ww)[EN_WW]) {
			const uint64_t g = blockIdx.x + (uint64_t) i * G;
			const uint64_t lo = i == 0 ? 0 : g - G + 1;
#pragma unroll
			for (int q = 0; q < EN_WW; q++) {
				const uint64_t j = lo + 32 * q + lane;
				ww[q] = j < g ? ld_relaxed(a.status + j) : LB_AGG;
			}
		};
		// reload the words that were still empty last time (non-blocking)
		auto repoll = [&](uint32_t i, uint64_t(&ww)[EN_WW]) {
			const uint64_t g = blockIdx.x + (uint64_t) i * G;
			const uint64_t lo = i == 0 ? 0 : g - G + 1;
#pragma unroll
			for (int q = 0; q < EN_WW; q++)
				if ((ww[q] >> 62) == 0) {
					ww[q] = ld_relaxed(a.status + lo + 32 * q + lane);
#ifdef PG_ROLE_TIMING
					atomicAdd((unsigned long long *) &a.hdr->aux[15], 1ull);
#endif
				}
		};
		auto complete = [&](const uint64_t(&ww)[EN_WW]) -> bool {
			bool ok = true;
#pragma unroll
			for (int q = 0; q < EN_WW; q++)
				ok = ok && (ww[q] >> 62) != 0;
			return __all_sync(0xffffffffu, ok);
		};
		// -> sum of the (complete) window; Wt = sum of its last t words (the tiles of the
		// same read before this one, when the read started inside the window)
		auto resolve = [&](uint32_t i, const uint64_t(&ww)[EN_WW], uint32_t t, uint64_t &Wt) -> uint64_t {
			const uint64_t g = blockIdx.x + (uint64_t) i * G;
			const uint64_t lo = i == 0 ? 0 : g - G + 1;
			uint64_t sum = 0, sumt = 0;
#pragma unroll
			for (int q = 0; q < EN_WW; q++) {
				const uint64_t j = lo + 32 * q + lane;
				sum += ww[q] & LB_VAL;
				sumt += j + t >= g ? ww[q] & LB_VAL : 0;	// j >= g reads as 0
			}
			Wt = warp_sum64(sumt);
			return warp_sum64(sum);
		};
#pragma unroll
		for (int u = 0; u < EN_PD; u++)
			if ((uint32_t) u < nt)
				issue(u, w[u]);
		uint64_t P = 0;		// exclusive prefix of the tile at hand
		uint32_t Lprev = 0;
		uint32_t r_prev = 0xffffffffu;
		uint64_t base_prev = 0;
		for (uint32_t i0 = 0; i0 < nt; i0 += EN_PD) {
#pragma unroll
			for (int u = 0; u < EN_PD; u++) {
				const uint32_t i = i0 + u;
				if (i >= nt)
					break;
				const uint32_t s = i % EN_STAGES;
				RT_BEGIN();
				mbar_wait(&sm.cnt_done[s], (i / EN_STAGES) & 1);
				RT_END(0);
				const EnTileInfo ti = sm.info[s];
				// start of the read's stream = kx0 + prefix at the read's first tile.  When
				// that tile lies inside the window (or is this tile) it follows from the
				// window words; else it was published by that tile, at least a row ago
				// (and the previous own tile of the same read already fetched it).
				const bool far = ti.t > g_window(i, G);
				uint64_t base = SENTINEL;
				if (far && ti.r == r_prev)
					base = base_prev;
				else if (far && lane == 0)
					base = ld_relaxed(a.out_offsets + ti.r);
				RT_BEGIN();
				// windows of this and the next EN_PD - 1 tiles are polled together, so the
				// round trips of one tile's late words overlap the others'
				while (!complete(w[u])) {
#pragma unroll
					for (int v = 0; v < EN_PD; v++) {
						const uint32_t tv = v >= u ? i0 + v : i0 + EN_PD + v;
						if (tv < nt)
							repoll(tv, w[v]);
					}
				}
				uint64_t Wt;
				const uint64_t W = resolve(i, w[u], ti.t, Wt);
				RT_END(1);
				RT_BEGIN();
				P = i == 0 ? W : P + Lprev + W;
				if (i + EN_PD < nt)
					issue(i + EN_PD, w[u]);
				RT_END(2);
				RT_BEGIN();
				const uint32_t lv = lane < EN_NB ? sm.blkL[s][lane] : 0u;
				uint32_t inc = lv;	// inclusive prefix over the blocks
#pragma unroll
				for (int o = 1; o < EN_NB; o <<= 1) {
					const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
					if (lane >= o)
						inc += y;
				}
				const uint32_t L = __shfl_sync(0xffffffffu, inc, EN_NB - 1);
				Lprev = L;
				if (!far) {
					base = ti.kx0 + P - Wt;
				} else if (ti.r != r_prev) {
					if (lane == 0)
						while (base == SENTINEL)
							base = ld_relaxed(a.out_offsets + ti.r);
					base = __shfl_sync(0xffffffffu, base, 0);
				}
				r_prev = ti.r;
				base_prev = base;
				// the tile's bytes and the read's frame + key section must fit
				const uint64_t n = ti.rend - ti.off0;
				const uint64_t kb = base + (FRAME ? 4 : 0);
				const bool fits = ti.kx1 + P + L <= a.out_cap &&
						  kb + (n + 8 / KEYBITS - 1) / (8 / KEYBITS) <= a.out_cap;
				if (lane < EN_NB)
					sm.blkP[s][lane] = fits ? ti.kx1 + P + inc - lv : SENTINEL;
				if (FRAME && fits && ti.t == 0 && lane < 4)	// press.c:1875-1879, 2035-2039
					a.out[base + lane] = (uint8_t) ((uint32_t) n >> (8 * lane));
				__syncwarp();	// ... before lane 0's releasing arrive
				RT_END(3);
				RT_BEGIN();
				if (lane == 0) {
					if (!fits)
						atomicOr(&a.hdr->error, ERR_CAPACITY);
					if (ti.t == 0)
						st_relaxed(a.out_offsets + ti.r, base);
					if (ti.flags >> 31)
						st_relaxed(a.out_offsets + a.nreads, __ldg(a.kx + a.nreads) + P + L);
					sm.kbase[s] = kb;
					mbar_arrive(&sm.scan_done[s]);
				}
				__syncwarp();
				RT_END(4);
			}
		}
		RT_FLUSH(3);
		return;
	}

	// ---- worker warps: team = tiles of one parity, block = warp index in the team -------------
	const int team = warp / EN_NB, bw = warp - team * EN_NB;
	const uint32_t nk = nt > (uint32_t) team ? (nt - team + EN_TEAMS - 1) / EN_TEAMS : 0;	// team tiles
	EnBlk R[EN_K];
	RT_DECL;
	const long long rt_start = clock64();
	for (uint32_t k0 = 0; k0 < nk + EN_K - 1; k0 += EN_K) {
#pragma unroll
		for (int u = 0; u < EN_K; u++) {
			const uint32_t k = k0 + u;
			if (k < nk) {		// phase A of team tile k
				const uint32_t i = team + k * EN_TEAMS, s = i % EN_STAGES;
				RT_BEGIN();
				mbar_wait(&sm.full[s], (i / EN_STAGES) & 1);
				RT_END(0);
				RT_BEGIN();
				en_phase_a(sm.info[s], sm.in[s], sm.blkL[s], bw, lane, R[u]);
				__syncwarp();
				if (lane == 0)
					mbar_arrive(&sm.cnt_done[s]);
				RT_END(1);
			}
			if (k >= EN_K - 1 && k - (EN_K - 1) < nk) {	// phase B of team tile k - (EN_K - 1)
				const uint32_t i = team + (k - (EN_K - 1)) * EN_TEAMS, s = i % EN_STAGES;
				RT_BEGIN();
				mbar_wait(&sm.scan_done[s], (i / EN_STAGES) & 1);
				RT_END(0);
				RT_BEGIN();
				const uint64_t P = sm.blkP[s][bw];
				if (P != SENTINEL)
					en_phase_b<KEYBITS>(a, sm.info[s], sm.in[s], P, sm.kbase[s], sm.blkL[s][bw],
							    sm.line[warp], bw, lane, R[(u + 1) % EN_K]);
				__syncwarp();
				if (lane == 0)
					mbar_arrive(&sm.empty[s]);
#ifdef PG_ROLE_TIMING
				if (R[(u + 1) % EN_K].fast) {
					RT_END(2);
				} else {
					RT_END(4);
					rt_acc[5]++;
				}
#endif
			}
		}
	}
#ifdef PG_ROLE_TIMING
	if (lane == 0)
		atomicAdd((unsigned long long *) &a.hdr->aux[12], (unsigned long long) (clock64() - rt_start));
#endif
	(void) rt_start;
	RT_FLUSH(9);
}

}  // namespace pg
