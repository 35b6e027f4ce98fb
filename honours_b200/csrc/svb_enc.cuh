// svb_enc.cuh -- fused zigzag-delta + streamvbyte ENCODE, CTA tiles staged by TMA.
//
//   KEYBITS = 1 : ONT svb16, 1 key bit / value   (svb16/encode_scalar.hpp:14-62; press.c:1678-1687)
//   KEYBITS = 2 : classical streamvbyte over zd widened to u32; zd < 65536 so only
//                 codes 0/1 occur (streamvbyte_encode.c:11-59; press.c:1590-1598)
//   FRAME       : the [n u32 LE] prefix of the pre-zstd VBZ stream (press.c:1875-1879, 2035-2039)
//
// Stream of read r:  [frame][keys: ceil(n / (8 / KEYBITS)) bytes][data bytes].
//
// One pass over HBM.  A tile is 8192 samples of ONE read cut on 16-byte
// boundaries of the sample array.  Per CTA (see pipe.cuh for the roles):
//   producer warp : tile metadata -> one cp.async.bulk per tile into a ring of
//                   EN_STAGES shared-memory stages (group before the tile and
//                   group after it included: previous sample / last key byte);
//   8 consumer warps, 1024 samples each (4 x LDS.128 per lane, lane interleaved):
//       COUNT(i+1): delta + zigzag on 16-bit pairs (VIADD.16x2 / PRMT), number of
//                   data bytes of the warp -> shared memory;
//       EMIT(i)   : same registers recomputed from shared memory, bytes placed at
//                   the tile's container position: aligned 64-bit stores for
//                   warps without a two-byte value (the common case on nanopore
//                   signal), per-sample placement through a staging line otherwise;
//   scan warp     : decoupled look-back over the per-tile byte counts (batch
//                   wide; the key sections are data independent, plan.cuh).
#pragma once

#include "common.cuh"
#include "pipe.cuh"
#include "svb_zd.cuh"

namespace pg {

constexpr int EN_TS = 8192;			// samples per tile
constexpr int EN_TG = EN_TS / 8;		// 16-byte groups per tile
constexpr int EN_CW = 8;			// consumer warps
constexpr int EN_WG = EN_TG / EN_CW;		// groups per consumer warp (128)
constexpr int EN_THREADS = (EN_CW + 2) * 32;	// + producer warp + scan warp
constexpr int EN_STAGES = 4;			// input ring depth
constexpr int EN_Q = 4;				// depth of the count / scan hand-off rings
constexpr int EN_STAGE_BYTES = (EN_TG + 2) * 16 + 96;	// slots 0 .. TG+1, padded to 128
constexpr int EN_LINE_BYTES = 2 * 8 * EN_WG + 16;	// per-warp staging line (general path)

struct EnTileInfo {
	uint64_t off0, rend;	// sample range of the tile's read
	uint64_t kx0, kx1;	// fixed-byte prefix before / through the read (plan.cuh)
	uint64_t k0;		// first 16-byte group of the tile
	uint32_t g, r, t, last;	// tile id, read id, tile index inside the read, last tile of the batch
};

struct EnSmem {
	alignas(128) uint8_t in[EN_STAGES][EN_STAGE_BYTES];
	alignas(16) uint8_t line[EN_CW][EN_LINE_BYTES];
	EnTileInfo info[EN_STAGES];
	uint64_t full[EN_STAGES], empty[EN_STAGES];
	uint64_t cnt_done[EN_Q], scan_done[EN_Q];
	uint64_t D[EN_Q], base[EN_Q];
	uint32_t warpL[EN_Q][EN_CW];
};

// the 4 groups of a lane: zd words, valid-sample masks, two-byte flags
struct EnRegs {
	uint32_t z[WC][4];
	uint32_t valid[WC], fl[WC];
	uint32_t fh;		// flags of the group after the warp's region (valid part)
	uint32_t high;		// nonzero iff one of this lane's values (or fh) needs two bytes
};

__device__ __forceinline__ uint4 en_group(const SvbEncArgs &a, const uint8_t *st, uint64_t k0, uint32_t gi,
					  uint64_t nfull)
{
	const uint64_t k = k0 + gi;
	if (k < nfull)
		return *reinterpret_cast<const uint4 *>(st + 16 * (gi + 1));
	return load_group(a.samples, a.nsamples, k);
}

// Everything EMIT needs and COUNT sums, from the staged tile.  One warp.
__device__ __forceinline__ void en_regs(const SvbEncArgs &a, const EnTileInfo &ti, const uint8_t *st,
					int w, int lane, EnRegs &R)
{
	const uint64_t nfull = a.nsamples >> 3;
	const uint64_t A = ti.off0 >> 3;
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	const uint32_t g0 = w * EN_WG;	// first group of the warp inside the tile
	uint4 x[WC];
#pragma unroll
	for (int c = 0; c < WC; c++)
		x[c] = en_group(a, st, ti.k0, g0 + 32 * c + lane, nfull);
	// sample before the warp's region: last halfword of slot g0 (group k0 + g0 - 1)
	const uint32_t prev0 = *reinterpret_cast<const uint16_t *>(st + 16 * g0 + 14);
	const uint4 xh = en_group(a, st, ti.k0, g0 + EN_WG, nfull);
#pragma unroll
	for (int c = 0; c < WC; c++) {
		const uint32_t last = x[c].w >> 16;
		uint32_t prev = __shfl_up_sync(0xffffffffu, last, 1);
		const uint32_t wrap = __shfl_sync(0xffffffffu, x[c > 0 ? c - 1 : 0].w >> 16, 31);
		if (lane == 0)
			prev = c > 0 ? wrap : prev0;
		const uint64_t k = ti.k0 + g0 + 32 * c + lane;
		zd_group(x[c], prev, (k == A) ? (int) m : -1, R.z[c]);
		const int lo = k == A ? (int) m : 0;
		const int64_t left = (int64_t) ti.rend - (int64_t) (8 * k);
		const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
		R.valid[c] = k >= A ? range_mask(lo, hi) : 0u;
		R.fl[c] = flags_group(R.z[c]) & R.valid[c];
	}
	uint32_t zh[4];
	zd_group(xh, __shfl_sync(0xffffffffu, x[WC - 1].w >> 16, 31), -1, zh);
	{
		const int64_t left = (int64_t) ti.rend - (int64_t) (8 * (ti.k0 + g0 + EN_WG));
		const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
		R.fh = flags_group(zh) & range_mask(0, hi);
	}
	R.high = R.fl[0] | R.fl[1] | R.fl[2] | R.fl[3] | R.fh;
}

template <int KEYBITS, bool FRAME>
__device__ __forceinline__ void en_emit(const SvbEncArgs &a, const EnTileInfo &ti, const EnRegs &R,
					 uint64_t D, uint64_t base, const uint32_t *warpL, uint8_t *line,
					 int w, int lane)
{
	constexpr uint32_t VPK = 8 / KEYBITS;
	const uint64_t n = ti.rend - ti.off0;
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	const uint32_t g0 = w * EN_WG;
	const uint64_t kw = ti.k0 + g0;			// first group of the warp
	// the warp's region lies inside the read and has no two-byte value
	const bool inside = 8 * kw >= ti.off0 && 8 * (kw + EN_WG + 1) <= ti.rend && !(ti.t == 0 && w == 0);
	const bool fast = inside && !__any_sync(0xffffffffu, R.high != 0);

	uint32_t woff = 0;
#pragma unroll
	for (int q = 0; q < EN_CW; q++)
		woff += q < w ? warpL[q] : 0u;
	const uint32_t L = warpL[w];
	const uint64_t P = ti.kx1 + D + woff;		// first data byte of this warp
	if (P + L > a.out_cap) {
		if (lane == 0)
			atomicOr(&a.hdr->error, ERR_CAPACITY);
		return;
	}
	uint8_t *S = a.out + P;
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(S) & 7);

	if (fast) {
		// chunk i = 32 c + lane holds exactly 8 data bytes; aligned word i of the output =
		// top s bytes of chunk i - 1 | low 8 - s bytes of chunk i
		uint64_t B[WC];
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint32_t w0 = __byte_perm(R.z[c][0], R.z[c][1], 0x6420);
			const uint32_t w1 = __byte_perm(R.z[c][2], R.z[c][3], 0x6420);
			B[c] = ((uint64_t) w1 << 32) | w0;
		}
		uint64_t *W = reinterpret_cast<uint64_t *>(S - s);
		if (s == 0) {
#pragma unroll
			for (int c = 0; c < WC; c++)
				W[32 * c + lane] = B[c];
		} else {
			const uint32_t sh = 8 * s;
#pragma unroll
			for (int c = 0; c < WC; c++) {
				uint64_t pb = __shfl_up_sync(0xffffffffu, B[c], 1);
				const uint64_t wrap = __shfl_sync(0xffffffffu, B[c > 0 ? c - 1 : 0], 31);
				if (lane == 0)
					pb = wrap;
				const uint64_t v = (pb >> (64 - sh)) | (B[c] << sh);
				if (c > 0 || lane > 0) {
					W[32 * c + lane] = v;
				} else {	// the low s bytes of word 0 belong to whoever precedes
					for (uint32_t b = s; b < 8; b++)
						S[b - s] = (uint8_t) (B[0] >> (8 * (b - s)));
				}
			}
			if (lane == 31) {	// word EN_WG: the last s bytes of the warp's data
				uint8_t *q = reinterpret_cast<uint8_t *>(W + EN_WG);
				const uint64_t v = B[WC - 1] >> (64 - sh);
				for (uint32_t b = 0; b < s; b++)
					q[b] = (uint8_t) (v >> (8 * b));
			}
		}
	} else {
		uint32_t run = 0;
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint32_t cnt = __popc(R.valid[c]) + __popc(R.fl[c]);
			const uint32_t inc = warp_inc_scan(cnt, lane);
			uint8_t *p = line + s + run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
			for (int e = 0; e < 8; e++) {
				if ((R.valid[c] >> e) & 1) {
					const uint32_t v = (R.z[c][e >> 1] >> (16 * (e & 1))) & 0xffffu;
					*p++ = (uint8_t) v;
					if ((R.fl[c] >> e) & 1)
						*p++ = (uint8_t) (v >> 8);
				}
			}
		}
		__syncwarp();
		warp_copy_out(line, S, L, lane);
		__syncwarp();
	}

	// ---- keys ---------------------------------------------------------------------
	const uint64_t nkeys = (n + VPK - 1) / VPK;
	if (base + (FRAME ? 4 : 0) + nkeys > a.out_cap) {
		if (lane == 0)
			atomicOr(&a.hdr->error, ERR_CAPACITY);
		return;
	}
	uint8_t *kout = a.out + base + (FRAME ? 4 : 0);
	if (FRAME && ti.t == 0 && w == 0 && lane < 4)
		a.out[base + lane] = (uint8_t) ((uint32_t) n >> (8 * lane));
	// read-relative 8-sample block of group k is k - A: block j = bits m..7 of group j's
	// flags + bits 0..m-1 of group j + 1's
	const uint64_t A = ti.off0 >> 3;
	if (kw + EN_WG <= A || 8 * kw >= ti.rend)
		return;			// warp entirely outside the read
	const uint64_t j0 = kw - A;	// kw >= A here except for the read's first warp (kw == A)
	if (fast) {
		warp_zero(kout + j0 * KEYBITS, EN_WG * KEYBITS, lane);
		return;
	}
#pragma unroll
	for (int c = 0; c < WC; c++) {
		uint32_t fnext = __shfl_down_sync(0xffffffffu, R.fl[c], 1);
		const uint32_t wrap = c + 1 < WC ? __shfl_sync(0xffffffffu, R.fl[c + 1 < WC ? c + 1 : c], 0) : R.fh;
		if (lane == 31)
			fnext = wrap;
		const uint64_t k = kw + 32 * c + lane;
		if (k < A)
			continue;
		const uint64_t j = k - A;
		const uint32_t f = ((R.fl[c] >> m) | (fnext << (8 - m))) & 0xffu;
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

// ----------------------------------------------------------------------------
// the kernel

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(EN_THREADS) svb_enc_kernel(const __grid_constant__ SvbEncArgs a)
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
			mbar_init(&sm.empty[s], EN_CW);
		}
		for (int q = 0; q < EN_Q; q++) {
			mbar_init(&sm.cnt_done[q], EN_CW);
			mbar_init(&sm.scan_done[q], 1);
		}
		mbar_fence_init();
	}
	__syncthreads();
	if (nt == 0)
		return;

	if (warp == EN_CW) {
		// ---- producer: metadata 32 tiles at a time, one bulk copy per tile ----------
		const uint64_t nfull = a.nsamples >> 3;
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
				me.k0 = (me.off0 >> 3) + (uint64_t) me.t * EN_TG;
				me.last = me.g == total - 1;
			}
			const uint32_t cnt = nt - i0 < 32 ? nt - i0 : 32;
			for (uint32_t l = 0; l < cnt; l++) {
				const uint32_t it = i0 + l, s = it % EN_STAGES;
				// the whole warp waits: a lane running ahead of its turn would see the
				// parity of a phase that has not started yet as "complete"
				if (it >= EN_STAGES)
					mbar_wait(&sm.empty[s], ((it / EN_STAGES) & 1) ^ 1);
				if (lane == (int) l) {
					sm.info[s] = me;
					// groups [kf, ke) -> slots kf - (k0 - 1) ...
					const uint64_t kf = me.k0 > 0 ? me.k0 - 1 : 0;
					uint64_t ke = me.k0 + EN_TG + 1;
					const uint64_t kr = (me.rend + 7) >> 3;
					ke = ke < kr ? ke : kr;
					ke = ke < nfull ? ke : nfull;
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
		return;
	}

	if (warp == EN_CW + 1) {
		// ---- scan warp: prefix of the data bytes, stream start of the read -----------
		for (uint32_t i = 0; i < nt; i++) {
			const uint32_t q = i % EN_Q, s = i % EN_STAGES;
			mbar_wait(&sm.cnt_done[q], (i / EN_Q) & 1);
			const EnTileInfo ti = sm.info[s];
			uint32_t L = lane < EN_CW ? sm.warpL[q][lane] : 0u;
#pragma unroll
			for (int o = 4; o; o >>= 1)
				L += __shfl_xor_sync(0xffffffffu, L, o);
			L = __shfl_sync(0xffffffffu, L, 0);
			const uint64_t D = scan_tile(a.status, a.status_blk, ti.g, 0, L);
			uint64_t base = 0;
			if (lane == 0) {
				if (ti.t == 0) {
					base = ti.kx0 + D;
					st_relaxed(a.out_offsets + ti.r, base);
				} else {	// published by the read's first tile (a lower tile)
					do {
						base = ld_relaxed(a.out_offsets + ti.r);
					} while (base == SENTINEL);
				}
				if (ti.last)
					st_relaxed(a.out_offsets + a.nreads, __ldg(a.kx + a.nreads) + D + L);
				sm.D[q] = D;
				sm.base[q] = base;
				mbar_arrive(&sm.scan_done[q]);
			}
			__syncwarp();
		}
		return;
	}

	// ---- consumers ----------------------------------------------------------------------
	auto count = [&](uint32_t j) {
		const uint32_t s = j % EN_STAGES, q = j % EN_Q;
		mbar_wait(&sm.full[s], (j / EN_STAGES) & 1);
		EnRegs R;
		en_regs(a, sm.info[s], sm.in[s], warp, lane, R);
		uint32_t cnt = 0;
#pragma unroll
		for (int c = 0; c < WC; c++)
			cnt += __popc(R.valid[c]) + __popc(R.fl[c]);
#pragma unroll
		for (int o = 16; o; o >>= 1)
			cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
		if (lane == 0) {
			sm.warpL[q][warp] = cnt;
			mbar_arrive(&sm.cnt_done[q]);
		}
	};
	count(0);
	for (uint32_t i = 0; i < nt; i++) {
		if (i + 1 < nt)
			count(i + 1);
		const uint32_t s = i % EN_STAGES, q = i % EN_Q;
		EnRegs R;
		en_regs(a, sm.info[s], sm.in[s], warp, lane, R);
		mbar_wait(&sm.scan_done[q], (i / EN_Q) & 1);
		en_emit<KEYBITS, FRAME>(a, sm.info[s], R, sm.D[q], sm.base[q], sm.warpL[q], sm.line[warp], warp, lane);
		__syncwarp();
		if (lane == 0)
			mbar_arrive(&sm.empty[s]);
	}
}

}  // namespace pg
