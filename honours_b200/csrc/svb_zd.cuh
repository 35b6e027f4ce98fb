// svb_zd.cuh -- fused zigzag-delta + streamvbyte kernels (warp tiles).
//
//   KEYBITS = 1 : ONT svb16, 1 key bit / value   (svb16/encode_scalar.hpp:14-62,
//                 decode_scalar.hpp:31-71; press.c:1678-1694)
//   KEYBITS = 2 : classical streamvbyte over zd widened to u32; because
//                 zd < 65536 only codes 0/1 occur (streamvbyte_encode.c:11-59,
//                 streamvbyte_decode.c:16-58; press.c:1585-1612)
//   FRAME       : the 4-byte [n u32 LE] prefix of the pre-zstd VBZ stream
//                 (press.c:1875-1879, 2035-2039)
//
// Stream of read r:  [frame][keys: ceil(n / (8 / KEYBITS)) bytes][data bytes].
//
// Both directions are ONE pass over HBM.  A tile is 1024 samples of one read,
// cut on 16-byte boundaries of the sample array, and is owned by one warp:
//   encode: 4 x LDG.128 per lane -> delta + zigzag on 16-bit pairs in registers
//           (VIADD.16x2 / PRMT) -> data-byte count -> decoupled look-back over
//           tiles gives the tile's byte position in the container -> bytes leave
//           as aligned 64-bit stores (a per-warp shared-memory line re-phases
//           them to the destination's alignment).
//   decode: key bytes -> byte count -> look-back (data position inside the
//           read) -> aligned 64-bit loads re-phased the same way -> unzigzag and
//           running sums on 16-bit pairs -> second look-back (value carry) ->
//           aligned 128-bit sample stores.
// Tiles with no two-byte value (the common case on nanopore signal: 4e-5 of the
// samples are exceptions) take a branch-free fast path; everything else --
// first / last tiles of a read, tiles with exceptions -- takes the general path
// with per-sample byte placement in the warp's staging line.
#pragma once

#include "common.cuh"

namespace pg {

constexpr int STAGE_BYTES = 2 * WT + 16;	// worst case: every value takes 2 bytes, + phase

struct SvbEncArgs {
	const uint4 *samples;		// 16-byte aligned
	uint64_t nsamples;		// samples readable behind `samples`
	const uint64_t *offsets;	// [nreads + 1]
	uint64_t nreads;
	uint8_t *out;			// 16-byte aligned
	uint64_t out_cap;
	uint64_t *out_offsets;		// [nreads + 1], pre-set to SENTINEL
	Hdr *hdr;
	const uint32_t *tile_start;
	const uint32_t *tile_read;
	const uint64_t *kx;
	uint64_t *status;		// look-back: data bytes, batch-wide
	uint64_t *status_blk;
};

// 16-byte group `k` of the sample array, zero filled past the end.
__device__ __forceinline__ uint4 load_group(const uint4 *samples, uint64_t nsamples, uint64_t k)
{
	if (8 * k + 8 <= nsamples)
		return __ldg(samples + k);
	uint32_t w[4] = {0, 0, 0, 0};
	const uint16_t *s = reinterpret_cast<const uint16_t *>(samples);
	for (int e = 0; e < 8; e++)
		if (8 * k + e < nsamples)
			w[e >> 1] |= (uint32_t) s[8 * k + e] << (16 * (e & 1));
	return make_uint4(w[0], w[1], w[2], w[3]);
}

// zd of the 8 samples in x given the sample before them; if zero_at >= 0 the
// predecessor of element zero_at is forced to 0 (first sample of a read,
// trans.c:221 "prev = 0").
__device__ __forceinline__ void zd_group(const uint4 &x, uint32_t prev_half, int zero_at,
					 uint32_t (&z)[4])
{
	uint32_t p[4];
	p[0] = __byte_perm(prev_half, x.x, 0x5410);
	p[1] = __funnelshift_l(x.x, x.y, 16);
	p[2] = __funnelshift_l(x.y, x.z, 16);
	p[3] = __funnelshift_l(x.z, x.w, 16);
	if (zero_at >= 0) {
#pragma unroll
		for (int j = 0; j < 4; j++)
			if ((zero_at >> 1) == j)
				p[j] &= (zero_at & 1) ? 0x0000ffffu : 0xffff0000u;
	}
	z[0] = zigzag16x2(sub16x2(x.x, p[0]));
	z[1] = zigzag16x2(sub16x2(x.y, p[1]));
	z[2] = zigzag16x2(sub16x2(x.z, p[2]));
	z[3] = zigzag16x2(sub16x2(x.w, p[3]));
}

__device__ __forceinline__ uint32_t flags_group(const uint32_t (&z)[4])
{
	return ge256x2(z[0]) | (ge256x2(z[1]) << 2) | (ge256x2(z[2]) << 4) | (ge256x2(z[3]) << 6);
}
// nonzero iff one of the 8 values needs two bytes
__device__ __forceinline__ uint32_t any_high(const uint32_t (&z)[4])
{
	return (z[0] | z[1] | z[2] | z[3]) & 0xff00ff00u;
}

// bits [lo, hi) of an 8-bit mask
__device__ __forceinline__ uint32_t range_mask(int lo, int hi)
{
	return hi > lo ? ((0xffu >> (8 - hi)) & (0xffu << lo)) & 0xffu : 0u;
}

// inclusive warp scan
__device__ __forceinline__ uint32_t warp_inc_scan(uint32_t x, int lane)
{
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
		if (lane >= o)
			x += y;
	}
	return x;
}

// Write `len` staged bytes to global `dst`; stage[] holds them at offset
// (dst & 7), so 8-byte words of shared and global memory line up.  One warp.
__device__ __forceinline__ void warp_copy_out(const uint8_t *stage, uint8_t *dst, uint32_t len, int lane)
{
	const uint32_t a = (uint32_t) (reinterpret_cast<uintptr_t>(dst) & 7);
	uint8_t *gbase = dst - a;
	const uint32_t end = a + len;
	const uint32_t nwords = (end + 7) >> 3;
	for (uint32_t q = lane; q < nwords; q += 32) {
		const uint32_t lo = q << 3, hi = lo + 8;
		if (lo >= a && hi <= end) {
			*reinterpret_cast<uint64_t *>(gbase + lo) =
				*reinterpret_cast<const uint64_t *>(stage + lo);
		} else {
			const uint32_t b0 = lo < a ? a : lo, b1 = hi > end ? end : hi;
			for (uint32_t b = b0; b < b1; b++)
				gbase[b] = stage[b];
		}
	}
}

// Zero `len` bytes at `dst` (any alignment) with 32-bit stores where possible.  One warp.
__device__ __forceinline__ void warp_zero(uint8_t *dst, uint32_t len, int lane)
{
	const uint32_t head = (uint32_t) ((4 - (reinterpret_cast<uintptr_t>(dst) & 3)) & 3);
	if (len <= head + 3) {
		if ((uint32_t) lane < len)
			dst[lane] = 0;
		return;
	}
	const uint32_t nw = (len - head) >> 2, tail = (len - head) & 3;
	uint32_t *w = reinterpret_cast<uint32_t *>(dst + head);
	for (uint32_t q = lane; q < nw; q += 32)
		w[q] = 0;
	if ((uint32_t) lane < head)
		dst[lane] = 0;
	if ((uint32_t) lane < tail)
		dst[head + 4 * nw + lane] = 0;
}

// ----------------------------------------------------------------------------
// encode

template <int KEYBITS, bool FRAME>
__device__ __forceinline__ void svb_encode_tile(const SvbEncArgs &a, const uint32_t g,
						 const uint32_t total_tiles, uint8_t *stage, const int lane)
{
	constexpr uint32_t VPK = 8 / KEYBITS;	// values per key byte
	const uint16_t *s16 = reinterpret_cast<const uint16_t *>(a.samples);

	const uint32_t r = __ldg(a.tile_read + g);
	const uint32_t t = g - __ldg(a.tile_start + r);
	const uint64_t off0 = __ldg(a.offsets + r), rend = __ldg(a.offsets + r + 1);
	const uint64_t kx0 = __ldg(a.kx + r), kx1 = __ldg(a.kx + r + 1);
	const uint64_t n = rend - off0;
	const uint64_t A = off0 >> 3;		// first 16-byte group of the read
	const uint32_t m = (uint32_t) (off0 & 7);
	const uint64_t k0 = A + (uint64_t) t * WG;

	// ---- load, zd -------------------------------------------------------------
	uint4 x[WC];
#pragma unroll
	for (int c = 0; c < WC; c++)
		x[c] = load_group(a.samples, a.nsamples, k0 + 32 * c + lane);
	uint4 xh = make_uint4(0, 0, 0, 0);	// the group after the tile (its flags feed the last key)
	uint32_t prev0 = 0;			// the sample before the tile
	if (lane == 0) {
		xh = load_group(a.samples, a.nsamples, k0 + WG);
		if (k0 > 0 && 8 * k0 - 1 < a.nsamples)
			prev0 = s16[8 * k0 - 1];
	}
	uint32_t z[WC][4];
#pragma unroll
	for (int c = 0; c < WC; c++) {
		const uint32_t last = x[c].w >> 16;
		uint32_t prev = __shfl_up_sync(0xffffffffu, last, 1);
		if (c > 0) {
			const uint32_t wrap = __shfl_sync(0xffffffffu, x[c > 0 ? c - 1 : 0].w >> 16, 31);
			if (lane == 0)
				prev = wrap;
		} else if (lane == 0) {
			prev = prev0;
		}
		zd_group(x[c], prev, (t == 0 && c == 0 && lane == 0) ? (int) m : -1, z[c]);
	}
	uint32_t zh[4];
	zd_group(xh, __shfl_sync(0xffffffffu, x[WC - 1].w >> 16, 31), -1, zh);
	// valid elements of the halo group
	uint32_t fh = 0;
	{
		const int64_t left = (int64_t) rend - (int64_t) (8 * (k0 + WG));
		const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
		if (lane == 0)
			fh = flags_group(zh) & range_mask(0, hi);
	}

	// ---- fast path test: tile and its halo inside the read, no 2-byte value -----
	uint32_t high = any_high(z[0]) | any_high(z[1]) | any_high(z[2]) | any_high(z[3]) | fh;
	const bool inside = t > 0 && 8 * (k0 + WG + 1) <= rend;
	const bool fast = inside && !__any_sync(0xffffffffu, high != 0);

	uint32_t valid[WC], fl[WC], excl[WC], L;
	if (fast) {
		L = WT;
	} else {
		uint32_t run = 0;
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint64_t k = k0 + 32 * c + lane;
			const int lo = k == A ? (int) m : 0;
			const int64_t left = (int64_t) rend - (int64_t) (8 * k);
			const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
			valid[c] = range_mask(lo, hi);
			fl[c] = flags_group(z[c]) & valid[c];
			const uint32_t cnt = __popc(valid[c]) + __popc(fl[c]);
			const uint32_t inc = warp_inc_scan(cnt, lane);
			excl[c] = run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
		}
		L = run;
	}

	// ---- global byte position of this tile's data --------------------------------
	const uint64_t D = scan_tile(a.status, a.status_blk, g, 0, L);
	uint64_t base = 0;
	if (lane == 0) {
		if (t == 0) {
			base = kx0 + D;
			st_relaxed(a.out_offsets + r, base);
		}
		if (g == total_tiles - 1)
			st_relaxed(a.out_offsets + a.nreads, __ldg(a.kx + a.nreads) + D + L);
	}
	const uint64_t P = kx1 + D;		// first data byte of this tile
	if (P + L > a.out_cap) {
		if (lane == 0)
			atomicOr(&a.hdr->error, ERR_CAPACITY);
		return;
	}
	uint8_t *S = a.out + P;
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(S) & 7);

	if (fast) {
		// every lane-chunk i = 32 c + lane holds exactly 8 data bytes; aligned word i of
		// the output = top s bytes of chunk i - 1 | low 8 - s bytes of chunk i
		uint64_t *line = reinterpret_cast<uint64_t *>(stage);	// line[1 + i] = bytes of chunk i
		uint64_t B[WC];
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint32_t w0 = __byte_perm(z[c][0], z[c][1], 0x6420);
			const uint32_t w1 = __byte_perm(z[c][2], z[c][3], 0x6420);
			B[c] = ((uint64_t) w1 << 32) | w0;
			line[1 + 32 * c + lane] = B[c];
		}
		__syncwarp();
		uint64_t *W = reinterpret_cast<uint64_t *>(S - s);
		if (s == 0) {
#pragma unroll
			for (int c = 0; c < WC; c++)
				W[32 * c + lane] = B[c];
		} else {
			const uint32_t sh = 8 * s;
#pragma unroll
			for (int c = 0; c < WC; c++) {
				const int i = 32 * c + lane;
				const uint64_t w = (line[i] >> (64 - sh)) | (B[c] << sh);
				if (i > 0) {
					W[i] = w;
				} else {	// the low s bytes of word 0 belong to the tile before
					for (uint32_t b = s; b < 8; b++)
						S[b - s] = (uint8_t) (B[0] >> (8 * (b - s)));
				}
			}
			if (lane == 31) {	// word WG: the last s bytes of the tile
				uint8_t *q = reinterpret_cast<uint8_t *>(W + WG);
				const uint64_t w = B[WC - 1] >> (64 - sh);
				for (uint32_t b = 0; b < s; b++)
					q[b] = (uint8_t) (w >> (8 * b));
			}
		}
		__syncwarp();
	} else {
#pragma unroll
		for (int c = 0; c < WC; c++) {
			uint8_t *p = stage + s + excl[c];
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
		__syncwarp();
		warp_copy_out(stage, S, L, lane);
		__syncwarp();
	}

	// ---- keys ---------------------------------------------------------------------
	if (t != 0) {	// the read's first tile published the stream start
		if (lane == 0) {
			do {
				base = ld_relaxed(a.out_offsets + r);
			} while (base == SENTINEL);
		}
	}
	base = __shfl_sync(0xffffffffu, base, 0);
	const uint64_t nkeys = (n + VPK - 1) / VPK;
	if (base + (FRAME ? 4 : 0) + nkeys > a.out_cap) {
		if (lane == 0)
			atomicOr(&a.hdr->error, ERR_CAPACITY);
		return;
	}
	uint8_t *kout = a.out + base + (FRAME ? 4 : 0);
	if (FRAME && t == 0 && lane < 4)
		a.out[base + lane] = (uint8_t) ((uint32_t) n >> (8 * lane));
	const uint64_t j0 = (uint64_t) t * WG;	// first read-relative 8-sample block of the tile
	if (fast) {
		warp_zero(kout + j0 * KEYBITS, WG * KEYBITS, lane);
		return;
	}
#pragma unroll
	for (int c = 0; c < WC; c++) {
		// block j = bits m..7 of this group's flags + bits 0..m-1 of the next group's
		uint32_t fnext = __shfl_down_sync(0xffffffffu, fl[c], 1);
		const uint32_t wrap = c + 1 < WC ? __shfl_sync(0xffffffffu, fl[c + 1 < WC ? c + 1 : c], 0)
						 : __shfl_sync(0xffffffffu, fh, 0);
		if (lane == 31)
			fnext = wrap;
		const uint64_t j = j0 + 32 * c + lane;
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

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(NT) svb_zd_encode_kernel(const __grid_constant__ SvbEncArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[NW][STAGE_BYTES];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t total_tiles = a.hdr->total_tiles;
	// Cooperative launch: every warp of the grid is resident, so the static
	// round-robin below starts tiles in index order (what the look-back wants)
	// without a ticket atomic, and a warp only ever waits on lower tiles.
	const uint32_t stride = gridDim.x * NW;
	for (uint32_t g = blockIdx.x * NW + warp; g < total_tiles; g += stride)
		svb_encode_tile<KEYBITS, FRAME>(a, g, total_tiles, s_stage[warp], lane);
}

// ----------------------------------------------------------------------------
// decode

struct SvbDecArgs {
	const uint8_t *in;		// container, 16-byte aligned
	uint64_t in_bytes;		// bytes readable behind `in`
	const uint64_t *in_offsets;	// [nreads + 1] stream starts
	const uint64_t *offsets;	// [nreads + 1] sample offsets of the output
	uint64_t nreads;
	int16_t *samples_out;		// 16-byte aligned
	Hdr *hdr;
	const uint32_t *tile_start;
	const uint32_t *tile_read;
	uint64_t *status_bytes;		// look-back 1: data bytes consumed inside the read
	uint64_t *status_bytes_blk;
	uint64_t *status_sum;		// look-back 2: sum of deltas inside the read
	uint64_t *status_sum_blk;
};

// Stage global bytes [src, src + len) into stage[] at offset (src & 7); aligned
// 8-byte words are fetched wherever they stay inside [base, limit).  One warp.
__device__ __forceinline__ void warp_copy_in(uint8_t *stage, const uint8_t *src, uint32_t len,
					     const uint8_t *base, const uint8_t *limit, int lane)
{
	const uint32_t a = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 7);
	const uint8_t *gbase = src - a;
	const uint32_t end = a + len;
	const uint32_t nwords = (end + 7) >> 3;
	for (uint32_t q = lane; q < nwords; q += 32) {
		const uint32_t lo = q << 3;
		const uint8_t *gp = gbase + lo;
		if (gp >= base && gp + 8 <= limit) {
			*reinterpret_cast<uint64_t *>(stage + lo) = __ldg(reinterpret_cast<const uint64_t *>(gp));
		} else {
			const uint32_t b0 = lo < a ? a : lo, b1 = lo + 8 > end ? end : lo + 8;
			for (uint32_t b = b0; b < b1; b++)
				stage[b] = gbase[b];
		}
	}
}

// 8-bit "two bytes" flags of read-relative block j (8 values), 0 past the end
template <int KEYBITS>
__device__ __forceinline__ uint32_t load_block_flags(const uint8_t *keys, uint64_t j, uint64_t n)
{
	if (8 * j >= n)
		return 0;
	const uint64_t left = n - 8 * j;
	uint32_t f;
	if (KEYBITS == 1) {
		f = keys[j];
	} else {
		uint32_t kk = keys[2 * j];
		if (left > 4)
			kk |= (uint32_t) keys[2 * j + 1] << 8;
		f = squeeze16(kk);
	}
	return left >= 8 ? f : f & (0xffu >> (8 - (uint32_t) left));
}

template <int KEYBITS, bool FRAME>
__device__ __forceinline__ void svb_decode_tile(const SvbDecArgs &a, const uint32_t g, uint8_t *stage,
						 const int lane)
{
	constexpr uint32_t VPK = 8 / KEYBITS;
	const uint32_t r = __ldg(a.tile_read + g);
	const uint32_t g0 = __ldg(a.tile_start + r);
	const uint32_t t = g - g0;
	const uint64_t off0 = __ldg(a.offsets + r), rend = __ldg(a.offsets + r + 1);
	const uint64_t n = rend - off0;
	const uint64_t sbase = __ldg(a.in_offsets + r);
	const uint64_t nkeys = (n + VPK - 1) / VPK;
	const uint8_t *keys = a.in + sbase + (FRAME ? 4 : 0);
	const uint8_t *data = keys + nkeys;
	const uint8_t *limit = a.in + a.in_bytes;
	const uint64_t A = off0 >> 3;
	const uint32_t m = (uint32_t) (off0 & 7);
	const uint64_t k0 = A + (uint64_t) t * WG;

	bool bad = sbase + (FRAME ? 4 : 0) + nkeys + n > a.in_bytes;	// at least 1 byte / value
	if (FRAME && !bad) {
		const uint32_t nn = keys[-4] | (keys[-3] << 8) | (keys[-2] << 16) | ((uint32_t) keys[-1] << 24);
		bad = nn != (uint32_t) n;	// press.c:2062 takes the count from the prefix
	}
	if (bad) {
		if (lane == 0) {
			atomicOr(&a.hdr->error, ERR_MALFORMED);
			st_relaxed(a.status_bytes + g, LB_INC);
			st_relaxed(a.status_sum + g, LB_INC);
			if ((g & 31) == 31) {
				st_relaxed(a.status_bytes_blk + (g >> 5), LB_INC);
				st_relaxed(a.status_sum_blk + (g >> 5), LB_INC);
			}
		}
		return;
	}

	// ---- keys -> per-group flags -------------------------------------------------
	// group k holds samples 8 (k - A) - m + e: its low m elements belong to block
	// k - A - 1, the rest to block k - A
	const uint64_t j0 = (uint64_t) t * WG;
	uint32_t F[WC];
#pragma unroll
	for (int c = 0; c < WC; c++)
		F[c] = load_block_flags<KEYBITS>(keys, j0 + 32 * c + lane, n);
	uint32_t Fb = 0;	// block before the tile
	if (lane == 0 && j0 > 0 && m > 0)
		Fb = load_block_flags<KEYBITS>(keys, j0 - 1, n);
	const bool inside = t > 0 && 8 * (k0 + WG) <= rend;
	const bool clean = !__any_sync(0xffffffffu, (F[0] | F[1] | F[2] | F[3] | Fb) != 0);

	uint32_t valid[WC], fl[WC], excl[WC], L;
#pragma unroll
	for (int c = 0; c < WC; c++) {
		uint32_t flo = __shfl_up_sync(0xffffffffu, F[c], 1);
		const uint32_t wrap = __shfl_sync(0xffffffffu, c > 0 ? F[c > 0 ? c - 1 : 0] : Fb, c > 0 ? 31 : 0);
		if (lane == 0)
			flo = wrap;
		const uint64_t k = k0 + 32 * c + lane;
		const int lo = k == A ? (int) m : 0;
		const int64_t left = (int64_t) rend - (int64_t) (8 * k);
		const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
		valid[c] = range_mask(lo, hi);
		fl[c] = (m ? ((flo >> (8 - m)) | (F[c] << m)) : F[c]) & valid[c];
	}
	const bool fast = inside && clean;
	if (fast) {
		L = WT;
	} else {
		uint32_t run = 0;
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint32_t cnt = __popc(valid[c]) + __popc(fl[c]);
			const uint32_t inc = warp_inc_scan(cnt, lane);
			excl[c] = run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
		}
		L = run;
	}

	// ---- look-back 1: data bytes of earlier tiles of this read ----------------------
	const uint64_t D = scan_tile(a.status_bytes, a.status_bytes_blk, g, g0, L);
	const uint8_t *src = data + D;
	if (src + L > limit) {
		if (lane == 0) {
			atomicOr(&a.hdr->error, ERR_MALFORMED);
			st_relaxed(a.status_sum + g, LB_INC);
			if ((g & 31) == 31)
				st_relaxed(a.status_sum_blk + (g >> 5), LB_INC);
		}
		return;
	}
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 7);

	// ---- bytes -> zd values, two per word -------------------------------------------
	uint32_t zz[WC][4];
	if (fast && src - s + 8 * (WG + 1) <= limit) {
		uint64_t *line = reinterpret_cast<uint64_t *>(stage);
		const uint64_t *W = reinterpret_cast<const uint64_t *>(src - s);
		uint64_t w[WC];
#pragma unroll
		for (int c = 0; c < WC; c++) {
			w[c] = __ldg(W + 32 * c + lane);
			line[32 * c + lane] = w[c];
		}
		if (lane == 31)
			line[WG] = s ? __ldg(W + WG) : 0;
		__syncwarp();
#pragma unroll
		for (int c = 0; c < WC; c++) {
			uint64_t B = w[c];
			if (s)
				B = (B >> (8 * s)) | (line[32 * c + lane + 1] << (64 - 8 * s));
			const uint32_t blo = (uint32_t) B, bhi = (uint32_t) (B >> 32);
			zz[c][0] = __byte_perm(blo, 0, 0x4140);
			zz[c][1] = __byte_perm(blo, 0, 0x4342);
			zz[c][2] = __byte_perm(bhi, 0, 0x4140);
			zz[c][3] = __byte_perm(bhi, 0, 0x4342);
		}
		__syncwarp();
	} else {
		if (fast) {	// clean tile at the very end of the buffer: general loads
#pragma unroll
			for (int c = 0; c < WC; c++)
				excl[c] = 8 * (32 * c + lane);
		}
		warp_copy_in(stage, src, L, a.in, limit, lane);
		__syncwarp();
#pragma unroll
		for (int c = 0; c < WC; c++) {
			const uint8_t *p = stage + s + excl[c];
			zz[c][0] = zz[c][1] = zz[c][2] = zz[c][3] = 0;
#pragma unroll
			for (int e = 0; e < 8; e++) {
				if ((valid[c] >> e) & 1) {
					uint32_t v = *p++;
					if ((fl[c] >> e) & 1)
						v |= (uint32_t) (*p++) << 8;
					zz[c][e >> 1] |= v << (16 * (e & 1));
				}
			}
		}
		__syncwarp();
	}

	// ---- unzigzag, running sums inside the group (mod 2^16 in each half) ------------
	uint32_t y[WC][4], T[WC];
#pragma unroll
	for (int c = 0; c < WC; c++) {
		uint32_t carry = 0;
#pragma unroll
		for (int q = 0; q < 4; q++) {
			const uint32_t d = unzigzag16x2(zz[c][q]);	// decode_scalar.hpp:10-13
			y[c][q] = add16x2(d * 0x00010001u, carry);	// (lo, lo + hi) + carry
			carry = __byte_perm(y[c][q], 0, 0x3232);
		}
		T[c] = carry;	// group total in both halves
	}
	// scan of the group totals in (chunk, lane) order, two chunks per register
	uint32_t p01 = __byte_perm(T[0], T[1], 0x5410), p23 = __byte_perm(T[2], T[3], 0x5410);
	const uint32_t o01 = p01, o23 = p23;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t y01 = __shfl_up_sync(0xffffffffu, p01, o);
		const uint32_t y23 = __shfl_up_sync(0xffffffffu, p23, o);
		if (lane >= o) {
			p01 = add16x2(p01, y01);
			p23 = add16x2(p23, y23);
		}
	}
	const uint32_t t01 = __shfl_sync(0xffffffffu, p01, 31), t23 = __shfl_sync(0xffffffffu, p23, 31);
	const uint32_t e01 = sub16x2(p01, o01), e23 = sub16x2(p23, o23);	// exclusive over lanes
	uint32_t vex[WC];
	vex[0] = e01 & 0xffffu;
	vex[1] = (e01 >> 16) + t01;
	vex[2] = e23 + t01 + (t01 >> 16);
	vex[3] = (e23 >> 16) + t01 + (t01 >> 16) + t23;
	const uint32_t S = (t01 + (t01 >> 16) + t23 + (t23 >> 16)) & 0xffffu;

	// ---- look-back 2: sum of deltas of earlier tiles of this read -------------------
	const uint64_t V = scan_tile(a.status_sum, a.status_sum_blk, g, g0, S);

	uint4 *out4 = reinterpret_cast<uint4 *>(a.samples_out);
	uint16_t *out16 = reinterpret_cast<uint16_t *>(a.samples_out);
#pragma unroll
	for (int c = 0; c < WC; c++) {
		const uint32_t cw = ((uint32_t) V + vex[c]) & 0xffffu;
		const uint32_t carry = cw * 0x00010001u;
		uint4 o;
		o.x = add16x2(y[c][0], carry);
		o.y = add16x2(y[c][1], carry);
		o.z = add16x2(y[c][2], carry);
		o.w = add16x2(y[c][3], carry);
		const uint64_t k = k0 + 32 * c + lane;
		if (valid[c] == 0xffu) {
			out4[k] = o;
		} else if (valid[c]) {
			const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
			for (int e = 0; e < 8; e++)
				if ((valid[c] >> e) & 1)
					out16[8 * k + e] = (uint16_t) (w[e >> 1] >> (16 * (e & 1)));
		}
	}
}

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(NT) svb_zd_decode_kernel(const __grid_constant__ SvbDecArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[NW][STAGE_BYTES];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t total_tiles = a.hdr->total_tiles;
	const uint32_t stride = gridDim.x * NW;	// cooperative launch, see the encode kernel
	for (uint32_t g = blockIdx.x * NW + warp; g < total_tiles; g += stride)
		svb_decode_tile<KEYBITS, FRAME>(a, g, s_stage[warp], lane);
}

}  // namespace pg
