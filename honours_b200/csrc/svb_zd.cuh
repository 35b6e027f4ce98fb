// svb_zd.cuh -- fused zigzag-delta + streamvbyte kernels.
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
// Encode is ONE pass over the samples: delta + zigzag in registers (two 16-bit
// lanes per word), key bits from a >=256 test, per-thread byte counts ->
// CTA scan -> decoupled look-back across tiles gives the global byte position
// of the tile's data; bytes are staged in shared memory pre-shifted to the
// destination's 16-byte phase and leave as aligned 128-bit stores.
#pragma once

#include "common.cuh"

namespace pg {

struct SvbEncArgs {
	const uint4 *samples;		// 16-byte aligned
	uint64_t nsamples;		// samples readable behind `samples`
	const uint64_t *offsets;	// [nreads + 1]
	uint64_t nreads;
	uint8_t *out;
	uint64_t out_cap;
	uint64_t *out_offsets;		// [nreads + 1], pre-set to SENTINEL
	Hdr *hdr;
	const uint32_t *tile_start;
	const uint32_t *tile_read;
	const uint64_t *kx;
	uint64_t *status;
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
	p[0] = (x.x << 16) | (prev_half & 0xffffu);
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

// Copy `len` staged bytes to global `dst`.  The staging buffer was filled at
// offset (dst & 15), so 16-byte chunks of shared and global memory line up.
__device__ __forceinline__ void copy_out_aligned(const uint8_t *s_buf, uint8_t *dst, uint32_t len)
{
	const uint32_t a = (uint32_t) (reinterpret_cast<uintptr_t>(dst) & 15);
	uint8_t *gbase = dst - a;
	const uint32_t end = a + len;
	const uint32_t nchunks = (end + 15) >> 4;
	for (uint32_t q = threadIdx.x; q < nchunks; q += NT) {
		uint32_t lo = q << 4, hi = lo + 16;
		if (lo >= a && hi <= end) {
			*reinterpret_cast<uint4 *>(gbase + lo) =
				*reinterpret_cast<const uint4 *>(s_buf + lo);
		} else {
			uint32_t b0 = lo < a ? a : lo, b1 = hi > end ? end : hi;
			for (uint32_t b = b0; b < b1; b++)
				gbase[b] = s_buf[b];
		}
	}
}

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(NT) svb_zd_encode_kernel(SvbEncArgs a)
{
	__shared__ __align__(16) uint8_t s_buf[2 * TILE + 32];
	__shared__ uint8_t s_flags[GROUPS + 8];
	__shared__ uint32_t s_warp[CHUNKS * NW + 1];
	__shared__ uint32_t s_ticket;
	__shared__ uint64_t s_D, s_base;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const uint16_t *s16 = reinterpret_cast<const uint16_t *>(a.samples);
	const uint32_t total_tiles = a.hdr->total_tiles;
	constexpr uint32_t VPK = 8 / KEYBITS;	// values per key byte

	for (;;) {
		__syncthreads();	// protects s_ticket / s_buf reuse
		if (tid == 0)
			s_ticket = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint32_t g = s_ticket;
		if (g >= total_tiles)
			break;

		const uint32_t r = a.tile_read[g];
		const uint32_t t = g - a.tile_start[r];
		const uint64_t off0 = a.offsets[r];
		const uint64_t n = a.offsets[r + 1] - off0;
		const uint64_t A = off0 >> 3;		// first 16-byte group of the read
		const uint32_t m = (uint32_t) (off0 & 7);
		const uint64_t vbase = A + (uint64_t) t * GROUPS;
		const uint64_t rend = off0 + n;

		// ---- load, zd, flags, byte counts ----------------------------------
		uint32_t z[CHUNKS][4], valid[CHUNKS], fl[CHUNKS], cnt[CHUNKS];
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			const uint64_t k = vbase + c * NT + tid;
			uint4 x = load_group(a.samples, a.nsamples, k);
			uint32_t prev = __shfl_up_sync(0xffffffffu, x.w >> 16, 1);
			if (lane == 0)
				prev = (k > 0 && 8 * k - 1 < a.nsamples) ? s16[8 * k - 1] : 0;
			zd_group(x, prev, k == A ? (int) m : -1, z[c]);
			// valid element range [lo, hi) of this group inside the read
			int lo = k == A ? (int) m : 0;
			int64_t left = (int64_t) rend - (int64_t) (8 * k);
			int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
			valid[c] = hi > lo ? ((0xffu >> (8 - hi)) & (0xffu << lo)) & 0xffu : 0;
			fl[c] = flags_group(z[c]) & valid[c];
			cnt[c] = __popc(valid[c]) + __popc(fl[c]);
			s_flags[c * NT + tid] = (uint8_t) fl[c];
			// flags of the group after the tile (needed when keys straddle groups)
			if (c == CHUNKS - 1 && tid == NT - 1) {
				uint4 xh = load_group(a.samples, a.nsamples, k + 1);
				uint32_t zh[4];
				zd_group(xh, x.w >> 16, -1, zh);
				int64_t lefth = (int64_t) rend - (int64_t) (8 * (k + 1));
				int hih = lefth >= 8 ? 8 : (lefth > 0 ? (int) lefth : 0);
				s_flags[GROUPS] = (uint8_t) (flags_group(zh) & (hih ? (0xffu >> (8 - hih)) : 0u));
			}
		}

		uint32_t excl[CHUNKS], L;
		block_scan<CHUNKS>(cnt, excl, L, s_warp);

		// ---- global byte position of this tile's data (warp 0) -------------
		if (warp == 0) {
			uint64_t D = 0;
			if (g > 0) {
				if (lane == 0)
					st_relaxed(a.status + g, LB_AGG | L);
				D = lookback(a.status, g, 0);
			}
			if (lane == 0) {
				st_relaxed(a.status + g, LB_INC | (D + L));
				uint64_t base;
				if (t == 0) {
					base = a.kx[r] + D;
					st_relaxed(a.out_offsets + r, base);
				} else {
					do {
						base = ld_relaxed(a.out_offsets + r);
					} while (base == SENTINEL);
				}
				if (g == total_tiles - 1)
					st_relaxed(a.out_offsets + a.nreads, a.kx[a.nreads] + D + L);
				s_D = D;
				s_base = base;
			}
		}
		__syncthreads();
		const uint64_t P = a.kx[r + 1] + s_D;	// first data byte of this tile
		const uint64_t base = s_base;
		const uint64_t nkeys = (n + VPK - 1) / VPK;
		const bool fits = base + (FRAME ? 4 : 0) + nkeys <= a.out_cap && P + L <= a.out_cap;
		if (!fits) {
			if (tid == 0)
				atomicOr(&a.hdr->error, ERR_CAPACITY);
			continue;
		}

		// ---- stage data bytes at the destination's 16-byte phase -----------
		const uint32_t ph = (uint32_t) (reinterpret_cast<uintptr_t>(a.out + P) & 15);
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			uint8_t *p = s_buf + ph + excl[c];
#pragma unroll
			for (int e = 0; e < 8; e++) {
				if ((valid[c] >> e) & 1) {
					uint32_t v = (z[c][e >> 1] >> (16 * (e & 1))) & 0xffffu;
					*p++ = (uint8_t) v;
					if ((fl[c] >> e) & 1)
						*p++ = (uint8_t) (v >> 8);
				}
			}
		}
		__syncthreads();
		copy_out_aligned(s_buf, a.out + P, L);

		// ---- keys -----------------------------------------------------------
		uint8_t *kout = a.out + base + (FRAME ? 4 : 0);
		if (FRAME && t == 0 && tid < 4)
			a.out[base + tid] = (uint8_t) ((uint32_t) n >> (8 * tid));
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			const uint32_t gi = c * NT + tid;
			// read-relative 8-sample block j is made of group gi (bits m..7)
			// and group gi + 1 (bits 0..m-1)
			const uint64_t j = (uint64_t) t * GROUPS + gi;
			uint32_t f = ((uint32_t) s_flags[gi] >> m) | ((uint32_t) s_flags[gi + 1] << (8 - m));
			f &= 0xffu;
			if (KEYBITS == 1) {
				if (8 * j < n)
					kout[j] = (uint8_t) f;
			} else {
				uint32_t kk = spread8(f);
				if (8 * j < n)
					kout[2 * j] = (uint8_t) kk;
				if (8 * j + 4 < n)
					kout[2 * j + 1] = (uint8_t) (kk >> 8);
			}
		}
	}
}

// ----------------------------------------------------------------------------
// decode

struct SvbDecArgs {
	const uint8_t *in;		// container
	uint64_t in_bytes;		// bytes readable behind `in`
	const uint64_t *in_offsets;	// [nreads + 1] stream starts
	const uint64_t *offsets;	// [nreads + 1] sample offsets of the output
	uint64_t nreads;
	int16_t *samples_out;		// 16-byte aligned
	Hdr *hdr;
	const uint32_t *tile_start;
	const uint32_t *tile_read;
	uint64_t *status_bytes;		// look-back 1: data bytes consumed inside the read
	uint64_t *status_sum;		// look-back 2: sum of deltas inside the read
};

// Copy `cnt` staged int16 to global `dst` (2-byte aligned); staging buffer was
// filled at element offset ((dst >> 1) & 7).
__device__ __forceinline__ void copy_out_samples(const uint16_t *s_out, int16_t *dst, uint32_t cnt)
{
	const uint32_t a = (uint32_t) ((reinterpret_cast<uintptr_t>(dst) >> 1) & 7);
	uint16_t *gbase = reinterpret_cast<uint16_t *>(dst) - a;
	const uint32_t end = a + cnt;
	const uint32_t nchunks = (end + 7) >> 3;
	for (uint32_t q = threadIdx.x; q < nchunks; q += NT) {
		uint32_t lo = q << 3, hi = lo + 8;
		if (lo >= a && hi <= end) {
			*reinterpret_cast<uint4 *>(gbase + lo) =
				*reinterpret_cast<const uint4 *>(s_out + lo);
		} else {
			uint32_t b0 = lo < a ? a : lo, b1 = hi > end ? end : hi;
			for (uint32_t b = b0; b < b1; b++)
				gbase[b] = s_out[b];
		}
	}
}

// Stage global bytes [src, src + len) into shared memory at offset (src & 15).
// [base, limit) is the readable container; whole 16-byte chunks are fetched
// wherever they stay inside it.
__device__ __forceinline__ void copy_in_aligned(uint8_t *s_buf, const uint8_t *src, uint32_t len,
						const uint8_t *base, const uint8_t *limit)
{
	const uint32_t a = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
	const uint8_t *gbase = src - a;
	const uint32_t end = a + len;
	const uint32_t nchunks = (end + 15) >> 4;
	for (uint32_t q = threadIdx.x; q < nchunks; q += NT) {
		const uint32_t lo = q << 4;
		const uint8_t *gp = gbase + lo;
		if (gp >= base && gp + 16 <= limit) {
			*reinterpret_cast<uint4 *>(s_buf + lo) = __ldg(reinterpret_cast<const uint4 *>(gp));
		} else {
			uint32_t b0 = lo < a ? a : lo, b1 = lo + 16 > end ? end : lo + 16;
			for (uint32_t b = b0; b < b1; b++)
				s_buf[b] = gbase[b];
		}
	}
}

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(NT) svb_zd_decode_kernel(SvbDecArgs a)
{
	__shared__ __align__(16) uint8_t s_buf[2 * TILE + 32];
	__shared__ __align__(16) uint16_t s_out[TILE + 16];
	__shared__ uint32_t s_warp[CHUNKS * NW + 1];
	__shared__ uint32_t s_ticket;
	__shared__ uint64_t s_D;
	__shared__ uint32_t s_V;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const uint32_t total_tiles = a.hdr->total_tiles;
	constexpr uint32_t VPK = 8 / KEYBITS;

	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_ticket = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint32_t g = s_ticket;
		if (g >= total_tiles)
			break;

		const uint32_t r = a.tile_read[g];
		const uint32_t g0 = a.tile_start[r];
		const uint32_t t = g - g0;
		const uint64_t off0 = a.offsets[r];
		uint64_t n = a.offsets[r + 1] - off0;
		const uint64_t sbase = a.in_offsets[r];
		const uint64_t nkeys = (n + VPK - 1) / VPK;
		const uint8_t *keys = a.in + sbase + (FRAME ? 4 : 0);
		const uint8_t *data = keys + nkeys;
		const uint8_t *limit = a.in + a.in_bytes;
		bool bad = sbase + (FRAME ? 4 : 0) + nkeys + n > a.in_bytes;	// at least 1 byte / value
		if (FRAME && !bad) {
			uint32_t nn = keys[-4] | (keys[-3] << 8) | (keys[-2] << 16) | ((uint32_t) keys[-1] << 24);
			bad = nn != (uint32_t) n;	// press.c:2062 takes the count from the prefix
		}
		if (bad) {
			if (tid == 0) {
				atomicOr(&a.hdr->error, ERR_MALFORMED);
				st_relaxed(a.status_bytes + g, LB_INC);
				st_relaxed(a.status_sum + g, LB_INC);
			}
			continue;
		}

		// ---- keys -> flags, byte counts ------------------------------------
		uint32_t fl[CHUNKS], nv[CHUNKS], cnt[CHUNKS];
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			const uint64_t j = (uint64_t) t * GROUPS + c * NT + tid;	// 8-sample block in the read
			int64_t left = (int64_t) n - (int64_t) (8 * j);
			nv[c] = left >= 8 ? 8 : (left > 0 ? (uint32_t) left : 0);
			uint32_t f = 0;
			if (nv[c]) {
				if (KEYBITS == 1) {
					f = keys[j];
				} else {
					uint32_t kk = keys[2 * j];
					if (nv[c] > 4)
						kk |= (uint32_t) keys[2 * j + 1] << 8;
					f = squeeze16(kk);
				}
				f &= 0xffu >> (8 - nv[c]);
			}
			fl[c] = f;
			cnt[c] = nv[c] + __popc(f);
		}
		uint32_t excl[CHUNKS], L;
		block_scan<CHUNKS>(cnt, excl, L, s_warp);

		// ---- look-back 1: data bytes of earlier tiles of this read ---------
		if (warp == 0) {
			uint64_t D = 0;
			if (t > 0) {
				if (lane == 0)
					st_relaxed(a.status_bytes + g, LB_AGG | L);
				D = lookback(a.status_bytes, g, g0);
			}
			if (lane == 0) {
				st_relaxed(a.status_bytes + g, LB_INC | (D + L));
				s_D = D;
			}
		}
		__syncthreads();
		const uint8_t *src = data + s_D;
		if (src + L > limit) {
			if (tid == 0) {
				atomicOr(&a.hdr->error, ERR_MALFORMED);
				st_relaxed(a.status_sum + g, LB_INC);
			}
			continue;
		}
		const uint32_t ph = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
		copy_in_aligned(s_buf, src, L, a.in, limit);
		__syncthreads();

		// ---- gather, unzigzag, local running sums ---------------------------
		uint32_t x[CHUNKS][8], tsum[CHUNKS];
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			const uint8_t *p = s_buf + ph + excl[c];
			uint32_t acc = 0;
#pragma unroll
			for (int e = 0; e < 8; e++) {
				uint32_t v = 0;
				if (e < (int) nv[c]) {
					v = *p++;
					if ((fl[c] >> e) & 1)
						v |= (uint32_t) (*p++) << 8;
				}
				uint32_t d = (v >> 1) ^ (0u - (v & 1u));	// unzigzag (decode_scalar.hpp:10-13)
				acc += d;
				x[c][e] = acc;
			}
			tsum[c] = acc & 0xffffu;
		}
		uint32_t vex[CHUNKS], S;
		block_scan<CHUNKS>(tsum, vex, S, s_warp);
		S &= 0xffffu;

		// ---- look-back 2: sum of deltas of earlier tiles of this read -------
		if (warp == 0) {
			uint64_t V = 0;
			if (t > 0) {
				if (lane == 0)
					st_relaxed(a.status_sum + g, LB_AGG | S);
				V = lookback(a.status_sum, g, g0);
			}
			if (lane == 0) {
				st_relaxed(a.status_sum + g, LB_INC | ((V + S) & 0xffffu));
				s_V = (uint32_t) V;
			}
		}
		__syncthreads();
		const uint32_t V = s_V;
		int16_t *dst = a.samples_out + off0 + (uint64_t) t * TILE;
		const uint32_t pho = (uint32_t) ((reinterpret_cast<uintptr_t>(dst) >> 1) & 7);
		const uint32_t ntile = (uint32_t) (n - (uint64_t) t * TILE < TILE ? n - (uint64_t) t * TILE : TILE);
#pragma unroll
		for (int c = 0; c < CHUNKS; c++) {
			const uint32_t i0 = (c * NT + tid) * 8;
			const uint32_t carry = V + vex[c];
#pragma unroll
			for (int e = 0; e < 8; e++)
				if (e < (int) nv[c])
					s_out[pho + i0 + e] = (uint16_t) (carry + x[c][e]);
		}
		__syncthreads();
		copy_out_samples(s_out, dst, ntile);
	}
}

}  // namespace pg
