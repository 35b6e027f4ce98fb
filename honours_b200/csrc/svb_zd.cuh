// svb_zd.cuh -- zigzag-delta and streamvbyte building blocks shared by the kernels.
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
// The kernels themselves live in svb_enc.cuh (encode) and svb_dec.cuh (decode).
#pragma once

#include "common.cuh"

namespace pg {

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
	uint64_t *status;		// per-tile data-byte counts, batch-wide (svb_enc.cuh)
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
	// whole 8-byte words by their lanes; the < 8 bytes of the first and of the last word by a lane each
	const uint32_t w0 = (a + 7) >> 3, w1 = end >> 3;
	if (w1 >= w0) {
		for (uint32_t q = w0 + lane; q < w1; q += 32)
			*reinterpret_cast<uint64_t *>(gbase + 8 * q) = *reinterpret_cast<const uint64_t *>(stage + 8 * q);
		const uint32_t nh = 8 * w0 - a, ntl = end - 8 * w1;
		if ((uint32_t) lane < nh)
			gbase[a + lane] = stage[a + lane];
		else if ((uint32_t) lane - 8 < ntl)
			gbase[8 * w1 + lane - 8] = stage[8 * w1 + lane - 8];
	} else if ((uint32_t) lane < len) {	// inside one word
		gbase[a + lane] = stage[a + lane];
	}
	(void) nwords;
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

}  // namespace pg
