// misc_codecs.cuh -- two further integer codecs on the zigzag-delta front end:
//
//   SVB0124_ZD  svb0124_zd_*_16, press.c:1616-1643 -> streamvbyte_0124_encode.c:8-59 (scalar form):
//               keys[ceil(n / 4)] (2-bit codes, value i at bit 2 (i & 3) of byte i >> 2: 0 = no byte
//               (value 0), 1 = one byte, 2 = two bytes, 3 = four bytes) || data bytes, over zd of
//               all n samples (the first against 0, trans.c:233-249)
//   UINT_ZD     uint_zd_*_16, press.c:639-687 -> uint_press_16 (489-507) -> uintx_press_core
//               (285-397): [first sample i16 LE][width u8] || the n - 1 values zd[1 ..] as `width`-bit
//               big-endian fields back to back (most significant bit first, zero padded to a byte),
//               width = uint_get_minbits(max zd) (press.c:461-471)
//
// Neither has a header whose length depends on the data, but both need one number per read before
// the first byte of the NEXT read can be placed (data-byte count / width), so encode is a size
// sweep, a scan, and the coding sweep.  One CTA per read (ticket), tiles of 4096 values cut at
// multiples of 4096 INSIDE the read (so 16 values per thread are 4 whole key bytes, and a tile of
// width-bit fields ends on a byte).  Global memory is touched by whole warps on neighbouring
// addresses only: samples come in through a padded shared-memory tile, streams leave (and enter)
// through a stage that mirrors the 16-byte words of the container.  Both payloads are written by
// the bit packer of the Huffman emitter (he_put, exfam.cuh): a svb0124 data section is the values
// themselves as "codes" of 0, 8 or 16 bits; a uint-zd payload is the bit-reversed fields as codes
// of `width` bits, with the bits of every byte reversed on the way out (most significant bit
// first, press.c:285-397).
#pragma once

#include "exfam.cuh"
#include "exfam_dec.cuh"

namespace pg {

constexpr int MT = 4096;			// values per tile, 16 per thread
constexpr int MC_XPAD = MT + 2 * (MT / 16);	// the tile's samples, two of padding after every 16
constexpr int MC_STAGE = 2 * MT + 64;		// stream bytes of a tile (<= 2 per value) + phase + slack

struct McArgs {
	int method;
	const int16_t *samples;		// encode in
	const uint64_t *offsets;	// [nreads + 1]
	uint64_t nreads;
	uint8_t *out;			// encode out
	uint64_t out_cap;
	uint64_t *out_offsets;		// [nreads + 1]
	uint64_t *r_size;		// per read: stream bytes
	uint32_t *r_aux;		// per read: UINT_ZD width
	Hdr *hdr;			// ticket / ticket2: next read of the size / coding sweep
	const uint8_t *in;		// decode in
	uint64_t in_bytes;
	const uint64_t *in_offsets;
	int16_t *samples_out;
};

__device__ __forceinline__ uint32_t mc_zigzag(int32_t d)	// trans.c:75-78 on the 16-bit difference
{
	const uint32_t x = (uint32_t) d & 0xffffu;
	return ((x << 1) ^ (0u - (x >> 15))) & 0xffffu;
}
__device__ __forceinline__ uint32_t mc_minbits(uint32_t mx)	// press.c:461-471
{
	return mx ? 32u - __clz(mx) : 0u;
}
// bit 15 / 31 = (low / high halfword != 0)
__device__ __forceinline__ uint32_t mc_nz16x2(uint32_t z)
{
	return (((z & 0x7fff7fffu) + 0x7fff7fffu) | z) & 0x80008000u;
}

// CTA-wide sum / max of one value per thread (two barriers)
__device__ __forceinline__ uint64_t mc_block_sum(uint64_t v, uint64_t *s_w)
{
	v = warp_sum64(v);
	__syncthreads();
	if ((threadIdx.x & 31) == 0)
		s_w[threadIdx.x >> 5] = v;
	__syncthreads();
	uint64_t t = 0;
#pragma unroll
	for (int w = 0; w < NW; w++)
		t += s_w[w];
	return t;
}

// size sweep: data bytes (SVB0124_ZD) / width (UINT_ZD) of every read -> r_size, r_aux.  The read's
// whole 16-byte sample groups are taken as 128-bit loads, eight zd values at a time on 16-bit pairs;
// the up to 8 samples in front of the first whole group and the up to 7 after the last one by a
// thread each.
__global__ void __launch_bounds__(NT) mc_size_kernel(const __grid_constant__ McArgs a)
{
	__shared__ uint64_t s_w[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
	const uint4 *s4 = reinterpret_cast<const uint4 *>(a.samples);
	const uint16_t *s16 = reinterpret_cast<const uint16_t *>(a.samples);
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t off0 = a.offsets[r], rend = a.offsets[r + 1], n = rend - off0;
		uint32_t cnt = 0, mx = 0;	// SVB0124: data bytes (of this thread: < 2^32); UINT_ZD: max value, both halves
		// one value: sample i of the read
		auto single = [&](uint64_t i) {
			const int32_t prev = i ? (int32_t) (int16_t) __ldg(s16 + off0 + i - 1) : 0;
			const uint32_t z = mc_zigzag((int32_t) (int16_t) __ldg(s16 + off0 + i) - prev);
			cnt += (z != 0) + (z >= 256u);
			if (i)
				mx = z > mx ? z : mx;
		};
		const uint64_t g0 = (off0 >> 3) + 1, g1 = rend >> 3;	// whole groups [g0, g1) past the read's first sample
		const uint64_t head = 8 * g0 - off0 < n ? 8 * g0 - off0 : n;
		if ((uint64_t) tid < head)
			single(tid);
		uint32_t mx2 = 0;
		for (uint64_t k = g0 + tid; k < g1; k += NT) {
			const uint4 x = __ldg(s4 + k);
			uint32_t z[4];
			zd_group(x, __ldg(s16 + 8 * k - 1), -1, z);
#pragma unroll
			for (int j = 0; j < 4; j++) {
				// (bits 15 / 31: value != 0; bits 14 / 30: value >= 256)
				cnt += __popc(mc_nz16x2(z[j]) | (mc_nz16x2(z[j] & 0xff00ff00u) >> 1));
				mx2 = __vmaxu2(mx2, z[j]);
			}
		}
		if (g1 >= g0) {
			const uint64_t t0 = 8 * g1 - off0;	// the samples after the last whole group
			if (t0 + tid < n && t0 >= head)
				single(t0 + tid);
		}
		mx2 = (mx2 & 0xffffu) > (mx2 >> 16) ? (mx2 & 0xffffu) : (mx2 >> 16);
		mx = mx2 > mx ? mx2 : mx;
		if (svb) {
			const uint64_t tot = mc_block_sum(cnt, s_w);
			if (tid == 0) {
				a.r_size[r] = (n + 3) / 4 + tot;
				a.r_aux[r] = 0;
			}
		} else {
			mx = __reduce_max_sync(0xffffffffu, mx);
			__syncthreads();
			if ((tid & 31) == 0)
				s_w[tid >> 5] = mx;
			__syncthreads();
#pragma unroll
			for (int w = 0; w < NW; w++)
				mx = s_w[w] > mx ? (uint32_t) s_w[w] : mx;
			const uint32_t bits = mc_minbits(mx);
			if (tid == 0) {
				// an empty read has no stream (the reference reads in[0] of nothing)
				a.r_size[r] = n ? 3 + ((n - 1) * bits + 7) / 8 : 0;
				a.r_aux[r] = bits;
			}
		}
	}
}

// The tile's samples base[t0 .. t0 + 4096) (zeros past nval) through shared memory: every warp loads
// 32 neighbouring samples at a time, every thread then finds its 16 in 8 words of its own (the
// padding puts the threads 9 words apart: no bank is met twice).  z: zd of the 16, two per word;
// prev_tile: the sample before the tile's first.  One barrier inside; the caller has one between
// this call's reads of s_x and the next call.  (Asking for the next tile's samples a tile ahead
// was measured slower: into registers 17.2, from L2 18.3, against 19.0 % of HBM peak for svb0124-zd encode.)
__device__ __forceinline__ void mc_tile_zd(const int16_t *base, uint64_t nval, uint64_t t0, uint32_t prev_tile, uint16_t *s_x,
					   uint32_t (&z)[8])
{
	const int tid = threadIdx.x;
	const uint16_t *b = reinterpret_cast<const uint16_t *>(base) + t0;
	const uint64_t left = nval - t0;
#pragma unroll
	for (int k = 0; k < MT / NT; k++) {
		const uint32_t i = k * NT + tid;
		s_x[i + 2 * (i >> 4)] = i < left ? __ldg(b + i) : (uint16_t) 0;
	}
	__syncthreads();
	const uint32_t *w = reinterpret_cast<const uint32_t *>(s_x) + 9 * tid;
	const uint4 xa = make_uint4(w[0], w[1], w[2], w[3]), xb = make_uint4(w[4], w[5], w[6], w[7]);
	const uint32_t prev = tid ? s_x[18 * tid - 3] : prev_tile;	// (sample 16 tid - 1 sits before two of padding)
	uint32_t za[4], zb[4];
	zd_group(xa, prev, -1, za);
	zd_group(xb, xa.w >> 16, -1, zb);
#pragma unroll
	for (int j = 0; j < 4; j++) {
		z[j] = za[j];
		z[4 + j] = zb[j];
	}
	// values past the end are no values
	const uint64_t j0 = 16ull * tid;
	if (j0 + 16 > left) {
		const uint32_t cnt = j0 < left ? (uint32_t) (left - j0) : 0u;
#pragma unroll
		for (int j = 0; j < 8; j++)
			z[j] &= (2 * j < (int) cnt ? 0xffffu : 0u) | (2 * j + 1 < (int) cnt ? 0xffff0000u : 0u);
	}
}

// The tile's stage -> the container.  The stage mirrors global memory from the 16-byte word g16 on;
// its bytes [fb, lb) are the tile's.  Whole words go out as they are, the others byte by byte (the
// neighbours' bytes belong to other tiles or reads).  REV: the bits of every byte are reversed on
// the way (uint-zd).  The stage is left zeroed.  One CTA.
template <bool REV>
__device__ __forceinline__ void mc_copy_out(uint8_t *s_stage, uint8_t *g16, uint32_t fb, uint32_t lb)
{
	const uint32_t nw = (lb + 15) >> 4;
	for (uint32_t w = threadIdx.x; w < nw; w += NT) {
		uint4 x = reinterpret_cast<const uint4 *>(s_stage)[w];
		reinterpret_cast<uint4 *>(s_stage)[w] = make_uint4(0, 0, 0, 0);
		if (REV) {
			x.x = __byte_perm(__brev(x.x), 0, 0x0123);
			x.y = __byte_perm(__brev(x.y), 0, 0x0123);
			x.z = __byte_perm(__brev(x.z), 0, 0x0123);
			x.w = __byte_perm(__brev(x.w), 0, 0x0123);
		}
		const uint32_t lo = w << 4, hi = lo + 16;
		if (lo >= fb && hi <= lb) {
			reinterpret_cast<uint4 *>(g16)[w] = x;
		} else {
			const uint32_t b0 = lo < fb ? fb : lo, b1 = hi > lb ? lb : hi;
			for (uint32_t b = b0; b < b1; b++) {
				const uint32_t j = b & 15, v = j < 4 ? x.x : j < 8 ? x.y : j < 12 ? x.z : x.w;
				g16[b] = (uint8_t) (v >> (8 * (j & 3)));
			}
		}
	}
}

// The thread's 16 codes (e[]: code << 8 | bits, see he_put) into the zeroed stage at bit bp.  A warp
// whose threads all bring a word's worth of bits passes the shared words along by shuffle; any
// other warp, and every warp at its two ends, ORs them in.
__device__ __forceinline__ void mc_put(const uint32_t (&e)[16], uint32_t bits, uint32_t bp, uint32_t stage_sa)
{
	const int lane = threadIdx.x & 31;
	const uint32_t fw = stage_sa + ((bp >> 3) & ~3u);
	uint32_t sa = fw, first;
	bool got;
	const uint32_t left = he_put<true>(e, 0u, bp, sa, first, got);
	if (__all_sync(0xffffffffu, bits >= 32)) {
		const uint32_t before = __shfl_up_sync(0xffffffffu, left, 1);
		if (lane)
			asm volatile("st.shared.u32 [%0], %1;" ::"r"(fw), "r"(first | before) : "memory");
		else
			sred_or(fw, first);
		if (lane == 31)
			sred_or(sa, left);
	} else {
		if (got)
			sred_or(fw, first);
		if (left)
			sred_or(sa, left);
	}
}

// coding sweep
#ifndef PG_MC_CTAS
#define PG_MC_CTAS 4
#endif
__global__ void __launch_bounds__(NT, PG_MC_CTAS) mc_encode_kernel(const __grid_constant__ McArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[MC_STAGE];
	__shared__ __align__(16) uint16_t s_x[MC_XPAD];
	__shared__ __align__(16) uint32_t s_keys[NT + 8];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
	if (a.hdr->error & ERR_CAPACITY)
		return;
	for (int w = tid; w < MC_STAGE / 16; w += NT)
		reinterpret_cast<uint4 *>(s_stage)[w] = make_uint4(0, 0, 0, 0);
	const uint32_t stage_sa = (uint32_t) __cvta_generic_to_shared(s_stage);
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket2, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t off0 = a.offsets[r], n = a.offsets[r + 1] - off0;
		if (n == 0)
			continue;
		const int16_t *s = a.samples + off0;
		uint8_t *dst = a.out + a.out_offsets[r];
		if (svb) {
			// streamvbyte_0124_encode.c:8-59: keys, then the values' bytes
			const uint64_t nk = (n + 3) / 4;
			uint8_t *data = dst + nk;
			uint64_t D = 0;		// data bytes written so far
			for (uint64_t t0 = 0; t0 < n; t0 += MT) {
				uint32_t z[8], e[16];
				mc_tile_zd(s, n, t0, t0 ? (uint32_t) (uint16_t) __ldg(s + t0 - 1) : 0u, s_x, z);
				uint32_t keys = 0;
#pragma unroll
				for (int j = 0; j < 8; j++) {
					// two values per word: code = (v != 0) + (v >= 256) (zd < 65536: never four bytes), as
					// bit 0 = one byte, bit 1 = two bytes, picked from bits 15 / 31 of the two tests
					const uint32_t t2 = mc_nz16x2(z[j] & 0xff00ff00u), t1 = mc_nz16x2(z[j]) & ~t2;
					keys |= (((t1 >> 15) & 1u) | ((t2 >> 14) & 2u) | ((t1 >> 29) & 4u) | ((t2 >> 28) & 8u)) << (4 * j);
					e[2 * j] = ((z[j] << 8) & 0xffff00u) | ((t1 >> 12) & 8u) | ((t2 >> 11) & 16u);
					e[2 * j + 1] = ((z[j] >> 8) & 0xffff00u) | ((t1 >> 28) & 8u) | ((t2 >> 27) & 16u);
				}
				s_keys[tid] = keys;
				const uint32_t nb = __popc(keys & 0x55555555u) + 2 * __popc(keys & 0xaaaaaaaau);
				uint32_t c1[1] = {nb}, e1[1], tot;
				block_scan<1>(c1, e1, tot, s_warp);
				uint8_t *g = data + D;
				const uint32_t s0 = (uint32_t) (reinterpret_cast<uintptr_t>(g) & 15);
				mc_put(e, 8 * nb, 8 * (s0 + e1[0]), stage_sa);
				__syncthreads();
				mc_copy_out<false>(s_stage, g - s0, s0, s0 + tot);
				const uint64_t left = n - t0;
				block_copy_out(reinterpret_cast<const uint8_t *>(s_keys), dst + t0 / 4,
					       (uint32_t) (((left < MT ? left : MT) + 3) / 4));	// (unused high bits of the last key byte: 0, :56)
				D += tot;
				__syncthreads();
			}
		} else {
			const uint32_t bits = a.r_aux[r];
			if (tid == 0) {	// press.c:661-672 (first sample), 489-507 (width byte)
				dst[0] = (uint8_t) (uint16_t) s[0];
				dst[1] = (uint8_t) ((uint16_t) s[0] >> 8);
				dst[2] = (uint8_t) bits;
			}
			uint8_t *pk = dst + 3;
			const uint64_t m = n - 1;	// values: zd of samples 1 ..
			for (uint64_t t0 = 0; t0 < m && bits; t0 += MT) {
				uint32_t z[8], e[16];
				mc_tile_zd(s + 1, m, t0, (uint32_t) (uint16_t) __ldg(s + t0), s_x, z);
				const uint64_t left = m - t0;
				const uint32_t cntt = left < MT ? (uint32_t) left : (uint32_t) MT;
				const uint32_t cnt = 16u * tid >= cntt ? 0u : (cntt - 16u * tid < 16u ? cntt - 16u * tid : 16u);
#pragma unroll
				for (int i = 0; i < 16; i++) {
					const uint32_t v = (z[i >> 1] >> (16 * (i & 1))) & 0xffffu;
					// the field's bits in stream order, first (most significant) bit lowest
					e[i] = (uint32_t) i < cnt ? ((__brev(v) >> (32 - bits)) << 8) | bits : 0u;
				}
				uint8_t *g = pk + t0 / 8 * bits;	// (4096 fields end on a byte)
				const uint32_t s0 = (uint32_t) (reinterpret_cast<uintptr_t>(g) & 15);
				mc_put(e, cnt * bits, 8 * s0 + 16u * bits * tid, stage_sa);
				__syncthreads();
				// the run's last byte is zero padded (press.c:386-390)
				mc_copy_out<true>(s_stage, g - s0, s0, s0 + (cntt * bits + 7) / 8);
				__syncthreads();
			}
		}
	}
}

// ---------------------------------------------------------------------------
// decode: stream -> samples (press.c:1631-1643 / 674-687: values, un-zigzag, running sum mod 2^16)

// stream bytes [g, g + len) -> the stage, from the 16-byte word holding g on (whole words, guarded at
// the container's ends).  REV: the bits of every byte reversed (uint-zd: fields then read LSB first).
template <bool REV>
__device__ __forceinline__ void mc_stage_in(uint8_t *s_stage, const uint8_t *g, uint32_t len, const uint8_t *lim_lo,
					    const uint8_t *lim_hi)
{
	const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(g) & 15);
	const uint8_t *b16 = g - al;
	const uint32_t nw = (al + len + 15) >> 4;
	for (uint32_t w = threadIdx.x; w < nw + 1; w += NT) {	// (+ 1: the windows read a word past the last byte)
		uint4 x = w < nw ? load_word_guarded(b16, w, lim_lo, lim_hi) : make_uint4(0, 0, 0, 0);
		if (REV) {
			x.x = __byte_perm(__brev(x.x), 0, 0x0123);
			x.y = __byte_perm(__brev(x.y), 0, 0x0123);
			x.z = __byte_perm(__brev(x.z), 0, 0x0123);
			x.w = __byte_perm(__brev(x.w), 0, 0x0123);
		}
		reinterpret_cast<uint4 *>(s_stage)[w] = x;
	}
}

// 64 stream bits from stage bit `pos` on
__device__ __forceinline__ void mc_window(uint32_t stage_sa, uint32_t pos, uint32_t &lo, uint32_t &hi)
{
	const uint32_t a = stage_sa + ((pos >> 3) & ~3u);
	uint32_t w0, w1, w2;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a));
	asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(a));
	asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(a));
	lo = __funnelshift_r(w0, w1, pos);	// (shift taken mod 32)
	hi = __funnelshift_r(w1, w2, pos);
}

__global__ void __launch_bounds__(NT, PG_MC_CTAS) mc_decode_kernel(const __grid_constant__ McArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[MC_STAGE + 16];
	__shared__ __align__(16) uint32_t s_out[MT / 2 + 8];	// the tile's samples on their way out
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ __align__(16) uint32_t s_warp2[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
	const uint32_t stage_sa = (uint32_t) __cvta_generic_to_shared(s_stage);
	const uint8_t *lim_hi = a.in + a.in_bytes;
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t off0 = a.offsets[r], n = a.offsets[r + 1] - off0;
		if (n == 0)
			continue;
		const uint64_t s0 = a.in_offsets[r], s1 = a.in_offsets[r + 1];
		const uint64_t nk = (n + 3) / 4;
		bool bad = s1 < s0 || s1 > a.in_bytes || s1 - s0 < (svb ? nk : 3);
		const uint8_t *src = a.in + s0;
		uint32_t bits = 0;
		uint32_t carry = 0;	// running sum so far (mod 2^16)
		if (!bad && !svb) {
			bits = src[2];
			bad = bits > 16 || s1 - s0 < 3 + ((n - 1) * bits + 7) / 8;
			carry = (uint32_t) src[0] | ((uint32_t) src[1] << 8);
		}
		if (bad) {
			if (tid == 0)
				atomicOr(&a.hdr->error, ERR_MALFORMED);
			continue;
		}
		int16_t *dst = a.samples_out + off0;
		const uint64_t ndata = s1 - s0 - (svb ? nk : 3);
		const uint8_t *data = src + (svb ? nk : 3);
		if (!svb && tid == 0)
			dst[0] = (int16_t) carry;	// press.c:679: the first sample as stored
		const uint64_t nval = svb ? n : n - 1;	// coded values; UINT_ZD value j is sample j + 1
		uint64_t D = 0;		// SVB0124: data bytes consumed
		bool overrun = false;
		// SVB0124: the thread's 4 key bytes of the tile at t0 (neighbouring threads, neighbouring bytes),
		// asked for one tile ahead
		auto load_keys = [&](uint64_t t0) {
			const uint64_t j0 = t0 + 16 * tid;
			const uint32_t cnt = j0 >= nval ? 0u : (nval - j0 < 16 ? (uint32_t) (nval - j0) : 16u);
			uint32_t k = 0;
			for (uint32_t q = 0; 4 * q < cnt; q++)
				k |= (uint32_t) __ldg(src + j0 / 4 + q) << (8 * q);
			return cnt < 16 ? k & ((1u << (2 * cnt)) - 1) : k;
		};
		uint32_t keys_next = svb ? load_keys(0) : 0u;
		for (uint64_t t0 = 0; t0 < nval && !overrun; t0 += MT) {
			const uint64_t j0 = t0 + 16 * tid;
			const uint32_t cntt = nval - t0 < MT ? (uint32_t) (nval - t0) : (uint32_t) MT;
			const uint32_t cnt = j0 >= nval ? 0u : (nval - j0 < 16 ? (uint32_t) (nval - j0) : 16u);
			uint32_t v[16];
			if (svb) {
				const uint32_t keys = keys_next;
				if (t0 + MT < nval)
					keys_next = load_keys(t0 + MT);
				// streamvbyte_0124_decode: 0, 1, 2 or 4 bytes
				const uint32_t k3 = keys & (keys >> 1) & 0x55555555u;
				const uint32_t nb = __popc(keys & 0x55555555u) + 2 * __popc(keys & 0xaaaaaaaau) + __popc(k3);
				uint32_t c1[1] = {nb}, e1[1], tot;
				block_scan<1>(c1, e1, tot, s_warp);
				if (D + tot > ndata) {	// (the same in every thread)
					overrun = true;
					break;
				}
				const uint8_t *g = data + D;
				if (g + tot + 32 * tid + 32 <= lim_hi)	// (where the next tile's bytes will be, near enough: asked from L2)
					prefetch_l2(g + tot + 32 * tid);
				if (tot > 2u * MT) {
					// four-byte values (never written for a 16-bit signal, but a valid stream): more than the
					// stage holds, so byte by byte from the container
					const uint8_t *p = g + e1[0];
#pragma unroll 1
					for (int e = 0; e < 16; e++) {
						const uint32_t code = (keys >> (2 * e)) & 3u, len = code == 3 ? 4u : code;
						uint32_t x = 0;
						for (uint32_t b = 0; b < len; b++)
							x |= (uint32_t) p[b] << (8 * b);
						p += len;
#pragma unroll
						for (int k = 0; k < 16; k++)	// (static indices: v stays in registers)
							if (k == e)
								v[k] = x;
					}
				} else {
					mc_stage_in<false>(s_stage, g, tot, a.in, lim_hi);
					__syncthreads();
					uint32_t pos = 8 * ((uint32_t) (reinterpret_cast<uintptr_t>(g) & 15) + e1[0]), lo = 0, hi = 0;
#pragma unroll
					for (int e = 0; e < 16; e++) {
						if (e % 2 == 0)
							mc_window(stage_sa, pos, lo, hi);
						const uint32_t code = (keys >> (2 * e)) & 3u;
						const uint32_t l = code == 3 ? 32u : 8 * code;
						v[e] = code == 3 ? lo : lo & ((1u << l) - 1);
						if (code == 3) {
							lo = hi;
							hi = 0;
						} else {
							lo = __funnelshift_r(lo, hi, l);
							hi >>= l;
						}
						pos += l;
					}
				}
				D += tot;
			} else {
				// 16 fields of `bits` bits from bit 16 * bits * (j0 / 16) of the payload on, most significant bit first
				const uint8_t *g = data + t0 / 8 * bits;
				if (g + 512 * bits + 32 * tid + 32 <= lim_hi)
					prefetch_l2(g + 512 * bits + 32 * tid);
				mc_stage_in<true>(s_stage, g, (cntt * bits + 7) / 8, a.in, lim_hi);
				__syncthreads();
				uint32_t pos = 8 * (uint32_t) (reinterpret_cast<uintptr_t>(g) & 15) + 16u * bits * tid, lo = 0, hi = 0;
#pragma unroll
				for (int e = 0; e < 16; e++) {
					if (e % 2 == 0)
						mc_window(stage_sa, pos, lo, hi);
					const uint32_t f = bits ? __brev(lo) >> (32 - bits) : 0u;	// (the field's first bit is its most significant)
					v[e] = (uint32_t) e < cnt ? f : 0u;
					lo = __funnelshift_r(lo, hi, bits);
					hi >>= bits;
					pos += bits;
				}
			}
			// un-zigzag the values' low 16 bits (trans.c:272-282 hands every value to unzigzag_one_16, which
			// takes a uint16_t: a four-byte value counts by its low half), running sum inside the thread on
			// 16-bit pairs, then over the CTA
			uint32_t y[8], cs = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) {
				const uint32_t d = unzigzag16x2((v[2 * j] & 0xffffu) | (v[2 * j + 1] << 16));	// (values past cnt are 0)
				y[j] = add16x2(d * 0x00010001u, cs);
				cs = __byte_perm(y[j], 0, 0x3232);
			}
			uint32_t c1[1] = {cs & 0xffffu}, e1[1], tot;
			block_scan<1>(c1, e1, tot, s_warp2);
			const uint32_t base = ((carry + e1[0]) & 0xffffu) * 0x00010001u;
			uint32_t o[8];
#pragma unroll
			for (int j = 0; j < 8; j++)
				o[j] = add16x2(y[j], base);
			reinterpret_cast<uint4 *>(s_out)[2 * tid] = make_uint4(o[0], o[1], o[2], o[3]);
			reinterpret_cast<uint4 *>(s_out)[2 * tid + 1] = make_uint4(o[4], o[5], o[6], o[7]);
			carry = (carry + tot) & 0xffffu;
			__syncthreads();
			block_copy_out(reinterpret_cast<const uint8_t *>(s_out), reinterpret_cast<uint8_t *>(dst + (svb ? t0 : t0 + 1)), 2 * cntt);
			__syncthreads();	// s_stage / s_out / s_warp / s_warp2 are written again by the next tile
		}
		if (overrun && tid == 0)
			atomicOr(&a.hdr->error, ERR_MALFORMED);
	}
}

}  // namespace pg
