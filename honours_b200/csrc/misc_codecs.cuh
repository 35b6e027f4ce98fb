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
// sweep, a scan, and the coding sweep.  These are coverage methods, written for clarity: one CTA
// per read (ticket), tiles of 4096 values cut at multiples of 4096 INSIDE the read (so 16 values
// per thread are 4 whole key bytes / 2 x width whole bytes), 16-bit sample loads.
#pragma once

#include "exfam.cuh"

namespace pg {

constexpr int MT = 4096;	// values per tile, 16 per thread

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

// size sweep: data bytes (SVB0124_ZD) / width (UINT_ZD) of every read -> r_size, r_aux
__global__ void __launch_bounds__(NT) mc_size_kernel(const __grid_constant__ McArgs a)
{
	__shared__ uint64_t s_w[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t off0 = a.offsets[r], n = a.offsets[r + 1] - off0;
		const int16_t *s = a.samples + off0;
		uint64_t acc = 0;	// SVB0124: data bytes; UINT_ZD: max value
		for (uint64_t i = tid; i < n; i += NT) {
			const int32_t prev = i ? (int32_t) __ldg(s + i - 1) : 0;
			const uint32_t z = mc_zigzag((int32_t) __ldg(s + i) - prev);
			if (svb)
				acc += (z != 0) + (z >= 256u);
			else if (i)
				acc = z > acc ? z : acc;
		}
		if (svb) {
			const uint64_t tot = mc_block_sum(acc, s_w);
			if (tid == 0) {
				a.r_size[r] = (n + 3) / 4 + tot;
				a.r_aux[r] = 0;
			}
		} else {
			uint32_t mx = __reduce_max_sync(0xffffffffu, (uint32_t) acc);
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

// the thread's 16 values of tile t: zd of samples i0 .. i0 + 15 (UINT_ZD: of samples i0 + 1 ..), 0 past the end
__device__ __forceinline__ void mc_load16(const int16_t *s, uint64_t n, uint64_t i0, bool uintzd, uint32_t (&z)[16],
					   uint32_t &cnt)
{
	const uint64_t first = uintzd ? i0 + 1 : i0;	// sample index of value 0
	int32_t prev = first && first - 1 < n ? (int32_t) __ldg(s + first - 1) : 0;
	cnt = 0;
#pragma unroll
	for (int e = 0; e < 16; e++) {
		z[e] = 0;
		if (first + e < n) {
			const int32_t x = (int32_t) __ldg(s + first + e);
			z[e] = mc_zigzag(x - prev);
			prev = x;
			cnt = e + 1;
		}
	}
}

// coding sweep
__global__ void __launch_bounds__(NT) mc_encode_kernel(const __grid_constant__ McArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[2 * MT + 64];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
	if (a.hdr->error & ERR_CAPACITY)
		return;
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
			const uint64_t nk = (n + 3) / 4;
			uint8_t *data = dst + nk;
			uint64_t D = 0;		// data bytes written so far
			for (uint64_t t0 = 0; t0 < n; t0 += MT) {
				const uint64_t i0 = t0 + 16 * tid;
				uint32_t z[16], cnt;
				mc_load16(s, n, i0, false, z, cnt);
				uint32_t nb = 0, keys = 0;
#pragma unroll
				for (int e = 0; e < 16; e++) {
					const uint32_t code = (uint32_t) e < cnt ? (z[e] != 0) + (z[e] >= 256u) : 0u;	// streamvbyte_0124_encode.c:8-29
					nb += code;
					keys |= code << (2 * e);
				}
				uint32_t c1[1] = {nb}, e1[1], tot;
				block_scan<1>(c1, e1, tot, s_warp);
				uint8_t *p = s_stage + e1[0];
#pragma unroll
				for (int e = 0; e < 16; e++) {
					if ((uint32_t) e < cnt && z[e]) {
						*p++ = (uint8_t) z[e];
						if (z[e] >= 256u)
							*p++ = (uint8_t) (z[e] >> 8);
					}
				}
				for (uint32_t q = 0; 4 * q < cnt; q++)	// key byte i >> 2; unused high bits 0 (:56)
					dst[i0 / 4 + q] = (uint8_t) (keys >> (8 * q));
				__syncthreads();
				block_copy_out(s_stage, data + D, tot);
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
			const uint64_t m = n - 1;	// values
			for (uint64_t t0 = 0; t0 < m && bits; t0 += MT) {
				const uint64_t j0 = t0 + 16 * tid;
				if (j0 >= m)
					continue;
				uint32_t z[16], cnt;
				mc_load16(s, n, j0, true, z, cnt);
				// 16 fields of `bits` bits, most significant bit first: 2 x bits whole bytes
				uint8_t *p = pk + j0 / 16 * (2ull * bits);
				uint64_t acc = 0;
				uint32_t have = 0;
#pragma unroll
				for (int e = 0; e < 16; e++) {
					if ((uint32_t) e < cnt) {
						acc = (acc << bits) | z[e];
						have += bits;
						while (have >= 8) {
							*p++ = (uint8_t) (acc >> (have - 8));
							have -= 8;
						}
					}
				}
				if (have)	// the run's last byte, zero padded (press.c:386-390)
					*p = (uint8_t) (acc << (8 - have));
			}
		}
	}
}

// decode: stream -> samples (press.c:1631-1643 / 674-687: values, un-zigzag, running sum mod 2^16)
__global__ void __launch_bounds__(NT) mc_decode_kernel(const __grid_constant__ McArgs a)
{
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ __align__(16) uint32_t s_warp2[NW];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const bool svb = a.method == PRESS_GPU_SVB0124_ZD;
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
		for (uint64_t t0 = 0; t0 < nval && !overrun; t0 += MT) {
			const uint64_t j0 = t0 + 16 * tid;
			const uint32_t cnt = j0 >= nval ? 0u : (nval - j0 < 16 ? (uint32_t) (nval - j0) : 16u);
			uint32_t v[16];
			if (svb) {
				uint32_t keys = 0, nb = 0;
				for (uint32_t q = 0; 4 * q < cnt; q++)
					keys |= (uint32_t) src[j0 / 4 + q] << (8 * q);
#pragma unroll
				for (int e = 0; e < 16; e++) {
					const uint32_t code = (uint32_t) e < cnt ? (keys >> (2 * e)) & 3u : 0u;
					nb += code == 3 ? 4u : code;	// streamvbyte_0124_decode: 0, 1, 2 or 4 bytes
				}
				uint32_t c1[1] = {nb}, e1[1], tot;
				block_scan<1>(c1, e1, tot, s_warp);
				if (D + tot > ndata) {	// (the same in every thread)
					overrun = true;
					break;
				}
				const uint8_t *p = data + D + e1[0];
#pragma unroll
				for (int e = 0; e < 16; e++) {
					const uint32_t code = (uint32_t) e < cnt ? (keys >> (2 * e)) & 3u : 0u;
					uint32_t x = 0;
					const uint32_t len = code == 3 ? 4u : code;
					for (uint32_t b = 0; b < len; b++)
						x |= (uint32_t) p[b] << (8 * b);
					p += len;
					v[e] = x;
				}
				D += tot;
			} else {
				// 16 fields of `bits` bits from byte 2 * bits * (j0 / 16) on, most significant bit first
				const uint8_t *p = data + j0 / 16 * (2ull * bits);
				uint64_t acc = 0;
				uint32_t have = 0;
#pragma unroll
				for (int e = 0; e < 16; e++) {
					v[e] = 0;
					if ((uint32_t) e < cnt && bits) {
						while (have < bits) {
							acc = (acc << 8) | *p++;
							have += 8;
						}
						v[e] = (uint32_t) (acc >> (have - bits)) & ((1u << bits) - 1);
						have -= bits;
					}
				}
			}
			// un-zigzag (on 32 bits, trans.c:272-282 / 16 bits, :251-258; the sum is kept mod 2^16)
			uint32_t run = 0, y[16];
#pragma unroll
			for (int e = 0; e < 16; e++) {
				const uint32_t d = (v[e] >> 1) ^ (0u - (v[e] & 1u));
				run = (run + ((uint32_t) e < cnt ? d : 0u)) & 0xffffu;
				y[e] = run;
			}
			uint32_t c1[1] = {run}, e1[1], tot;
			block_scan<1>(c1, e1, tot, s_warp2);
			const uint32_t base = carry + e1[0];
			int16_t *o = dst + (svb ? j0 : j0 + 1);
#pragma unroll
			for (int e = 0; e < 16; e++)
				if ((uint32_t) e < cnt)
					o[e] = (int16_t) (uint16_t) (base + y[e]);
			carry = (carry + tot) & 0xffffu;
			__syncthreads();	// s_warp / s_warp2 are written again by the next tile
		}
		if (overrun && tid == 0)
			atomicOr(&a.hdr->error, ERR_MALFORMED);
	}
}

}  // namespace pg
