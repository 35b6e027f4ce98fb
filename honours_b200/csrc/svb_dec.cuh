// svb_dec.cuh -- fused streamvbyte + inverse zigzag-delta DECODE, one pass over HBM.
//
//   KEYBITS = 1 : ONT svb16 (svb16/decode_scalar.hpp:31-71; press.c:1689-1694)
//   KEYBITS = 2 : classical streamvbyte, codes 0/1 only (streamvbyte_decode.c:16-58; press.c:1600-1612)
//   FRAME       : [n u32 LE] prefix of the pre-zstd VBZ stream (press.c:1887-1910, 2047-2070)
//
// The reference's decoder carries two values from one sample to the next: the
// data pointer (advanced by 1 + key bit) and the running sum of the deltas.  Here
// one CTA owns one READ (reads are handed out by an atomic ticket) and walks it
// tile by tile (4096 samples), carrying both values itself, so no tile ever waits
// for another CTA and nothing is read twice:
//
//   keys of tile t+2   -> registers        (their address does not depend on the data)
//   keys of tile t+1   -> byte counts -> block scan -> the tile's data range is known
//   data of tile t+1   -> registers        (aligned 128-bit loads, issued before ...)
//   tile t             -> smem stage -> values -> unzigzag -> block scan + carry ->
//                         aligned 128-bit stores (tiles are cut on 16-byte boundaries
//                         of the output array)
//
// so the loads of a tile are in flight for a whole tile time before they are
// needed.  1184 reads are in flight on a B200 (8 CTAs per SM).
#pragma once

#include "common.cuh"
#include "exfam_dec.cuh"

namespace pg {

constexpr int DT = 4096;		// samples per tile
constexpr int DPT = DT / NT;		// samples per thread (16)

struct SvbDec2Args {
	const uint8_t *in;		// container, 16-byte aligned
	uint64_t in_bytes;
	const uint64_t *in_offsets;	// [nreads + 1]
	const uint64_t *offsets;	// [nreads + 1]
	uint64_t nreads;
	int16_t *samples_out;
	Hdr *hdr;			// ticket = next read
};

// Key bits of the 16 values with read indices i0 .. i0 + 15 (i0 > -8), requested as two
// aligned 32-bit words and turned into "two bytes" flags one tile later (so nothing waits
// on the loads).  svb16: bit i & 7 of key byte i >> 3; svb32: 2-bit code at bit 2 (i & 3)
// of key byte i >> 2, codes 0/1 only (zd < 65536).  Bits of indices < 0 read as 0, bits
// past the key section are garbage and masked by the caller's valid mask.
struct KeyRaw {
	uint32_t w0, w1, sh, skip;
};

// interior tiles of a read whose key section ends >= 8 bytes before the container does:
// i0 >= 0 and both words are readable, so no guard is needed
template <int KEYBITS>
__device__ __forceinline__ KeyRaw dec_keys_issue_fast(const uint8_t *keys, uint64_t i0)
{
	KeyRaw k;
	k.skip = 0;
	const uint8_t *p = keys + (KEYBITS == 1 ? i0 >> 3 : i0 >> 2);
	const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(p) & 3);
	const uint32_t *pw = reinterpret_cast<const uint32_t *>(p - al);
	k.sh = 8 * al + (KEYBITS == 1 ? (uint32_t) (i0 & 7) : 2 * (uint32_t) (i0 & 3));
	k.w0 = __ldg(pw);
	k.w1 = __ldg(pw + 1);
	return k;
}

template <int KEYBITS>
__device__ __forceinline__ KeyRaw dec_keys_issue(const uint8_t *keys, int64_t i0, const uint8_t *lo, const uint8_t *hi)
{
	KeyRaw k;
	k.skip = i0 < 0 ? (uint32_t) (-i0) : 0u;
	const uint64_t ic = i0 < 0 ? 0 : (uint64_t) i0;
	const uint8_t *p = keys + (KEYBITS == 1 ? ic >> 3 : ic >> 2);
	const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(p) & 3);
	const uint8_t *pw = p - al;
	k.sh = 8 * al + (KEYBITS == 1 ? (uint32_t) (ic & 7) : 2 * (uint32_t) (ic & 3));
	if (pw >= lo && pw + 8 <= hi) {
		k.w0 = __ldg(reinterpret_cast<const uint32_t *>(pw));
		k.w1 = __ldg(reinterpret_cast<const uint32_t *>(pw) + 1);
	} else {	// the last bytes of the container
		k.w0 = k.w1 = 0;
		for (int b = 0; b < 8; b++)
			if (pw + b >= lo && pw + b < hi)
				(b < 4 ? k.w0 : k.w1) |= (uint32_t) pw[b] << (8 * (b & 3));
	}
	return k;
}

template <int KEYBITS>
__device__ __forceinline__ uint32_t dec_keys_flags(const KeyRaw &k)
{
	const uint32_t c = __funnelshift_r(k.w0, k.w1, k.sh);	// sh <= 31
	if (KEYBITS == 1)
		return (c << k.skip) & 0xffffu;
	return ((squeeze16(c) | (squeeze16(c >> 16) << 8)) << k.skip) & 0xffffu;
}

#ifndef PG_DEC_CTAS
#define PG_DEC_CTAS 5
#endif
template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(NT, PG_DEC_CTAS) svb_dec_kernel(const __grid_constant__ SvbDec2Args a)
{
	// the tile's data bytes, double buffered: tile t + 1 arrives by cp.async (no registers held,
	// no stage stores) while tile t is decoded
	__shared__ __align__(16) uint8_t s_in2[2][2 * DT + 64];
	__shared__ __align__(16) uint32_t s_warp[NW];	// byte-count scans
	__shared__ __align__(16) uint32_t s_warp2[NW];	// value scans
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x;
	const uint8_t *limit = a.in + a.in_bytes;
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t off0 = __ldg(a.offsets + r), n = __ldg(a.offsets + r + 1) - off0;
		const uint64_t sbase = __ldg(a.in_offsets + r), send = __ldg(a.in_offsets + r + 1);
		const uint64_t nkeys = KEYBITS == 1 ? (n + 7) / 8 : (n + 3) / 4;
		const uint64_t fixed = (FRAME ? 4 : 0) + nkeys;
		bool bad = send < sbase || send > a.in_bytes || send - sbase < fixed + n;	// >= 1 byte per value
		if (FRAME && !bad && n) {	// press.c:2062 takes the count from the prefix
			const uint8_t *p = a.in + sbase;
			bad = (p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t) p[3] << 24)) != (uint32_t) n;
		}
		if (bad) {
			if (tid == 0)
				atomicOr(&a.hdr->error, ERR_MALFORMED);
			continue;
		}
		if (n == 0)
			continue;
		const uint8_t *keys = a.in + sbase + (FRAME ? 4 : 0);
		const uint8_t *data = keys + nkeys;
		const uint64_t ndata = send - sbase - fixed;
		// tiles are cut on 16-byte boundaries of the OUTPUT (aligned 128-bit stores): element e
		// of thread tid in tile t is read index t * DT + 16 tid + e - m
		const uint32_t m = (uint32_t) (off0 & 7);
		const uint64_t A = off0 >> 3;
		const uint32_t ntiles = (uint32_t) ((m + n + DT - 1) / DT);
		uint4 *out4 = reinterpret_cast<uint4 *>(a.samples_out);
		uint16_t *out16 = reinterpret_cast<uint16_t *>(a.samples_out);

		auto first_of = [&](uint32_t t) -> int64_t { return (int64_t) t * DT + tid * DPT - (int64_t) m; };
		// tiles [t_lo, t_hi) are interior: every element is a sample of the read
		const uint32_t t_lo = m ? 1u : 0u, t_hi = (uint32_t) ((m + n) / DT);
		auto interior_of = [&](uint32_t t) -> bool { return t >= t_lo && t < t_hi; };
		auto valid_of = [&](uint32_t t) -> uint32_t {
			if (interior_of(t))
				return 0xffffu;
			const int64_t i0 = first_of(t), left = (int64_t) n - i0;
			const int lo = i0 < 0 ? (int) (-i0 < 16 ? -i0 : 16) : 0;
			const int hi = left >= 16 ? 16 : (left > 0 ? (int) left : 0);
			return hi > lo ? (0xffffu >> (16 - hi)) & (0xffffu << lo) & 0xffffu : 0u;
		};
		// data bytes of a tile -> stage `buf`: 16-byte words u, u + 256, (u + 512)
		auto issue_data = [&](uint32_t buf, uint64_t D, uint32_t L) {
			const uint8_t *src = data + D;
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
			const uint8_t *b16 = src - al;
			const uint32_t nw = (al + L + 15) >> 4;
			uint4 *dst = reinterpret_cast<uint4 *>(s_in2[buf]);
			if (b16 + 16ull * nw <= limit) {	// (b16 >= a.in: the container is 16-byte aligned)
#pragma unroll
				for (int j = 0; j < 3; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						cp_async16(dst + w, reinterpret_cast<const uint4 *>(b16) + w);
				}
			} else {
#pragma unroll
				for (int j = 0; j < 3; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						dst[w] = load_word_guarded(b16, w, a.in, limit);
				}
			}
			cp_async_commit();
		};

		const bool keys_safe = keys + nkeys + 8 <= limit;
		auto keys_of = [&](uint32_t t) -> KeyRaw {
			if (keys_safe && t >= 1)	// (tile 0 may start before the read: i0 < 0)
				return dec_keys_issue_fast<KEYBITS>(keys, (uint64_t) first_of(t));
			return dec_keys_issue<KEYBITS>(keys, first_of(t), a.in, limit);
		};
		// ---- prologue: tile 0 scanned and its data requested, keys of tile 1 requested
		uint32_t va_cur = valid_of(0);
		uint32_t fl_cur = dec_keys_flags<KEYBITS>(dec_keys_issue<KEYBITS>(keys, first_of(0), a.in, limit)) & va_cur;
		KeyRaw k_next = {0, 0, 0, 0};
		if (ntiles > 1)
			k_next = keys_of(1);
		uint32_t ex_cur, L_cur;
		{
			uint32_t c1[1] = {(uint32_t) (__popc(va_cur) + __popc(fl_cur))}, e1[1];
			block_scan<1>(c1, e1, L_cur, s_warp);
			ex_cur = e1[0];
		}
		uint64_t D = 0;
		uint32_t carry = 0;
		bool overrun = L_cur > ndata;
		if (!overrun)
			issue_data(0, 0, L_cur);
		__syncthreads();	// s_warp is reused by the first scan-ahead

		for (uint32_t t = 0; t < ntiles && !overrun; t++) {
			// ---- data of tile t: in stage t & 1 once this thread's copies have landed and the
			// barrier below has passed (same 16-byte phase as in global memory)
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(data + D) & 15);
			const uint8_t *s_in = s_in2[t & 1];
			cp_async_wait_all();
			// ---- scan ahead: tile t + 1's byte counts, then request its data and tile t + 2's keys
			uint32_t fl_n = 0, va_n = 0, ex_n = 0, L_n = 0;
			if (t + 1 < ntiles) {
				va_n = valid_of(t + 1);
				fl_n = dec_keys_flags<KEYBITS>(k_next) & va_n;
				// (the barrier inside either branch also publishes s_in)
				if (!__syncthreads_or(fl_n != 0 || va_n != 0xffffu)) {
					ex_n = 16 * tid;	// 4096 one-byte values: no scan needed
					L_n = DT;
				} else {
					uint32_t c1[1] = {(uint32_t) (__popc(va_n) + __popc(fl_n))}, e1[1];
					block_scan<1>(c1, e1, L_n, s_warp);
					ex_n = e1[0];
				}
				if (D + L_cur + L_n > ndata)
					overrun = true;
				else
					issue_data((t + 1) & 1, D + L_cur, L_n);
				if (t + 2 < ntiles)
					k_next = keys_of(t + 2);
			} else {
				__syncthreads();
			}
			// ---- bytes -> zd values, two per word
			uint32_t zz[8];
			const uint32_t p = al + ex_cur;
			if (L_cur == DT && va_cur == 0xffffu && interior_of(t)) {
				// full tile without a two-byte value: the thread's 16 bytes sit at stage[al + 16 tid]
				stage_bytes16(s_in, al, tid, zz);
			} else if (va_cur == 0xffffu && fl_cur == 0) {
				// 16 one-byte values at an arbitrary stage offset
				const uint32_t *w32 = reinterpret_cast<const uint32_t *>(s_in) + (p >> 2);
				const uint32_t sh = (p & 3) * 8;
				uint32_t w[5];
#pragma unroll
				for (int j = 0; j < 5; j++)
					w[j] = w32[j];
#pragma unroll
				for (int j = 0; j < 4; j++) {
					const uint32_t v = __funnelshift_r(w[j], w[j + 1], sh);
					zz[2 * j] = __byte_perm(v, 0, 0x4140);
					zz[2 * j + 1] = __byte_perm(v, 0, 0x4342);
				}
			} else {
				const uint8_t *q = s_in + p;
#pragma unroll
				for (int j = 0; j < 8; j++)
					zz[j] = 0;
#pragma unroll
				for (int e = 0; e < 16; e++) {
					if ((va_cur >> e) & 1) {
						uint32_t v = *q++;
						if ((fl_cur >> e) & 1)
							v |= (uint32_t) (*q++) << 8;
						zz[e >> 1] |= v << (16 * (e & 1));
					}
				}
			}
			// ---- unzigzag, running sum (mod 2^16 in each half): decode_scalar.hpp:10-13, 52-65
			uint32_t y[8], c = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) {
				const uint32_t d = unzigzag16x2(zz[j]);
				y[j] = add16x2(d * 0x00010001u, c);
				c = __byte_perm(y[j], 0, 0x3232);
			}
			uint32_t c1[1] = {c & 0xffffu}, e1[1], tot;
			block_scan<1>(c1, e1, tot, s_warp2);
			const uint32_t cw = ((e1[0] + carry) & 0xffffu) * 0x00010001u;
			uint32_t o[8];
#pragma unroll
			for (int j = 0; j < 8; j++)
				o[j] = add16x2(y[j], cw);
			const uint64_t ka = A + (uint64_t) t * (DT / 8) + 2 * tid;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t vm = (va_cur >> (8 * h)) & 0xffu;
				if (vm == 0xffu) {
					out4[ka + h] = make_uint4(o[4 * h], o[4 * h + 1], o[4 * h + 2], o[4 * h + 3]);
				} else if (vm) {
#pragma unroll
					for (int e = 0; e < 8; e++)
						if ((vm >> e) & 1)
							out16[8 * (ka + h) + e] = (uint16_t) (o[4 * h + (e >> 1)] >> (16 * (e & 1)));
				}
			}
			carry = (carry + tot) & 0xffffu;
			D += L_cur;
			fl_cur = fl_n;
			va_cur = va_n;
			ex_cur = ex_n;
			L_cur = L_n;
			// (no barrier here: every read of s_in above precedes the value scan's barrier, every
			//  read of s_warp / s_warp2 is followed by the next tile's first barrier before the
			//  array is written again)
		}
		if (overrun && tid == 0)
			atomicOr(&a.hdr->error, ERR_MALFORMED);
	}
}

}  // namespace pg
