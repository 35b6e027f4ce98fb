// svb_dec.cuh -- fused streamvbyte + inverse zigzag-delta DECODE, one pass over HBM.
//
//   KEYBITS = 1 : ONT svb16 (svb16/decode_scalar.hpp:31-71; press.c:1689-1694)
//   KEYBITS = 2 : classical streamvbyte, codes 0/1 only (streamvbyte_decode.c:16-58; press.c:1600-1612)
//   FRAME       : [n u32 LE] prefix of the pre-zstd VBZ stream (press.c:1887-1910, 2047-2070)
//
// The reference's decoder carries two values from one sample to the next: the
// data pointer (advanced by 1 + key bit) and the running sum of the deltas.  Here
// one CTA owns one READ (reads are handed out by an atomic ticket) and walks it
// step by step (8192 samples), carrying both values itself, so no tile ever waits
// for another CTA and nothing is read twice:
//
//   keys of step t+2   -> registers        (their address does not depend on the data)
//   keys of step t+1   -> byte counts -> block scan -> the step's data range is known
//   data of step t+1   -> shared memory by cp.async (no registers held)
//   step t             -> stage -> values -> unzigzag -> block scan + carry -> aligned
//                         128-bit stores (steps are cut on 16-byte boundaries of the
//                         output array)
//
// so the loads of a step are in flight for a whole step time before they are
// needed.  592 reads are in flight on a B200 (4 CTAs per SM).
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

// ---------------------------------------------------------------------------
// A step is TWO 4096-sample subtiles (thread tid holds 16 samples of each): one "anything but
// one-byte values?" vote, one byte-count scan (both counts packed in one word) and one value scan
// (both 16-bit sums packed, added per half) per 8192 samples, the two subtiles' dependency chains
// interleave, and a subtile of one-byte values is un-zigzagged four values per word before it is
// widened.  (r01's kernel did all of this per 4096 samples: 57 % of HBM peak; this one 64 %.)

constexpr int DT2 = 2 * DT;	// samples per step

template <int KEYBITS, bool FRAME>
#ifndef PG_DEC_CTAS
#define PG_DEC_CTAS 4
#endif
__global__ void __launch_bounds__(NT, PG_DEC_CTAS) svb_dec2_kernel(const __grid_constant__ SvbDec2Args a)
{
	// the step's data bytes, double buffered: step t + 1 arrives by cp.async while step t is decoded
	__shared__ __align__(16) uint8_t s_in2[2][2 * DT2 + 64];
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
		// steps are cut on 16-byte boundaries of the OUTPUT (aligned 128-bit stores): element e of
		// thread tid in subtile h of step t is read index t * DT2 + h * DT + 16 tid + e - m
		const uint32_t m = (uint32_t) (off0 & 7);
		const uint64_t A = off0 >> 3;
		const uint32_t nsteps = (uint32_t) ((m + n + DT2 - 1) / DT2);
		uint4 *out4 = reinterpret_cast<uint4 *>(a.samples_out);
		uint16_t *out16 = reinterpret_cast<uint16_t *>(a.samples_out);
		// steps [t_lo, t_hi) are interior: every element is a sample of the read
		const uint32_t t_lo = m ? 1u : 0u, t_hi = (uint32_t) ((m + n) / DT2);
		const bool keys_safe = keys + nkeys + 8 <= limit;

		auto first_of = [&](uint32_t t, int h) -> int64_t {
			return (int64_t) t * DT2 + h * DT + tid * DPT - (int64_t) m;
		};
		// valid masks of both subtiles (low / high 16 bits)
		auto valid_of = [&](uint32_t t) -> uint32_t {
			if (t >= t_lo && t < t_hi)
				return 0xffffffffu;
			uint32_t v = 0;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const int64_t i0 = first_of(t, h), left = (int64_t) n - i0;
				const int lo = i0 < 0 ? (int) (-i0 < 16 ? -i0 : 16) : 0;
				const int hi = left >= 16 ? 16 : (left > 0 ? (int) left : 0);
				const uint32_t vh = hi > lo ? (0xffffu >> (16 - hi)) & (0xffffu << lo) & 0xffffu : 0u;
				v |= vh << (16 * h);
			}
			return v;
		};
		auto keys_of = [&](uint32_t t, int h) -> KeyRaw {
			const int64_t i0 = first_of(t, h);
			if (keys_safe && i0 >= 0)
				return dec_keys_issue_fast<KEYBITS>(keys, (uint64_t) i0);
			return dec_keys_issue<KEYBITS>(keys, i0, a.in, limit);
		};
		// data bytes of a step -> stage `buf`
		auto issue_data = [&](uint32_t buf, uint64_t D, uint32_t L) {
			const uint8_t *src = data + D;
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
			const uint8_t *b16 = src - al;
			const uint32_t nw = (al + L + 15) >> 4;
			uint4 *dst = reinterpret_cast<uint4 *>(s_in2[buf]);
			if (b16 + 16ull * nw <= limit) {	// (b16 >= a.in: the container is 16-byte aligned)
#pragma unroll
				for (int j = 0; j < 5; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						cp_async16(dst + w, reinterpret_cast<const uint4 *>(b16) + w);
				}
			} else {
#pragma unroll
				for (int j = 0; j < 5; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						dst[w] = load_word_guarded(b16, w, a.in, limit);
				}
			}
			cp_async_commit();
		};
		// flags of both subtiles from their raw key words, masked by the valid bits
		auto flags_of = [&](const KeyRaw &ka, const KeyRaw &kb, uint32_t va) -> uint32_t {
			return (dec_keys_flags<KEYBITS>(ka) | (dec_keys_flags<KEYBITS>(kb) << 16)) & va;
		};
		// byte counts of both subtiles -> (exclusive offsets in the step's data, total); uniform
		// fast exit when the whole step is one-byte values
		auto count_scan = [&](uint32_t fl, uint32_t va, uint32_t &exa, uint32_t &exb, uint32_t &L) {
			if (!__syncthreads_or(fl != 0 || va != 0xffffffffu)) {
				exa = 16 * tid;
				exb = DT + 16 * tid;
				L = DT2;
				return;
			}
			const uint32_t c = (uint32_t) (__popc(va & 0xffffu) + __popc(fl & 0xffffu)) |
					   ((uint32_t) (__popc(va >> 16) + __popc(fl >> 16)) << 16);
			uint32_t e, tot;
			block_scan_pair(c, e, tot, s_warp);
			exa = e & 0xffffu;
			exb = (tot & 0xffffu) + (e >> 16);
			L = (tot & 0xffffu) + (tot >> 16);
		};

		// ---- prologue: step 0 scanned and its data requested, keys of step 1 requested
		uint32_t va_cur = valid_of(0);
		uint32_t fl_cur;
		{
			const KeyRaw k0a = dec_keys_issue<KEYBITS>(keys, first_of(0, 0), a.in, limit);
			const KeyRaw k0b = dec_keys_issue<KEYBITS>(keys, first_of(0, 1), a.in, limit);
			fl_cur = flags_of(k0a, k0b, va_cur);
		}
		KeyRaw kna = {0, 0, 0, 0}, knb = {0, 0, 0, 0};
		if (nsteps > 1) {
			kna = keys_of(1, 0);
			knb = keys_of(1, 1);
		}
		uint32_t exa_cur, exb_cur, L_cur;
		count_scan(fl_cur, va_cur, exa_cur, exb_cur, L_cur);
		uint64_t D = 0;
		uint32_t carry = 0;
		bool overrun = L_cur > ndata;
		if (!overrun)
			issue_data(0, 0, L_cur);
		__syncthreads();	// s_warp is reused by the first scan-ahead

		for (uint32_t t = 0; t < nsteps && !overrun; t++) {
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(data + D) & 15);
			const uint8_t *s_in = s_in2[t & 1];
			cp_async_wait_all();
			// ---- scan ahead: step t + 1's byte counts, then request its data and step t + 2's keys
			uint32_t fl_n = 0, va_n = 0, exa_n = 0, exb_n = 0, L_n = 0;
			if (t + 1 < nsteps) {
				va_n = valid_of(t + 1);
				fl_n = flags_of(kna, knb, va_n);
				count_scan(fl_n, va_n, exa_n, exb_n, L_n);	// (its barrier also publishes s_in)
				if (D + L_cur + L_n > ndata)
					overrun = true;
				else
					issue_data((t + 1) & 1, D + L_cur, L_n);
				if (t + 2 < nsteps) {
					kna = keys_of(t + 2, 0);
					knb = keys_of(t + 2, 1);
				}
			} else {
				__syncthreads();
			}
			// ---- bytes -> deltas, two per word, for both subtiles
			const bool plain = L_cur == DT2 && va_cur == 0xffffffffu && t >= t_lo && t < t_hi;
			uint32_t y[2][8], sums = 0;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t vah = (va_cur >> (16 * h)) & 0xffffu, flh = (fl_cur >> (16 * h)) & 0xffffu;
				const uint32_t p = al + (h ? exb_cur : exa_cur);
				uint32_t d[8];
				if (plain) {
					// whole step of one-byte values: the thread's 16 bytes sit at stage[al + h DT + 16 tid]
					uint32_t w[4];
					stage_words_aligned(s_in + h * DT, al, tid, w);
					unzigzag_bytes_x16(w, d);
				} else if (vah == 0xffffu && flh == 0) {
					uint32_t w[4];
					stage_words_any(s_in, p, w);
					unzigzag_bytes_x16(w, d);
				} else if (vah == 0xffffu && (flh & (flh - 1)) == 0) {
					// ONE two-byte value among the thread's 16, at element x (one step in three has such a
					// thread at NA12878's rate, and the element loop below holds its warp, and with it the
					// CTA's next barrier, for 16 dependent shared-memory loads): 17 bytes in a row, those
					// after element x moved down by one, the value's 16 bits put in its place
					const uint32_t x = __ffs(flh) - 1;
					const uint32_t *w32 = reinterpret_cast<const uint32_t *>(s_in) + (p >> 2);
					const uint32_t sh = (p & 3) * 8;
					uint32_t u[6], b[5], r[4];
#pragma unroll
					for (int j = 0; j < 6; j++)
						u[j] = w32[j];
#pragma unroll
					for (int j = 0; j < 5; j++)
						b[j] = __funnelshift_r(u[j], u[j + 1], sh);
#pragma unroll
					for (int j = 0; j < 4; j++) {
						const uint32_t dn = __funnelshift_r(b[j], b[j + 1], 8);	// byte i = byte i + 1
						const uint32_t m = x + 1 <= 4u * j ? 0u : x + 1 >= 4u * j + 4 ? 0xffffffffu : (1u << (8 * (x + 1 - 4u * j))) - 1;
						r[j] = (b[j] & m) | (dn & ~m);
					}
					unzigzag_bytes_x16(r, d);
					const uint32_t q = x >> 2;
					const uint32_t wl = q == 0 ? b[0] : q == 1 ? b[1] : q == 2 ? b[2] : b[3];
					const uint32_t wh = q == 0 ? b[1] : q == 1 ? b[2] : q == 2 ? b[3] : b[4];
					const uint32_t zx = __funnelshift_r(wl, wh, 8 * (x & 3)) & 0xffffu;
					const uint32_t dv = (((zx >> 1) ^ (0u - (zx & 1u))) & 0xffffu) * 0x00010001u;	// trans.c:80-83
#pragma unroll
					for (int j = 0; j < 8; j++) {	// (masks, not an indexed store: d stays in registers)
						const uint32_t hit = (x >> 1) == (uint32_t) j ? ((x & 1) ? 0xffff0000u : 0x0000ffffu) : 0u;
						d[j] = (d[j] & ~hit) | (dv & hit);
					}
				} else {
					const uint8_t *q = s_in + p;
					uint32_t zz[8];
#pragma unroll
					for (int j = 0; j < 8; j++)
						zz[j] = 0;
#pragma unroll
					for (int e = 0; e < 16; e++) {
						if ((vah >> e) & 1) {
							uint32_t v = *q++;
							if ((flh >> e) & 1)
								v |= (uint32_t) (*q++) << 8;
							zz[e >> 1] |= v << (16 * (e & 1));
						}
					}
#pragma unroll
					for (int j = 0; j < 8; j++)
						d[j] = unzigzag16x2(zz[j]);
				}
				// running sum inside the thread (mod 2^16 in each half): decode_scalar.hpp:52-65
				uint32_t c = 0;
#pragma unroll
				for (int j = 0; j < 8; j++) {
					y[h][j] = add16x2(d[j] * 0x00010001u, c);
					c = __byte_perm(y[h][j], 0, 0x3232);
				}
				sums |= (c & 0xffffu) << (16 * h);
			}
			uint32_t e2, tot2;
			block_scan_pair(sums, e2, tot2, s_warp2);
			const uint32_t cwa = ((e2 + carry) & 0xffffu) * 0x00010001u;
			const uint32_t cwb = (((e2 >> 16) + tot2 + carry) & 0xffffu) * 0x00010001u;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t cw = h ? cwb : cwa;
				const uint64_t ka = A + (uint64_t) t * (DT2 / 8) + h * (DT / 8) + 2 * tid;
#pragma unroll
				for (int g = 0; g < 2; g++) {
					const uint32_t vm = (va_cur >> (16 * h + 8 * g)) & 0xffu;
					const uint32_t o0 = add16x2(y[h][4 * g], cw), o1 = add16x2(y[h][4 * g + 1], cw);
					const uint32_t o2 = add16x2(y[h][4 * g + 2], cw), o3 = add16x2(y[h][4 * g + 3], cw);
					if (vm == 0xffu) {
						out4[ka + g] = make_uint4(o0, o1, o2, o3);
					} else if (vm) {
						const uint32_t o[4] = {o0, o1, o2, o3};
#pragma unroll
						for (int e = 0; e < 8; e++)
							if ((vm >> e) & 1)
								out16[8 * (ka + g) + e] = (uint16_t) (o[e >> 1] >> (16 * (e & 1)));
					}
				}
			}
			carry = (carry + (tot2 & 0xffffu) + (tot2 >> 16)) & 0xffffu;
			D += L_cur;
			fl_cur = fl_n;
			va_cur = va_n;
			exa_cur = exa_n;
			exb_cur = exb_n;
			L_cur = L_n;
			// (no barrier here: every read of s_in above precedes the value scan's barrier, every
			//  read of s_warp / s_warp2 is followed by the next step's first barrier before the
			//  array is written again)
		}
		if (overrun && tid == 0)
			atomicOr(&a.hdr->error, ERR_MALFORMED);
	}
}

}  // namespace pg
