// exfam_dec.cuh -- decode of the exception-coded methods (see exfam.cuh for the
// formats) and of the static Huffman payload.
//
//   ex_parse   (read)  header -> n, q, zd0, payload position; both exception lists
//                      decoded into a pool (sorted positions, values)
//   huff_decode(read)  SHUFF only: bitstream -> one-byte symbols in a scratch payload
//                      (self-synchronising parallel decode, see below)
//   ex_decode2 (read)  one CTA walks one read step by step (two 4096-sample subtiles): payload
//                      + exceptions -> unzigzag -> running sum (trans.c:260-270) carried by the
//                      CTA from step to step -> << q -> samples; the next step's payload bytes
//                      are requested before the current step is processed
#pragma once

#include "exfam.cuh"

namespace pg {

struct ExDecArgs {
	int method;
	const uint8_t *in;		// container, 16-byte aligned
	uint64_t in_bytes;
	const uint64_t *in_offsets;	// [nreads + 1]
	const uint64_t *offsets;	// [nreads + 1] sample offsets of the output
	uint64_t nreads;
	int16_t *samples_out;		// 16-byte aligned
	Hdr *hdr;			// aux[0] = pool entries used, ticket = next read (huff_decode)
	const uint32_t *tile_start, *tile_read;
	uint32_t *r_nex, *r_q, *r_zd0, *r_pool;
	uint64_t *r_pay;		// byte offset of the payload / bitstream in `in`; SENTINEL: bad read
	uint2 *pool;
	uint64_t pool_cap;
	uint8_t *paybuf;		// SHUFF: symbols of read r at paybuf + offsets[r]
	const uint32_t *huff_dec;	// 4096 two-symbol entries: see hd_run
	const uint16_t *huff_sub;	// second level: [k][1 << huff_sub_bits]
	uint32_t huff_sub_bits, huff_minlen, huff_nsub;
	const uint16_t *sidecar;	// SHUFF, optional: see huff_fused.cuh (null: self-synchronising decode)
};

// 16-byte word `w` of the 16-byte aligned region starting at `base16`, bytes outside [lo, hi) read as 0
__device__ __forceinline__ uint4 load_word_guarded(const uint8_t *base16, uint32_t w, const uint8_t *lo, const uint8_t *hi)
{
	const uint8_t *p = base16 + 16ull * w;
	if (p >= lo && p + 16 <= hi)
		return __ldg(reinterpret_cast<const uint4 *>(p));
	uint32_t v[4] = {0, 0, 0, 0};
	for (int b = 0; b < 16; b++)
		if (p + b >= lo && p + b < hi)
			v[b >> 2] |= (uint32_t) p[b] << (8 * (b & 3));
	return make_uint4(v[0], v[1], v[2], v[3]);
}

// 16 bytes at stage[al + 16 tid ...) (al < 16, the same in every thread) as zd values,
// two per word: two conflict-free 128-bit loads and a byte shift
__device__ __forceinline__ void stage_bytes16(const uint8_t *stage, uint32_t al, int tid, uint32_t (&zz)[8])
{
	const uint4 a = reinterpret_cast<const uint4 *>(stage)[tid], b = reinterpret_cast<const uint4 *>(stage)[tid + 1];
	uint32_t u0, u1, u2, u3, u4;
	switch (al >> 2) {
	case 0: u0 = a.x; u1 = a.y; u2 = a.z; u3 = a.w; u4 = b.x; break;
	case 1: u0 = a.y; u1 = a.z; u2 = a.w; u3 = b.x; u4 = b.y; break;
	case 2: u0 = a.z; u1 = a.w; u2 = b.x; u3 = b.y; u4 = b.z; break;
	default: u0 = a.w; u1 = b.x; u2 = b.y; u3 = b.z; u4 = b.w; break;
	}
	const uint32_t sh = (al & 3) * 8;
	const uint32_t v0 = __funnelshift_r(u0, u1, sh), v1 = __funnelshift_r(u1, u2, sh);
	const uint32_t v2 = __funnelshift_r(u2, u3, sh), v3 = __funnelshift_r(u3, u4, sh);
	zz[0] = __byte_perm(v0, 0, 0x4140); zz[1] = __byte_perm(v0, 0, 0x4342);
	zz[2] = __byte_perm(v1, 0, 0x4140); zz[3] = __byte_perm(v1, 0, 0x4342);
	zz[4] = __byte_perm(v2, 0, 0x4140); zz[5] = __byte_perm(v2, 0, 0x4342);
	zz[6] = __byte_perm(v3, 0, 0x4140); zz[7] = __byte_perm(v3, 0, 0x4342);
}

__device__ __forceinline__ uint64_t get_le(const uint8_t *p, int nbytes)
{
	uint64_t v = 0;
	for (int b = 0; b < nbytes; b++)
		v |= (uint64_t) p[b] << (8 * b);
	return v;
}

// streamvbyte list decode by one warp (streamvbyte_decode.c:16-58 / svb16 decode_scalar.hpp:31-71);
// sink(k, active, value) is called by all lanes for every chunk of 32 values.
// Returns the bytes consumed; sets bad when the list does not fit in `avail`.
template <int KB, class F>
__device__ __forceinline__ uint32_t warp_svb_decode(const uint8_t *src, uint64_t avail, uint32_t cnt,
						    F sink, int lane, bool &bad)
{
	const uint32_t nkeys = KB == 2 ? (cnt + 3) / 4 : (cnt + 7) / 8;
	if (nkeys > avail) {
		bad = true;
		return 0;
	}
	uint64_t run = nkeys;
	for (uint32_t k0 = 0; k0 < cnt; k0 += 32) {
		const uint32_t k = k0 + lane;
		const bool act = k < cnt;
		uint32_t len = 0;
		if (act)
			len = KB == 2 ? ((src[k >> 2] >> (2 * (k & 3))) & 3u) + 1u : ((src[k >> 3] >> (k & 7)) & 1u) + 1u;
		const uint32_t inc = warp_inc_scan(len, lane);
		const uint64_t o = run + inc - len;
		const bool over = act && o + len > avail;
		if (__any_sync(0xffffffffu, over)) {
			bad = true;
			return 0;
		}
		uint32_t val = 0;
		for (uint32_t b = 0; b < len; b++)
			val |= (uint32_t) src[o + b] << (8 * b);
		sink(k, act, val);
		run += __shfl_sync(0xffffffffu, inc, 31);
	}
	return (uint32_t) run;
}

// uint_depress_16 / uint_depress_32 by one warp (press.c:509-528, 566-585): [width u8] + big-endian
// bit fields (see warp_bitpack_list); sink as in warp_svb_decode.  Returns the bytes consumed.
template <class F>
__device__ __forceinline__ uint32_t warp_bitpack_decode(const uint8_t *src, uint64_t avail, uint32_t cnt,
							F sink, int lane, bool &bad)
{
	if (avail < 1) {
		bad = true;
		return 0;
	}
	const uint32_t bits = src[0];
	const uint64_t nbytes = ((uint64_t) cnt * bits + 7) >> 3;
	if (bits > 32 || 1 + nbytes > avail) {
		bad = true;
		return 0;
	}
	for (uint32_t k0 = 0; k0 < cnt; k0 += 32) {
		const uint32_t k = k0 + lane;
		const bool act = k < cnt;
		uint32_t val = 0;
		if (act && bits) {
			const uint64_t bo = (uint64_t) k * bits, b0 = bo >> 3;
			uint64_t w = 0;		// bytes b0 .. b0 + 4 of the run, first byte on top
			for (uint32_t i = 0; i < 5; i++)
				if (b0 + i < nbytes)
					w |= (uint64_t) src[1 + b0 + i] << (8 * (7 - i));
			val = (uint32_t) ((w << (bo & 7)) >> (64 - bits));
		}
		sink(k, act, val);
	}
	return (uint32_t) (1 + nbytes);
}

__global__ void __launch_bounds__(NT) ex_parse_kernel(const __grid_constant__ ExDecArgs a)
{
	const uint64_t r = ((uint64_t) blockIdx.x * NT + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (r >= a.nreads)
		return;
	const uint64_t n = a.offsets[r + 1] - a.offsets[r];
	const uint64_t s0 = a.in_offsets[r], s1 = a.in_offsets[r + 1];
	const bool exz = a.method == PRESS_GPU_EX_ZD, vbe = a.method == PRESS_GPU_VBE21_ZD ||
							    a.method == PRESS_GPU_SHUFF_VBE21_ZD;
	const uint32_t fx = ex_fixed_bytes(a.method);
	if (n == 0) {	// no stream exists for an empty read (the reference underflows, press.c:3426)
		if (lane == 0) {
			a.r_pay[r] = SENTINEL;
			a.r_nex[r] = 0;
		}
		return;
	}
	bool bad = s1 < s0 || s1 > a.in_bytes || s1 - s0 < fx || n > 0xffffffffull;
	uint32_t q = 0, zd0 = 0, nex = 0;
	const uint64_t m = n - 1;
	if (!bad) {
		const uint8_t *p = a.in + s0;
		if (exz) {	// ex_zd.c:459-515
			bad = p[0] != 0 || get_le(p + 1, 8) != n || p[9] > 5;
			q = p[9];
			p += 10;
		}
		zd0 = (uint32_t) get_le(p, 2);
		nex = (uint32_t) get_le(p + 2, 4);
		bad = bad || nex > m;
	}
	uint64_t off = s0 + fx;
	uint64_t pbase = 0;
	if (!bad && nex) {
		if (lane == 0)
			pbase = atomicAdd((unsigned long long *) &a.hdr->aux[0], (unsigned long long) nex);
		pbase = __shfl_sync(0xffffffffu, pbase, 0);
		if (pbase + nex > a.pool_cap) {
			if (lane == 0)
				atomicOr(&a.hdr->error, ERR_SCRATCH);
			bad = true;
		}
	}
	if (!bad && nex) {
		uint2 *pl = a.pool + pbase;
		bool order_bad = false;
		if (vbe) {	// press.c:2738-2748: raw lists
			if (off + 6ull * nex > s1) {
				bad = true;
			} else {
				const uint8_t *pp = a.in + off, *pv = pp + 4ull * nex;
				for (uint32_t k = lane; k < nex; k += 32) {
					const uint32_t pos = (uint32_t) get_le(pp + 4ull * k, 4);
					const uint32_t prv = k ? (uint32_t) get_le(pp + 4ull * (k - 1), 4) : 0u;
					order_bad |= pos >= m || (k && pos <= prv);
					pl[k] = make_uint2(pos, (uint32_t) get_le(pv + 2ull * k, 2));
				}
				off += 6ull * nex;
			}
		} else if (nex == 1) {	// press.c:3323-3330 / ex_zd.c:264-272
			const uint32_t vb = exz ? 4 : 2;
			if (off + 4 + vb > s1) {
				bad = true;
			} else {
				if (lane == 0) {
					const uint32_t pos = (uint32_t) get_le(a.in + off, 4);
					order_bad = pos >= m;
					pl[0] = make_uint2(pos, ((uint32_t) get_le(a.in + off + 4, vb) + 256u) & 0xffffu);
				}
				off += 4 + vb;
			}
		} else {
			// positions: svb32 gaps, p[k] = p[k - 1] + gap + 1 (trans.c:186-197)
			uint64_t carry = 0;
			auto pos_sink = [&](uint32_t k, bool act, uint32_t val) {
				uint64_t inc = act ? (uint64_t) val + 1 : 0;
#pragma unroll
				for (int o = 1; o < 32; o <<= 1) {
					const uint64_t y = __shfl_up_sync(0xffffffffu, inc, o);
					if (lane >= o)
						inc += y;
				}
				const uint64_t pos = carry + inc - 1;
				if (act) {
					order_bad |= pos >= m;
					pl[k].x = (uint32_t) pos;
				}
				carry += __shfl_sync(0xffffffffu, inc, 31);
			};
			auto val_sink = [&](uint32_t k, bool act, uint32_t val) {
				if (act)
					pl[k].y = (val + 256u) & 0xffffu;	// stored into uint16: wraps (press.c:3336)
			};
			for (int pass = 0; pass < 2 && !bad; pass++) {
				if (off + 4 > s1) {
					bad = true;
					break;
				}
				const uint32_t len = (uint32_t) get_le(a.in + off, 4);
				off += 4;
				if (off + len > s1) {
					bad = true;
					break;
				}
				uint32_t used;
				if (ex_is_bitpacked(a.method))	// press.c:2912-2938
					used = pass == 0 ? warp_bitpack_decode(a.in + off, len, nex, pos_sink, lane, bad)
							 : warp_bitpack_decode(a.in + off, len, nex, val_sink, lane, bad);
				else if (pass == 0)
					used = warp_svb_decode<2>(a.in + off, len, nex, pos_sink, lane, bad);
				else if (exz)
					used = warp_svb_decode<2>(a.in + off, len, nex, val_sink, lane, bad);
				else
					used = warp_svb_decode<1>(a.in + off, len, nex, val_sink, lane, bad);
				if (exz && used != len)	// ex_zd.c:222-232, 252-262
					bad = true;
				off += len;
			}
		}
		bad = bad || __any_sync(0xffffffffu, order_bad);
	}
	if (!bad) {
		if (ex_is_huff(a.method)) {
			// [count u32 BE] + bitstream; huffman.c:696-711 memread fails when no byte follows
			bad = off + 4 >= s1;
			if (!bad) {
				const uint8_t *p = a.in + off;
				const uint32_t cnt = ((uint32_t) p[0] << 24) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 8) | p[3];
				bad = cnt != m - nex;
				off += 4;
			}
		} else {
			bad = s1 - off != m - nex;
		}
	}
	if (lane == 0) {
		if (bad)
			atomicOr(&a.hdr->error, ERR_MALFORMED);
		a.r_pay[r] = bad ? SENTINEL : off;
		a.r_nex[r] = nex;
		a.r_q[r] = q;
		a.r_zd0[r] = zd0;
		a.r_pool[r] = (uint32_t) pbase;
	}
}

// ---------------------------------------------------------------------------
// static Huffman decode (huffman.c:1219-1265: walk the tree bit by bit from bit 0
// of every byte until `count` symbols are out).
//
// One CTA per read (reads handed out by an atomic ticket), 4 KB of bitstream per
// step, one 128-bit sub-block per thread.  A thread does not know where the first
// code of its sub-block starts, so it guesses "at bit 0", decodes, and hands the
// position where it left its sub-block to the next thread; a thread whose entry
// point changed decodes again.  Prefix codes re-synchronise after a few symbols,
// so two or three rounds settle all 256 sub-blocks; the loop runs until nothing
// changes, which makes the result exact whatever the data (worst case: serial).
// The entry point into the next 4 KB step is carried by the CTA itself.

constexpr int HD_SUB = 128;			// bits per sub-block
constexpr int HD_CHUNK = NT * HD_SUB / 8;	// bytes per step
constexpr int HD_SLOT = HD_SUB / 4;		// symbols a sub-block can hold when no code is shorter than 4 bits
constexpr int HD_SYM = NT * HD_SLOT;		// symbol stage
constexpr int HD_L1 = 12;			// first-level table: 12 bits

// Decode the codes of one sub-block that start at bit `pos` (relative to the sub-block)
// or later; returns the position of the first code of the NEXT sub-block (>= HD_SUB) and
// the number of symbols in cnt.  MODE 0 counts only; 1 stores symbol i at dst[i * STRIDE]
// (the caller knows they fit); 2 stores those with i < lim.  Table levels, no tree walk
// (a walk taken by one lane stalls the other 31):
//   lut[w & 4095]: codes of <= 12 bits: symbol | symbol2 << 8 | length2 << 16 | length << 24,
//                  where symbol2 is the NEXT code when it also fits in the 12 bits
//                  (length2 = 0 if not) -- two symbols per lookup on most of the stream;
//                  longer codes: bit 31 | k, and sub[k][(w >> 12) & mask] = length << 8 | symbol.
// The staged bitstream is transposed: word j of thread t's sub-block at sw[j * NT + t],
// j = 0..4 (row 4 = the next sub-block's first word), so whatever words the lanes of a warp
// are at, they read 32 different banks, and the second word of a window is one row further.
constexpr int HD_ROW = NT;

// a code longer than the first-level table (rare; kept out of line so that its instructions
// do not ride, predicated off, through every turn of the decode loops)
__device__ __noinline__ uint32_t hd_long(uint32_t e, uint32_t w, const uint16_t *sub, uint32_t sub_bits)
{
	const uint32_t e1 = __ldg(sub + (((e & 0x7fffu) << sub_bits) | ((w >> HD_L1) & ((1u << sub_bits) - 1))));
	return (e1 & 0xffu) | ((e1 >> 8) << 24);
}

__device__ __forceinline__ uint32_t hd_lookup(uint32_t swt_sa, uint32_t lut_sa, const uint16_t *sub, uint32_t sub_bits,
					      uint32_t pos)
{
	// shared-window addresses by hand: word (pos >> 5) of the thread's column and the next row
	const uint32_t wa = swt_sa + ((pos & 0x60u) << 5);
	uint32_t w0, w1, e;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(wa));
	asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(w1) : "r"(wa), "n"(HD_ROW * 4));
	const uint32_t w = __funnelshift_r(w0, w1, pos);
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(lut_sa + ((w & ((1u << HD_L1) - 1)) << 2)));
	if (__builtin_expect((int32_t) e < 0, 0))
		e = hd_long(e, w, sub, sub_bits);
	return e;
}

// The slot pass (tables with >= 4-bit codes): turn i of the loop leaves its one or two
// symbols as a 16-bit pair at pairs[i * NT] (i is the same in every lane, so the lanes'
// stores fall into different banks) and bit i of `two` says whether the second one counts.
__device__ __forceinline__ uint32_t hd_run_pairs(uint32_t swt_sa, uint32_t lut_sa, const uint16_t *sub, uint32_t sub_bits,
						 uint32_t pos, uint32_t &turns, uint32_t &two, uint16_t *pairs)
{
	uint32_t i = 0, bit = 1u, m = 0;
	while (pos < (uint32_t) HD_SUB) {
		const uint32_t e = hd_lookup(swt_sa, lut_sa, sub, sub_bits, pos);
		pairs[i * NT] = (uint16_t) e;
		pos += e >> 24;
		const uint32_t l2 = (e >> 16) & 0xffu;
		if (l2 && pos < (uint32_t) HD_SUB) {	// the second code starts inside the sub-block as well
			pos += l2;
			m |= bit;
		}
		bit <<= 1;
		i++;
	}
	turns = i;
	two = m;
	return pos;
}

// The generic pass: MODE 0 counts the symbols only; 2 stores symbol i at dst[i] for i < lim.
template <int MODE>
__device__ __forceinline__ uint32_t hd_run(uint32_t swt_sa, uint32_t lut_sa, const uint16_t *sub, uint32_t sub_bits,
					   uint32_t pos, uint32_t &cnt, uint8_t *dst, uint32_t lim)
{
	uint32_t c = 0;
	while (pos < (uint32_t) HD_SUB) {
		const uint32_t e = hd_lookup(swt_sa, lut_sa, sub, sub_bits, pos);
		if (MODE == 2 && c < lim)
			dst[c] = (uint8_t) e;
		c++;
		pos += e >> 24;
		const uint32_t l2 = (e >> 16) & 0xffu;
		if (l2 && pos < (uint32_t) HD_SUB) {
			if (MODE == 2 && c < lim)
				dst[c] = (uint8_t) (e >> 8);
			c++;
			pos += l2;
		}
	}
	cnt = c;
	return pos;
}

__global__ void __launch_bounds__(NT, 5) huff_decode_kernel(const __grid_constant__ ExDecArgs a)
{
	__shared__ uint32_t s_lut[1 << HD_L1];
	__shared__ uint32_t s_exit[2][NT];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read;
	// the step's symbols, in order, before they leave as aligned 128-bit stores (a byte
	// store per symbol costs one L2 transaction each: that alone bounded the kernel)
	__shared__ __align__(16) uint8_t s_sym[HD_SYM + 32];
	// with no code shorter than 4 bits every pass leaves the sub-block's symbols in the
	// thread's own slots (hd_run_pairs: at most 32 turns), so nothing is decoded again once
	// the entry points have settled.  (Tables with shorter codes take counting passes and one
	// more storing pass.)
	__shared__ __align__(16) uint16_t s_pair[HD_SLOT * NT];
	const int tid = threadIdx.x;
	const bool slots = a.huff_minlen >= 4;
	const uint16_t *sub = a.huff_sub;
	const uint32_t sub_bits = a.huff_sub_bits;
	uint16_t *mypairs = s_pair + tid;
	// the staged bitstream (5 * HD_ROW words) lives in whichever of the two is idle while
	// codes are decoded: the symbol stage with slots, the slots without
	uint32_t *s_w = reinterpret_cast<uint32_t *>(slots ? s_sym : reinterpret_cast<uint8_t *>(s_pair));
	const uint32_t swt_sa = (uint32_t) __cvta_generic_to_shared(s_w + tid);
	const uint32_t lut_sa = (uint32_t) __cvta_generic_to_shared(s_lut);
	for (int i = tid; i < (1 << HD_L1); i += NT)
		s_lut[i] = a.huff_dec[i];
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t bs0 = a.r_pay[r];
		if (bs0 == SENTINEL)
			continue;
		const uint64_t off0 = a.offsets[r], n = a.offsets[r + 1] - off0;
		const uint64_t nbytes = a.in_offsets[r + 1] - bs0;
		const int64_t count = (int64_t) (n - 1 - a.r_nex[r]);
		uint8_t *dst = a.paybuf + off0;
		const uint8_t *src = a.in + bs0, *limit = a.in + a.in_bytes;
		uint32_t entry = 0;	// bit offset of the first code inside the step
		int64_t done = 0;	// symbols out so far
		uint64_t steps = 0, rounds = 0;
		// Every thread fetches the 20 bytes of its own sub-block (+ the next one's first word)
		// straight from global memory, one step ahead, as six aligned words; its column of the
		// transposed stage is private, so staging needs no barrier.  Bytes past the end of the
		// bitstream (the next read's stream, or zeros outside the container) only ever decode
		// into symbols beyond `count`, which are dropped.
		uint32_t g[6];
		const uint32_t gal = (uint32_t) (reinterpret_cast<uintptr_t>(src + 16 * tid) & 3);
		auto fetch = [&](uint64_t cb) {
			const uint8_t *s = src + cb + 16 * tid - gal;	// 4-byte aligned
			if (s >= a.in && s + 24 <= limit) {
				const uint32_t *gw = reinterpret_cast<const uint32_t *>(s);
#pragma unroll
				for (int j = 0; j < 6; j++)
					g[j] = __ldg(gw + j);
			} else {
#pragma unroll
				for (int j = 0; j < 6; j++) {
					uint32_t v = 0;
					for (int b = 0; b < 4; b++) {
						const uint8_t *q = s + 4 * j + b;
						if (q >= a.in && q < limit)
							v |= (uint32_t) *q << (8 * b);
					}
					g[j] = v;
				}
			}
		};
		fetch(0);
		for (uint64_t cb = 0; cb < nbytes && done < count; cb += HD_CHUNK) {
			const uint64_t left = nbytes - cb;
#pragma unroll
			for (int j = 0; j < 5; j++)
				s_w[j * HD_ROW + tid] = __funnelshift_r(g[j], g[j + 1], 8 * gal);
			if (cb + HD_CHUNK < nbytes)
				fetch(cb + HD_CHUNK);
			const uint32_t cbits = left < HD_CHUNK ? (uint32_t) left * 8 : HD_CHUNK * 8;
			const uint32_t nsub = (cbits + HD_SUB - 1) / HD_SUB;
			const bool act = (uint32_t) tid < nsub;
			uint32_t cur = tid == 0 ? entry : 0u, cnt = 0, two = 0, ex = 0;	// (cnt: turns with slots)
			if (act) {
				if (slots)
					ex = hd_run_pairs(swt_sa, lut_sa, sub, sub_bits, cur, cnt, two, mypairs) - HD_SUB;
				else
					ex = hd_run<0>(swt_sa, lut_sa, sub, sub_bits, cur, cnt, nullptr, 0) - HD_SUB;
			}
			s_exit[0][tid] = ex;
			__syncthreads();
			uint32_t rd = 0;
			// Rounds until nothing moves (exits double buffered: one barrier per round).  Inside a
			// round every warp settles its own 32 sub-blocks by shuffles -- a lane whose left
			// neighbour's exit moved decodes again at once -- so only the exit of a warp's last
			// lane, the one the next warp reads from shared memory, can ask for another round:
			// that exit is right after the first pass in all but ~0.5 % of the sub-blocks.
			const int lane = tid & 31;
			for (;;) {
				rd++;
				uint32_t e = tid ? s_exit[(rd - 1) & 1][tid - 1] : entry;
				const uint32_t ex0 = ex;
				for (;;) {
					const bool need = act && e != cur;
					if (!__any_sync(0xffffffffu, need))
						break;
					if (need) {
						cur = e;
						if (slots)
							ex = hd_run_pairs(swt_sa, lut_sa, sub, sub_bits, cur, cnt, two, mypairs) - HD_SUB;
						else
							ex = hd_run<0>(swt_sa, lut_sa, sub, sub_bits, cur, cnt, nullptr, 0) - HD_SUB;
					}
					const uint32_t left_ex = __shfl_up_sync(0xffffffffu, ex, 1);
					if (lane)
						e = left_ex;
				}
				s_exit[rd & 1][tid] = ex;
				if (!__syncthreads_or(lane == 31 && ex != ex0))
					break;
			}
			const uint32_t nsym = slots ? cnt + __popc(two) : cnt;
			uint32_t c1[1] = {act ? nsym : 0u}, ex1[1], tot;
			block_scan<1>(c1, ex1, tot, s_warp);
			if (tot <= HD_SYM) {
				if (act && nsym) {
					if (slots) {
						uint8_t *o = s_sym + ex1[0];
						for (uint32_t i = 0; i < cnt; i++) {
							const uint32_t pr = mypairs[i * NT];
							*o++ = (uint8_t) pr;
							if ((two >> i) & 1u)
								*o++ = (uint8_t) (pr >> 8);
						}
					} else {
						uint32_t dummy;
						(void) hd_run<2>(swt_sa, lut_sa, sub, sub_bits, cur, dummy, s_sym + ex1[0], nsym);
					}
				}
				__syncthreads();
				const int64_t room = count - done;
				block_copy_out(s_sym, dst + done, (uint32_t) ((int64_t) tot < room ? tot : room));
			} else if (act && nsym) {	// a table with very short codes: more symbols than the stage holds
				uint32_t dummy;
				const int64_t first = done + ex1[0], room = count - first;
				(void) hd_run<2>(swt_sa, lut_sa, sub, sub_bits, cur, dummy, dst + first,
						 room > 0 ? (uint32_t) (room < 0xffffffffll ? room : 0xffffffffll) : 0u);
			}
			entry = s_exit[rd & 1][nsub - 1];
			done += tot;
			steps++, rounds += rd;
			__syncthreads();
		}
		if (tid == 0) {	// statistics: steps and re-synchronisation rounds (press_gpu_debug_aux)
			atomicAdd((unsigned long long *) &a.hdr->aux[1], (unsigned long long) steps);
			atomicAdd((unsigned long long *) &a.hdr->aux[2], (unsigned long long) rounds);
			if (done < count)
				atomicOr(&a.hdr->error, ERR_MALFORMED);	// bitstream shorter than its symbol count
		}
	}
}

// ---------------------------------------------------------------------------
// payload + exceptions -> samples, TWO 4096-sample subtiles per step (thread tid holds 16 samples
// of each): one exception vote and one value scan (both 16-bit sums packed, added per half) per
// 8192 samples, the subtiles' chains interleave, and a subtile without exceptions is un-zigzagged
// four values per word before it is widened (cf. svb_dec.cuh)

constexpr int XT2 = 2 * XT;

// two 16-bit exclusive scans over the CTA in one (low / high half, each mod 2^16); one barrier
__device__ __forceinline__ void block_scan_pair(uint32_t v, uint32_t &excl, uint32_t &total, uint32_t *s_warp)
{
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t x = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
		if (lane >= o)
			x = add16x2(x, y);
	}
	if (lane == 31)
		s_warp[warp] = x;
	__syncthreads();
	const uint32_t t = s_warp[lane & 7];
	uint32_t s = t;
#pragma unroll
	for (int o = 1; o < 8; o <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
		if ((lane & 7) >= o)
			s = add16x2(s, y);
	}
	const uint32_t before = __shfl_sync(0xffffffffu, sub16x2(s, t), warp);
	total = __shfl_sync(0xffffffffu, s, 7);
	excl = add16x2(before, sub16x2(x, v));
}

// 16 one-byte zigzag values (four per word) -> deltas, two per word, sign extended
__device__ __forceinline__ void unzigzag_bytes_x16(const uint32_t (&w)[4], uint32_t (&d)[8])
{
#pragma unroll
	for (int j = 0; j < 4; j++) {
		const uint32_t u = ((w[j] >> 1) & 0x7f7f7f7fu) ^ ((w[j] & 0x01010101u) * 0xffu);	// trans.c:80-83 per byte
		asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d[2 * j]) : "r"(u), "r"(0u), "r"(0x9180u));
		asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d[2 * j + 1]) : "r"(u), "r"(0u), "r"(0xb3a2u));
	}
}

// four words holding the 16 bytes at stage[off .. off + 16), off = al + 16 tid with al the same in
// every thread (two conflict-free 128-bit loads), or at any offset (five 32-bit loads)
__device__ __forceinline__ void stage_words_aligned(const uint8_t *stage, uint32_t al, int tid, uint32_t (&w)[4])
{
	const uint4 a = reinterpret_cast<const uint4 *>(stage)[tid], b = reinterpret_cast<const uint4 *>(stage)[tid + 1];
	uint32_t u0, u1, u2, u3, u4;
	switch (al >> 2) {
	case 0: u0 = a.x; u1 = a.y; u2 = a.z; u3 = a.w; u4 = b.x; break;
	case 1: u0 = a.y; u1 = a.z; u2 = a.w; u3 = b.x; u4 = b.y; break;
	case 2: u0 = a.z; u1 = a.w; u2 = b.x; u3 = b.y; u4 = b.z; break;
	default: u0 = a.w; u1 = b.x; u2 = b.y; u3 = b.z; u4 = b.w; break;
	}
	const uint32_t sh = (al & 3) * 8;
	w[0] = __funnelshift_r(u0, u1, sh);
	w[1] = __funnelshift_r(u1, u2, sh);
	w[2] = __funnelshift_r(u2, u3, sh);
	w[3] = __funnelshift_r(u3, u4, sh);
}
__device__ __forceinline__ void stage_words_any(const uint8_t *stage, uint32_t p, uint32_t (&w)[4])
{
	const uint32_t *w32 = reinterpret_cast<const uint32_t *>(stage) + (p >> 2);
	const uint32_t sh = (p & 3) * 8;
	uint32_t u[5];
#pragma unroll
	for (int j = 0; j < 5; j++)
		u[j] = w32[j];
#pragma unroll
	for (int j = 0; j < 4; j++)
		w[j] = __funnelshift_r(u[j], u[j + 1], sh);
}


__global__ void __launch_bounds__(NT, 4) ex_decode2_kernel(const __grid_constant__ ExDecArgs a)
{
	// the step's payload bytes, double buffered: step t + 1 arrives by cp.async while step t is decoded
	__shared__ __align__(16) uint8_t s_stage2[2][XT2 + 64];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_cnt[NW];
	__shared__ uint32_t s_read, s_e[2];
	__shared__ uint2 s_ex[NT];	// the next 256 exceptions of the read (position, value)
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const bool shuff = ex_is_huff(a.method);
	const uint8_t *lim_lo = shuff ? a.paybuf : a.in;
	const uint8_t *lim_hi = shuff ? a.paybuf + ((a.offsets[a.nreads] + 15) & ~15ull) : a.in + a.in_bytes;
	uint4 *out4 = reinterpret_cast<uint4 *>(a.samples_out);
	uint16_t *out16 = reinterpret_cast<uint16_t *>(a.samples_out);
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket2, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint64_t pay0 = a.r_pay[r];
		if (pay0 == SENTINEL)
			continue;
		const uint64_t off0 = __ldg(a.offsets + r), rend = __ldg(a.offsets + r + 1), n = rend - off0;
		const uint32_t m = (uint32_t) (off0 & 7);
		const uint64_t A = off0 >> 3;
		const uint32_t nt4 = (uint32_t) ((m + n + XT - 1) / XT);	// 4096-sample subtiles of the read
		const uint32_t nsteps = (nt4 + 1) / 2;
		const uint32_t nex = a.r_nex[r], q = a.r_q[r], zd0 = a.r_zd0[r];
		const uint2 *ex = a.pool + a.r_pool[r];
		const uint8_t *pbase = shuff ? a.paybuf + off0 : a.in + pay0;
		const uint32_t nm1 = (uint32_t) (n - 1);
		// v-indices (sample index - 1) of the read before subtile tt: 0 for the first, everything
		// from the end on, tt * XT - m - 1 in between
		auto vb = [&](uint32_t tt) { return tt == 0 ? 0u : tt >= nt4 ? nm1 : tt * (uint32_t) XT - m - 1u; };
		const bool safe = pbase - 16 >= lim_lo && pbase + (uint64_t) nm1 + 48 <= lim_hi;
		// payload bytes of a step -> stage `buf`
		auto issue = [&](uint32_t buf, uint32_t j_lo, uint32_t j_hi, uint32_t e_lo) {
			const uint8_t *src = pbase + (j_lo - e_lo);
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
			const uint32_t nw = (al + (j_hi - j_lo) + 15) >> 4;	// (a few bytes more than needed when
			const uint8_t *b16 = src - al;				//  the step holds exceptions)
			uint4 *dst = reinterpret_cast<uint4 *>(s_stage2[buf]);
			const bool fastld = safe || (b16 >= lim_lo && b16 + 16ull * nw <= lim_hi);
#pragma unroll
			for (int j = 0; j < 3; j++) {
				const uint32_t w = tid + j * NT;
				if (w < nw && (j < 2 || tid == 0)) {
					if (fastld)
						cp_async16(dst + w, reinterpret_cast<const uint4 *>(b16) + w);
					else
						dst[w] = load_word_guarded(b16, w, lim_lo, lim_hi);
				}
			}
			cp_async_commit();
		};
		const uint32_t t_hi4 = (uint32_t) ((m + n) / XT);	// subtiles [1, t_hi4) lie inside the read
		uint2 exreg = make_uint2(0xffffffffu, 0u);
		auto issue_ex = [&](uint32_t e0) {
			exreg = e0 + tid < nex ? ex[e0 + tid] : make_uint2(0xffffffffu, 0u);
		};
		uint32_t e_lo = 0, carry = 0;
		issue(0, vb(0), vb(2), 0);
		if (nex)
			issue_ex(0);
		for (uint32_t t = 0; t < nsteps; t++) {
			const uint32_t j0 = vb(2 * t), j1 = vb(2 * t + 1), j2 = vb(2 * t + 2);
			const uint32_t p_lo = j0 - e_lo;	// payload index of the step's first byte
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(pbase + p_lo) & 15);
			const uint8_t *s_stage = s_stage2[t & 1];	// complete once the barrier below has passed
			cp_async_wait_all();
			uint32_t e_mid = e_lo, e_hi = e_lo;
			const uint2 *xs = ex;	// where exceptions [e_lo, e_hi) are read from
			if (!nex) {
				__syncthreads();
			} else {
				s_ex[tid] = exreg;
				// staged exceptions before the end of each subtile (one barrier for both counts)
				const uint32_t c1 = __popc(__ballot_sync(0xffffffffu, exreg.x < j1));
				const uint32_t c2 = __popc(__ballot_sync(0xffffffffu, exreg.x < j2));
				if (lane == 0)
					s_cnt[warp] = c1 | (c2 << 16);
				__syncthreads();
				uint32_t k = 0;
#pragma unroll
				for (int w = 0; w < NW; w++)
					k += s_cnt[w];
				const uint32_t k1 = k & 0xffffu, k2 = k >> 16;
				if (k2 < NT) {
					e_mid = e_lo + k1;
					e_hi = e_lo + k2;
					xs = s_ex - e_lo;
				} else {	// more than 255 in one step: lower_bound over the whole list
					if (tid < 2) {
						const uint32_t key = tid ? j2 : j1;
						uint32_t lo = e_lo, hi = nex;
						while (lo < hi) {
							const uint32_t mid = (lo + hi) >> 1;
							if (ex[mid].x < key)
								lo = mid + 1;
							else
								hi = mid;
						}
						s_e[tid] = lo;
					}
					__syncthreads();
					e_mid = s_e[0];
					e_hi = s_e[1];
				}
			}
			if (t + 1 < nsteps) {
				issue((t + 1) & 1, j2, vb(2 * t + 4), e_hi);
				if (nex)
					issue_ex(e_hi);
			}
			uint32_t y[2][8], sums = 0, valid2 = 0;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t tt = 2 * t + h;
				const uint32_t el = h ? e_mid : e_lo, eh = h ? e_hi : e_mid, jl = h ? j1 : j0;
				const uint64_t ka = A + (uint64_t) tt * XG + 2 * tid;
				const int64_t f = (int64_t) tt * XT - m;		// read index of the subtile's first element
				const bool interior = tt >= 1 && tt < t_hi4;
				const uint32_t valid = interior ? 0xffffu
								: tt < nt4 ? group_valid(ka, off0, rend) | (group_valid(ka + 1, off0, rend) << 8) : 0u;
				valid2 |= valid << (16 * h);
				const int64_t i0 = f + 16 * tid;
				uint32_t d[8];
				const uint32_t rel = (jl - el) - p_lo;	// stage offset of the subtile's first payload byte
				if (interior && eh == el) {
					// no exception in the subtile: the thread's 16 payload bytes sit at stage[al + rel + 16 tid]
					uint32_t w[4];
					if ((rel & 15) == 0)
						stage_words_aligned(s_stage + rel, al, tid, w);
					else
						stage_words_any(s_stage, al + rel + 16 * tid, w);
					unzigzag_bytes_x16(w, d);
				} else {
					uint32_t zz[8];
#pragma unroll
					for (int j = 0; j < 8; j++)
						zz[j] = 0;
					bool bytes = false;
					if (valid) {
						const int64_t jf = i0 + (__ffs(valid) - 1) - 1;	// v-index of the first valid element (-1: zd0)
						uint32_t e = el;
						if (eh > el) {
							uint32_t lo = el, hi = eh;
							const int64_t key = jf < 0 ? 0 : jf;
							while (lo < hi) {
								const uint32_t mid = (lo + hi) >> 1;
								if ((int64_t) xs[mid].x < key)
									lo = mid + 1;
								else
									hi = mid;
							}
							e = lo;
						}
						uint32_t nextpos = e < eh ? xs[e].x : 0xffffffffu;
						if (valid == 0xffffu && jf >= 0 && (uint64_t) nextpos >= (uint64_t) jf + 16) {
							// no exception among the thread's 16 values: 16 payload bytes in a row
							uint32_t w[4];
							stage_words_any(s_stage, al + ((uint32_t) jf - e - p_lo), w);
							unzigzag_bytes_x16(w, d);
							bytes = true;
						} else if (valid == 0xffffu && jf >= 0 && (e + 1 >= eh || (uint64_t) xs[e + 1].x >= (uint64_t) jf + 16)) {
							// ONE exception among the thread's 16 values, at element x: 15 payload bytes in a row,
							// those from x on moved up by one, the exception's value put in the gap (at rates
							// around 1e-3 this is the thread every other warp has: the element loop below would
							// hold the warp for some 200 instructions.  The same for up to three exceptions in a
							// loop was measured: 40.9 against 43.1 % of HBM peak at 1e-3, 25.2 against 23.9 at 1e-2)
							const uint32_t x = nextpos - (uint32_t) jf;
							uint32_t w[4], r[4];
							stage_words_any(s_stage, al + ((uint32_t) jf - e - p_lo), w);
#pragma unroll
							for (int j = 0; j < 4; j++) {
								const uint32_t up = __funnelshift_l(j ? w[j - 1] : 0u, w[j], 8);	// byte i = payload byte i - 1
								const uint32_t m = x <= 4u * j ? 0u : x >= 4u * j + 4 ? 0xffffffffu : (1u << (8 * (x - 4u * j))) - 1;
								r[j] = (w[j] & m) | (up & ~m);
							}
							unzigzag_bytes_x16(r, d);
							const uint32_t v = xs[e].y, dv = (((v >> 1) ^ (0u - (v & 1u))) & 0xffffu) * 0x00010001u;	// trans.c:80-83
#pragma unroll
							for (int j = 0; j < 8; j++) {	// (masks, not an indexed store: d stays in registers)
								const uint32_t hit = (x >> 1) == (uint32_t) j ? ((x & 1) ? 0xffff0000u : 0x0000ffffu) : 0u;
								d[j] = (d[j] & ~hit) | (dv & hit);
							}
							bytes = true;
						} else {
							const uint8_t *st = s_stage + al;
#pragma unroll
							for (int el16 = 0; el16 < 16; el16++) {
								if ((valid >> el16) & 1) {
									const int64_t j = i0 + el16 - 1;
									uint32_t z;
									if (j < 0) {
										z = zd0;
									} else if ((uint32_t) j == nextpos) {
										z = xs[e].y;
										e++;
										nextpos = e < eh ? xs[e].x : 0xffffffffu;
									} else {
										z = st[(uint32_t) j - e - p_lo];
									}
									zz[el16 >> 1] |= (z & 0xffffu) << (16 * (el16 & 1));
								}
							}
						}
					}
					if (!bytes) {
#pragma unroll
						for (int j = 0; j < 8; j++)
							d[j] = unzigzag16x2(zz[j]);
					}
				}
				// running sum inside the thread (mod 2^16 in each half), then over the CTA
				uint32_t c = 0;
#pragma unroll
				for (int j = 0; j < 8; j++) {
					y[h][j] = add16x2(d[j] * 0x00010001u, c);
					c = __byte_perm(y[h][j], 0, 0x3232);
				}
				sums |= (c & 0xffffu) << (16 * h);
			}
			uint32_t e2, tot2;
			block_scan_pair(sums, e2, tot2, s_warp);
			const uint32_t cwa = ((e2 + carry) & 0xffffu) * 0x00010001u;
			const uint32_t cwb = (((e2 >> 16) + tot2 + carry) & 0xffffu) * 0x00010001u;
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t cw = h ? cwb : cwa;
				const uint64_t ka = A + (uint64_t) (2 * t + h) * XG + 2 * tid;
#pragma unroll
				for (int g = 0; g < 2; g++) {
					const uint32_t vm = (valid2 >> (16 * h + 8 * g)) & 0xffu;
					uint32_t o[4];
#pragma unroll
					for (int j = 0; j < 4; j++) {
						o[j] = add16x2(y[h][4 * g + j], cw);
						if (q)
							o[j] = shl16x2(o[j], q);	// ex_zd.c:396-401
					}
					if (vm == 0xffu) {
						out4[ka + g] = make_uint4(o[0], o[1], o[2], o[3]);
					} else if (vm) {
#pragma unroll
						for (int e = 0; e < 8; e++)
							if ((vm >> e) & 1)
								out16[8 * (ka + g) + e] = (uint16_t) (o[e >> 1] >> (16 * (e & 1)));
					}
				}
			}
			carry = (carry + (tot2 & 0xffffu) + (tot2 >> 16)) & 0xffffu;
			e_lo = e_hi;
			// (no barrier here: the stage and s_ex are read before the scan's barrier and rewritten
			//  after it; s_warp / s_cnt are rewritten after the next step's first barrier)
		}
	}
}

}  // namespace pg
