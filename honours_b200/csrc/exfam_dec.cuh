// exfam_dec.cuh -- decode of the exception-coded methods (see exfam.cuh for the
// formats) and of the static Huffman payload.
//
//   ex_parse   (read)  header -> n, q, zd0, payload position; both exception lists
//                      decoded into a pool (sorted positions, values)
//   huff_decode(read)  SHUFF only: bitstream -> one-byte symbols in a scratch payload
//                      (self-synchronising parallel decode, see below)
//   ex_decode  (read)  one CTA walks one read tile by tile: payload + exceptions ->
//                      unzigzag -> running sum (trans.c:260-270) carried by the CTA from
//                      tile to tile -> << q -> samples; the next tile's payload bytes
//                      are requested before the current tile is processed
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
// payload + exceptions -> samples

__global__ void __launch_bounds__(NT, 6) ex_decode_kernel(const __grid_constant__ ExDecArgs a)
{
	// the tile's payload bytes, double buffered: tile t + 1 arrives by cp.async (no registers
	// held, no stage stores) while tile t is decoded
	__shared__ __align__(16) uint8_t s_stage2[2][XT + 64];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read, s_e;
	__shared__ uint2 s_ex[NT];	// the next 256 exceptions of the read (position, value)
	const int tid = threadIdx.x;
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
		const uint32_t nt = (uint32_t) ((m + n + XT - 1) / XT);
		const uint32_t nex = a.r_nex[r], q = a.r_q[r], zd0 = a.r_zd0[r];
		const uint2 *ex = a.pool + a.r_pool[r];
		const uint8_t *pbase = shuff ? a.paybuf + off0 : a.in + pay0;

		// v-indices of the read before tile t (ex_v_before), in 32 bits: 0 for the first tile,
		// everything for t >= nt, t * XT - m - 1 in between (below n - 1 there)
		const uint32_t nm1 = (uint32_t) (n - 1);
		auto vb = [&](uint32_t t) { return t == 0 ? 0u : t >= nt ? nm1 : t * (uint32_t) XT - m - 1u; };
		// may every 16-byte word the walk can touch be loaded without a look at the limits?
		const bool safe = pbase - 16 >= lim_lo && pbase + (uint64_t) nm1 + 48 <= lim_hi;
		// payload bytes of a tile -> stage `buf` (16-byte word tid, and word 256 by thread 0)
		auto issue = [&](uint32_t buf, uint32_t j_lo, uint32_t j_hi, uint32_t e_lo) {
			const uint8_t *src = pbase + (j_lo - e_lo);
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
			const uint32_t nw = (al + (j_hi - j_lo) + 15) >> 4;	// (a few bytes more than needed when
			const uint8_t *b16 = src - al;				//  the tile holds exceptions)
			uint4 *dst = reinterpret_cast<uint4 *>(s_stage2[buf]);
			if (safe || (b16 >= lim_lo && b16 + 16ull * nw <= lim_hi)) {
				if ((uint32_t) tid < nw)
					cp_async16(dst + tid, reinterpret_cast<const uint4 *>(b16) + tid);
				if (tid == 0 && nw > NT)
					cp_async16(dst + NT, reinterpret_cast<const uint4 *>(b16) + NT);
			} else {
				if ((uint32_t) tid < nw)
					dst[tid] = load_word_guarded(b16, tid, lim_lo, lim_hi);
				if (tid == 0 && nw > NT)
					dst[NT] = load_word_guarded(b16, NT, lim_lo, lim_hi);
			}
			cp_async_commit();
		};
		const uint32_t t_hi = (uint32_t) ((m + n) / XT);	// tiles [1, t_hi) lie inside the read
		// the read's next 256 exceptions travel with the payload: requested one tile ahead,
		// staged in shared memory, so the per-tile and per-thread searches never go to HBM
		uint2 exreg = make_uint2(0xffffffffu, 0u);
		auto issue_ex = [&](uint32_t e0) {
			exreg = e0 + tid < nex ? ex[e0 + tid] : make_uint2(0xffffffffu, 0u);
		};
		uint32_t e_lo = 0, carry = 0;
		uint32_t j_lo = 0, j_hi = vb(1), j_nx = vb(2);	// v-indices before tiles t, t + 1, t + 2
		issue(0, j_lo, j_hi, 0);
		if (nex)
			issue_ex(0);
		for (uint32_t t = 0; t < nt; t++, j_lo = j_hi, j_hi = j_nx, j_nx = vb(t + 2)) {
			const uint32_t p_lo = j_lo - e_lo;
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(pbase + p_lo) & 15);
			const uint8_t *s_stage = s_stage2[t & 1];	// complete once the barrier below has passed
			cp_async_wait_all();
			uint32_t e_hi = 0;
			const uint2 *xs = ex;	// where exceptions [e_lo, e_hi) are read from
			if (!nex) {
				__syncthreads();
			} else {
				s_ex[tid] = exreg;
				// exceptions before the end of this tile = staged positions below j_hi
				const uint32_t k = (uint32_t) __syncthreads_count(exreg.x < j_hi);
				if (k < NT) {
					e_hi = e_lo + k;
					xs = s_ex - e_lo;
				} else {	// more than 255 in one tile: lower_bound over the whole list
					if (tid == 0) {
						uint32_t lo = e_lo, hi = nex;
						while (lo < hi) {
							const uint32_t mid = (lo + hi) >> 1;
							if (ex[mid].x < j_hi)
								lo = mid + 1;
							else
								hi = mid;
						}
						s_e = lo;
					}
					__syncthreads();
					e_hi = s_e;
				}
			}
			if (t + 1 < nt) {
				issue((t + 1) & 1, j_hi, j_nx, e_hi);
				if (nex)
					issue_ex(e_hi);
			}
			// ---- zd values of the thread's 16 elements
			const uint64_t ka = A + (uint64_t) t * XG + 2 * tid;
			const int64_t f = (int64_t) t * XT - m;		// read index of the tile's first element
			const bool interior = t >= 1 && t < t_hi;
			const uint32_t valid = interior ? 0xffffu
							: group_valid(ka, off0, rend) | (group_valid(ka + 1, off0, rend) << 8);
			const int64_t i0 = f + 16 * tid;
			uint32_t zz[8];
			const uint8_t *st = s_stage + al;
			if (interior && e_hi == e_lo) {
				// no exception in the tile: the thread's 16 payload bytes sit at stage[al + 16 tid]
				stage_bytes16(s_stage, al, tid, zz);
			} else {
#pragma unroll
				for (int j = 0; j < 8; j++)
					zz[j] = 0;
				if (valid) {
					const int64_t jf = i0 + (__ffs(valid) - 1) - 1;	// v-index of the first valid element (-1: zd0)
					uint32_t e = e_lo;
					if (e_hi > e_lo) {
						uint32_t lo = e_lo, hi = e_hi;
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
					uint32_t nextpos = e < e_hi ? xs[e].x : 0xffffffffu;
					if (valid == 0xffffu && jf >= 0 && (uint64_t) nextpos >= (uint64_t) jf + 16) {
						// no exception among the thread's 16 values: 16 payload bytes in a row
						const uint32_t pp = al + ((uint32_t) jf - e - p_lo);
						const uint32_t *w32 = reinterpret_cast<const uint32_t *>(s_stage) + (pp >> 2);
						const uint32_t sh = (pp & 3) * 8;
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
					} else
#pragma unroll
					for (int el = 0; el < 16; el++) {
						if ((valid >> el) & 1) {
							const int64_t j = i0 + el - 1;
							uint32_t z;
							if (j < 0) {
								z = zd0;
							} else if ((uint32_t) j == nextpos) {
								z = xs[e].y;
								e++;
								nextpos = e < e_hi ? xs[e].x : 0xffffffffu;
							} else {
								z = st[(uint32_t) j - e - p_lo];
							}
							zz[el >> 1] |= (z & 0xffffu) << (16 * (el & 1));
						}
					}
				}
			}
			// ---- unzigzag, running sum inside the thread (mod 2^16 in each half), then over the CTA
			uint32_t y[8], c = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) {
				const uint32_t d = unzigzag16x2(zz[j]);
				y[j] = add16x2(d * 0x00010001u, c);
				c = __byte_perm(y[j], 0, 0x3232);
			}
			uint32_t c1[1] = {c & 0xffffu}, ex1[1], tot;
			block_scan<1>(c1, ex1, tot, s_warp);
			const uint32_t cw = ((ex1[0] + carry) & 0xffffu) * 0x00010001u;
			uint32_t o[8];
#pragma unroll
			for (int j = 0; j < 8; j++) {
				o[j] = add16x2(y[j], cw);
				if (q)
					o[j] = shl16x2(o[j], q);	// ex_zd.c:396-401
			}
#pragma unroll
			for (int h = 0; h < 2; h++) {
				const uint32_t vm = (valid >> (8 * h)) & 0xffu;
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
			e_lo = e_hi;
			// (no barrier here: the stage and s_ex are read before the scan's barrier and
			//  rewritten after it; s_warp is rewritten after the next tile's first barrier)
		}
	}
}

}  // namespace pg
