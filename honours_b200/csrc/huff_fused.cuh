// huff_fused.cuh -- static-Huffman decode WITH the encoder's side-car: bits -> samples in one
// kernel (huffman.c:1219-1265 + press.c:4455-4498 + trans.c:260-270 fused).
//
// The reference's bitstream carries a symbol count and nothing else, so a parallel decoder has
// no entry points (huff_decode_kernel, exfam_dec.cuh, finds them by speculation).  When the
// stream was written by this library, press leaves beside the container one uint16 per 32
// payload symbols: the low 16 bits of the bit offset of payload symbol 32 g inside the read's
// bitstream (entries of read r at index (offsets[r] >> 5) + r).  Differences of neighbouring
// entries are the groups' bit lengths (<= 32 x 24 bits), so a block scan over 256 entries gives
// every thread the exact start of its own group of 32 codes:
//
//   entries of step c + 2 -> registers
//   entries of step c + 1 -> lengths -> block scan -> bit range of the step -> cp.async of
//                            those bitstream bytes into the idle half of a double buffer
//   step c                -> every thread decodes its 32 codes (one table lookup per code, the
//                            window slides over three staged words held in registers) and
//                            leaves 32 symbol bytes in a 16 KB ring; a thread whose decode does
//                            not end where the next group starts flags the side-car as wrong
//   4096-sample tiles     -> symbols from the ring (+ the read's exceptions) -> unzigzag ->
//                            running sum carried by the CTA -> << q -> 128-bit stores, exactly as
//                            ex_decode_kernel does from a staged byte payload
//
// so no symbol goes through global memory and nothing is decoded twice.  One CTA per read
// (ticket), like the other decoders.
#pragma once

#include "exfam_dec.cuh"

namespace pg {

constexpr int HF_GROUP = 32;			// payload symbols per side-car entry
constexpr int HF_CHUNK = NT * HF_GROUP;		// symbols per decode step (8192)
constexpr int HF_RING = 16384;			// symbol ring, bytes (power of two, >= 2 steps)
constexpr int HF_BITS = 12288;			// bitstream bytes staged per step on the fast path (12 bits / symbol)
constexpr int HF_BITS_ALLOC = HF_BITS + 160;	// + what a decode running on garbage can over-read (26 words)

// a code longer than the first-level table: sub-table k = e & 0x7fff, indexed by the bits after the first 12
__device__ __noinline__ uint32_t hf_long(uint32_t e, uint32_t w, const uint16_t *sub, uint32_t sub_bits)
{
	const uint32_t e1 = __ldg(sub + (((e & 0x7fffu) << sub_bits) | ((w >> HD_L1) & ((1u << sub_bits) - 1))));
	return e1;	// length << 8 | symbol
}

// Decode 32 codes starting at stage bit `bitpos` (the stage is an array of 32-bit words in shared
// memory); symbol i goes to byte i of out[].  Returns the bit position after the 32nd code;
// endv = the position after code lim - 1 (LAST only: groups at the end of a read hold fewer
// symbols; what is decoded past them is never used).
template <bool LAST>
__device__ __forceinline__ uint32_t hf_decode32(const uint32_t *bits32, uint32_t bitpos, const uint16_t *lut,
						 const uint16_t *sub, uint32_t sub_bits, uint32_t lim, uint32_t (&out)[8],
						 uint32_t &endv)
{
	uint32_t wi = bitpos >> 5, pos = bitpos;
	uint32_t w0 = bits32[wi], w1 = bits32[wi + 1], w2 = bits32[wi + 2];
	uint32_t ev = bitpos;
#pragma unroll
	for (int j = 0; j < 8; j++)
		out[j] = 0;
#pragma unroll
	for (int i = 0; i < 32; i++) {
		const uint32_t win = __funnelshift_r(w0, w1, pos & 31u);
		uint32_t e = lut[win & ((1u << HD_L1) - 1)];
		if (__builtin_expect((e & 0x8000u) != 0, 0))
			e = hf_long(e, win, sub, sub_bits);
		out[i >> 2] |= (e & 0xffu) << (8 * (i & 3));
		pos += e >> 8;
		if (LAST && (uint32_t) i < lim)
			ev = pos;
		if ((pos >> 5) != wi) {		// (a code has <= 24 bits: at most one word is crossed)
			wi++;
			w0 = w1;
			w1 = w2;
			w2 = bits32[wi + 2];
		}
	}
	endv = ev;
	return pos;
}

// The same from global memory, code by code, for steps whose bits do not fit the stage (a run
// of very long codes): slow, exact.  `lim` symbols go straight to the ring; returns the bit
// position after them.
__device__ __noinline__ uint64_t hf_decode_global(const uint8_t *src, uint64_t nbytes, uint64_t bit0,
						  const uint16_t *lut, const uint16_t *sub, uint32_t sub_bits,
						  uint32_t lim, uint8_t *ring, uint32_t ring_off)
{
	uint64_t pos = bit0;
#pragma unroll 1
	for (uint32_t i = 0; i < lim; i++) {
		const uint64_t b0 = pos >> 3;
		uint64_t w64 = 0;
#pragma unroll 1
		for (int b = 0; b < 5; b++)	// 5 bytes cover 7 + 24 bits
			if (b0 + b < nbytes)
				w64 |= (uint64_t) src[b0 + b] << (8 * b);
		const uint32_t w = (uint32_t) (w64 >> (pos & 7));
		uint32_t e = lut[w & ((1u << HD_L1) - 1)];
		if (e & 0x8000u)
			e = hf_long(e, w, sub, sub_bits);
		ring[(ring_off + i) & (HF_RING - 1)] = (uint8_t) e;
		pos += (e >> 8) ? (e >> 8) : 1u;	// (a complete code has no zero length; never loop in place)
	}
	return pos;
}

// 16 bytes of the ring starting at payload index p + 16 tid (p the same in every thread), as zd
// values two per word: two 128-bit loads and a byte shift (cf. stage_bytes16)
__device__ __forceinline__ void ring_bytes16(const uint8_t *ring, uint32_t p, int tid, uint32_t (&zz)[8])
{
	const uint32_t al = p & 15u;
	const uint32_t i0 = ((p >> 4) + tid) & (HF_RING / 16 - 1), i1 = (i0 + 1) & (HF_RING / 16 - 1);
	const uint4 a = reinterpret_cast<const uint4 *>(ring)[i0], b = reinterpret_cast<const uint4 *>(ring)[i1];
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
