// huff_fused.cuh -- static-Huffman decode WITH the encoder's side-car: bits -> samples in one
// kernel (huffman.c:1219-1265 + press.c:4455-4498 + trans.c:260-270 fused).
//
// The reference's bitstream carries a symbol count and nothing else, so a parallel decoder has
// no entry points (huff_decode_kernel, exfam_dec.cuh, finds them by speculation).  When the
// stream was written by this library, press leaves beside the container one uint16 per 16
// payload symbols: the low 16 bits of the bit offset of payload symbol 16 j inside the read's
// bitstream (entries of read r at index (offsets[r] >> 4) + r).  Differences of neighbouring
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

constexpr int HF_GROUP = 32;			// payload symbols per thread and step
constexpr int HF_SUB = 16;			// payload symbols per side-car entry
constexpr int HF_CHUNK = NT * HF_GROUP;		// symbols per decode step (8192)
constexpr int HF_RING = 16384;			// symbol ring, bytes (power of two, >= 2 steps)
constexpr int HF_BITS = 8192;			// bitstream bytes staged per step on the fast path (8 bits / symbol)
constexpr int HF_L1 = 13;			// first-level table bits (the device tables of exfam_dec.cuh are 12 + sub-tables)
constexpr int HF_BITS_ALLOC = HF_BITS + 160;	// + what a decode running on garbage can over-read (26 words)

// Codes longer than the first-level table (0.5 % of the NA12878 symbols) go through a second
// level kept in shared memory: l2[k][next HF_L2 bits] = length << 8 | symbol for codes of up to
// 12 + HF_L2 bits, 0 where the code is longer still (third level: the global sub-tables).
constexpr int HF_L2 = 6;
constexpr int HF_L2_TABLES = 24;	// prefixes with a second level in shared memory (NA12878: 20)

__device__ __noinline__ uint32_t hf_long(uint32_t e, uint32_t w, const uint16_t *sub, uint32_t sub_bits)
{
	const uint32_t e1 = __ldg(sub + (((e & 0x7fffu) << sub_bits) | ((w >> HD_L1) & ((1u << sub_bits) - 1))));
	return e1;	// length << 8 | symbol
}

// two signed bytes -> two 16-bit lanes, sign extended (prmt's sign-replicate selectors)
__device__ __forceinline__ uint32_t sext8x2(uint32_t v, uint32_t sel)
{
	uint32_t r;
	asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(0u), "r"(sel));
	return r;
}
// zigzag symbol -> its delta as a signed byte (trans.c:80-83 on a one-byte value)
__device__ __forceinline__ uint32_t hf_delta8(uint32_t sym)
{
	return ((sym >> 1) ^ (0u - (sym & 1u))) & 0xffu;
}

// 63 stream bits from stage position pos on (positions name the bit before a code)
__device__ __forceinline__ void hf_window(uint32_t bits_sa, uint32_t pos, uint32_t &lo, uint32_t &hi)
{
	const uint32_t a = bits_sa + 4 * (pos >> 5);
	uint32_t w0, w1, w2;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a));
	asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(a));
	asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(a));
	lo = __funnelshift_r(w0, w1, pos);	// (shift taken mod 32)
	hi = __funnelshift_r(w1, w2, pos);
}

// a code longer than the first-level table, out of line (inline, or with the 16 rounds rolled into
// a loop of 4, the kernel is no faster: 16.6 % / 21.7 % against 21.4 %; inline and unrolled it misses
// the instruction cache, no_inst 30 % of the stall samples; a 14-bit first level at 3 CTAs/SM: 19.2 %): fresh window, second level (shared) or
// third (global), fresh window after it -> (entry, lo, hi, pos)
__device__ __noinline__ uint4 hf_long_code(uint32_t e, uint32_t pos, uint32_t bits_sa, const uint16_t *l2, uint32_t nl2,
					   const uint16_t *sub, uint32_t sub_bits)
{
	uint32_t lo, hi;
	hf_window(bits_sa, pos, lo, hi);
	const uint32_t k = e & 0x7fffu, win = lo >> 1;
	e = k < nl2 ? l2[(k << HF_L2) | ((win >> HD_L1) & ((1u << HF_L2) - 1))] : 0u;
	if (e == 0) {	// third level; raw symbols there
		const uint32_t e1 = __ldg(sub + ((k << sub_bits) | ((win >> HD_L1) & ((1u << sub_bits) - 1))));
		e = (e1 & 0xff00u) | hf_delta8(e1 & 0xffu);
	}
	pos += e >> 8;
	hf_window(bits_sa, pos, lo, hi);
	return make_uint4(e, lo, hi, pos);
}

// A decode chain: 63 bits of the stream from position `pos` on live in (lo, hi); positions name
// the bit BEFORE the code, so lo & 0x3ffe is the 13-bit window times the table's entry size.
// The table entries are length << 8 | DELTA (the symbol already un-zigzagged, a signed byte).
// A code of the first level takes <= 13 bits, so after a refill four codes can be decoded
// without another look at memory; longer codes (rare) refill before and after themselves.
struct HfChain {
	uint32_t lo, hi, pos;
	__device__ __forceinline__ void refill(uint32_t bits_sa) { hf_window(bits_sa, pos, lo, hi); }
	// lut_sa: shared-memory address of the first-level table
	__device__ __forceinline__ uint32_t step(uint32_t bits_sa, uint32_t lut_sa, const uint16_t *l2, uint32_t nl2,
						 const uint16_t *sub, uint32_t sub_bits)
	{
		int32_t es;	// sign-extending load: the long-code marker (bit 15) becomes the sign
		asm volatile("ld.shared.s16 %0, [%1];" : "=r"(es) : "r"(lut_sa + (lo & ((2u << HF_L1) - 2u))));
		const uint32_t e = (uint32_t) es;
		if (__builtin_expect(es < 0, 0)) {
			const uint4 r = hf_long_code(e, pos, bits_sa, l2, nl2, sub, sub_bits);
			lo = r.y;
			hi = r.z;
			pos = r.w;
			return r.x;
		}
		const uint32_t len = e >> 8;
		pos += len;
		lo = __funnelshift_r(lo, hi, len);
		hi >>= len;
		return e;
	}
};

// Decode the thread's 32 codes as two independent chains of 16 (the side-car has an entry every
// 16 symbols: twice the instruction-level parallelism of one chain of 32).  (Four chains of 8 on a
// side-car entry every 8 symbols were measured: 16.3 % of HBM peak at 4 CTAs/SM, where they spill,
// 17.7 % at 3 without spills, against 21.4 %: the kernel gains from resident warps, not from longer
// independent instruction streams, and the encoder pays for the second entry per thread.)  Delta i goes to byte
// i of out[].  enda / endb = positions after the 16th code of each chain; LAST: the positions
// after codes lima - 1 / limb - 1 instead (the read's last groups hold fewer symbols; what is
// decoded past them is never used).
template <bool LAST>
__device__ __forceinline__ void hf_decode2x16(uint32_t bits_sa, uint32_t pa, uint32_t pb, uint32_t lut,
					      const uint16_t *l2, uint32_t nl2, const uint16_t *sub, uint32_t sub_bits,
					      uint32_t lima, uint32_t limb, uint32_t (&out)[8], uint32_t &enda,
					      uint32_t &endb)
{
	HfChain A, B;
	A.pos = pa;
	B.pos = pb;
	uint32_t eva = pa, evb = pb;
#pragma unroll
	for (int j = 0; j < 8; j++)
		out[j] = 0;
#pragma unroll
	for (int i = 0; i < 16; i++) {
		if (i % 4 == 0) {
			A.refill(bits_sa);
			B.refill(bits_sa);
		}
		const uint32_t sel = (i & 3) == 0 ? 0x3214 : (i & 3) == 1 ? 0x3240 : (i & 3) == 2 ? 0x3410 : 0x4210;
		const uint32_t ea = A.step(bits_sa, lut, l2, nl2, sub, sub_bits);
		const uint32_t eb = B.step(bits_sa, lut, l2, nl2, sub, sub_bits);
		out[i >> 2] = __byte_perm(out[i >> 2], ea, sel);
		out[4 + (i >> 2)] = __byte_perm(out[4 + (i >> 2)], eb, sel);
		if (LAST && (uint32_t) i < lima)
			eva = A.pos;
		if (LAST && (uint32_t) i < limb)
			evb = B.pos;
	}
	enda = LAST ? eva : A.pos;
	endb = LAST ? evb : B.pos;
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
		uint32_t e = lut[w & ((1u << HF_L1) - 1)];	// length << 8 | delta
		if (e & 0x8000u) {
			e = hf_long(e, w, sub, sub_bits);	// length << 8 | symbol
			e = (e & 0xff00u) | hf_delta8(e & 0xffu);
		}
		ring[(ring_off + i) & (HF_RING - 1)] = (uint8_t) e;
		pos += (e >> 8) ? (e >> 8) : 1u;	// (a complete code has no zero length; never loop in place)
	}
	return pos;
}

// 16 bytes of the ring (deltas, signed) starting at payload index p + 16 tid (p the same in every
// thread), two per word sign extended to 16 bits: two 128-bit loads and a byte shift (cf. stage_bytes16)
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
	zz[0] = sext8x2(v0, 0x9180); zz[1] = sext8x2(v0, 0xb3a2);
	zz[2] = sext8x2(v1, 0x9180); zz[3] = sext8x2(v1, 0xb3a2);
	zz[4] = sext8x2(v2, 0x9180); zz[5] = sext8x2(v2, 0xb3a2);
	zz[6] = sext8x2(v3, 0x9180); zz[7] = sext8x2(v3, 0xb3a2);
}

struct HfSmem {		// dynamic shared memory (> 48 KB: opt-in at launch)
	alignas(16) uint8_t ring[HF_RING];		// decoded symbols, by payload index mod HF_RING
	alignas(16) uint8_t bits[2][HF_BITS_ALLOC];	// the step's bitstream bytes, double buffered
	alignas(16) uint16_t l2[HF_L2_TABLES << HF_L2];	// second level (see hf_decode32)
	alignas(16) uint2 ex[NT];			// the next 256 exceptions of the read (position, value)
	alignas(16) uint32_t warp[NW];			// bit-length scans
	alignas(16) uint32_t warp2[NW];			// value scans
	uint32_t read, e;
};

__global__ void __launch_bounds__(NT, 4) huff_fused_decode_kernel(const __grid_constant__ ExDecArgs a)
{
	extern __shared__ __align__(16) uint8_t hf_smem_raw[];
	__shared__ __align__(16) uint16_t s_lut[1 << HF_L1];	// first level, one symbol per entry (static: immediate addressing)
	HfSmem &sm = *reinterpret_cast<HfSmem *>(hf_smem_raw);
	uint8_t *s_ring = sm.ring;
	uint8_t (*s_bits)[HF_BITS_ALLOC] = sm.bits;
	uint32_t *s_warp = sm.warp, *s_warp2 = sm.warp2;
	uint32_t &s_read = sm.read, &s_e = sm.e;
	uint2 *s_ex = sm.ex;
	const int tid = threadIdx.x;
	const uint16_t *sub = a.huff_sub;
	const uint32_t sub_bits = a.huff_sub_bits;
	// first level of HF_L1 bits from the device tables (12-bit first level: symbol | symbol2 << 8 |
	// length2 << 16 | length << 24, or bit 31 | k, + sub-tables sub[k][bits after the first 12])
	for (int i = tid; i < (1 << HF_L1); i += NT) {
		const uint32_t e = a.huff_dec[i & ((1 << HD_L1) - 1)];
		uint32_t v;
		if (!(e >> 31)) {
			v = hf_delta8(e & 0xffu) | ((e >> 24) << 8);
		} else {
			const uint32_t k = e & 0x7fffu, x = (uint32_t) i >> HD_L1;	// the HF_L1 - 12 further bits, zeros above
			const uint32_t e1 = sub_bits >= (uint32_t) (HF_L1 - HD_L1) ? sub[(k << sub_bits) | x]
										   : sub[(k << sub_bits) | (x & ((1u << sub_bits) - 1))];
			v = (e1 >> 8) <= (uint32_t) HF_L1 ? (e1 & 0xff00u) | hf_delta8(e1 & 0xffu) : 0x8000u | k;
		}
		s_lut[i] = (uint16_t) v;
	}
	// second level from the global sub-tables: entry (k, x) = sub[k][x with zeros above] when that code
	// is no longer than 12 + HF_L2 bits (then every completion of x decodes to it), else 0
	const uint32_t nl2 = a.huff_nsub < (uint32_t) HF_L2_TABLES ? a.huff_nsub : (uint32_t) HF_L2_TABLES;
	for (uint32_t i = tid; i < (nl2 << HF_L2); i += NT) {
		const uint32_t k = i >> HF_L2, x = i & ((1u << HF_L2) - 1);
		uint32_t e1 = sub[(k << sub_bits) | (sub_bits >= (uint32_t) HF_L2 ? x : x & ((1u << sub_bits) - 1))];
		if ((e1 >> 8) > (uint32_t) (HD_L1 + HF_L2))
			e1 = 0;
		else
			e1 = (e1 & 0xff00u) | hf_delta8(e1 & 0xffu);
		sm.l2[i] = (uint16_t) e1;
	}
	// (laundered through an opaque move: the compiler otherwise recomputes the window address,
	//  S2R + LEA, at every table lookup)
	uint32_t bits_sa0 = (uint32_t) __cvta_generic_to_shared(sm.bits[0]);
	uint32_t lut_sa = (uint32_t) __cvta_generic_to_shared(s_lut);
	asm volatile("mov.u32 %0, %0;" : "+r"(bits_sa0));
	asm volatile("mov.u32 %0, %0;" : "+r"(lut_sa));
	const uint8_t *limit = a.in + a.in_bytes;
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
		const uint32_t nm1 = (uint32_t) (n - 1);
		const uint32_t msym = nm1 - nex;			// payload symbols (ex_parse checked the count field)
		const uint8_t *src = a.in + pay0;			// the read's bitstream
		const uint64_t nbytes = __ldg(a.in_offsets + r + 1) - pay0, nbits = 8 * nbytes;
		const uint16_t *sc = a.sidecar + (off0 >> 4) + r;
		const uint32_t ngroups = (msym + HF_GROUP - 1) / HF_GROUP, nsteps = (ngroups + NT - 1) / NT;
		const uint32_t nent = (msym + HF_SUB - 1) / HF_SUB;	// side-car entries of the read

		// ---- decode steps ------------------------------------------------------------------
		// raw side-car entries of the step after the staged one: the thread's two halves and the next
		// thread's first (an entry past the read's last: length unknown, taken as 0)
		uint32_t en0 = 0, en1 = 0, en2 = 0;
		auto load_entries = [&](uint32_t s) {
			const uint32_t j = 2 * (s * NT + tid);
			en0 = j < nent ? __ldg(sc + j) : 0u;
			en1 = j + 1 < nent ? __ldg(sc + j + 1) : en0;
			en2 = j + 2 < nent ? __ldg(sc + j + 2) : en1;
		};
		// bit range [b_lo, b_hi) of the bitstream -> stage `buf`; returns false when it does not fit
		auto issue_bits = [&](uint32_t buf, uint64_t b_lo, uint64_t b_hi) -> bool {
			const uint64_t byte0 = b_lo >> 3, L = ((b_hi + 7) >> 3) - byte0;
			const uint8_t *s0 = src + byte0;
			const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(s0) & 15);
			if (al + L > (uint64_t) HF_BITS - 16)
				return false;
			const uint32_t nw = (uint32_t) ((al + L + 15) >> 4);
			const uint8_t *b16 = s0 - al;
			uint4 *dst = reinterpret_cast<uint4 *>(s_bits[buf]) + 1;	// (16 bytes of lead: see hf_decode32)
			if (b16 + 16ull * nw <= limit) {	// (b16 >= a.in: the container is 16-byte aligned)
#pragma unroll
				for (int j = 0; j < (HF_BITS / 16 + NT - 1) / NT; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						cp_async16(dst + w, reinterpret_cast<const uint4 *>(b16) + w);
				}
			} else {
#pragma unroll
				for (int j = 0; j < (HF_BITS / 16 + NT - 1) / NT; j++) {
					const uint32_t w = tid + j * NT;
					if (w < nw)
						dst[w] = load_word_guarded(b16, w, a.in, limit);
				}
			}
			cp_async_commit();
			return true;
		};
		// lengths of step s from the entries in registers -> (exclusive offset, total), bit range staged
		uint32_t ex_cur = 0, len_cur = 0, tot_cur = 0;	// of the step to decode next (len: first half | second half << 16)
		bool fast_cur = true;
		uint64_t B = 0;			// bit position of the step to decode next
		uint32_t c = 0, decoded = 0;	// next step, payload symbols in the ring so far
		auto clampb = [&](uint64_t b) { return b < nbits ? b : nbits; };
		if (nsteps) {
			load_entries(0);
			len_cur = ((en1 - en0) & 0xffffu) | ((en2 - en1) << 16);
			uint32_t c1[1] = {(len_cur & 0xffffu) + (len_cur >> 16)}, e1[1];
			block_scan<1>(c1, e1, tot_cur, s_warp);
			ex_cur = e1[0];
			fast_cur = issue_bits(0, 0, nsteps == 1 ? nbits : clampb(tot_cur));
			if (nsteps > 1)
				load_entries(1);
			__syncthreads();	// s_warp is reused by the first scan-ahead
		}
		auto decode_step = [&]() {
			const uint32_t buf = c & 1;
			cp_async_wait_all();
			uint32_t ex_n = 0, len_n = 0, tot_n = 0;
			bool fast_n = true;
			if (c + 1 < nsteps) {	// scan ahead: step c + 1's lengths, its bits requested, step c + 2's entries
				len_n = ((en1 - en0) & 0xffffu) | ((en2 - en1) << 16);
				uint32_t c1[1] = {(len_n & 0xffffu) + (len_n >> 16)}, e1[1];
				block_scan<1>(c1, e1, tot_n, s_warp);	// (its barrier also publishes the staged bits)
				ex_n = e1[0];
				const uint64_t b_lo = clampb(B + tot_cur);
				fast_n = issue_bits(buf ^ 1, b_lo, c + 2 == nsteps ? nbits : clampb(B + tot_cur + tot_n));
				if (c + 2 < nsteps)
					load_entries(c + 2);
			} else {
				__syncthreads();
			}
			const uint32_t g = c * NT + tid;
			const bool act = g < ngroups;
			const uint32_t left = act ? msym - HF_GROUP * g : 0u;	// symbols of the read from this group on
			const uint32_t lima = left < HF_SUB ? left : HF_SUB;
			const uint32_t limb = left < HF_SUB ? 0u : (left - HF_SUB < HF_SUB ? left - HF_SUB : HF_SUB);
			const uint32_t la = len_cur & 0xffffu, lb = len_cur >> 16;
			const uint64_t Bc = clampb(B);
			bool wrong = false;
			if (act && fast_cur) {
				// the staged range starts at the byte holding bit Bc, in its 16-byte phase
				const uint32_t al = (uint32_t) (reinterpret_cast<uintptr_t>(src + (Bc >> 3)) & 15);
				// (+ 128: the 16 bytes of lead; - 1: positions name the bit before the code)
				const uint32_t base = 127u + 8 * al + (uint32_t) (Bc & 7);	// stage position of bitstream bit Bc
				uint32_t pa = base + ex_cur, pb = pa + la;
				if (pa > 8u * HF_BITS)
					pa = 8u * HF_BITS;	// (a side-car that does not belong to this stream)
				if (pb > 8u * HF_BITS)
					pb = 8u * HF_BITS;
				const uint32_t bsa = bits_sa0 + buf * (uint32_t) HF_BITS_ALLOC;
				uint32_t sym[8], enda, endb;
				if (c + 1 == nsteps)
					hf_decode2x16<true>(bsa, pa, pb, lut_sa, sm.l2, nl2, sub, sub_bits, lima, limb, sym, enda, endb);
				else
					hf_decode2x16<false>(bsa, pa, pb, lut_sa, sm.l2, nl2, sub, sub_bits, lima, limb, sym, enda, endb);
				uint4 *rd = reinterpret_cast<uint4 *>(s_ring + ((HF_GROUP * g) & (HF_RING - 1)));
				rd[0] = make_uint4(sym[0], sym[1], sym[2], sym[3]);
				rd[1] = make_uint4(sym[4], sym[5], sym[6], sym[7]);
				// every half must end where the next one starts; the read's last one inside the bitstream
				if (left > HF_SUB)
					wrong = enda != pb;
				else
					wrong = (uint64_t) (enda - base) + Bc > nbits;
				if (left > HF_GROUP)
					wrong = wrong || endb - pb != lb;
				else if (left > HF_SUB)
					wrong = wrong || (uint64_t) (endb - base) + Bc > nbits;
			} else if (act) {
				const uint64_t b0 = clampb(Bc + ex_cur);
				const uint64_t e = hf_decode_global(src, nbytes, b0, s_lut, sub, sub_bits, lima + limb, s_ring, HF_GROUP * g);
				wrong = left > HF_GROUP ? e - b0 != (uint64_t) la + lb : e > nbits;
			}
			if (wrong)
				atomicOr(&a.hdr->error, ERR_MALFORMED);
			B += tot_cur;
			c++;
			decoded = (uint64_t) c * HF_CHUNK < msym ? c * (uint32_t) HF_CHUNK : msym;
			ex_cur = ex_n;
			len_cur = len_n;
			tot_cur = tot_n;
			fast_cur = fast_n;
			__syncthreads();	// the symbols are in the ring; s_warp may be written again
		};

		// ---- tiles of 4096 samples: symbols + exceptions -> samples (cf. ex_decode_kernel) ----------
		auto vb = [&](uint32_t tt) { return tt == 0 ? 0u : tt >= nt ? nm1 : tt * (uint32_t) XT - m - 1u; };
		const uint32_t t_hi = (uint32_t) ((m + n) / XT);	// tiles [1, t_hi) lie inside the read
		uint2 exreg = make_uint2(0xffffffffu, 0u);
		auto issue_ex = [&](uint32_t e0) {
			exreg = e0 + tid < nex ? ex[e0 + tid] : make_uint2(0xffffffffu, 0u);
		};
		uint32_t e_lo = 0, carry = 0;
		uint32_t j_lo = 0, j_hi = vb(1);	// v-indices before tiles t, t + 1
		if (nex)
			issue_ex(0);
		const uint32_t *ring32 = reinterpret_cast<const uint32_t *>(s_ring);
		for (uint32_t t = 0; t < nt; t++, j_lo = j_hi, j_hi = vb(t + 1)) {
			const uint32_t p_lo = j_lo - e_lo;	// payload index of the tile's first symbol
			uint32_t e_hi = 0;
			const uint2 *xs = ex;	// where exceptions [e_lo, e_hi) are read from
			if (!nex) {
				__syncthreads();
			} else {
				s_ex[tid] = exreg;
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
			// the symbols of this tile: payload indices [p_lo, j_hi - e_hi)
			const uint32_t need = j_hi - e_hi < msym ? j_hi - e_hi : msym;
			while (decoded < need)
				decode_step();
			if (nex && t + 1 < nt) {
				// (every thread has read s_ex / exreg above; the searches below read s_ex before the
				//  value scan's barrier, the next tile rewrites it after that barrier)
				issue_ex(e_hi);
			}
			const uint64_t ka = A + (uint64_t) t * XG + 2 * tid;
			const int64_t f = (int64_t) t * XT - m;		// read index of the tile's first element
			const bool interior = t >= 1 && t < t_hi;
			const uint32_t valid = interior ? 0xffffu
							: group_valid(ka, off0, rend) | (group_valid(ka + 1, off0, rend) << 8);
			const int64_t i0 = f + 16 * tid;
			uint32_t zz[8];
			if (interior && e_hi == e_lo) {
				ring_bytes16(s_ring, p_lo, tid, zz);
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
						// no exception among the thread's 16 values: 16 symbols in a row
						const uint32_t pp = (uint32_t) jf - e;	// payload index of the first one
						const uint32_t sh = (pp & 3) * 8;
						uint32_t w[5];
#pragma unroll
						for (int j = 0; j < 5; j++)
							w[j] = ring32[((pp >> 2) + j) & (HF_RING / 4 - 1)];
#pragma unroll
						for (int j = 0; j < 4; j++) {
							const uint32_t v = __funnelshift_r(w[j], w[j + 1], sh);
							zz[2 * j] = sext8x2(v, 0x9180);
							zz[2 * j + 1] = sext8x2(v, 0xb3a2);
						}
					} else if (valid == 0xffffu && jf >= 0 && (e + 1 >= e_hi || (uint64_t) xs[e + 1].x >= (uint64_t) jf + 16)) {
						// ONE exception among the thread's 16 values, at element x: 15 symbols in a row, those
						// from x on moved up by one, the exception's delta put in the gap (cf. ex_decode2_kernel)
						const uint32_t x = nextpos - (uint32_t) jf;
						const uint32_t pp = (uint32_t) jf - e;	// payload index of the first one
						const uint32_t sh = (pp & 3) * 8;
						uint32_t w[5], b[4];
#pragma unroll
						for (int j = 0; j < 5; j++)
							w[j] = ring32[((pp >> 2) + j) & (HF_RING / 4 - 1)];
#pragma unroll
						for (int j = 0; j < 4; j++)
							b[j] = __funnelshift_r(w[j], w[j + 1], sh);
#pragma unroll
						for (int j = 0; j < 4; j++) {
							const uint32_t up = __funnelshift_l(j ? b[j - 1] : 0u, b[j], 8);	// byte i = symbol i - 1
							const uint32_t m = x <= 4u * j ? 0u : x >= 4u * j + 4 ? 0xffffffffu : (1u << (8 * (x - 4u * j))) - 1;
							const uint32_t r = (b[j] & m) | (up & ~m);
							zz[2 * j] = sext8x2(r, 0x9180);
							zz[2 * j + 1] = sext8x2(r, 0xb3a2);
						}
						const uint32_t v = xs[e].y, dv = (((v >> 1) ^ (0u - (v & 1u))) & 0xffffu) * 0x00010001u;	// trans.c:80-83
#pragma unroll
						for (int j = 0; j < 8; j++) {	// (masks, not an indexed store: zz stays in registers)
							const uint32_t hit = (x >> 1) == (uint32_t) j ? ((x & 1) ? 0xffff0000u : 0x0000ffffu) : 0u;
							zz[j] = (zz[j] & ~hit) | (dv & hit);
						}
					} else
#pragma unroll
					for (int el = 0; el < 16; el++) {
						if ((valid >> el) & 1) {
							const int64_t j = i0 + el - 1;
							uint32_t z;	// the element's delta, 16 bits
							if (j < 0) {
								z = (zd0 >> 1) ^ (0u - (zd0 & 1u));	// trans.c:80-83
							} else if ((uint32_t) j == nextpos) {
								z = (xs[e].y >> 1) ^ (0u - (xs[e].y & 1u));
								e++;
								nextpos = e < e_hi ? xs[e].x : 0xffffffffu;
							} else {
								z = (uint32_t) (int32_t) (int8_t) s_ring[((uint32_t) j - e) & (HF_RING - 1)];
							}
							zz[el >> 1] |= (z & 0xffffu) << (16 * (el & 1));
						}
					}
				}
			}
			// ---- unzigzag, running sum inside the thread (mod 2^16 in each half), then over the CTA
			uint32_t y[8], cs = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) {
				y[j] = add16x2(zz[j] * 0x00010001u, cs);	// (zz: deltas already, see the decode table)
				cs = __byte_perm(y[j], 0, 0x3232);
			}
			uint32_t c1[1] = {cs & 0xffffu}, ex1[1], tot;
			block_scan<1>(c1, ex1, tot, s_warp2);
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
		}
	}
}

}  // namespace pg
