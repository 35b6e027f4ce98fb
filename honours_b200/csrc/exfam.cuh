// exfam.cuh -- the exception-coded methods: one-byte zigzag-delta payload with the
// values > 255 ("exceptions") pulled out into a header.
//
//   VBE21_ZD        press.c:2679-2765, 3411-3448   raw u32 positions / u16 values
//   VBSSE21_ZD      press.c:3186-3360, 3543-3580   svb32 position gaps / svb16 values - 256
//   EX_ZD           ex_zd.c:9-515, press.c:8461-8501   qts shift, svb32 gaps / svb32 values - 256
//   SHUFF_VBE21_ZD  press.c:4409-4498, huffman.c:848-887   VBE21 header + static Huffman payload
//
// Every stream is header first, and the header's length depends on ALL exceptions
// of the read, so no byte of a read's payload can be placed before the whole read
// has been seen: encode is two sweeps over the samples with one scan in between.
// In both sweeps one CTA owns one READ (reads are handed out by an atomic ticket)
// and walks it tile by tile, carrying the running values (exception index, payload
// position, Huffman bit position and the bits of the last partial byte) itself, so
// no CTA ever waits for another one; the next tile's samples are requested before
// the current tile is processed.
//
//   ex_count   (read)  zd in registers -> #exceptions, Huffman bits, OR of the samples
//                      (ex-zd: q = min(5, ctz(OR)), ex_zd.c:358-381; reads with q > 0
//                      are walked again on the shifted signal); exceptions captured in
//                      order into a pool; list sizes -> header length, stream size
//   ex_offsets (batch) exclusive scan of the sizes -> out_offsets
//   ex_emit    (read)  header + exception lists, then zd again -> payload bytes /
//                      Huffman bits to their final place
//
// Tile = 4096 samples of one read cut on 16-byte boundaries of the sample array
// (aligned 128-bit loads), 16 contiguous samples per thread.
#pragma once

#include "common.cuh"
#include "svb_zd.cuh"

namespace pg {

constexpr int XT = 4096;		// samples per tile
constexpr int XG = XT / 8;		// 16-byte groups per tile
constexpr uint32_t X_FAIL = 0xffffffffu;	// r_hdr of a read that has no stream (empty, ex-zd cap)
#ifndef PG_PF_COUNT
#define PG_PF_COUNT 3	// sweep 1: tiles ahead of the walk asked from L2
#endif
#ifndef PG_PF_EMIT
#define PG_PF_EMIT 3	// sweep 2
#endif
constexpr int X_STAGE = 24 * XT / 8 + 64;	// payload stage: 4096 codes of the longest length accepted
						// (24 bits, press_gpu_set_huffman_table), + phase + slack

enum : uint32_t { ERR_SCRATCH = 8 };	// exception pool too small (press_gpu_set_exception_capacity)

// payload = static Huffman bitstream (else: the low bytes themselves)
__host__ __device__ __forceinline__ bool ex_is_huff(int method)
{
	return method == PRESS_GPU_SHUFF_VBE21_ZD || method == PRESS_GPU_SHUFF_VBSSE21_ZD ||
	       method == PRESS_GPU_SHUFF_VBBE21_ZD;
}
// exception lists bit-packed (vbbe21) instead of streamvbyte coded
__host__ __device__ __forceinline__ bool ex_is_bitpacked(int method)
{
	return method == PRESS_GPU_VBBE21_ZD || method == PRESS_GPU_SHUFF_VBBE21_ZD;
}

struct ExArgs {
	int method;
	const uint4 *samples;		// encode in
	uint64_t nsamples;
	const uint64_t *offsets;	// [nreads + 1] sample offsets
	uint64_t nreads;
	uint8_t *out;			// encode out (16-byte aligned)
	uint64_t out_cap;
	uint64_t *out_offsets;		// [nreads + 1]
	Hdr *hdr;			// aux[0] = pool entries handed out (in chunks, see ex_count_kernel)
	const uint32_t *tile_start, *tile_read;
	uint32_t *t_nex, *t_pool, *t_eb;	// per (tile, warp): exceptions in the slice, pool position, prefix
	uint32_t *r_nex, *r_q, *r_hdr, *r_pool;			// per read
	uint64_t *r_size, *r_bits;
	uint2 *pool;			// (position in v = sample index - 1, value), slice by slice as captured
	uint2 *pool2;			// the same, one contiguous run per read (Hdr aux[3] = entries used)
	uint64_t pool_cap;
	const uint32_t *huff_enc;	// 256 x (code | length << 24)
	uint16_t *sidecar;		// HUFF, optional: see huff_fused.cuh (null: none)
	uint8_t *lowb;			// byte payloads: low bytes of the zd values of every tile that lies inside
					// a read, as sweep 1 computed them (1 B/sample; see ex_lowb_ptr)
};

// Sweep 1 leaves the 16 low bytes of a thread's zd values where sweep 2 finds them with one aligned
// 128-bit load: tile t of the read whose first 16-byte sample group is A, thread tid.  (Only slices
// (512 samples, one warp) wholly inside a read are kept: they share no sample group with another read; the 8-byte
// phase of A is taken out so that the address is 16-byte aligned.)
__device__ __forceinline__ uint4 *ex_lowb_ptr(uint8_t *lowb, uint64_t A, uint32_t t, int tid)
{
	return reinterpret_cast<uint4 *>(lowb + 8 * (A + (A & 1))) + (uint64_t) t * (XT / 16) + tid;
}

// per-halfword arithmetic shift right (ex_zd.c:383-394 on int16)
__device__ __forceinline__ uint32_t sar16x2(uint32_t x, uint32_t q)
{
	const uint32_t m = (0xffffu >> q) * 0x00010001u;
	const uint32_t t = ((x ^ 0x80008000u) >> q) & m;
	return __vsub2(t, (0x8000u >> q) * 0x00010001u);
}
__device__ __forceinline__ uint32_t shl16x2(uint32_t x, uint32_t q)
{
	return (x << q) & (((0xffffu << q) & 0xffffu) * 0x00010001u);
}

struct ExTile {
	uint32_t r, t, ntiles;
	uint64_t off0, rend, k0;	// read's sample range, tile's first 16-byte group
	uint32_t m;			// off0 & 7
};

__device__ __forceinline__ void ex_tile_info(const uint32_t *tile_start, const uint32_t *tile_read,
					     const uint64_t *offsets, uint32_t g, ExTile &T)
{
	T.r = __ldg(tile_read + g);
	const uint32_t g0 = __ldg(tile_start + T.r);
	T.t = g - g0;
	T.ntiles = __ldg(tile_start + T.r + 1) - g0;
	T.off0 = __ldg(offsets + T.r);
	T.rend = __ldg(offsets + T.r + 1);
	T.m = (uint32_t) (T.off0 & 7);
	T.k0 = (T.off0 >> 3) + (uint64_t) T.t * XG;
}

// number of v-indices (sample index - 1) of the read that lie before tile t
__device__ __forceinline__ uint32_t ex_v_before(uint32_t t, uint32_t m, uint64_t n)
{
	const int64_t first = (int64_t) t * XT - m;	// read index of the tile's first element
	int64_t j = first - 1;
	if (j < 0)
		j = 0;
	const int64_t mm = n ? (int64_t) n - 1 : 0;
	return (uint32_t) (j < mm ? j : mm);
}

// The thread's 16 samples of the tile as zd values (two per word), with masks:
// valid = element inside the read, first = the read's first sample (stored apart
// as zd0), exc = exception (value > 255, not first).  i0 = read index of element 0.
struct ExVals {
	uint32_t z[8];
	uint32_t valid, first, exc;
	int64_t i0;
	uint32_t orv;		// OR of the valid raw samples (both halves folded)
};

// the thread's 32 bytes of a tile, requested ...
struct ExRaw {
	uint4 xa, xb;
	uint32_t prev0;		// lane 0: the sample before the thread's first one
};

__device__ __forceinline__ void ex_issue(const uint4 *samples, uint64_t nsamples, uint64_t k0, ExRaw &w)
{
	const int tid = threadIdx.x;
	const uint64_t ka = k0 + 2 * tid;
	w.xa = load_group(samples, nsamples, ka);
	w.xb = load_group(samples, nsamples, ka + 1);
	w.prev0 = 0;
	if ((tid & 31) == 0 && ka > 0 && 8 * ka - 1 < nsamples)
		w.prev0 = __ldg(reinterpret_cast<const uint16_t *>(samples) + 8 * ka - 1);
}

// ... and turned into zd values one tile later
__device__ __forceinline__ void ex_compute(const ExRaw &w, const ExTile &T, uint32_t q, ExVals &v)
{
	const int tid = threadIdx.x, lane = tid & 31;
	const uint64_t ka = T.k0 + 2 * tid;
	uint4 xa = w.xa, xb = w.xb;
	uint32_t prev = __shfl_up_sync(0xffffffffu, xb.w >> 16, 1);
	if (lane == 0)
		prev = w.prev0;
	v.valid = group_valid(ka, T.off0, T.rend) | (group_valid(ka + 1, T.off0, T.rend) << 8);
	{
		// OR of the valid samples: mask out elements outside the read
		uint32_t x[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w}, o = 0;
#pragma unroll
		for (int j = 0; j < 8; j++) {
			const uint32_t mk = ((v.valid >> (2 * j)) & 1u ? 0xffffu : 0u) |
					    ((v.valid >> (2 * j + 1)) & 1u ? 0xffff0000u : 0u);
			o |= x[j] & mk;
		}
		v.orv = (o | (o >> 16)) & 0xffffu;
	}
	if (q) {
		xa.x = sar16x2(xa.x, q); xa.y = sar16x2(xa.y, q); xa.z = sar16x2(xa.z, q); xa.w = sar16x2(xa.w, q);
		xb.x = sar16x2(xb.x, q); xb.y = sar16x2(xb.y, q); xb.z = sar16x2(xb.z, q); xb.w = sar16x2(xb.w, q);
		prev = sar16x2(prev, q) & 0xffffu;
	}
	v.i0 = (int64_t) (8 * ka) - (int64_t) T.off0;
	const bool holds_first = T.t == 0 && tid == 0;
	v.first = holds_first ? 1u << T.m : 0u;
	uint32_t za[4], zb[4];
	zd_group(xa, prev, holds_first ? (int) T.m : -1, za);
	zd_group(xb, xa.w >> 16, -1, zb);
#pragma unroll
	for (int j = 0; j < 4; j++) {
		v.z[j] = za[j];
		v.z[4 + j] = zb[j];
	}
	const uint32_t f = flags_group(za) | (flags_group(zb) << 8);
	v.exc = f & v.valid & ~v.first;
}

// Interior tile (every element is a sample of the read past its first one): no
// masks.  Returns nonzero iff one of the thread's values is an exception; orw = OR
// of the raw sample words (for qts).  The instruction count of this path bounds
// the encode sweeps: they are issue-bound, not HBM-bound, when it is long.
__device__ __forceinline__ uint32_t ex_compute_fast(const ExRaw &w, uint32_t q, ExVals &v, uint32_t &orw)
{
	uint4 xa = w.xa, xb = w.xb;
	uint32_t prev = __shfl_up_sync(0xffffffffu, xb.w >> 16, 1);
	if ((threadIdx.x & 31) == 0)
		prev = w.prev0;
	orw = xa.x | xa.y | xa.z | xa.w | xb.x | xb.y | xb.z | xb.w;
	if (q) {
		xa.x = sar16x2(xa.x, q); xa.y = sar16x2(xa.y, q); xa.z = sar16x2(xa.z, q); xa.w = sar16x2(xa.w, q);
		xb.x = sar16x2(xb.x, q); xb.y = sar16x2(xb.y, q); xb.z = sar16x2(xb.z, q); xb.w = sar16x2(xb.w, q);
		prev = sar16x2(prev, q) & 0xffffu;
	}
	uint32_t za[4], zb[4];
	zd_group(xa, prev, -1, za);
	zd_group(xb, xa.w >> 16, -1, zb);
#pragma unroll
	for (int j = 0; j < 4; j++) {
		v.z[j] = za[j];
		v.z[4 + j] = zb[j];
	}
	v.valid = 0xffffu;
	v.first = 0;
	v.orv = 0;
	const uint32_t hi = (any_high(za) | any_high(zb));
	v.exc = hi ? (flags_group(za) | (flags_group(zb) << 8)) : 0u;
	return hi;
}

// sum of the code lengths of the thread's non-exception values (lut: 256 lengths, one byte each:
// symbols s and s + 128 are the only ones that meet in a bank)
__device__ __forceinline__ uint32_t ex_code_bits(const ExVals &v, uint32_t ne, const uint8_t *lut)
{
	uint32_t bits = 0;
	if (ne == 0xffffu) {
#pragma unroll
		for (int j = 0; j < 8; j++)
			bits += (uint32_t) lut[v.z[j] & 0xffffu] + lut[v.z[j] >> 16];
	} else {
#pragma unroll
		for (int e = 0; e < 16; e++)
			if ((ne >> e) & 1)
				bits += lut[(v.z[e >> 1] >> (16 * (e & 1))) & 0xffffu];
	}
	return bits;
}

__device__ __forceinline__ uint32_t ex_elem(const ExVals &v, int e)
{
	return (v.z[e >> 1] >> (16 * (e & 1))) & 0xffffu;
}

// ---------------------------------------------------------------------------
// exception lists of a read, addressed by their index k inside the read

// the exceptions of one read, in order (a run of pool2)
struct ExList {
	const uint2 *p;
	__device__ __forceinline__ uint2 at(uint32_t k) const { return p[k]; }
};

// streamvbyte coding of `cnt` values by one warp: KB = 2 classical svb32
// (streamvbyte_encode.c:11-59), KB = 1 svb16 without zd (svb16/encode_scalar.hpp:14-62).
// Returns the byte length; writes when WRITE.
template <int KB, bool WRITE, class F>
__device__ __forceinline__ uint32_t warp_svb_list(F value, uint32_t cnt, uint8_t *dst, int lane)
{
	const uint32_t nkeys = KB == 2 ? (cnt + 3) / 4 : (cnt + 7) / 8;
	uint32_t run = 0;
	for (uint32_t k0 = 0; k0 < cnt; k0 += 32) {
		const uint32_t k = k0 + lane;
		const bool act = k < cnt;
		const uint32_t val = act ? value(k) : 0u;
		uint32_t len = 0;
		if (act)
			len = KB == 2 ? 1u + (val >= 256u) + (val >= 65536u) + (val >= (1u << 24))
				      : 1u + ((val & 0xffffu) >= 256u);
		const uint32_t inc = warp_inc_scan(len, lane);
		if (WRITE) {
			uint8_t *p = dst + nkeys + run + inc - len;
			for (uint32_t b = 0; b < len; b++)
				p[b] = (uint8_t) (val >> (8 * b));
			uint32_t kk = act ? (len - 1) << (KB == 2 ? 2 * (lane & 3) : (lane & 7)) : 0u;
			kk |= __shfl_xor_sync(0xffffffffu, kk, 1);
			kk |= __shfl_xor_sync(0xffffffffu, kk, 2);
			if (KB == 1)
				kk |= __shfl_xor_sync(0xffffffffu, kk, 4);
			if (act && (lane & (KB == 2 ? 3 : 7)) == 0)
				dst[KB == 2 ? k >> 2 : k >> 3] = (uint8_t) kk;
		}
		run += __shfl_sync(0xffffffffu, inc, 31);
	}
	return nkeys + run;
}

// uint_press_16 / uint_press_32 of `cnt` values by one warp (press.c:489-507, 546-564 ->
// uintx_press_core 285-397): [width u8] then the values as `width`-bit big-endian fields back to
// back, most significant bit first, zero padded to a byte; width = uint_get_minbits(max)
// (press.c:461-471).  Returns the byte length; writes when WRITE (a lane per output byte).
template <bool WRITE, class F>
__device__ __forceinline__ uint32_t warp_bitpack_list(F value, uint32_t cnt, uint8_t *dst, int lane)
{
	uint32_t mx = 0;
	for (uint32_t k = lane; k < cnt; k += 32)
		mx = max(mx, value(k));
	mx = __reduce_max_sync(0xffffffffu, mx);
	const uint32_t bits = mx ? 32u - __clz(mx) : 0u;
	const uint64_t tot = (uint64_t) cnt * bits;
	const uint32_t nbytes = (uint32_t) ((tot + 7) >> 3);
	if (WRITE) {
		if (lane == 0)
			dst[0] = (uint8_t) bits;
		for (uint32_t b = lane; b < nbytes; b += 32) {
			// the fields overlapping bits [8b, 8b + 8) of the packed run
			const uint64_t g0 = 8ull * b;
			uint32_t k = (uint32_t) (g0 / bits);
			uint32_t byte = 0;
			for (uint64_t f0 = (uint64_t) k * bits; f0 < g0 + 8 && k < cnt; f0 += bits, k++) {
				// field k covers bits [f0, f0 + bits); bit j of the byte (7 = first) is run bit g0 + 7 - j
				const int64_t sh = (int64_t) (f0 + bits) - (int64_t) (g0 + 8);	// field's low bits past the byte
				const uint64_t v = value(k);
				byte |= (uint32_t) (sh >= 0 ? v >> sh : v << -sh) & 0xffu;
			}
			dst[1 + b] = (uint8_t) byte;
		}
	}
	return 1 + nbytes;
}

__device__ __forceinline__ void put_le(uint8_t *p, uint64_t v, int nbytes)
{
	for (int b = 0; b < nbytes; b++)
		p[b] = (uint8_t) (v >> (8 * b));
}

// header of one read past the fixed fields; returns its length.  One warp.
template <bool WRITE>
__device__ __forceinline__ uint32_t ex_lists(int method, const ExList &L, uint32_t nex, uint8_t *dst, int lane)
{
	auto gap = [&](uint32_t k) {
		const uint32_t p = L.at(k).x;
		return k ? p - L.at(k - 1).x - 1 : p;	// trans.c:129-149
	};
	auto val256 = [&](uint32_t k) { return L.at(k).y - 256u; };
	if (method == PRESS_GPU_VBE21_ZD || method == PRESS_GPU_SHUFF_VBE21_ZD) {
		if (WRITE)
			for (uint32_t k = lane; k < nex; k += 32) {
				const uint2 e = L.at(k);
				put_le(dst + 4 * (uint64_t) k, e.x, 4);
				put_le(dst + 4 * (uint64_t) nex + 2 * (uint64_t) k, e.y, 2);
			}
		return 6 * nex;
	}
	if (nex == 0)
		return 0;
	const bool ex = method == PRESS_GPU_EX_ZD;
	if (nex == 1) {		// press.c:3269-3274 (u32, i16) / ex_zd.c:146-153 (u32, u32)
		if (WRITE && lane == 0) {
			const uint2 e = L.at(0);
			put_le(dst, e.x, 4);
			put_le(dst + 4, e.y - 256u, ex ? 4 : 2);
		}
		return ex ? 8 : 6;
	}
	const bool bp = ex_is_bitpacked(method);	// press.c:2822-2866
	const uint32_t lp = bp ? warp_bitpack_list<WRITE>(gap, nex, dst + 4, lane) : warp_svb_list<2, WRITE>(gap, nex, dst + 4, lane);
	if (WRITE && lane == 0)
		put_le(dst, lp, 4);
	uint8_t *d2 = dst + 4 + lp;
	auto val16 = [&](uint32_t k) { return val256(k) & 0xffffu; };
	const uint32_t lv = bp ? warp_bitpack_list<WRITE>(val16, nex, d2 + 4, lane)
			  : ex ? warp_svb_list<2, WRITE>(val256, nex, d2 + 4, lane)
			       : warp_svb_list<1, WRITE>(val256, nex, d2 + 4, lane);
	if (WRITE && lane == 0)
		put_le(d2, lv, 4);
	return 8 + lp + lv;
}

__device__ __forceinline__ uint32_t ex_fixed_bytes(int method)
{
	return method == PRESS_GPU_EX_ZD ? 16u : 6u;	// [ver][n u64][q] + [zd0 u16][nex u32]
}

// exclusive scan of the stream sizes -> out_offsets.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) ex_offsets_kernel(const uint64_t *size, uint64_t nreads,
							  uint64_t *out_offsets, uint64_t out_cap, Hdr *hdr)
{
	__shared__ uint64_t s_w[33];
	__shared__ uint64_t s_carry;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (tid == 0)
		s_carry = 0;
	__syncthreads();
	constexpr int IT = 4;
	for (uint64_t base = 0; base < nreads; base += 1024 * IT) {
		const uint64_t r0 = base + (uint64_t) tid * IT;
		uint64_t v[IT], s = 0;
#pragma unroll
		for (int i = 0; i < IT; i++) {
			v[i] = r0 + i < nreads ? size[r0 + i] : 0;
			s += v[i];
		}
		uint64_t inc = s;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint64_t y = __shfl_up_sync(0xffffffffu, inc, o);
			if (lane >= o)
				inc += y;
		}
		if (lane == 31)
			s_w[warp] = inc;
		__syncthreads();
		if (warp == 0) {
			uint64_t x = s_w[lane], o0 = x;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
				if (lane >= o)
					x += y;
			}
			s_w[lane] = x - o0;
			if (lane == 31)
				s_w[32] = x;
		}
		__syncthreads();
		uint64_t e = s_carry + s_w[warp] + inc - s;
#pragma unroll
		for (int i = 0; i < IT; i++) {
			if (r0 + i < nreads)
				out_offsets[r0 + i] = e;
			e += v[i];
		}
		__syncthreads();
		if (tid == 0)
			s_carry += s_w[32];
		__syncthreads();
	}
	if (tid == 0) {
		out_offsets[nreads] = s_carry;
		if (s_carry > out_cap)
			atomicOr(&hdr->error, ERR_CAPACITY);
	}
}

// ---------------------------------------------------------------------------
// sweep 2: payload

// Copy `len` bytes staged at stage[0 .. len) (8-byte aligned base, 24 bytes of
// readable slack) to global dst (any alignment) with aligned 16-byte stores.  One CTA.
__device__ __forceinline__ void block_copy_out(const uint8_t *stage, uint8_t *dst, uint32_t len)
{
	const uint32_t a = (uint32_t) (reinterpret_cast<uintptr_t>(dst) & 15);
	uint8_t *gb = dst - a;
	const uint32_t end = a + len;
	const uint32_t nw = (end + 15) >> 4;
	const uint64_t *s64 = reinterpret_cast<const uint64_t *>(stage);
	for (uint32_t w = threadIdx.x; w < nw; w += NT) {
		const uint32_t lo = w << 4, hi = lo + 16;
		if (lo >= a && hi <= end) {
			const uint32_t o = lo - a, qd = o >> 3, sh = (o & 7) * 8;
			const uint64_t x0 = s64[qd], x1 = s64[qd + 1], x2 = s64[qd + 2];
			uint2 l2, h2;
			uint64_t l = x0, h = x1;
			if (sh) {
				l = (x0 >> sh) | (x1 << (64 - sh));
				h = (x1 >> sh) | (x2 << (64 - sh));
			}
			l2 = make_uint2((uint32_t) l, (uint32_t) (l >> 32));
			h2 = make_uint2((uint32_t) h, (uint32_t) (h >> 32));
			*reinterpret_cast<uint4 *>(gb + lo) = make_uint4(l2.x, l2.y, h2.x, h2.y);
		} else {
			const uint32_t b0 = lo < a ? a : lo, b1 = hi > end ? end : hi;
			for (uint32_t b = b0; b < b1; b++)
				gb[b] = stage[b - a];
		}
	}
}

// per-read geometry of the walk
__device__ __forceinline__ bool ex_read_tile0(const ExArgs &a, uint64_t r, ExTile &T, uint32_t &g0)
{
	T.r = (uint32_t) r;
	T.off0 = __ldg(a.offsets + r);
	T.rend = __ldg(a.offsets + r + 1);
	T.m = (uint32_t) (T.off0 & 7);
	g0 = __ldg(a.tile_start + r);
	T.ntiles = __ldg(a.tile_start + r + 1) - g0;
	T.t = 0;
	T.k0 = T.off0 >> 3;
	return T.rend > T.off0;
}

// ---------------------------------------------------------------------------
// sweep 1: counts, exception capture, stream size

#ifndef PG_COUNT_CTAS
#define PG_COUNT_CTAS 4
#endif
template <bool HUFF>
__global__ void __launch_bounds__(NT, PG_COUNT_CTAS) ex_count_kernel(const __grid_constant__ ExArgs a)
{
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read, s_or, s_nex, s_base2;
	__shared__ unsigned long long s_bits;
	__shared__ uint8_t s_len[HUFF ? 256 : 4];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (HUFF)
		s_len[tid] = (uint8_t) (a.huff_enc[tid] >> 24);
	// pool chunk of a warp: 1/16 of the pool at most is held in unfinished chunks; none when the pool
	// is small or has its full size (then every entry may be needed: the retry of the host paths)
	uint32_t chunk = 0, wleft = 0;
	unsigned long long wbase = 0;
	if (a.pool_cap <= a.nsamples) {
		const uint64_t c = a.pool_cap / (16ull * gridDim.x * NW);
		chunk = c >= 16 ? (uint32_t) (c < 1024 ? c : 1024) : 0u;
	}
	for (;;) {
		__syncthreads();
		if (tid == 0) {
			s_read = atomicAdd(&a.hdr->ticket, 1u);
			s_or = 0;
			s_nex = 0;
			s_bits = 0;
		}
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		ExTile T;
		uint32_t g0;
		if (!ex_read_tile0(a, r, T, g0)) {	// empty read: no stream (the reference underflows, press.c:3426)
			if (tid == 0) {
				a.r_nex[r] = 0;
				a.r_q[r] = 0;
				a.r_hdr[r] = X_FAIL;
				a.r_size[r] = 0;
			}
			continue;
		}
		const uint64_t n = T.rend - T.off0, A = T.off0 >> 3;
		const uint32_t nt = T.ntiles;
		const uint64_t rec0 = (uint64_t) g0 * NW;
		uint32_t q = 0, bits_thr = 0;
		for (int pass = 0; pass < 2; pass++) {
			// the walk: no CTA barrier, every warp runs through its 512-sample slice of every tile
			uint32_t nex_thr = 0, or_thr = 0;
			bits_thr = 0;
			ExRaw W;
			ex_issue(a.samples, a.nsamples, A, W);
#pragma unroll(HUFF ? 1 : 2)	// (the Huffman variant's body is long: unrolled twice it misses the instruction cache)
			for (uint32_t t = 0; t < nt; t++) {
				T.t = t;
				T.k0 = A + (uint64_t) t * XG;
				const ExRaw C = W;
				if (t + 1 < nt)
					ex_issue(a.samples, a.nsamples, T.k0 + XG, W);
				if (t + PG_PF_COUNT < nt)
					prefetch_l2(a.samples + T.k0 + PG_PF_COUNT * XG + 2 * tid);
				ExVals v;
				// (the warp's 512-sample slice lies inside the read, past its first sample)
				const int64_t f = (int64_t) t * XT - T.m + 512 * warp;
				if (f >= 1 && f + 512 <= (int64_t) n) {
					uint32_t orw;
					(void) ex_compute_fast(C, q, v, orw);
					or_thr |= orw;
					v.i0 = (int64_t) t * XT - T.m + 16 * tid;
					// (ex-zd: the pass with the final q comes last)
					*ex_lowb_ptr(a.lowb, A, t, tid) =
							make_uint4(__byte_perm(v.z[0], v.z[1], 0x6420), __byte_perm(v.z[2], v.z[3], 0x6420),
								   __byte_perm(v.z[4], v.z[5], 0x6420), __byte_perm(v.z[6], v.z[7], 0x6420));
				} else {
					ex_compute(C, T, q, v);
					or_thr |= v.orv;
				}
				if (HUFF)
					bits_thr += ex_code_bits(v, v.valid & ~v.exc & ~v.first, s_len);
				uint32_t slice_nex = 0, base = 0;
				if (__any_sync(0xffffffffu, v.exc)) {	// capture, in order
					const uint32_t cnt = __popc(v.exc);
					const uint32_t inc = warp_inc_scan(cnt, lane);
					slice_nex = __shfl_sync(0xffffffffu, inc, 31);
					// pool space: from the warp's own chunk (one global atomic per `chunk` entries: at
					// exception rates >= 1e-3 an atomic per slice on one address was the whole sweep's
					// bound), else asked for directly
					unsigned long long b64 = wbase;
					if (slice_nex <= wleft) {
						wbase += slice_nex;
						wleft -= slice_nex;
					} else {
						// (a new chunk gives up the old one's rest, < slice_nex < chunk / 2 entries)
						const bool refill = 2 * slice_nex < chunk;
						const uint32_t take = refill ? chunk : slice_nex;
						if (lane == 0)
							b64 = atomicAdd((unsigned long long *) &a.hdr->aux[0], (unsigned long long) take);
						b64 = __shfl_sync(0xffffffffu, b64, 0);
						if (refill) {
							const bool fits = b64 + take <= a.pool_cap;
							wbase = b64 + slice_nex;
							wleft = fits ? take - slice_nex : 0u;
						}
					}
					base = (uint32_t) b64;
					if (b64 + slice_nex <= a.pool_cap) {
						uint2 *p = a.pool + b64 + inc - cnt;
						if (__reduce_max_sync(0xffffffffu, cnt) <= 2) {
							// few exceptions: one turn per exception, its value picked by selects
							// (v.z stays in registers)
							for (uint32_t mk = v.exc; mk; mk &= mk - 1) {
								const uint32_t e = __ffs(mk) - 1, j = e >> 1;
								const uint32_t lo4 = (j & 2) ? ((j & 1) ? v.z[3] : v.z[2]) : ((j & 1) ? v.z[1] : v.z[0]);
								const uint32_t hi4 = (j & 2) ? ((j & 1) ? v.z[7] : v.z[6]) : ((j & 1) ? v.z[5] : v.z[4]);
								const uint32_t w = (j & 4) ? hi4 : lo4;
								*p++ = make_uint2((uint32_t) (v.i0 + e - 1), (w >> (16 * (e & 1))) & 0xffffu);
							}
						} else {
#pragma unroll
							for (int e = 0; e < 16; e++)	// (static indices: v.z stays in registers)
								if ((v.exc >> e) & 1)
									*p++ = make_uint2((uint32_t) (v.i0 + e - 1), ex_elem(v, e));
						}
					} else if (lane == 0) {
						atomicOr(&a.hdr->error, ERR_SCRATCH);
					}
					nex_thr += cnt;
				}
				if (lane == 0) {
					a.t_nex[rec0 + (uint64_t) t * NW + warp] = slice_nex;
					a.t_pool[rec0 + (uint64_t) t * NW + warp] = base;
				}
			}
			nex_thr = __reduce_add_sync(0xffffffffu, nex_thr);
			or_thr = __reduce_or_sync(0xffffffffu, (or_thr | (or_thr >> 16)) & 0xffffu);
			if (lane == 0) {
				if (nex_thr)
					atomicAdd(&s_nex, nex_thr);
				if (or_thr)
					atomicOr(&s_or, or_thr);
			}
			__syncthreads();
			if (a.method != PRESS_GPU_EX_ZD || pass == 1)
				break;
			// ex_zd.c:358-381: q from the OR of all samples (a zero sample never lowers q; cap 5)
			q = __ffs(s_or | 32u) - 1;
			if (q == 0)
				break;
			__syncthreads();	// everybody has read s_or / s_nex
			if (tid == 0)
				s_nex = 0;
			__syncthreads();
		}
		if (HUFF) {
			const uint32_t wb = __reduce_add_sync(0xffffffffu, bits_thr);
			if (lane == 0)
				atomicAdd(&s_bits, (unsigned long long) wb);
		}
		const uint32_t nex = s_nex;
		// exclusive prefix of the slice counts -> t_eb (all records of the read, in order), and
		// the read's exceptions gathered into one contiguous run of pool2
		const uint32_t nrec = nt * NW;
		if (nex) {
			if (tid == 0)
				s_base2 = (uint32_t) atomicAdd((unsigned long long *) &a.hdr->aux[3], (unsigned long long) nex);
			uint32_t run = 0;
			for (uint32_t b0 = 0; b0 < nrec; b0 += NT) {
				__syncthreads();	// records written above / s_warp of the previous round / s_base2
				const uint64_t base2 = s_base2;
				const uint32_t i = b0 + tid;
				const uint32_t c = i < nrec ? a.t_nex[rec0 + i] : 0u;
				uint32_t c1[1] = {c}, e1[1], tot;
				block_scan<1>(c1, e1, tot, s_warp);
				if (i < nrec)
					a.t_eb[rec0 + i] = run + e1[0];
				if (c && base2 + nex <= a.pool_cap && (uint64_t) a.t_pool[rec0 + i] + c <= a.pool_cap) {
					const uint2 *src = a.pool + a.t_pool[rec0 + i];
					uint2 *dst2 = a.pool2 + base2 + run + e1[0];
					for (uint32_t j = 0; j < c; j++)
						dst2[j] = src[j];
				}
				run += tot;
			}
			if (tid == 0 && (uint64_t) s_base2 + nex > a.pool_cap)
				atomicOr(&a.hdr->error, ERR_SCRATCH);
		} else {
			for (uint32_t i = tid; i < nrec; i += NT)
				a.t_eb[rec0 + i] = 0;
		}
		__syncthreads();
		if (tid == 0) {
			a.r_nex[r] = nex;
			a.r_q[r] = q;
			a.r_pool[r] = nex ? s_base2 : 0u;
			a.r_bits[r] = HUFF ? s_bits : 0ull;
		}
	}
}

// header length and stream size of every read.  One warp per read.
__global__ void __launch_bounds__(NT) ex_size_kernel(const __grid_constant__ ExArgs a)
{
	const uint64_t r = ((uint64_t) blockIdx.x * NT + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (r >= a.nreads)
		return;
	const uint64_t n = a.offsets[r + 1] - a.offsets[r];
	uint32_t hdr = X_FAIL;
	uint64_t size = 0;
	if (n > 0 && !(a.hdr->error & ERR_SCRATCH)) {
		const uint32_t nex = a.r_nex[r];
		const ExList L = {a.pool2 + a.r_pool[r]};
		hdr = ex_fixed_bytes(a.method) + ex_lists<false>(a.method, L, nex, nullptr, lane);
		if ((a.method == PRESS_GPU_SHUFF_VBSSE21_ZD || a.method == PRESS_GPU_SHUFF_VBBE21_ZD) && hdr - 2 > 0xffffu) {
			// press.c:4753, 4521: the header length lives in a uint16_t there; longer ones wrap and
			// the reference cannot read its own stream back.  Reported like the ex-zd cap.
			if (lane == 0)
				atomicOr(&a.hdr->error, ERR_EXZD_CAP);
			hdr = X_FAIL;
		} else if (ex_is_huff(a.method)) {
			hdr += 4;	// [count u32 BE], huffman.c:1199-1203
			size = hdr + (a.r_bits[r] + 7) / 8;
		} else {
			size = hdr + (n - 1 - nex);
		}
		if (a.method == PRESS_GPU_EX_ZD && size > 2 * n + 1024) {	// ex_zd.c:411 scratch cap
			if (lane == 0)
				atomicOr(&a.hdr->error, ERR_EXZD_CAP);
			hdr = X_FAIL;
			size = 0;
		}
	}
	if (lane == 0) {
		a.r_hdr[r] = hdr;
		a.r_size[r] = size;
	}
}

// header and exception lists of every read at their final place.  One warp per read.
__global__ void __launch_bounds__(NT) ex_header_kernel(const __grid_constant__ ExArgs a)
{
	const uint64_t r = ((uint64_t) blockIdx.x * NT + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (r >= a.nreads || (a.hdr->error & (ERR_CAPACITY | ERR_SCRATCH)))
		return;
	if (a.r_hdr[r] == X_FAIL)
		return;
	const uint64_t off0 = a.offsets[r], n = a.offsets[r + 1] - off0;
	const uint32_t nex = a.r_nex[r];
	const uint32_t q = a.method == PRESS_GPU_EX_ZD ? a.r_q[r] : 0u;
	uint8_t *dst = a.out + a.out_offsets[r];
	if (lane == 0) {
		const int16_t x0 = reinterpret_cast<const int16_t *>(a.samples)[off0];
		const uint16_t s0 = (uint16_t) (int16_t) (x0 >> q);
		const uint16_t zd0 = (uint16_t) ((uint16_t) (s0 << 1) ^ (uint16_t) (0 - (s0 >> 15)));	// trans.c:75-78
		uint8_t *p = dst;
		if (a.method == PRESS_GPU_EX_ZD) {
			*p++ = 0;		// ex_zd.c:420-423
			put_le(p, n, 8);	// ex_zd.c:425-428
			p += 8;
			*p++ = (uint8_t) q;	// ex_zd.c:436-439
		}
		put_le(p, zd0, 2);
		put_le(p + 2, nex, 4);
	}
	const ExList L = {a.pool2 + a.r_pool[r]};
	const uint32_t fx = ex_fixed_bytes(a.method);
	const uint32_t ll = ex_lists<true>(a.method, L, nex, dst + fx, lane);
	if (ex_is_huff(a.method) && lane < 4) {
		const uint32_t cnt = (uint32_t) (n - 1 - nex);
		dst[fx + ll + lane] = (uint8_t) (cnt >> (8 * (3 - lane)));	// htonl, huffman.c:1199
	}
}

// ---------------------------------------------------------------------------
// sweep 2, Huffman payloads (huffman.c:848-887: codes back to back, first bit of the stream
// in bit 0 of byte 0).  One CTA per read, a tile of 4096 samples per step:
//   1. every thread looks its 16 symbols up (one table word per symbol: code << 8 | length) and
//      adds up the lengths -- the thread's bit count, known before a single bit is placed -- and
//      the bit count in front of the one symbol that opens a side-car group;
//   2. one block scan of the bit counts gives every thread its place in the tile's bitstream;
//   3. the thread appends its codes, at the stage's own word alignment, to a register and stores
//      every word it completes; the stage's words are written once, with plain stores, except
//      the one word a thread shares with the thread before (see he_put);
//   4. the stage goes out as aligned 128-bit words and is left zeroed; the bits of the last,
//      partial byte are carried into the next tile.
// A shared-memory atomic costs about 2 cycles per LANE on this part: one ORed word per thread and
// tile is as much LSU time as four table lookups, which is why the words shared by neighbours go
// through a shuffle instead (an earlier form with two atomicOr per thread spent half its time there).

// step 1.  ALL: the thread's 16 elements are payload symbols.  `tgt`: index (among the thread's
// symbols) of the one whose bit offset is wanted in `cap`.  The lengths are added as whole table
// words: the low byte of a sum of up to 8 entries is the sum of their lengths (8 x 24 < 256).
template <bool ALL>
__device__ __forceinline__ uint32_t he_lookup(const uint32_t (&qv)[4], uint32_t ne, uint32_t tgt, const uint32_t (&lut)[256],
					       uint32_t (&e)[16], uint32_t &cap)
{
#pragma unroll
	for (int i = 0; i < 16; i++) {
		e[i] = lut[(qv[i >> 2] >> (8 * (i & 3))) & 0xffu];
		if (!ALL && !((ne >> i) & 1))
			e[i] = 0;
	}
	uint32_t ps[8], qs[4];
#pragma unroll
	for (int k = 0; k < 8; k++)
		ps[k] = e[2 * k] + e[2 * k + 1];
#pragma unroll
	for (int j = 0; j < 4; j++)
		qs[j] = ps[2 * j] + ps[2 * j + 1];
	const uint32_t o0 = (qs[0] + qs[1]) & 0xffu, o1 = (qs[2] + qs[3]) & 0xffu;
	if (ALL) {
		// (tgt is the same in every thread of a tile without exceptions)
		uint32_t c = 0;
		switch (tgt & 7u) {
		case 0: c = 0; break;
		case 1: c = tgt & 8u ? e[8] : e[0]; break;
		case 2: c = tgt & 8u ? ps[4] : ps[0]; break;
		case 3: c = tgt & 8u ? ps[4] + e[10] : ps[0] + e[2]; break;
		case 4: c = tgt & 8u ? qs[2] : qs[0]; break;
		case 5: c = tgt & 8u ? qs[2] + e[12] : qs[0] + e[4]; break;
		case 6: c = tgt & 8u ? qs[2] + ps[6] : qs[0] + ps[2]; break;
		default: c = tgt & 8u ? qs[2] + ps[6] + e[14] : qs[0] + ps[2] + e[6]; break;
		}
		cap = (c & 0xffu) + (tgt & 8u ? o0 : 0u);
	} else {
		uint32_t cnt = 0, nb = 0;
#pragma unroll
		for (int i = 0; i < 16; i++) {
			if (((ne >> i) & 1) && cnt++ == tgt)
				cap = nb;
			nb += e[i] & 0xffu;
		}
	}
	return o0 + o1;
}

// step 3.  The thread's codes appended to a0 (what its first word holds already below bit
// pos & 31), every completed word but the first stored at sa onwards (sa: shared-memory address of
// the word holding bit pos; it follows the words).  The first completed word comes back in `first`
// (got: there was one), the bits after the last completed word are the return value.
// EDGE = false: 8 codes are 32 bits at least, so the first word is complete after the eighth.
template <bool EDGE>
__device__ __forceinline__ uint32_t he_put(const uint32_t (&e)[16], uint32_t a0, uint32_t pos, uint32_t &sa, uint32_t &first,
					    bool &got)
{
	got = false;
	first = 0;
#pragma unroll
	for (int i = 0; i < 16; i++) {
		const uint32_t c = e[i] >> 8;
		a0 |= __funnelshift_l(0u, c, pos);	// (shifts by pos & 31)
		const uint32_t a1 = __funnelshift_l(c, 0u, pos);
		const uint32_t np = pos + e[i];	// (only bits 0..5 of a position are ever looked at: the code above bit 8 may ride along)
		if ((np ^ pos) & 32u) {		// a code is 24 bits at most: one word boundary at most
			if (EDGE || i < 8) {
				if (got)
					asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa), "r"(a0) : "memory");
				else
					first = a0;
				got = true;
			} else {
				asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa), "r"(a0) : "memory");
			}
			sa += 4;
			a0 = a1;
		}
		pos = np;
	}
	return a0;
}

__device__ __forceinline__ void sred_or(uint32_t saddr, uint32_t v)	// OR into a shared-memory word, no result wanted
{
	asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}

#ifndef PG_HE_CTAS
#define PG_HE_CTAS 5	// (47 registers; 4: 21.1, 5 and 6: 21.3 % of HBM peak)
#endif
__global__ void __launch_bounds__(NT, PG_HE_CTAS) huff_emit_kernel(const __grid_constant__ ExArgs a)
{
	__shared__ __align__(16) uint8_t s_stage[X_STAGE];
	__shared__ uint32_t s_lut[256];
	__shared__ __align__(16) uint32_t s_warp[NW];
	__shared__ uint32_t s_read, s_carry;
	const int tid = threadIdx.x;
	if (a.hdr->error & (ERR_CAPACITY | ERR_SCRATCH))
		return;
	const uint32_t x = a.huff_enc[tid];	// code | length << 24 -> code << 8 | length
	s_lut[tid] = (x << 8) | (x >> 24);
	for (int w = tid; w < X_STAGE / 16; w += NT)
		reinterpret_cast<uint4 *>(s_stage)[w] = make_uint4(0, 0, 0, 0);
	if (tid == 0)
		s_carry = 0;
	// the plain-store path wants 8 codes to be 32 bits at least (he_put)
	const bool tab_ok = __syncthreads_and((x >> 24) >= 4);
	const uint32_t stage_sa = (uint32_t) __cvta_generic_to_shared(s_stage);
	uint32_t *sw = reinterpret_cast<uint32_t *>(s_stage);
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = atomicAdd(&a.hdr->ticket2, 1u);
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		const uint32_t hdr = a.r_hdr[r];
		if (hdr == X_FAIL)
			continue;
		ExTile T;
		uint32_t g0;
		(void) ex_read_tile0(a, r, T, g0);
		const uint64_t n = T.rend - T.off0, A = T.off0 >> 3;
		const uint32_t nt = T.ntiles;
		uint8_t *pay = a.out + a.out_offsets[r] + hdr;	// first payload byte of the read
		uint16_t *sc = a.sidecar ? a.sidecar + (T.off0 >> 4) + r : nullptr;
		uint64_t pos = 0;	// payload bits written so far
		// (the bits of the last, still partial byte wait in s_carry: 0 at every read's start)
		// exceptions in the tile's eight slices, as sweep 1 counted them: a tile without any inside
		// the read ("all": every element a payload symbol) takes the path without masks and without the
		// symbol-count scan, and starts from the low bytes sweep 1 left (16 bytes per thread instead of
		// 32 of samples, no zd); the others start from the samples again
		const uint4 *rec = reinterpret_cast<const uint4 *>(a.t_nex + (uint64_t) g0 * NW);
		auto tile_all = [&](uint32_t t, const uint4 &ra, const uint4 &rb) {
			const int64_t f = (int64_t) t * XT - T.m;
			return f >= 1 && f + XT <= (int64_t) n && !(ra.x | ra.y | ra.z | ra.w | rb.x | rb.y | rb.z | rb.w);
		};
		// one tile ahead in registers: the low bytes or the samples; records two tiles ahead
		ExRaw W;
		uint4 Lw = make_uint4(0, 0, 0, 0);
		bool all = tile_all(0, __ldg(rec), __ldg(rec + 1)), all_next = false;	// (tile 0 holds the first sample: never all)
		ex_issue(a.samples, a.nsamples, A, W);
		uint4 ra = make_uint4(1, 0, 0, 0), rb = ra;
		if (nt > 1) {
			ra = __ldg(rec + 2);
			rb = __ldg(rec + 3);
		}
		uint32_t psym = 0;	// payload symbols written so far (side-car index)
		for (uint32_t t = 0; t < nt; t++, all = all_next) {
			T.t = t;
			T.k0 = A + (uint64_t) t * XG;
			const ExRaw C = W;
			const uint4 Lc = Lw;
			all_next = t + 1 < nt && tile_all(t + 1, ra, rb);
			if (t + 1 < nt) {
				if (all_next)
					Lw = __ldg(ex_lowb_ptr(a.lowb, A, t + 1, tid));
				else
					ex_issue(a.samples, a.nsamples, T.k0 + XG, W);
			}
			if (t + 2 < nt) {
				ra = __ldg(rec + 2 * (t + 2));
				rb = __ldg(rec + 2 * (t + 2) + 1);
			}
			if (t + PG_PF_COUNT + 1 < nt)	// (not the read's last tile: its low bytes may lie past the scratch)
				prefetch_l2(ex_lowb_ptr(a.lowb, A, t + PG_PF_COUNT, tid));
			uint32_t ne = 0xffffu, qv[4] = {Lc.x, Lc.y, Lc.z, Lc.w};
			if (!all) {
				const int64_t f = (int64_t) t * XT - T.m;
				ExVals v;
				if (f >= 1 && f + XT <= (int64_t) n) {
					uint32_t orw;
					(void) ex_compute_fast(C, 0, v, orw);
				} else {
					ex_compute(C, T, 0, v);
				}
				ne = v.valid & ~v.exc & ~v.first;
				qv[0] = __byte_perm(v.z[0], v.z[1], 0x6420);
				qv[1] = __byte_perm(v.z[2], v.z[3], 0x6420);
				qv[2] = __byte_perm(v.z[4], v.z[5], 0x6420);
				qv[3] = __byte_perm(v.z[6], v.z[7], 0x6420);
			}
			// payload index of the thread's first symbol
			uint32_t si = psym + 16 * tid, nsym = XT;
			if (!all) {
				uint32_t c1[1] = {(uint32_t) __popc(ne)}, e1[1];
				block_scan<1>(c1, e1, nsym, s_warp);
				si = psym + e1[0];
				__syncthreads();	// (s_warp is written again below)
			}
			psym += nsym;
			// 1. table words, the thread's bit count
			const uint32_t tgt = (0u - si) & 15u;
			const bool plain = all && tab_ok;	// (the same in every thread)
			uint32_t cap = 0xffffffffu, e[16];
			const uint32_t len = all ? he_lookup<true>(qv, ne, tgt, s_lut, e, cap) : he_lookup<false>(qv, ne, tgt, s_lut, e, cap);
			// 2. place in the tile's bitstream.  The stage mirrors global memory from the 16-byte
			// word holding the tile's first bit
			uint32_t l1[1] = {len}, x1[1], tot;
			block_scan<1>(l1, x1, tot, s_warp);
			const uint64_t G = 8 * (uint64_t) (pay - a.out) + pos;
			const uint32_t s0 = (uint32_t) (G & 127), e0 = s0 + tot;
			if (sc && cap != 0xffffffffu)	// side-car: the bit offset (in the read's bitstream) of every 16th symbol
				sc[(si + tgt) >> 4] = (uint16_t) (pos + x1[0] + cap);
			// 3. codes -> stage
			const uint32_t bp = s0 + x1[0];
			const uint32_t fw = stage_sa + ((bp >> 3) & ~3u);	// the thread's first word
			uint32_t sa = fw, first;
			bool got;
			// the tile's first word opens with the carried byte (written before the scan's barrier,
			// rewritten after the next one)
			const uint32_t carried = tid == 0 ? s_carry << (8 * ((s0 >> 3) & 3)) : 0u;
			if (plain) {
				// every thread has 64 bits or more: its first word is shared with the thread before and
				// nobody else, whose bits there are what that thread has left over at its end.  Across
				// warps (and only there) the two halves are ORed into the zeroed stage.
				const uint32_t left = he_put<false>(e, carried, bp, sa, first, got);
				const uint32_t before = __shfl_up_sync(0xffffffffu, left, 1);
				if (tid & 31)
					asm volatile("st.shared.u32 [%0], %1;" ::"r"(fw), "r"(first | before) : "memory");
				else
					sred_or(fw, first);
				if ((tid & 31) == 31)
					sred_or(sa, left);
			} else {
				// exceptions, or an end of the read: nothing is known about the neighbours' lengths, so
				// the thread's first and last words are ORed in (the words in between are its own)
				const uint32_t left = he_put<true>(e, carried, bp, sa, first, got);
				if (got)
					sred_or(fw, first);
				if (left)
					sred_or(sa, left);
			}
			__syncthreads();
			// 4. bytes [fb, lb) are complete: the first one thanks to the carried bits, the last
			// partial one only at the end of the stream (zero padded, huffman.c:886)
			const bool lastt = t + 1 == nt;
			uint8_t *gb = a.out + ((G >> 7) << 4);	// global address of stage byte 0
			const uint32_t fb = s0 >> 3, lb = (e0 >> 3) + ((lastt && (e0 & 7)) ? 1 : 0);
			const uint32_t cw = e0 >> 7;	// the 16-byte word holding the partial byte
			for (uint32_t w = tid; w <= cw; w += NT) {
				const uint4 x = reinterpret_cast<const uint4 *>(s_stage)[w];
				if (w == cw)
					s_carry = (!lastt && (e0 & 7)) ? s_stage[e0 >> 3] : 0u;
				const uint32_t lo = w << 4, hi = lo + 16;
				if (lo >= fb && hi <= lb) {
					reinterpret_cast<uint4 *>(gb)[w] = x;
				} else {
					const uint32_t b0 = lo < fb ? fb : lo, b1 = hi > lb ? lb : hi;
					for (uint32_t b = b0; b < b1; b++)
						gb[b] = s_stage[b];
				}
				reinterpret_cast<uint4 *>(s_stage)[w] = make_uint4(0, 0, 0, 0);
			}
			// (no barrier here: the next tile's scan has one between these stage writes and its ORs)
			pos += tot;
		}
	}
}

// sweep 2 for the byte payloads (vbe21 / vbsse21 / ex-zd): no CTA barrier in the walk.
// A warp knows where its 512-sample slice goes from the record prefix alone
// (values before the slice minus exceptions before the slice), so the eight warps
// of the CTA run through the read independently.  A slice without exception inside
// the read is 32 x 16 low bytes as sweep 1 left them (ex_lowb_ptr): one aligned
// 128-bit load per lane, re-phased to the destination's 16-byte alignment with
// shuffles and stored as aligned 128-bit words; anything else is computed again
// from the samples and goes through a per-warp staging line.

// 32 lanes x 16 payload bytes (qv) -> D[0, 512), any alignment
__device__ __forceinline__ void ex_store_slice(const uint32_t (&qv)[4], uint8_t *D, int lane)
{
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(D) & 15);
	if (s == 0) {
		reinterpret_cast<uint4 *>(D)[lane] = make_uint4(qv[0], qv[1], qv[2], qv[3]);
		return;
	}
	uint32_t c[8];
#pragma unroll
	for (int j = 0; j < 4; j++) {
		c[j] = __shfl_up_sync(0xffffffffu, qv[j], 1);
		c[4 + j] = qv[j];
	}
	// aligned word `lane` = bytes [16 - s, 32 - s) of (previous lane's, own) 32 bytes
	const uint32_t sh = 16 - s, bs = (sh & 3) * 8;
	uint32_t u0, u1, u2, u3, u4;
	switch (sh >> 2) {
	case 0: u0 = c[0]; u1 = c[1]; u2 = c[2]; u3 = c[3]; u4 = c[4]; break;
	case 1: u0 = c[1]; u1 = c[2]; u2 = c[3]; u3 = c[4]; u4 = c[5]; break;
	case 2: u0 = c[2]; u1 = c[3]; u2 = c[4]; u3 = c[5]; u4 = c[6]; break;
	default: u0 = c[3]; u1 = c[4]; u2 = c[5]; u3 = c[6]; u4 = c[7]; break;
	}
	uint8_t *Wd = D - s;
	if (lane > 0)
		reinterpret_cast<uint4 *>(Wd)[lane] = make_uint4(__funnelshift_r(u0, u1, bs), __funnelshift_r(u1, u2, bs),
								  __funnelshift_r(u2, u3, bs), __funnelshift_r(u3, u4, bs));
	// head: the first 16 - s bytes of lane 0; tail: the last s bytes of lane 31
	uint32_t h[4], tl[4];
#pragma unroll
	for (int j = 0; j < 4; j++) {
		h[j] = __shfl_sync(0xffffffffu, qv[j], 0);
		tl[j] = __shfl_sync(0xffffffffu, qv[j], 31);
	}
	if ((uint32_t) lane < sh) {
		const uint32_t wv = lane < 4 ? h[0] : lane < 8 ? h[1] : lane < 12 ? h[2] : h[3];
		D[lane] = (uint8_t) (wv >> (8 * (lane & 3)));
	}
	if ((uint32_t) lane < s) {
		const uint32_t b = sh + lane;	// byte of lane 31's 16
		const uint32_t wv = b < 4 ? tl[0] : b < 8 ? tl[1] : b < 12 ? tl[2] : tl[3];
		Wd[512 + lane] = (uint8_t) (wv >> (8 * (b & 3)));
	}
}

// the bytes of qv selected by `ne`, lane after lane -> D (through the warp's staging line)
__device__ __forceinline__ void ex_store_slice_packed(const uint32_t (&qv)[4], uint32_t ne, uint8_t *D, uint8_t *line, int lane)
{
	const uint32_t cnt = __popc(ne);
	const uint32_t inc = warp_inc_scan(cnt, lane);
	const uint32_t L = __shfl_sync(0xffffffffu, inc, 31);
	uint8_t *p = line + (reinterpret_cast<uintptr_t>(D) & 7) + inc - cnt;
#pragma unroll
	for (int e = 0; e < 16; e++)
		if ((ne >> e) & 1)
			*p++ = (uint8_t) ((qv[e >> 2] >> (8 * (e & 3))));
	__syncwarp();
	warp_copy_out(line, D, L, lane);
	__syncwarp();
}

#ifndef PG_EMIT_RING
#define PG_EMIT_RING 4	// tiles of low bytes per warp on their way through shared memory (4 / 6 / 8: the same)
#endif
// four words holding the 16 bytes at stage[16 i + al .. + 16), 0 < al < 16 the same in every lane
// (two conflict-free 128-bit loads)
__device__ __forceinline__ void stage_words16(const uint8_t *stage, uint32_t al, int i, uint32_t (&w)[4])
{
	const uint4 a = reinterpret_cast<const uint4 *>(stage)[i], b = reinterpret_cast<const uint4 *>(stage)[i + 1];
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
#ifndef PG_EMIT_CTAS
#define PG_EMIT_CTAS 5	// (48 registers: nothing of the slices ahead is held in them; 4: 52.4, 5: 53.4, 6: 53.2, 8: 45.8 %)
#endif
__global__ void __launch_bounds__(NT, PG_EMIT_CTAS) ex_emit_bytes_kernel(const __grid_constant__ ExArgs a)
{
	__shared__ __align__(16) uint8_t s_line[NW][512 + 16];
	__shared__ __align__(16) uint8_t s_ring[NW][PG_EMIT_RING][512];
	__shared__ uint32_t s_read;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (a.hdr->error & (ERR_CAPACITY | ERR_SCRATCH))
		return;
	uint8_t *line = s_line[warp];
	uint32_t next = tid == 0 ? atomicAdd(&a.hdr->ticket2, 1u) : 0u;	// (the ticket after this one is asked for a read early)
	for (;;) {
		__syncthreads();
		if (tid == 0)
			s_read = next;
		__syncthreads();
		const uint64_t r = s_read;
		if (r >= a.nreads)
			break;
		if (tid == 0)
			next = atomicAdd(&a.hdr->ticket2, 1u);
		ExTile T;
		uint32_t g0;
		(void) ex_read_tile0(a, r, T, g0);
		const uint32_t hdr = __ldg(a.r_hdr + r);
		const uint32_t q = a.method == PRESS_GPU_EX_ZD ? __ldg(a.r_q + r) : 0u;
		const uint64_t oo = a.out_offsets[r];
		if (hdr == X_FAIL)
			continue;
		const uint64_t n = T.rend - T.off0, A = T.off0 >> 3;
		const uint32_t nt = T.ntiles;
		const uint64_t rec0 = (uint64_t) g0 * NW + warp;
		uint8_t *pay = a.out + oo + hdr;	// first payload byte of the read
		const int64_t mm = (int64_t) n - 1;
		// a slice at one of the read's ends: samples again, element masks
		auto edge_finish = [&](uint32_t t, const ExRaw &C, uint32_t eb) {
			const int64_t f = (int64_t) t * XT - T.m + 512 * warp;	// read index of the slice's first element
			if (f >= (int64_t) n)
				return;
			T.t = t;
			T.k0 = A + (uint64_t) t * XG;
			// values of the read before this slice, minus the exceptions among them
			const int64_t vb = f - 1 < 0 ? 0 : (f - 1 < mm ? f - 1 : mm);
			ExVals v;
			ex_compute(C, T, q, v);
			const uint32_t qv[4] = {__byte_perm(v.z[0], v.z[1], 0x6420), __byte_perm(v.z[2], v.z[3], 0x6420),
						__byte_perm(v.z[4], v.z[5], 0x6420), __byte_perm(v.z[6], v.z[7], 0x6420)};
			ex_store_slice_packed(qv, v.valid & ~v.exc & ~v.first, pay + ((uint64_t) vb - eb), line, lane);
		};
		auto edge_slice = [&](uint32_t t) {
			if ((int64_t) t * XT - T.m + 512 * warp >= (int64_t) n)
				return;
			ExRaw C;
			ex_issue(a.samples, a.nsamples, A + (uint64_t) t * XG, C);
			edge_finish(t, C, __ldg(a.t_eb + rec0 + (uint64_t) t * NW));
		};
		// the samples of tiles t_in and t_in + 1 (the read's last slices of this warp) asked from L2
		auto prefetch_tail = [&](uint32_t t_in) {
#pragma unroll
			for (int k = 0; k < 2; k++) {
				const uint64_t g = A + (uint64_t) (t_in + k) * XG + 2 * tid;
				if (t_in + k < nt && 8 * g + 16 <= a.nsamples)
					prefetch_l2(a.samples + g);
			}
		};
		// tiles [1, t_in): the warp's slice lies inside the read
		const uint64_t span = T.m + n;
		const uint32_t t_in = span >= 512u * (warp + 1) ? (uint32_t) ((span - 512u * (warp + 1)) / XT) + 1 : 0u;
		// everything the start of the read needs is asked for at once: tile 0's samples, the
		// first records, the first low bytes (L2)
		ExRaw E0;
		ex_issue(a.samples, a.nsamples, A, E0);
		const uint32_t eb0 = __ldg(a.t_eb + rec0);
		if (t_in > 1) {
			// records two tiles ahead, low bytes PG_EMIT_RING tiles ahead (shared memory; the first ones
			// also asked from L2 at once); every address advances by a constant.  A slice's low bytes are
			// read from the ring at the DESTINATION's 16-byte phase s, so that lane j holds the aligned
			// output word j, and one byte for each of the < 16 bytes either side.  A slice with an
			// exception is computed again from its samples (asked from L2 when its record arrives).
			const uint8_t *lp = reinterpret_cast<const uint8_t *>(ex_lowb_ptr(a.lowb, A, 1, 32 * warp));
			const uint4 *sp = a.samples + A + XG + 2 * tid;
			const uint32_t *pe = a.t_eb + rec0 + NW;
			const ptrdiff_t dnx = a.t_nex - a.t_eb;
			uint8_t *Db = pay + (XT - T.m + 512 * warp - 1);
			uint32_t eb1 = __ldg(pe), nx1 = __ldg(pe + dnx), eb2 = 0, nx2 = 0;
			if (t_in > 2) {
				eb2 = __ldg(pe + NW);
				nx2 = __ldg(pe + dnx + NW);
			}
#pragma unroll
			for (int k = 0; k < PG_PF_EMIT; k++)
				if (t_in > 1 + k)
					prefetch_l2(lp + k * XT + 16 * lane);
			prefetch_tail(t_in);
			edge_finish(0, E0, eb0);
			// the warp's low bytes, PG_EMIT_RING tiles ahead, through a ring in shared memory (cp.async: no
			// register waits for them, and the destination's phase is dealt with when they are read)
			const uint8_t *lp0 = lp;
			auto ring_issue = [&](uint32_t t) {
				if (t < t_in)
					cp_async16(s_ring[warp][t % PG_EMIT_RING] + 16 * lane, lp0 + (uint64_t) (t - 1) * XT + 16 * lane);
				cp_async_commit();
			};
#pragma unroll
			for (int k = 1; k <= PG_EMIT_RING; k++)
				ring_issue(k);
			if (nx1)
				prefetch_l2(sp);
#pragma unroll 1
			for (uint32_t t = 1; t < t_in; t++) {
				const uint32_t eb = eb1, nx = nx1;
				uint8_t *D = Db - eb;
				const uint4 *spc = sp;
				eb1 = eb2;
				nx1 = nx2;
				sp += XG;
				pe += NW;
				Db += XT;
				if (t + 1 < t_in) {
					if (nx1)
						prefetch_l2(sp);
					if (t + 2 < t_in) {
						eb2 = __ldg(pe + NW);
						nx2 = __ldg(pe + dnx + NW);
					}
				}
				asm volatile("cp.async.wait_group %0;" ::"n"(PG_EMIT_RING - 1) : "memory");
				__syncwarp();
				if (nx == 0) {
					const uint8_t *S = s_ring[warp][t % PG_EMIT_RING];
					const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(D) & 15);
					if (s == 0) {
						reinterpret_cast<uint4 *>(D)[lane] = reinterpret_cast<const uint4 *>(S)[lane];
					} else {
						if (lane > 0) {	// aligned word `lane` = slice bytes [16 lane - s, 16 lane - s + 16)
							uint32_t w[4];
							stage_words16(S - 16, 16 - s, lane, w);
							reinterpret_cast<uint4 *>(D - s)[lane] = make_uint4(w[0], w[1], w[2], w[3]);
						}
						if (lane < 16)	// the 16 - s bytes before the first aligned word, the s bytes after the last
							D[(uint32_t) lane < 16 - s ? lane : 496 + lane] = S[(uint32_t) lane < 16 - s ? lane : 496 + lane];
					}
				} else {
					ExRaw R;
					R.xa = __ldg(spc);
					R.xb = __ldg(spc + 1);
					R.prev0 = lane == 0 ? __ldg(reinterpret_cast<const uint16_t *>(spc) - 1) : 0u;
					ExVals v;
					uint32_t orw;
					(void) ex_compute_fast(R, q, v, orw);
					const uint32_t qv[4] = {__byte_perm(v.z[0], v.z[1], 0x6420), __byte_perm(v.z[2], v.z[3], 0x6420),
								__byte_perm(v.z[4], v.z[5], 0x6420), __byte_perm(v.z[6], v.z[7], 0x6420)};
					ex_store_slice_packed(qv, ~v.exc & 0xffffu, D, line, lane);
				}
				__syncwarp();	// (every lane has read the stage)
				ring_issue(t + PG_EMIT_RING);
			}
			cp_async_wait_all();
			__syncwarp();
		} else {
			edge_finish(0, E0, eb0);
		}
		for (uint32_t t = t_in > 1 ? t_in : 1u; t < nt; t++)
			edge_slice(t);
	}
}

}  // namespace pg
