// svb_enc.cuh -- fused zigzag-delta + streamvbyte ENCODE: one persistent CTA per
// SM, warp-specialised dataflow over a shared-memory ring filled by TMA.
//
//   KEYBITS = 1 : ONT svb16, 1 key bit / value   (svb16/encode_scalar.hpp:14-62; press.c:1678-1687)
//   KEYBITS = 2 : classical streamvbyte over zd widened to u32; zd < 65536 so only
//                 codes 0/1 occur (streamvbyte_encode.c:11-59; press.c:1590-1598)
//   FRAME       : the [n u32 LE] prefix of the pre-zstd VBZ stream (press.c:1875-1879, 2035-2039)
//
// Stream of read r:  [frame][keys: ceil(n / (8 / KEYBITS)) bytes][data bytes].
//
// One pass over HBM.  A tile is 8192 samples of ONE read cut on 16-byte
// boundaries of the sample array; tiles are dealt round-robin to the CTAs.
// Every tile walks through the ring stage it was loaded into:
//
//   producer warp   tile metadata -> ONE cp.async.bulk (the tile, the group before
//                   it for the previous sample, the group after it for the last key)
//   12 A warps      pull (tile, 1024-sample block) jobs: 4 x LDS.128 per lane, delta +
//                   zigzag on 16-bit pairs (VIADD.16x2 / PRMT).  No two-byte value
//                   in the block (the common case on nanopore signal): its 1024
//                   data bytes are packed into the middle of the block's own staged
//                   samples and the byte count (1024) is posted; else the exact
//                   count is posted and the samples stay as they are
//   12 B warps      pull (tile, block) jobs once the tile's container position is
//                   known: aligned 64-bit stores of the packed bytes re-phased to
//                   the destination, or (blocks with two-byte values, first / last
//                   blocks of a read) per-sample placement through a staging line
//   publish warp    the tile's byte count -> global status word, as soon as it is known
//   window warp +   batch-wide prefix of the byte counts (the key sections are data
//   prefix warp     independent, plan.cuh) by a sliding window: P(g) = P(g - G) +
//                   the G aggregates in between, G = grid size; the window holds
//                   other CTAs' tiles only and is fetched several tiles ahead by a
//                   bulk copy; then the tile's position, the read's key pointer,
//                   capacity checks, frame
//
// A runs ahead of B by up to the ring depth, so the look-back latency (a few L2
// round trips) and the occasional slow general-path block never hold up the
// counts other CTAs are waiting for.
#pragma once

#include "common.cuh"
#include "pipe.cuh"
#include "svb_zd.cuh"

namespace pg {

constexpr int EN_TS = 8192;			// samples per tile
constexpr int EN_TG = EN_TS / 8;		// 16-byte groups per tile
constexpr int EN_NB = 8;			// 1024-sample blocks per tile
constexpr int EN_NA = 12;			// A warps
constexpr int EN_NBW = 12;			// B warps
constexpr int EN_EG = EN_TG / EN_NB;		// groups per block (128)
constexpr int EN_EC = EN_EG / 32;		// groups per lane (4)
constexpr int EN_W_B = EN_NA;			// first B warp
constexpr int EN_W_SCAN = EN_NA + EN_NBW;	// prefix warp; + 1: publish warp
constexpr int EN_W_PROD = EN_W_SCAN + 3;	// (+ 2: window warp)
constexpr int EN_THREADS = (EN_W_PROD + 1) * 32;
#ifndef PG_EN_STAGES
#define PG_EN_STAGES 13
#endif
constexpr int EN_STAGES = PG_EN_STAGES;	// ring depth: tiles in flight per CTA.  Throughput follows it (6 / 9 / 12 stages:
					// 38 / 51 / 59 % of HBM peak before phase A packed every block, r02): the kernel is bound by tile lifetime x ring size
constexpr int EN_STAGE_BYTES = (EN_TG + 2) * 16 + 96;	// slots 0 .. TG+1, padded to 128
constexpr int EN_WS = 4;				// window sums in flight between the window and the prefix warp
constexpr int EN_WW = 5;				// window words per lane: grid <= 32 * EN_WW + 1
constexpr int EN_MAX_GRID = 32 * EN_WW + 1;
constexpr int EN_WIN_WORDS = 32 * EN_WW + 4;	// window + alignment slack
constexpr int EN_PACK_OFF = 512;			// packed bytes of a fast block: middle of its 2 KB of samples
// A block that is not fast (a two-byte value, a read boundary) is packed by phase A as well when
// its bytes fit between the block's first and last group, which the neighbouring blocks read and
// which therefore stay intact: data bytes from EN_SLOW_DATA on, the block's key bytes at
// EN_SLOW_KEYS.  (Phase B then only copies; before r02 it redid the block sample by sample while
// holding the tile's ring stage: 6800 cycles per such block, 0.69 of them per tile.)
constexpr int EN_SLOW_DATA = 16;
constexpr int EN_SLOW_KEYS = 16 * EN_EG - 16 - 256;
constexpr int EN_SLOW_MAX = EN_SLOW_KEYS - EN_SLOW_DATA;	// data bytes a packed slow block can hold (1760)

struct EnTileInfo {
	uint64_t off0, rend;	// sample range of the tile's read
	uint64_t kx0, kx1;	// fixed-byte prefix before / through the read (plan.cuh)
	uint64_t k0;		// first 16-byte group of the tile
	uint32_t g, r, t;	// tile id, read id, tile index inside the read
	uint32_t flags;		// bit b < 8: block b and the group after it lie inside the read, past its
				// first group; bit 31: last tile of the batch
};

struct EnSmem {
	alignas(128) uint8_t in[EN_STAGES][EN_STAGE_BYTES];
	EnTileInfo info[EN_STAGES];
	uint64_t full[EN_STAGES], cnt_done[EN_STAGES], pub_done[EN_STAGES], scan_done[EN_STAGES], empty[EN_STAGES];
	uint64_t tileP[EN_STAGES];		// container position of the tile's data (prefix warp);
						// SENTINEL: the tile does not fit, write nothing
	uint32_t blkOff[EN_STAGES][EN_NB];	// data bytes of the tile's earlier blocks (publish warp)
	uint32_t tileL[EN_STAGES];		// data bytes of the tile (publish warp)
	uint64_t kbase[EN_STAGES];		// container position of the read's key section
	uint32_t blkL[EN_STAGES][EN_NB];	// data bytes of each 1024-sample block (phase A)
	uint32_t blkF[EN_STAGES][EN_NB];	// phase A left the block 1: packed at EN_PACK_OFF (fast), 2: packed at
						// EN_SLOW_DATA with its keys at EN_SLOW_KEYS, 0: as loaded (phase B codes it)
	uint32_t ajob, bjob;			// next (tile, block) job of the A / B warps
	uint64_t wsum_done[EN_WS], wsum_free[EN_WS];
	uint2 wsum[EN_WS];				// window sum, sum of its last t words
};

// word holding the sample before each of x's words' low halves
__device__ __forceinline__ void prev_words(const uint4 &x, uint32_t prev_half, uint32_t (&p)[4])
{
	p[0] = __byte_perm(prev_half, x.x, 0x5410);
	p[1] = __funnelshift_l(x.x, x.y, 16);
	p[2] = __funnelshift_l(x.y, x.z, 16);
	p[3] = __funnelshift_l(x.z, x.w, 16);
}

// valid-sample mask of group k of a read [off0, rend)
__device__ __forceinline__ uint32_t group_valid(uint64_t k, uint64_t off0, uint64_t rend)
{
	const int64_t lo64 = (int64_t) off0 - (int64_t) (8 * k);
	const int64_t left = (int64_t) rend - (int64_t) (8 * k);
	const int lo = lo64 > 0 ? (lo64 < 8 ? (int) lo64 : 8) : 0;
	const int hi = left >= 8 ? 8 : (left > 0 ? (int) left : 0);
	return range_mask(lo, hi);
}

// ---------------------------------------------------------------------------
// worker warps: one 1024-sample block of a tile.  The code a warp runs for the
// common block is kept short and straight (the kernel is instruction-fetch
// bound otherwise: five roles share one SM's instruction cache); everything
// rare sits in compact, not unrolled, out-of-line routines.

// zd words of group gi of the staged tile (gi = -1 ... EN_TG: group before / after
// included); the sample before the group comes straight from shared memory
__device__ __forceinline__ void en_group_zd(const EnTileInfo &ti, const uint8_t *st, int gi, uint32_t (&z)[4])
{
	const uint4 x = *reinterpret_cast<const uint4 *>(st + 16 * (gi + 1));
	const uint32_t prev = *reinterpret_cast<const uint16_t *>(st + 16 * (gi + 1) - 2);
	// the read's first sample is coded against 0 (trans.c:221)
	const bool first = ti.t == 0 && gi == 0;
	zd_group(x, prev, first ? (int) (ti.off0 & 7) : -1, z);
}

// Phase A of a block that is not fast: count, and when the bytes fit (EN_SLOW_MAX) pack data and
// key bytes into the block's own stage area.  All four columns are loaded before the first byte
// is stored.  Returns the data-byte count; packed = false leaves the block as loaded.
template <int KEYBITS>
__device__ __noinline__ uint32_t en_pack_slow(const EnTileInfo &ti, uint8_t *st, int bw, int lane, bool &packed)
{
	const uint64_t kw = ti.k0 + bw * EN_EG;
	packed = true;
	if (8 * kw >= ti.rend)
		return 0;	// block past the end of the read
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	uint32_t z[EN_EC][4], valid[EN_EC], fl[EN_EC];
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		en_group_zd(ti, st, bw * EN_EG + 32 * c + lane, z[c]);
		valid[c] = group_valid(kw + 32 * c + lane, ti.off0, ti.rend);
		fl[c] = flags_group(z[c]) & valid[c];
	}
	uint32_t fafter;	// flags of the group after the block (its first key bits belong to our last key byte)
	{
		uint32_t zn[4];
		en_group_zd(ti, st, bw * EN_EG + EN_EG, zn);
		fafter = flags_group(zn) & group_valid(kw + EN_EG, ti.off0, ti.rend);
	}
	__syncwarp();	// every lane holds its samples: the stage may be overwritten
	const uint32_t sa = (__popc(valid[0]) + __popc(fl[0])) | ((__popc(valid[1]) + __popc(fl[1])) << 16);
	const uint32_t sb = (__popc(valid[2]) + __popc(fl[2])) | ((__popc(valid[3]) + __popc(fl[3])) << 16);
	const uint32_t ia = warp_inc_scan(sa, lane), ib = warp_inc_scan(sb, lane);
	const uint32_t ta = __shfl_sync(0xffffffffu, ia, 31), tb = __shfl_sync(0xffffffffu, ib, 31);
	const uint32_t t0 = ta & 0xffffu, t1 = ta >> 16, t2 = tb & 0xffffu, t3 = tb >> 16;
	const uint32_t L = t0 + t1 + t2 + t3;
	if (L > (uint32_t) EN_SLOW_MAX) {
		packed = false;
		return L;
	}
	const uint32_t off[EN_EC] = {(ia & 0xffffu) - (sa & 0xffffu), t0 + (ia >> 16) - (sa >> 16),
				     t0 + t1 + (ib & 0xffffu) - (sb & 0xffffu), t0 + t1 + t2 + (ib >> 16) - (sb >> 16)};
	uint8_t *blk = st + 16 * (bw * EN_EG + 1);
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		uint8_t *p = blk + EN_SLOW_DATA + off[c];
#pragma unroll
		for (int e = 0; e < 8; e++) {
			const uint32_t v = (z[c][e >> 1] >> (16 * (e & 1))) & 0xffffu;
			if ((valid[c] >> e) & 1) {
				*p++ = (uint8_t) v;
				if ((fl[c] >> e) & 1)
					*p++ = (uint8_t) (v >> 8);
			}
		}
		// keys: read-relative key block j of group k is k - A: bits m..7 of group j's flags +
		// bits 0..m-1 of group j + 1's (the next lane's, the next column's first, the group after)
		uint32_t fnext = __shfl_down_sync(0xffffffffu, fl[c], 1);
		const uint32_t wrap = __shfl_sync(0xffffffffu, fl[c < EN_EC - 1 ? c + 1 : c], 0);
		if (lane == 31)
			fnext = c < EN_EC - 1 ? wrap : fafter;
		const uint32_t f = ((fl[c] >> m) | (fnext << (8 - m))) & 0xffu;
		if (KEYBITS == 1)
			blk[EN_SLOW_KEYS + 32 * c + lane] = (uint8_t) f;
		else
			reinterpret_cast<uint16_t *>(blk + EN_SLOW_KEYS)[32 * c + lane] = (uint16_t) spread8(f);
	}
	return L;
}

// Phase A: byte count of the block -> blkL[bw].  A fast block (inside the read, no
// two-byte value) also leaves its 1024 data bytes packed in the middle of its own
// 2 KB of staged samples: the first and last group of the block, which the
// neighbouring blocks read, stay intact.
template <int KEYBITS>
__device__ __forceinline__ void en_phase_a(const EnTileInfo &ti, uint8_t *st, uint32_t *blkL, uint32_t *blkF,
					   int bw, int lane)
{
	bool fast = (ti.flags >> bw) & 1u;	// block and the group after it inside the read
	if (fast) {
		const uint32_t g0 = bw * EN_EG;
		uint4 x[EN_EC];
#pragma unroll
		for (int c = 0; c < EN_EC; c++)
			x[c] = *reinterpret_cast<const uint4 *>(st + 16 * (g0 + 32 * c + lane + 1));
		const uint32_t prev0 = *reinterpret_cast<const uint16_t *>(st + 16 * g0 + 14);
		const uint4 xh = *reinterpret_cast<const uint4 *>(st + 16 * (g0 + EN_EG + 1));
		uint32_t z[EN_EC][4], zh[4];
#pragma unroll
		for (int c = 0; c < EN_EC; c++) {
			uint32_t prev = __shfl_up_sync(0xffffffffu, x[c].w >> 16, 1);
			if (c > 0) {
				const uint32_t wrap = __shfl_sync(0xffffffffu, x[c > 0 ? c - 1 : 0].w >> 16, 31);
				if (lane == 0)
					prev = wrap;
			} else if (lane == 0) {
				prev = prev0;
			}
			zd_group(x[c], prev, -1, z[c]);
		}
		zd_group(xh, __shfl_sync(0xffffffffu, x[EN_EC - 1].w >> 16, 31), -1, zh);
		uint32_t high = zh[0] | zh[1] | zh[2] | zh[3];
#pragma unroll
		for (int c = 0; c < EN_EC; c++)
			high |= z[c][0] | z[c][1] | z[c][2] | z[c][3];
		fast = !__any_sync(0xffffffffu, (high & 0xff00ff00u) != 0);	// also orders the loads
		if (fast) {
			uint2 *pk = reinterpret_cast<uint2 *>(st + 16 * (g0 + 1) + EN_PACK_OFF);
#pragma unroll
			for (int c = 0; c < EN_EC; c++)
				pk[32 * c + lane] = make_uint2(__byte_perm(z[c][0], z[c][1], 0x6420),
							       __byte_perm(z[c][2], z[c][3], 0x6420));
		}
	}
	uint32_t cnt = 8 * EN_EG, mode = 1;
	if (!fast) {
		bool packed;
		cnt = en_pack_slow<KEYBITS>(ti, st, bw, lane, packed);
		mode = packed ? 2u : 0u;
	}
	if (lane == 0) {
		blkF[bw] = mode;
		blkL[bw] = cnt;
	}
}

// Phase B of a block packed by en_pack_slow: L data bytes from the stage to S (any alignment) as
// aligned 64-bit stores (bytes at the edges), then the block's key bytes.
template <int KEYBITS>
__device__ __noinline__ void en_emit_packed(const SvbEncArgs &a, const EnTileInfo &ti, const uint8_t *st, uint64_t P,
					    uint64_t kbase, uint32_t L, int bw, int lane)
{
	const uint64_t kw = ti.k0 + bw * EN_EG;
	if (8 * kw >= ti.rend)
		return;			// block past the end of the read
	const uint8_t *blk = st + 16 * (bw * EN_EG + 1);
	uint8_t *S = a.out + P;
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(S) & 7), sh = 8 * s;
	const uint64_t *pk = reinterpret_cast<const uint64_t *>(blk + EN_SLOW_DATA);
	uint8_t *Wb = S - s;
	const uint32_t end = s + L, nwords = (end + 7) >> 3;
	for (uint32_t q = lane; q < nwords; q += 32) {
		// aligned word q = packed bytes [8 q - s, 8 q - s + 8)  (pk[-1] is the block's first group: masked below)
		const uint64_t hi = pk[q], lo = s ? pk[(int) q - 1] : 0ull;
		const uint64_t v = s ? (lo >> (64 - sh)) | (hi << sh) : hi;
		const uint32_t b0 = q << 3, b1 = b0 + 8;
		if (b0 >= s && b1 <= end) {
			*reinterpret_cast<uint64_t *>(Wb + b0) = v;
		} else {
			const uint32_t f = b0 < s ? s : b0, l = b1 > end ? end : b1;
			for (uint32_t b = f; b < l; b++)
				Wb[b] = (uint8_t) (v >> (8 * (b - b0)));
		}
	}
	// keys: group 32 c + lane of the block is key block j of the read (svb16: one byte, svb32: two)
	const uint64_t n = ti.rend - ti.off0;
	const uint64_t jw = (uint64_t) ti.t * EN_TG + bw * EN_EG;
	uint8_t *kout = a.out + kbase;
#pragma unroll
	for (int c = 0; c < EN_EC; c++) {
		const uint64_t j = jw + 32 * c + lane;
		if (KEYBITS == 1) {
			if (8 * j < n)
				kout[j] = blk[EN_SLOW_KEYS + 32 * c + lane];
		} else {
			const uint32_t kk = reinterpret_cast<const uint16_t *>(blk + EN_SLOW_KEYS)[32 * c + lane];
			if (8 * j < n)
				kout[2 * j] = (uint8_t) kk;
			if (8 * j + 4 < n)
				kout[2 * j + 1] = (uint8_t) (kk >> 8);
		}
	}
}

// Last resort of phase B, for blocks en_pack_slow could not pack (more than EN_SLOW_MAX data
// bytes: most values take two bytes, i.e. not a nanopore signal): per-sample placement straight
// into the container, key bytes from the two-byte flags.  Group by group, nothing unrolled.
template <int KEYBITS>
__device__ __noinline__ void en_emit_slow(const SvbEncArgs &a, const EnTileInfo &ti, const uint8_t *st,
					  uint64_t P, uint64_t kbase, int bw, int lane)
{
	const uint64_t kw = ti.k0 + bw * EN_EG;
	if (8 * kw >= ti.rend)
		return;			// block past the end of the read
	const uint64_t n = ti.rend - ti.off0;
	const uint32_t m = (uint32_t) (ti.off0 & 7);
	const uint64_t jw = (uint64_t) ti.t * EN_TG + bw * EN_EG;	// read-relative group of the block
	uint8_t *S = a.out + P, *kout = a.out + kbase;
	uint32_t run = 0;
#pragma unroll 1
	for (int c = 0; c < EN_EC; c++) {
		const int gi = bw * EN_EG + 32 * c + lane;
		uint32_t z[4], zn[4];
		en_group_zd(ti, st, gi, z);
		const uint32_t valid = group_valid(kw + 32 * c + lane, ti.off0, ti.rend);
		const uint32_t fl = flags_group(z) & valid;
		const uint32_t cnt = __popc(valid) + __popc(fl);
		const uint32_t inc = warp_inc_scan(cnt, lane);
		uint8_t *p = S + run + inc - cnt;
		run += __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
		for (int e = 0; e < 8; e++) {
			const uint32_t v = (z[e >> 1] >> (16 * (e & 1))) & 0xffffu;
			if ((valid >> e) & 1) {
				*p++ = (uint8_t) v;
				if ((fl >> e) & 1)
					*p++ = (uint8_t) (v >> 8);
			}
		}
		// keys: the read-relative 8-sample block of group k is k - A; key block j = bits
		// m..7 of group j's flags + bits 0..m-1 of group j + 1's
		uint32_t fnext = __shfl_down_sync(0xffffffffu, fl, 1);
		en_group_zd(ti, st, bw * EN_EG + 32 * c + 32, zn);	// the group after lane 31's
		if (lane == 31)
			fnext = flags_group(zn) & group_valid(kw + 32 * c + 32, ti.off0, ti.rend);
		const uint64_t j = jw + 32 * c + lane;
		const uint32_t f = ((fl >> m) | (fnext << (8 - m))) & 0xffu;
		if (KEYBITS == 1) {
			if (8 * j < n)
				kout[j] = (uint8_t) f;
		} else {
			const uint32_t kk = spread8(f);
			if (8 * j < n)
				kout[2 * j] = (uint8_t) kk;
			if (8 * j + 4 < n)
				kout[2 * j + 1] = (uint8_t) (kk >> 8);
		}
	}
}

// Phase B of a fast block: its packed bytes to the container, key bytes zero.
template <int KEYBITS>
__device__ __forceinline__ void en_emit_fast(const SvbEncArgs &a, const EnTileInfo &ti, const uint8_t *st,
					     uint64_t P, uint64_t kbase, int bw, int lane)
{
	const uint64_t jw = (uint64_t) ti.t * EN_TG + bw * EN_EG;	// read-relative group of the block
	uint8_t *S = a.out + P;
	const uint32_t s = (uint32_t) (reinterpret_cast<uintptr_t>(S) & 7);
	uint64_t B[EN_EC];
	const uint64_t *pk = reinterpret_cast<const uint64_t *>(st + 16 * (bw * EN_EG + 1) + EN_PACK_OFF);
#pragma unroll
	for (int c = 0; c < EN_EC; c++)
		B[c] = pk[32 * c + lane];
	// chunk i = 32 c + lane holds exactly 8 data bytes; aligned word i of the output =
	// top s bytes of chunk i - 1 | low 8 - s bytes of chunk i
	uint64_t *W = reinterpret_cast<uint64_t *>(S - s);
	if (s == 0) {
#pragma unroll
		for (int c = 0; c < EN_EC; c++)
			W[32 * c + lane] = B[c];
	} else {
		const uint32_t sh = 8 * s;
#pragma unroll
		for (int c = 0; c < EN_EC; c++) {
			uint64_t pb = __shfl_up_sync(0xffffffffu, B[c], 1);
			if (c > 0) {
				const uint64_t wrap = __shfl_sync(0xffffffffu, B[c > 0 ? c - 1 : 0], 31);
				if (lane == 0)
					pb = wrap;
			}
			const uint64_t v = (pb >> (64 - sh)) | (B[c] << sh);
			if (c > 0 || lane > 0)
				W[32 * c + lane] = v;
		}
		// partial words at both ends, one byte per lane: the low 8 - s bytes of chunk 0
		// and the top s bytes of the last chunk
		const uint64_t b0 = __shfl_sync(0xffffffffu, B[0], 0);
		const uint64_t bl = __shfl_sync(0xffffffffu, B[EN_EC - 1], 31) >> (64 - sh);
		if ((uint32_t) lane < 8 - s)
			S[lane] = (uint8_t) (b0 >> (8 * lane));
		if ((uint32_t) lane < s)
			reinterpret_cast<uint8_t *>(W + EN_EG)[lane] = (uint8_t) (bl >> (8 * lane));
	}
	warp_zero(a.out + kbase + jw * KEYBITS, EN_EG * KEYBITS, lane);
}

// number of window words of the CTA's i-th tile (tiles before it that are not its own)
__device__ __forceinline__ uint32_t g_window(uint32_t i, uint32_t G)
{
	return i == 0 ? blockIdx.x : G - 1;
}

// ----------------------------------------------------------------------------
// the kernel

// role timing (cycles summed over warps and CTAs into hdr->aux[], lane 0 only);
// compiled in with -DPG_ROLE_TIMING, see tools/role_timing.py
#ifdef PG_ROLE_TIMING
#define RT_DECL long long rt_t0 = 0, rt_acc[6] = {0, 0, 0, 0, 0, 0}
#define RT_BEGIN() (rt_t0 = clock64())
#define RT_END(k) (rt_acc[k] += clock64() - rt_t0)
#define RT_FLUSH(b)                                                                          \
	do {                                                                                 \
		if (lane == 0)                                                               \
			for (int k_ = 0; k_ < 6; k_++)                                       \
				atomicAdd((unsigned long long *) &a.hdr->aux[(b) + k_],      \
					  (unsigned long long) rt_acc[k_]);                  \
	} while (0)
#else
#define RT_DECL
#define RT_BEGIN()
#define RT_END(k)
#define RT_FLUSH(b)
#endif

template <int KEYBITS, bool FRAME>
__global__ void __launch_bounds__(EN_THREADS, 1) svb_enc_kernel(const __grid_constant__ SvbEncArgs a)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	EnSmem &sm = *reinterpret_cast<EnSmem *>(smem_raw);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t total = a.hdr->total_tiles;
	const uint32_t G = gridDim.x;
	// tiles of this CTA: blockIdx.x + i * G, i < nt
	const uint32_t nt = blockIdx.x < total ? (total - blockIdx.x + G - 1) / G : 0;

	if (threadIdx.x == 0) {
		for (int s = 0; s < EN_STAGES; s++) {
			mbar_init(&sm.full[s], 1);
			mbar_init(&sm.cnt_done[s], EN_NB);
			mbar_init(&sm.pub_done[s], 1);
			mbar_init(&sm.scan_done[s], 1);
			mbar_init(&sm.empty[s], EN_NB);
		}
		for (int u = 0; u < EN_WS; u++) {
			mbar_init(&sm.wsum_done[u], 1);
			mbar_init(&sm.wsum_free[u], 1);
		}
		sm.ajob = 0;
		sm.bjob = 0;
		mbar_fence_init();
	}
	__syncthreads();
	if (nt == 0)
		return;

	if (warp == EN_W_PROD) {
		// ---- producer: metadata 32 tiles at a time, one bulk copy per tile ----------
		const uint64_t nfull = a.nsamples >> 3;
		RT_DECL;
		for (uint32_t i0 = 0; i0 < nt; i0 += 32) {
			const uint32_t i = i0 + lane;
			EnTileInfo me = {};
			if (i < nt) {
				me.g = blockIdx.x + i * G;
				me.r = __ldg(a.tile_read + me.g);
				me.t = me.g - __ldg(a.tile_start + me.r);
				me.off0 = __ldg(a.offsets + me.r);
				me.rend = __ldg(a.offsets + me.r + 1);
				me.kx0 = __ldg(a.kx + me.r);
				me.kx1 = __ldg(a.kx + me.r + 1);
				const uint64_t A = me.off0 >> 3;
				me.k0 = A + (uint64_t) me.t * EN_TG;
				uint32_t fl = me.g == total - 1 ? 0x80000000u : 0u;
				for (int b = 0; b < EN_NB; b++) {
					const uint64_t kw = me.k0 + (uint64_t) b * EN_EG;
					if (kw > A && 8 * (kw + EN_EG + 1) <= me.rend)
						fl |= 1u << b;
				}
				me.flags = fl;
			}
			const uint32_t cnt = nt - i0 < 32 ? nt - i0 : 32;
			for (uint32_t l = 0; l < cnt; l++) {
				const uint32_t it = i0 + l, s = it % EN_STAGES;
				// the whole warp waits: a lane running ahead of its turn would see the
				// parity of a phase that has not started yet as "complete"
				RT_BEGIN();
				if (it >= EN_STAGES)
					mbar_wait(&sm.empty[s], ((it / EN_STAGES) & 1) ^ 1);
				RT_END(0);
				if (lane == (int) l) {
					sm.info[s] = me;
					// groups [kf, ke) of the sample array -> slots kf - (k0 - 1) ...
					const uint64_t kf = me.k0 > 0 ? me.k0 - 1 : 0;
					uint64_t kr = (me.rend + 7) >> 3;	// groups holding samples of the read
					uint64_t ke = me.k0 + EN_TG + 1;
					ke = ke < kr ? ke : kr;
					if (ke > nfull) {	// the array's last, partial group: by hand, zero filled
						const uint4 v = load_group(a.samples, a.nsamples, nfull);
						*reinterpret_cast<uint4 *>(sm.in[s] + 16 * (nfull + 1 - me.k0)) = v;
						ke = nfull;
					}
					if (ke > kf) {
						const uint32_t bytes = (uint32_t) (ke - kf) * 16;
						mbar_arrive_expect_tx(&sm.full[s], bytes);
						tma_load_1d(sm.in[s] + 16 * (kf + 1 - me.k0), a.samples + kf, bytes,
							    &sm.full[s]);
					} else {
						mbar_arrive(&sm.full[s]);
					}
				}
				__syncwarp();
			}
		}
		RT_FLUSH(0);
		return;
	}

	if (warp == EN_W_SCAN + 1) {
		// ---- publish warp: the tile's byte count goes out the moment it is known; the
		// blocks' offsets inside the tile for the B warps ---------------------------------
		for (uint32_t i = 0; i < nt; i++) {
			const uint32_t s = i % EN_STAGES;
			mbar_wait(&sm.cnt_done[s], (i / EN_STAGES) & 1);
			const uint32_t lv = lane < EN_NB ? sm.blkL[s][lane] : 0u;
			uint32_t inc = lv;	// inclusive prefix over the blocks
#pragma unroll
			for (int o = 1; o < EN_NB; o <<= 1) {
				const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
				if (lane >= o)
					inc += y;
			}
			if (lane < EN_NB)
				sm.blkOff[s][lane] = inc - lv;
			if (lane == EN_NB - 1) {
				st_relaxed(a.status + blockIdx.x + (uint64_t) i * G, LB_AGG | inc);
				sm.tileL[s] = inc;
			}
			__syncwarp();
			if (lane == EN_NB - 1)
				mbar_arrive(&sm.pub_done[s]);
		}
		return;
	}

	if (warp == EN_W_SCAN + 2) {
		// ---- window warp: for every tile of the CTA the sum of the G - 1 aggregates between
		// it and the CTA's previous tile (other CTAs' tiles only).  Plain relaxed loads, five
		// words per lane, requested the moment the CTA's own count of that tile is out: the
		// other CTAs run the same round at about the same time, so a window fetched earlier
		// (a bulk copy several tiles ahead was tried) is stale and has to be fetched again --
		// one more L2 round trip on the chain every emit of the CTA waits for.
		RT_DECL;
		uint64_t w[EN_WW];
		auto request = [&](uint32_t i) {	// window of the CTA's i-th tile: words [g - wl, g)
			const uint64_t g = blockIdx.x + (uint64_t) i * G;
			const uint32_t wl = g_window(i, G);
			const uint64_t *p = a.status + (g - wl);
#pragma unroll
			for (int q = 0; q < EN_WW; q++) {
				const uint32_t j = 32 * q + lane;
				w[q] = j < wl ? ld_relaxed(p + j) : LB_AGG;
			}
		};
		RT_BEGIN();
		mbar_wait(&sm.pub_done[0], 0);
		RT_END(0);
		request(0);
		for (uint32_t i = 0; i < nt; i++) {
			const uint32_t s = i % EN_STAGES, v = i % EN_WS;
			RT_BEGIN();
			const uint32_t t = sm.info[s].t;
			const uint32_t wl = g_window(i, G);
			uint32_t W, Wt;
			for (;;) {
				uint32_t sum = 0, sumt = 0;
				bool ok = true;
#pragma unroll
				for (int q = 0; q < EN_WW; q++) {
					const uint32_t j = 32 * q + lane;	// word j of the window
					ok = ok && (w[q] >> 62) != 0;
					sum += (uint32_t) w[q];			// a tile holds < 2^15 bytes
					sumt += j + t >= wl ? (uint32_t) w[q] : 0u;
				}
				if (__all_sync(0xffffffffu, ok)) {
					W = __reduce_add_sync(0xffffffffu, sum);
					Wt = __reduce_add_sync(0xffffffffu, sumt);
					break;
				}
				request(i);	// some counts were not out yet: fetch the window again
#ifdef PG_ROLE_TIMING
				if (lane == 0)
					atomicAdd((unsigned long long *) &a.hdr->aux[24], 1ull);
#endif
			}
			RT_END(1);
			RT_BEGIN();
			if (i + 1 < nt) {	// the next tile's count is out (its info with it): request its window
				mbar_wait(&sm.pub_done[(i + 1) % EN_STAGES], ((i + 1) / EN_STAGES) & 1);
				request(i + 1);
			}
			RT_END(0);
			RT_BEGIN();
			if (i >= EN_WS)		// the prefix warp has taken the sums of tile i - EN_WS
				mbar_wait(&sm.wsum_free[v], ((i / EN_WS) & 1) ^ 1);
			if (lane == 0) {
				sm.wsum[v] = make_uint2(W, Wt);
				mbar_arrive(&sm.wsum_done[v]);
			}
			__syncwarp();
			RT_END(2);
		}
		RT_FLUSH(21);
		return;
	}

	if (warp == EN_W_SCAN) {
		// ---- prefix warp: P(g) = P(g - G) + L(g - G) + window sum.  Nothing here waits for
		// another CTA's prefix, only (through the window warp) for its counts.
		RT_DECL;
		uint64_t P = 0;		// exclusive prefix of the tile at hand
		uint32_t Lprev = 0;
		uint32_t r_prev = 0xffffffffu;
		uint64_t base_prev = 0;
		for (uint32_t i = 0; i < nt; i++) {
			const uint32_t s = i % EN_STAGES, u = i % EN_WS;
			RT_BEGIN();
			mbar_wait(&sm.pub_done[s], (i / EN_STAGES) & 1);
			RT_END(0);
			RT_BEGIN();
			const EnTileInfo &ti = sm.info[s];
			const uint32_t t = ti.t, r = ti.r;
			const uint32_t L = sm.tileL[s];
			// start of the read's stream = kx0 + prefix at the read's first tile.  When
			// that tile lies inside the window (or is this tile) it follows from the
			// window words; else it was published by that tile, at least a row ago
			// (and the previous own tile of the same read already fetched it).
			const bool far = t > g_window(i, G);
			uint64_t base = SENTINEL;
			if (far && r == r_prev)
				base = base_prev;
			else if (far && lane == 0)
				base = ld_relaxed(a.out_offsets + r);
			mbar_wait(&sm.wsum_done[u], (i / EN_WS) & 1);
			const uint2 ws = sm.wsum[u];
			__syncwarp();
			if (lane == 0)
				mbar_arrive(&sm.wsum_free[u]);
			RT_END(1);
			RT_BEGIN();
			P = i == 0 ? ws.x : P + Lprev + ws.x;
			Lprev = L;
			if (!far) {
				base = ti.kx0 + P - ws.y;
			} else if (r != r_prev) {
				if (lane == 0)
					while (base == SENTINEL)
						base = ld_relaxed(a.out_offsets + r);
				base = __shfl_sync(0xffffffffu, base, 0);
			}
			r_prev = r;
			base_prev = base;
			if (lane == 0) {
				// the tile's bytes and the read's frame + key section must fit
				const uint64_t n = ti.rend - ti.off0;
				const uint64_t kb = base + (FRAME ? 4 : 0);
				const uint64_t tp = ti.kx1 + P;
				const bool fits = tp + L <= a.out_cap &&
						  kb + (n + 8 / KEYBITS - 1) / (8 / KEYBITS) <= a.out_cap;
				if (!fits)
					atomicOr(&a.hdr->error, ERR_CAPACITY);
				if (FRAME && fits && t == 0) {	// press.c:1875-1879, 2035-2039
#pragma unroll
					for (int b = 0; b < 4; b++)
						a.out[base + b] = (uint8_t) ((uint32_t) n >> (8 * b));
				}
				if (t == 0)
					st_relaxed(a.out_offsets + r, base);
				if (ti.flags >> 31)
					st_relaxed(a.out_offsets + a.nreads, __ldg(a.kx + a.nreads) + P + L);
				sm.tileP[s] = fits ? tp : SENTINEL;
				sm.kbase[s] = kb;
				mbar_arrive(&sm.scan_done[s]);
			}
			__syncwarp();
			RT_END(2);
		}
		RT_FLUSH(3);
		return;
	}

	const uint32_t njobs = nt * EN_NB;
	if (warp < EN_W_B) {
		// ---- A warps -------------------------------------------------------------------------
		RT_DECL;
		for (;;) {
			uint32_t j = 0;
			if (lane == 0)
				j = atomicAdd(&sm.ajob, 1u);
			j = __shfl_sync(0xffffffffu, j, 0);
			if (j >= njobs)
				break;
			const uint32_t i = j / EN_NB, bw = j - i * EN_NB, s = i % EN_STAGES;
			RT_BEGIN();
			// the stage's previous tile is done (so this wait names the right phase even
			// if bulk copies complete out of order), then this tile has landed
			if (i >= EN_STAGES)
				mbar_wait(&sm.empty[s], ((i / EN_STAGES) & 1) ^ 1);
			mbar_wait(&sm.full[s], (i / EN_STAGES) & 1);
			RT_END(0);
			RT_BEGIN();
			en_phase_a<KEYBITS>(sm.info[s], sm.in[s], sm.blkL[s], sm.blkF[s], bw, lane);
			__syncwarp();
			if (lane == 0)
				mbar_arrive(&sm.cnt_done[s]);
			RT_END(1);
		}
		RT_FLUSH(9);
		return;
	}

	// ---- B warps ---------------------------------------------------------------------------------
	RT_DECL;
	for (;;) {
		uint32_t j = 0;
		if (lane == 0)
			j = atomicAdd(&sm.bjob, 1u);
		j = __shfl_sync(0xffffffffu, j, 0);
		if (j >= njobs)
			break;
		const uint32_t i = j / EN_NB, bw = j - i * EN_NB, s = i % EN_STAGES;
		RT_BEGIN();
		mbar_wait(&sm.scan_done[s], (i / EN_STAGES) & 1);
		RT_END(0);
		RT_BEGIN();
		const uint64_t P = sm.tileP[s];
		const uint32_t mode = sm.blkF[s][bw];
		const bool fast = mode == 1;
		if (P != SENTINEL) {
			if (fast)
				en_emit_fast<KEYBITS>(a, sm.info[s], sm.in[s], P + sm.blkOff[s][bw], sm.kbase[s], bw, lane);
			else if (mode == 2)
				en_emit_packed<KEYBITS>(a, sm.info[s], sm.in[s], P + sm.blkOff[s][bw], sm.kbase[s],
							sm.blkL[s][bw], bw, lane);
			else
				en_emit_slow<KEYBITS>(a, sm.info[s], sm.in[s], P + sm.blkOff[s][bw], sm.kbase[s],
						      bw, lane);
		}
		__syncwarp();
		if (lane == 0)
			mbar_arrive(&sm.empty[s]);
#ifdef PG_ROLE_TIMING
		if (fast) {
			RT_END(1);
		} else {
			RT_END(2);
			rt_acc[3]++;
		}
#endif
	}
	RT_FLUSH(15);
}

}  // namespace pg
