// plan.cuh -- per-batch plan: tile map and the data-independent byte prefix.
//
// Everything here depends only on the read lengths (d_offsets), never on the
// samples: how many tiles each read gets, which read a tile belongs to, and
// the running sum of each read's FIXED stream bytes (framing + key section,
// svb16.h:11-15 / streamvbyte_encode.c:77), so that the single-pass kernels
// only have to scan the data-dependent part.
#pragma once

#include "common.cuh"

namespace pg {

enum PlanTiling : int {
	TILING_ALIGNED = 0	// tiles start on 16-byte (8-sample) boundaries of d_samples
};

struct PlanArgs {
	const uint64_t *offsets;	// [nreads + 1] sample offsets
	uint64_t nreads;
	int tiling;
	uint32_t tile;			// samples per tile
	uint32_t frame;			// fixed bytes before the key section (0, or 4 for VBZ)
	uint32_t keydiv;		// values per key byte: 8 (svb16), 4 (svb32), 0 = none
	uint32_t fixed;			// further fixed bytes per read (e.g. 6 for vbe21's zd0 + nex)
	uint32_t *tile_start;		// [nreads + 1] out: first tile of each read
	uint64_t *kx;			// [nreads + 1] out: exclusive prefix of fixed bytes
	Hdr *hdr;
};

__device__ __forceinline__ uint32_t plan_tiles(uint64_t off, uint64_t n, int tiling, uint32_t tile)
{
	uint64_t span = (off & 7) + n;
	(void) tiling;
	uint64_t t = (span + tile - 1) / tile;
	return t ? (uint32_t) t : 1u;	// an empty read still owns one (empty) tile
}

// One CTA of 1024 threads, 4 reads per thread per sweep.
__global__ void __launch_bounds__(1024) plan_scan_kernel(PlanArgs a)
{
	__shared__ uint64_t s_t[33], s_k[33];
	__shared__ uint64_t s_carry_t, s_carry_k;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (tid == 0) {
		s_carry_t = 0;
		s_carry_k = 0;
	}
	__syncthreads();
	constexpr int IT = 4;
	for (uint64_t base = 0; base < a.nreads; base += 1024 * IT) {
		uint64_t r0 = base + (uint64_t) tid * IT;
		uint64_t tl[IT], kb[IT], st = 0, sk = 0;
#pragma unroll
		for (int i = 0; i < IT; i++) {
			uint64_t r = r0 + i;
			tl[i] = kb[i] = 0;
			if (r < a.nreads) {
				uint64_t off = a.offsets[r], n = a.offsets[r + 1] - off;
				tl[i] = plan_tiles(off, n, a.tiling, a.tile);
				kb[i] = a.frame + a.fixed +
					(a.keydiv ? (n + a.keydiv - 1) / a.keydiv : 0);
			}
			st += tl[i];
			sk += kb[i];
		}
		uint64_t it = st, ik = sk;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint64_t yt = __shfl_up_sync(0xffffffffu, it, o);
			uint64_t yk = __shfl_up_sync(0xffffffffu, ik, o);
			if (lane >= o) {
				it += yt;
				ik += yk;
			}
		}
		if (lane == 31) {
			s_t[warp] = it;
			s_k[warp] = ik;
		}
		__syncthreads();
		if (warp == 0) {
			uint64_t xt = s_t[lane], xk = s_k[lane], ot = xt, ok = xk;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				uint64_t yt = __shfl_up_sync(0xffffffffu, xt, o);
				uint64_t yk = __shfl_up_sync(0xffffffffu, xk, o);
				if (lane >= o) {
					xt += yt;
					xk += yk;
				}
			}
			s_t[lane] = xt - ot;
			s_k[lane] = xk - ok;
			if (lane == 31) {
				s_t[32] = xt;
				s_k[32] = xk;
			}
		}
		__syncthreads();
		uint64_t et = s_carry_t + s_t[warp] + it - st;
		uint64_t ek = s_carry_k + s_k[warp] + ik - sk;
#pragma unroll
		for (int i = 0; i < IT; i++) {
			uint64_t r = r0 + i;
			if (r < a.nreads) {
				a.tile_start[r] = (uint32_t) et;
				a.kx[r] = ek;
			}
			et += tl[i];
			ek += kb[i];
		}
		__syncthreads();
		if (tid == 0) {
			s_carry_t += s_t[32];
			s_carry_k += s_k[32];
		}
		__syncthreads();
	}
	if (tid == 0) {
		a.tile_start[a.nreads] = (uint32_t) s_carry_t;
		a.kx[a.nreads] = s_carry_k;
		a.hdr->total_tiles = (uint32_t) s_carry_t;
		a.hdr->fixed_total = s_carry_k;
	}
}

// tile -> read map: one warp per read.
__global__ void __launch_bounds__(256) plan_fill_kernel(const uint32_t *tile_start,
							 uint64_t nreads, uint32_t *tile_read)
{
	uint64_t r = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (r >= nreads)
		return;
	uint32_t t0 = tile_start[r], t1 = tile_start[r + 1];
	for (uint32_t t = t0 + lane; t < t1; t += 32)
		tile_read[t] = (uint32_t) r;
}

}  // namespace pg
