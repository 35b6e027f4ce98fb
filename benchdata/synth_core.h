// synth_core.h -- synthetic NA12878-like reads (the bench / test workload of
// SURVEY.md 8(d)); integer-only, so the host and device generators agree bit
// for bit and any shard of a batch can be regenerated anywhere.  NOT part of the
// product library: compiled into benchdata/libsynth_gpu.so (nvcc, the device
// generator) and benchdata/libsynth_host.so (g++ only, the host twin; the CPU
// reference arm of bench.py loads nothing else besides oracle/).
//
//   sample 0 of read r          = x0[r]
//   sample i >= 1               : h  = mix(key(seed, r) + i * GOLD)
//                                 zd = exception ? 256 + 12*clz-ish + jitter
//                                                : inverse-CDF of the 256-bin one-byte-zd histogram
//                                                  (press/gen_huffman.c:11)
//                                 d  = unzigzag(zd)                         (trans.c:80-83)
//                                 S_i = S_{i-1} + d   (unbounded walk, only S mod 4094 matters)
//                                 x_i = fold(S_i) into [0, 2047]  (triangle wave: an 11-bit ADC
//                                       never leaves its range; folding flips the sign of a delta,
//                                       which leaves the |delta| histogram alone)
//   stored sample               = x_i << shift[r]        (shift 2 gives the q = 2 reads of ex-zd)
//
// The walk is a prefix sum, so one warp per read generates 256 samples per
// iteration with one warp scan; reads are handed out by an atomic ticket.
#pragma once

#include <stdint.h>

#include <thread>
#include <vector>

#ifndef __CUDACC__
#define __host__
#define __device__
#define __forceinline__ inline
#endif

namespace pg {

constexpr uint32_t SYN_PERIOD = 4094;	// 2 * 2047
constexpr uint64_t SYN_GOLD = 0x9E3779B97F4A7C15ull;

__host__ __device__ __forceinline__ uint64_t syn_mix(uint64_t z)
{
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t syn_key(uint64_t seed, uint64_t r)
{
	return syn_mix(seed ^ syn_mix(r + SYN_GOLD));
}

// number of leading zeros of a non-zero 32-bit word, host and device
__host__ __device__ __forceinline__ uint32_t syn_clz(uint32_t v)
{
#ifdef __CUDA_ARCH__
	return (uint32_t) __clz((int) v);
#else
	return (uint32_t) __builtin_clz(v);
#endif
}

// zd symbol of sample i (i >= 1).  cdf[k] = floor(2^32 * P(sym <= k)), cdf[255] = 2^32 - 1.
__host__ __device__ __forceinline__ uint32_t syn_zd(uint64_t key, uint64_t i, const uint32_t *cdf,
						    uint32_t p_ex_u32)
{
	const uint64_t h = syn_mix(key + i * SYN_GOLD);
	const uint32_t u1 = (uint32_t) h, u2 = (uint32_t) (h >> 32);
	if (u1 < p_ex_u32)	// exception: 256 + 12 * Geom(1/2) + U[0, 12)
		return 256u + 12u * syn_clz(u2 | 1u) + (((u2 >> 1) & 0xffu) * 12u >> 8);
	uint32_t lo = 0, hi = 255;	// smallest k with u2 <= cdf[k]
	while (lo < hi) {
		uint32_t mid = (lo + hi) >> 1;
		if (u2 <= cdf[mid])
			hi = mid;
		else
			lo = mid + 1;
	}
	return lo;
}

__host__ __device__ __forceinline__ int32_t syn_unzigzag(uint32_t z)
{
	return (int32_t) (z >> 1) ^ -(int32_t) (z & 1u);
}
// residue of a (possibly negative) walk position, in [0, 4094)
__host__ __device__ __forceinline__ uint32_t syn_mod(int64_t s)
{
	int64_t m = s % (int64_t) SYN_PERIOD;
	return (uint32_t) (m < 0 ? m + SYN_PERIOD : m);
}
__host__ __device__ __forceinline__ uint32_t syn_fold(uint32_t m)
{
	return m <= 2047u ? m : SYN_PERIOD - m;
}

#ifdef __CUDACC__
struct SynthArgs {
	int16_t *samples;
	const uint64_t *offsets;
	uint64_t nreads;
	const int16_t *x0;
	const uint8_t *shift;	// may be null
	uint32_t cdf[256];
	uint32_t p_ex_u32;
	uint64_t seed;
	uint64_t first_read;	// batch index of offsets' read 0 (shards of one batch agree)
	unsigned long long *ticket;
};

__global__ void __launch_bounds__(256) synth_kernel(const __grid_constant__ SynthArgs a)
{
	__shared__ uint32_t s_cdf[256];
	__shared__ uint16_t s_stage[8][256];
	s_cdf[threadIdx.x] = a.cdf[threadIdx.x];
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint16_t *stage = s_stage[warp];
	for (;;) {
		unsigned long long r = 0;
		if (lane == 0)
			r = atomicAdd(a.ticket, 1ull);
		r = __shfl_sync(0xffffffffu, r, 0);
		if (r >= a.nreads)
			break;
		const uint64_t off = a.offsets[r], n = a.offsets[r + 1] - off;
		if (n == 0)
			continue;
		const uint64_t key = syn_key(a.seed, a.first_read + r);
		const uint32_t sh = a.shift ? a.shift[r] : 0;
		uint32_t carry = syn_mod((int64_t) a.x0[r]);	// S_{i0 - 1} mod 4094
		uint16_t *dst = reinterpret_cast<uint16_t *>(a.samples) + off;
		for (uint64_t i0 = 0; i0 < n; i0 += 256) {
			// lane owns samples i0 + 8 * lane .. + 7
			int32_t d[8], acc = 0;
#pragma unroll
			for (int e = 0; e < 8; e++) {
				const uint64_t i = i0 + 8 * lane + e;
				int32_t v = 0;
				if (i >= 1 && i < n)
					v = syn_unzigzag(syn_zd(key, i, s_cdf, a.p_ex_u32));
				acc += v;
				d[e] = acc;
			}
			int32_t inc = acc;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				int32_t y = __shfl_up_sync(0xffffffffu, inc, o);
				if (lane >= o)
					inc += y;
			}
			const int32_t before = inc - acc;	// deltas of lower lanes; |.| < 2^24
			const int32_t total = __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
			for (int e = 0; e < 8; e++) {
				uint32_t m = syn_mod((int64_t) carry + before + d[e]);
				stage[8 * lane + e] = (uint16_t) (syn_fold(m) << sh);
			}
			__syncwarp();
			const uint32_t cnt = n - i0 < 256 ? (uint32_t) (n - i0) : 256u;
			for (uint32_t j = lane; j < cnt; j += 32)
				dst[i0 + j] = stage[j];
			__syncwarp();
			carry = syn_mod((int64_t) carry + total);
		}
	}
}

#endif  // __CUDACC__

// Host twin of the kernel: same arithmetic, sample by sample.  The inverse-CDF
// search is short-cut by a 64 Ki-entry table on the top 16 bits of the draw
// (exact: buckets that straddle a threshold fall back to the search), and the
// reads are split over the host's hardware threads.
inline void synth_host_range(int16_t *samples, const uint64_t *offsets, uint64_t r0, uint64_t r1,
			     const int16_t *x0, const uint8_t *shift, const uint32_t *cdf,
			     const uint16_t *lut, uint32_t p_ex_u32, uint64_t seed, uint64_t first_read)
{
	for (uint64_t r = r0; r < r1; r++) {
		const uint64_t off = offsets[r], n = offsets[r + 1] - off;
		if (!n)
			continue;
		const uint64_t key = syn_key(seed, first_read + r);
		const uint32_t sh = shift ? shift[r] : 0;
		int32_t m = (int32_t) syn_mod((int64_t) x0[r]);
		samples[off] = (int16_t) (uint16_t) (syn_fold((uint32_t) m) << sh);
		for (uint64_t i = 1; i < n; i++) {
			const uint64_t h = syn_mix(key + i * SYN_GOLD);
			uint32_t z = lut[h >> 48];
			if ((uint32_t) h < p_ex_u32 || z > 255)
				z = syn_zd(key, i, cdf, p_ex_u32);
			m += syn_unzigzag(z);	// |delta| < 4094: one conditional wrap each way
			m = m < 0 ? m + (int32_t) SYN_PERIOD : m;
			m = m >= (int32_t) SYN_PERIOD ? m - (int32_t) SYN_PERIOD : m;
			samples[off + i] = (int16_t) (uint16_t) (syn_fold((uint32_t) m) << sh);
		}
	}
}

inline void synth_host(int16_t *samples, const uint64_t *offsets, uint64_t nreads, const int16_t *x0,
		       const uint8_t *shift, const uint32_t *cdf, uint32_t p_ex_u32, uint64_t seed,
		       uint64_t first_read)
{
	std::vector<uint16_t> lut(65536);
	for (uint32_t b = 0; b < 65536; b++) {
		// symbols of the smallest and largest draw of bucket b
		uint32_t lo_u = b << 16, hi_u = lo_u | 0xffffu, k0 = 0, k1 = 0;
		while (k0 < 255 && lo_u > cdf[k0]) k0++;
		k1 = k0;
		while (k1 < 255 && hi_u > cdf[k1]) k1++;
		lut[b] = k0 == k1 ? (uint16_t) k0 : (uint16_t) 0xffff;
	}
	unsigned nt = std::thread::hardware_concurrency();
	if (nt < 1) nt = 1;
	if (nt > 64) nt = 64;
	if (nreads < 4 * nt) nt = 1;
	const uint64_t total = offsets[nreads] - offsets[0];
	std::vector<std::thread> th;
	uint64_t r0 = 0;
	for (unsigned t = 0; t < nt; t++) {	// sample-balanced contiguous ranges
		uint64_t r1 = r0;
		const uint64_t want = offsets[0] + total / nt * (t + 1);
		if (t == nt - 1)
			r1 = nreads;
		else
			while (r1 < nreads && offsets[r1 + 1] <= want) r1++;
		if (nt == 1)
			synth_host_range(samples, offsets, r0, r1, x0, shift, cdf, lut.data(), p_ex_u32, seed, first_read);
		else
			th.emplace_back(synth_host_range, samples, offsets, r0, r1, x0, shift, cdf, lut.data(),
					p_ex_u32, seed, first_read);
		r0 = r1;
	}
	for (auto &x : th)
		x.join();
}

// 256 histogram counts -> cumulative thresholds on the u32 range
inline void synth_cdf(const uint64_t *hist256, uint32_t *cdf)
{
	unsigned __int128 total = 0, cum = 0;
	for (int k = 0; k < 256; k++)
		total += hist256[k];
	for (int k = 0; k < 256; k++) {
		cum += hist256[k];
		unsigned __int128 t = total ? (cum << 32) / total : ((unsigned __int128) (k + 1) << 24);
		cdf[k] = t ? (uint32_t) (t - 1) : 0u;
	}
	cdf[255] = 0xffffffffu;
}

}  // namespace pg
