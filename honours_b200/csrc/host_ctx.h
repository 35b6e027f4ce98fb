// host_ctx.h -- host-side context shared by the launch code (no device code).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace pg {

struct ScratchLayout {
	size_t hdr, status, status2, status_blk, status2_blk, zero_bytes, tile_start, kx, tile_read, extra, total;
	// exception-coded methods (exfam.cuh): per-tile / per-read tables, Huffman payload, exception pool
	size_t x_t[6], x_r[4], x_r8[1], x_paybuf, x_pool;
	uint64_t x_pool_cap;	// entries
};

// pool_frac: exception pool entries per sample (< 0: the library default)
ScratchLayout scratch_layout(int method, uint64_t nreads, uint64_t nsamples, double pool_frac = -1);
// the layout of a scratch buffer of `bytes` bytes: the pool takes what is left
ScratchLayout scratch_layout_for(int method, uint64_t nreads, uint64_t nsamples, size_t bytes);
uint64_t max_tiles(uint64_t nreads, uint64_t nsamples);

// A device buffer that only ever grows (high-water mark), as the per-read
// drop-ins are called once per read with varying sizes (press/test.h:27-32).
struct GrowBuf {
	void *p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap)
			return 0;
		if (p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
		size_t want = bytes + bytes / 4 + 4096;
		if (cudaMalloc(&p, want) != cudaSuccess)
			return -1;
		cap = want;
		return 0;
	}
	void release()
	{
		if (p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
	}
};

struct HostCtx {
	bool ready = false;
	int device = 0;
	int sms = 0;
	cudaStream_t stream = nullptr;
	GrowBuf samples, offsets, out, out_offsets, scratch, misc;
	void release()
	{
		samples.release();
		offsets.release();
		out.release();
		out_offsets.release();
		scratch.release();
		misc.release();
	}
};

}  // namespace pg
