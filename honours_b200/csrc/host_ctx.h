// host_ctx.h -- host-side context shared by the launch code (no device code).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace pg {

struct ScratchLayout {
	size_t hdr, status, zero_bytes, tile_start, kx, tile_read, total;
	// exception-coded methods (exfam.cuh): per-tile / per-read tables, Huffman payload, exception pool
	size_t x_t[6], x_r[4], x_r8[2], x_paybuf, x_pool, x_pool2;
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

// One stage of the host-buffer pipeline (host_pipe.inc): its own stream, device
// buffers and pinned staging for the chunk's offset tables.
struct PipeSlot {
	cudaStream_t st = nullptr;
	cudaEvent_t ev = nullptr;
	GrowBuf samples, offsets, in_offsets, out, out_offsets, scratch;
	uint64_t *h_tab = nullptr;	// pinned: chunk-relative offsets (two tables) + the scratch header
	size_t h_tab_cap = 0;
	int reserve_tab(size_t words)
	{
		if (words <= h_tab_cap)
			return 0;
		if (h_tab)
			cudaFreeHost(h_tab);
		h_tab = nullptr;
		h_tab_cap = 0;
		if (cudaHostAlloc((void **) &h_tab, (words + 1024) * 8, cudaHostAllocDefault) != cudaSuccess)
			return -1;
		h_tab_cap = words + 1024;
		return 0;
	}
	void release()
	{
		samples.release(); offsets.release(); in_offsets.release(); out.release();
		out_offsets.release(); scratch.release();
		if (h_tab)
			cudaFreeHost(h_tab);
		h_tab = nullptr;
		h_tab_cap = 0;
		if (ev)
			cudaEventDestroy(ev);
		if (st)
			cudaStreamDestroy(st);
		ev = nullptr;
		st = nullptr;
	}
};

constexpr int PIPE_SLOTS = 6;	// chunks in flight per direction: deep enough that a copy queued behind the OTHER
				// pipeline's traffic (duplex use) does not stall this one's uploads

struct HostCtx {
	bool ready = false;
	int device = 0;
	int sms = 0;
	cudaStream_t stream = nullptr;
	GrowBuf samples, offsets, out, out_offsets, scratch, misc;
	PipeSlot slot[2][PIPE_SLOTS];	// [0]: press_host pipeline, [1]: depress_host pipeline -- one thread may
					// compress a batch while another decompresses an earlier one (full duplex link)
	void release()
	{
		for (int i = 0; i < PIPE_SLOTS; i++) {
			slot[0][i].release();
			slot[1][i].release();
		}
		samples.release();
		offsets.release();
		out.release();
		out_offsets.release();
		scratch.release();
		misc.release();
	}
};

}  // namespace pg
