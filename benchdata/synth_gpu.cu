// synth_gpu.cu -- C ABI of the device generator (benchdata/libsynth_gpu.so).  Bench / test
// support only: the product library (honours_b200/libpress_gpu.so) does not contain it.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include "synth_core.h"

using namespace pg;

namespace {
unsigned long long *g_ticket = nullptr;
int g_sms = 0, g_per_sm = 0;
}

extern "C" int synth_gpu_device(int16_t *d_samples, const uint64_t *d_offsets, uint64_t nreads,
				const int16_t *d_x0, const uint8_t *d_shift, const uint64_t *hist256,
				uint32_t p_ex_u32, uint64_t seed, uint64_t first_read, void *stream)
{
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream
	if (!g_ticket) {
		int dev = 0;
		cudaDeviceProp prop;
		if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess ||
		    cudaMalloc((void **) &g_ticket, 256) != cudaSuccess ||
		    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_per_sm, synth_kernel, 256, 0) != cudaSuccess) {
			fprintf(stderr, "synth_gpu: no usable CUDA device (%s)\n", cudaGetErrorString(cudaGetLastError()));
			return -1;
		}
		g_sms = prop.multiProcessorCount;
	}
	SynthArgs a;
	a.samples = d_samples;
	a.offsets = d_offsets;
	a.nreads = nreads;
	a.x0 = d_x0;
	a.shift = d_shift;
	synth_cdf(hist256, a.cdf);
	a.p_ex_u32 = p_ex_u32;
	a.seed = seed;
	a.first_read = first_read;
	a.ticket = g_ticket;
	if (cudaMemsetAsync(g_ticket, 0, 8, st) != cudaSuccess)
		return -1;
	synth_kernel<<<(g_per_sm > 0 ? g_per_sm : 1) * g_sms, 256, 0, st>>>(a);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ---------------------------------------------------------------------------
// Host link probe (bench support): what plain pinned cudaMemcpyAsync does on this box, one
// copy per `chunk_bytes`, so that the end-to-end numbers of bench.py have a measured ceiling.
// out[0] = H2D alone, out[1] = D2H alone, out[2] / out[3] = H2D / D2H while both run (GB/s).
extern "C" int bench_link_probe(size_t total_bytes, size_t chunk_bytes, int reps, double out[4])
{
	uint8_t *h_a = nullptr, *h_b = nullptr, *d_a = nullptr, *d_b = nullptr;
	cudaStream_t s1 = nullptr, s2 = nullptr;
	cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr};
	int rc = -1;
	auto up = [&]() {
		for (size_t o = 0; o < total_bytes; o += chunk_bytes) {
			const size_t n = total_bytes - o < chunk_bytes ? total_bytes - o : chunk_bytes;
			cudaMemcpyAsync(d_a + o, h_a + o, n, cudaMemcpyHostToDevice, s1);
		}
	};
	auto down = [&]() {
		for (size_t o = 0; o < total_bytes; o += chunk_bytes) {
			const size_t n = total_bytes - o < chunk_bytes ? total_bytes - o : chunk_bytes;
			cudaMemcpyAsync(h_b + o, d_b + o, n, cudaMemcpyDeviceToHost, s2);
		}
	};
	if (cudaHostAlloc((void **) &h_a, total_bytes, cudaHostAllocDefault) != cudaSuccess ||
	    cudaHostAlloc((void **) &h_b, total_bytes, cudaHostAllocDefault) != cudaSuccess ||
	    cudaMalloc((void **) &d_a, total_bytes) != cudaSuccess || cudaMalloc((void **) &d_b, total_bytes) != cudaSuccess ||
	    cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking) != cudaSuccess)
		goto done;
	for (int i = 0; i < 4; i++)
		if (cudaEventCreate(&e[i]) != cudaSuccess)
			goto done;
	memset(h_a, 1, total_bytes);
	memset(h_b, 2, total_bytes);
	cudaMemset(d_b, 3, total_bytes);
	up();
	down();
	cudaDeviceSynchronize();
	{
		float ms = 0;
		const double gb = (double) total_bytes * reps / 1e9;
		cudaEventRecord(e[0], s1);
		for (int r = 0; r < reps; r++)
			up();
		cudaEventRecord(e[1], s1);
		cudaEventSynchronize(e[1]);
		cudaEventElapsedTime(&ms, e[0], e[1]);
		out[0] = gb / (ms * 1e-3);
		cudaEventRecord(e[2], s2);
		for (int r = 0; r < reps; r++)
			down();
		cudaEventRecord(e[3], s2);
		cudaEventSynchronize(e[3]);
		cudaEventElapsedTime(&ms, e[2], e[3]);
		out[1] = gb / (ms * 1e-3);
		cudaEventRecord(e[0], s1);
		cudaEventRecord(e[2], s2);
		for (int r = 0; r < reps; r++) {
			up();
			down();
		}
		cudaEventRecord(e[1], s1);
		cudaEventRecord(e[3], s2);
		cudaEventSynchronize(e[1]);
		cudaEventSynchronize(e[3]);
		cudaEventElapsedTime(&ms, e[0], e[1]);
		out[2] = gb / (ms * 1e-3);
		cudaEventElapsedTime(&ms, e[2], e[3]);
		out[3] = gb / (ms * 1e-3);
	}
	rc = cudaGetLastError() == cudaSuccess ? 0 : -1;
done:
	for (int i = 0; i < 4; i++)
		if (e[i])
			cudaEventDestroy(e[i]);
	if (s1) cudaStreamDestroy(s1);
	if (s2) cudaStreamDestroy(s2);
	if (h_a) cudaFreeHost(h_a);
	if (h_b) cudaFreeHost(h_b);
	if (d_a) cudaFree(d_a);
	if (d_b) cudaFree(d_b);
	return rc;
}
