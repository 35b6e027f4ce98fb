// synth_gpu.cu -- C ABI of the device generator (benchdata/libsynth_gpu.so).  Bench / test
// support only: the product library (honours_b200/libpress_gpu.so) does not contain it.
#include <cuda_runtime.h>
#include <stdio.h>

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
