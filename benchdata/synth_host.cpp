// synth_host.cpp -- C ABI of the host twin of the generator (benchdata/libsynth_host.so,
// plain g++: no CUDA).  Bench / test support only.
#include "synth_core.h"

extern "C" int synth_host_fill(int16_t *h_samples, const uint64_t *h_offsets, uint64_t nreads,
			       const int16_t *h_x0, const uint8_t *h_shift, const uint64_t *hist256,
			       uint32_t p_ex_u32, uint64_t seed, uint64_t first_read)
{
	uint32_t cdf[256];
	pg::synth_cdf(hist256, cdf);
	pg::synth_host(h_samples, h_offsets, nreads, h_x0, h_shift, cdf, p_ex_u32, seed, first_read);
	return 0;
}
