// press_gpu.cu -- C ABI of libpress_gpu.so (see include/press_gpu.h).
//
// Host side only: context, scratch layout, launch sequences, the per-read
// drop-ins (press.h signatures) built on the batch path, and the host-buffer
// batch path.  All device code lives in the .cuh files next to this one.
#include "../../include/press_gpu.h"

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "plan.cuh"
#include "svb_zd.cuh"
#include "svb_enc.cuh"
#include "exfam.cuh"
#include "exfam_dec.cuh"
#include "huff_fused.cuh"
#include "misc_codecs.cuh"
#include "svb_dec.cuh"
#include "host_ctx.h"

using namespace pg;

// ---------------------------------------------------------------------------
// context

namespace {

char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};
std::once_flag g_once;
HostCtx g_ctx;

void set_err(const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
	fprintf(stderr, "press_gpu: %s\n", g_err);
}

#define CK(call)                                                                      \
	do {                                                                          \
		cudaError_t e_ = (call);                                              \
		if (e_ != cudaSuccess) {                                              \
			set_err("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
				__FILE__, __LINE__);                                  \
			return -1;                                                    \
		}                                                                     \
	} while (0)

// event brackets around the main kernels (press_gpu_profile_*)
struct EvPair { cudaEvent_t a, b; int kind; };
bool g_prof = false;
std::vector<EvPair> g_ev_pending, g_ev_free;

EvPair *prof_begin(int kind, cudaStream_t st)
{
	if (!g_prof)
		return nullptr;
	EvPair p;
	if (!g_ev_free.empty()) {
		p = g_ev_free.back();
		g_ev_free.pop_back();
	} else if (cudaEventCreate(&p.a) != cudaSuccess || cudaEventCreate(&p.b) != cudaSuccess) {
		return nullptr;
	}
	p.kind = kind;
	cudaEventRecord(p.a, st);
	g_ev_pending.push_back(p);
	return &g_ev_pending.back();
}
void prof_end(EvPair *p, cudaStream_t st)
{
	if (p)
		cudaEventRecord(p->b, st);
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int ensure_init()
{
	if (g_ctx.ready) {
		// a host thread other than the one that called press_gpu_init starts on device 0: the
		// library's streams and buffers live on g_ctx.device
		int d = -1;
		if (cudaGetDevice(&d) != cudaSuccess || d != g_ctx.device)
			CK(cudaSetDevice(g_ctx.device));
		return 0;
	}
	return press_gpu_init(-1);
}

}  // namespace

// scratch layout -------------------------------------------------------------
namespace pg {

// upper bound on the tiles of a batch: the svb encoder's 8192-sample tiles (the
// exception-coded methods size their own 4096-sample tile tables below)
uint64_t max_tiles(uint64_t nreads, uint64_t nsamples)
{
	return nsamples / EN_TS + 2 * nreads + 2;
}

static double default_pool_frac();

ScratchLayout scratch_layout(int method, uint64_t nreads, uint64_t nsamples, double pool_frac)
{
	ScratchLayout l = {};
	const uint64_t mt = max_tiles(nreads, nsamples);
	size_t o = 0;
	l.hdr = o;		o += sizeof(Hdr);
	if (method == PRESS_GPU_SVB0124_ZD || method == PRESS_GPU_UINT_ZD) {	// misc_codecs.cuh: per-read size and width
		l.zero_bytes = o;
		l.x_r8[0] = o;		o += align_up((nreads + 1) * 8, 256);
		l.x_r[0] = o;		o += align_up((nreads + 1) * 4, 256);
		l.total = o;
		return l;
	}
	if (method >= PRESS_GPU_VBE21_ZD) {
		const uint64_t xt = nsamples / 4096 + 2 * nreads + 2;	// tiles of 4096 samples
		l.zero_bytes = o;	// only the header is cleared
		l.tile_start = o;	o += align_up((nreads + 1) * 4, 256);
		l.kx = o;		o += align_up((nreads + 1) * 8, 256);
		l.tile_read = o;	o += align_up(xt * 4, 256);
		for (int i = 0; i < 6; i++) {	// [0..2]: 8 records (one per warp) per tile
			l.x_t[i] = o;
			if (i < 3)
				o += align_up(xt * 8 * 4, 256);
		}
		for (int i = 0; i < 4; i++) {
			l.x_r[i] = o;
			o += align_up((nreads + 1) * 4, 256);
		}
		l.x_r8[0] = o;		o += align_up((nreads + 1) * 8, 256);
		l.x_r8[1] = o;		o += align_up((nreads + 1) * 8, 256);
		// 1 B/sample: Huffman methods, decode without a side-car: the payload between bit decode and
		// sample decode; byte-payload methods, encode: the low bytes sweep 1 leaves for sweep 2
		l.x_paybuf = o;
		o += align_up(nsamples + 128, 256);
		l.x_pool = o;
		if (pool_frac < 0)
			pool_frac = default_pool_frac();
		// (ex-zd reads with q > 0 are captured twice, once per pass of ex_count_kernel, and the
		//  first capture's space is not given back: its pool may need 2 entries per sample)
		const uint64_t cap_max = (method == PRESS_GPU_EX_ZD ? 2 : 1) * (nsamples + 1);
		uint64_t cap = (uint64_t) ((double) nsamples * pool_frac) + 4096;
		if (cap > cap_max)
			cap = cap_max;
		l.x_pool_cap = cap;
		l.x_pool2 = o + align_up(cap * 8, 256);	// two pools: as captured, and gathered per read
		o += 2 * align_up(cap * 8, 256);
		l.total = o;
		return l;
	}
	l.status = o;		o += align_up(mt * 8, 256);
	l.zero_bytes = o;	// [0, zero_bytes) is cleared before every batch call
	l.tile_start = o;	o += align_up((nreads + 1) * 4, 256);
	l.kx = o;		o += align_up((nreads + 1) * 8, 256);
	l.tile_read = o;	o += align_up(mt * 4, 256);
	l.total = o;
	return l;
}

ScratchLayout scratch_layout_for(int method, uint64_t nreads, uint64_t nsamples, size_t bytes)
{
	ScratchLayout l = scratch_layout(method, nreads, nsamples, 0.0);
	if (method >= PRESS_GPU_VBE21_ZD && method <= PRESS_GPU_SHUFF_VBBE21_ZD && bytes > l.total) {
		uint64_t cap = l.x_pool_cap + (bytes - l.total) / 16 / 256 * 256;	// the two pools share what is left
		const uint64_t cap_max = (method == PRESS_GPU_EX_ZD ? 2 : 1) * (nsamples + 1);
		if (cap > cap_max)
			cap = cap_max;
		l.x_pool_cap = cap;
		l.x_pool2 = l.x_pool + align_up(cap * 8, 256);
		l.total = l.x_pool + 2 * align_up(cap * 8, 256);
	}
	return l;
}

}  // namespace pg

extern "C" {

int press_gpu_init(int device)
{
	if (g_ctx.ready)
		return 0;
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		set_err("no CUDA device (%s); libpress_gpu has no CPU fallback",
			e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
		return -1;
	}
	if (device >= 0)
		CK(cudaSetDevice(device));
	CK(cudaGetDevice(&g_ctx.device));
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, g_ctx.device));
	if (prop.major < 10) {
		set_err("device %d is sm_%d%d; this library is built for sm_100a only", g_ctx.device,
			prop.major, prop.minor);
		return -1;
	}
	g_ctx.sms = prop.multiProcessorCount;
	CK(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
	g_ctx.ready = true;
	return 0;
}

void press_gpu_shutdown(void)
{
	if (!g_ctx.ready)
		return;
	g_ctx.release();
	cudaStreamDestroy(g_ctx.stream);
	g_ctx.ready = false;
}

const char *press_gpu_last_error(void) { return g_err; }
const char *press_gpu_version(void) { return "press_gpu 0.1 sm_100a"; }
uint64_t press_gpu_launch_count(void) { return g_launches.load(); }
int press_gpu_method_implemented(int method)
{
	return method >= PRESS_GPU_SVB16_ZD && method < PRESS_GPU_NMETHODS;
}

uint64_t press_gpu_batch_bound(int method, uint64_t nreads, uint64_t nsamples)
{
	switch (method) {
	case PRESS_GPU_SVB16_ZD: return 2 * nsamples + nsamples / 8 + nreads + 16;
	case PRESS_GPU_VBZ_SVB16_ZD: return 2 * nsamples + nsamples / 8 + 5 * nreads + 16;
	case PRESS_GPU_SVB_ZD: return 2 * nsamples + nsamples / 4 + nreads + 16;
	case PRESS_GPU_VBZ_SVB_ZD: return 2 * nsamples + nsamples / 4 + 5 * nreads + 16;
	/* every sample an exception: 6 B (vbe21), <= 1.25 + 2.125 B (vbsse21) per sample */
	case PRESS_GPU_VBE21_ZD: return 6 * nsamples + 16;
	case PRESS_GPU_SHUFF_VBE21_ZD: return 6 * nsamples + 4 * nreads + 16;
	/* lists as vbsse21, 3-byte codes at most for the rest */
	case PRESS_GPU_SHUFF_VBSSE21_ZD: return 4 * nsamples + 20 * nreads + 16;
	/* bit-packed lists: <= 4 + 2 bytes per exception + two width bytes and two lengths */
	case PRESS_GPU_VBBE21_ZD: return 6 * nsamples + 16 * nreads + 16;
	case PRESS_GPU_SHUFF_VBBE21_ZD: return 6 * nsamples + 20 * nreads + 16;
	case PRESS_GPU_VBSSE21_ZD: return 4 * nsamples + 16 * nreads + 16;
	case PRESS_GPU_EX_ZD: return 2 * nsamples + 1024 * nreads + 16;	/* ex_zd.c:411 cap */
	case PRESS_GPU_SVB0124_ZD: return 2 * nsamples + nsamples / 4 + nreads + 16;
	case PRESS_GPU_UINT_ZD: return 2 * nsamples + 3 * nreads + 16;
	default: break;
	}
	return 0;
}

size_t press_gpu_batch_scratch_bytes(int method, uint64_t nreads, uint64_t nsamples)
{
	return scratch_layout(method, nreads, nsamples).total;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// launch sequences

namespace {

template <typename K>
int persistent_grid(K kernel, int *grid)
{
	int per_sm = 0;
	CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, NT, 0));
	if (per_sm < 1)
		per_sm = 1;
	*grid = per_sm * g_ctx.sms;
	return 0;
}

int run_plan(const ScratchLayout &l, uint8_t *scr, const uint64_t *d_offsets, uint64_t nreads,
	     int tiling, uint32_t tile, uint32_t frame, uint32_t keydiv, uint32_t fixed, cudaStream_t st)
{
	CK(cudaMemsetAsync(scr, 0, l.zero_bytes, st));
	PlanArgs pa;
	pa.offsets = d_offsets;
	pa.nreads = nreads;
	pa.tiling = tiling;
	pa.tile = tile;
	pa.frame = frame;
	pa.keydiv = keydiv;
	pa.fixed = fixed;
	pa.tile_start = reinterpret_cast<uint32_t *>(scr + l.tile_start);
	pa.kx = reinterpret_cast<uint64_t *>(scr + l.kx);
	pa.hdr = reinterpret_cast<Hdr *>(scr + l.hdr);
	plan_scan_kernel<<<1, 1024, 0, st>>>(pa);
	g_launches++;
	if (nreads) {
		unsigned blocks = (unsigned) ((nreads * 32 + 255) / 256);
		plan_fill_kernel<<<blocks, 256, 0, st>>>(pa.tile_start, nreads,
							 reinterpret_cast<uint32_t *>(scr + l.tile_read));
		g_launches++;
	}
	CK(cudaGetLastError());
	return 0;
}

template <int KEYBITS, bool FRAME>
int launch_svb_encode(const SvbEncArgs &a, cudaStream_t st)
{
	auto kernel = svb_enc_kernel<KEYBITS, FRAME>;
	const size_t smem = sizeof(EnSmem);
	static int grid = 0;	// per instantiation
	if (!grid) {
		CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		int per_sm = 0;
		CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, EN_THREADS, smem));
		if (per_sm < 1) {
			set_err("svb_enc_kernel does not fit on an SM (%zu B shared)", smem);
			return -1;
		}
		grid = per_sm * g_ctx.sms;
		if (grid > EN_MAX_GRID)
			grid = EN_MAX_GRID;	// the prefix window is sized for one CTA per SM
	}
	EvPair *ev = prof_begin(0, st);
	void *params[] = {const_cast<SvbEncArgs *>(&a)};
	CK(cudaLaunchCooperativeKernel((const void *) kernel, dim3(grid), dim3(EN_THREADS), params, smem, st));
	prof_end(ev, st);
	g_launches++;
	CK(cudaGetLastError());
	return 0;
}

template <int KEYBITS, bool FRAME>
int launch_svb_decode(const SvbDec2Args &a, cudaStream_t st)
{
	const uint64_t cap = (uint64_t) g_ctx.sms * PG_DEC_CTAS;	// (resident CTAs; reads go by ticket)
	const unsigned grid = (unsigned) (a.nreads < cap ? (a.nreads ? a.nreads : 1) : cap);
	EvPair *ev = prof_begin(1, st);
	svb_dec2_kernel<KEYBITS, FRAME><<<grid, NT, 0, st>>>(a);
	prof_end(ev, st);
	g_launches++;
	CK(cudaGetLastError());
	return 0;
}

bool is_svb(int m) { return m >= PRESS_GPU_SVB16_ZD && m <= PRESS_GPU_VBZ_SVB_ZD; }

}  // namespace

// side-car of the call in progress (press_gpu_batch_*_sidecar; null: none).  thread_local: the
// host-buffer pipelines of two threads never see each other's.
thread_local uint8_t *g_sidecar = nullptr;

#include "ex_host.inc"

namespace pg {
static double default_pool_frac() { return g_pool_frac; }
}

extern "C" {

int press_gpu_batch_press(int method, const int16_t *d_samples, const uint64_t *d_offsets,
			  uint64_t nreads, uint64_t nsamples, uint8_t *d_out, uint64_t out_capacity,
			  uint64_t *d_out_offsets, void *d_scratch, size_t scratch_bytes, void *stream)
{
	if (ensure_init())
		return -1;
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream
	const ScratchLayout l = scratch_layout_for(method, nreads, nsamples, scratch_bytes);
	if (scratch_bytes < l.total) {
		set_err("scratch too small: %zu < %zu", scratch_bytes, l.total);
		return -1;
	}
	if ((reinterpret_cast<uintptr_t>(d_samples) & 15) || (reinterpret_cast<uintptr_t>(d_scratch) & 255)) {
		set_err("d_samples must be 16-byte and d_scratch 256-byte aligned");
		return -1;
	}
	if (l.x_pool_cap >= (1ull << 32)) {
		set_err("batch too large for the exception pool (positions are 32-bit): split it");
		return -1;
	}
	if (nsamples / 4096 + 2 * nreads + 2 >= (1ull << 32)) {
		set_err("batch too large: split it (tile ids are 32-bit)");
		return -1;
	}
	uint8_t *scr = static_cast<uint8_t *>(d_scratch);
	CK(cudaMemsetAsync(d_out_offsets, 0xff, (nreads + 1) * sizeof(uint64_t), st));

	if (is_svb(method)) {
		const bool frame = method == PRESS_GPU_VBZ_SVB16_ZD || method == PRESS_GPU_VBZ_SVB_ZD;
		const bool k2 = method == PRESS_GPU_SVB_ZD || method == PRESS_GPU_VBZ_SVB_ZD;
		if (run_plan(l, scr, d_offsets, nreads, TILING_ALIGNED, EN_TS, frame ? 4 : 0, k2 ? 4 : 8, 0, st))
			return -1;
		SvbEncArgs a;
		a.samples = reinterpret_cast<const uint4 *>(d_samples);
		a.nsamples = nsamples;
		a.offsets = d_offsets;
		a.nreads = nreads;
		a.out = d_out;
		a.out_cap = out_capacity;
		a.out_offsets = d_out_offsets;
		a.hdr = reinterpret_cast<Hdr *>(scr + l.hdr);
		a.tile_start = reinterpret_cast<const uint32_t *>(scr + l.tile_start);
		a.tile_read = reinterpret_cast<const uint32_t *>(scr + l.tile_read);
		a.kx = reinterpret_cast<const uint64_t *>(scr + l.kx);
		a.status = reinterpret_cast<uint64_t *>(scr + l.status);
		if (!k2 && !frame) return launch_svb_encode<1, false>(a, st);
		if (!k2 && frame) return launch_svb_encode<1, true>(a, st);
		if (k2 && !frame) return launch_svb_encode<2, false>(a, st);
		return launch_svb_encode<2, true>(a, st);
	}
	if (method == PRESS_GPU_SVB0124_ZD || method == PRESS_GPU_UINT_ZD)
		return mc_batch_press(method, d_samples, d_offsets, nreads, d_out, out_capacity, d_out_offsets, l, scr, st);
	if (method >= PRESS_GPU_VBE21_ZD && method < PRESS_GPU_NMETHODS)
		return ex_batch_press(method, d_samples, d_offsets, nreads, nsamples, d_out, out_capacity,
				      d_out_offsets, l, scr, st);
	set_err("batch_press: method %d not implemented", method);
	return -1;
}

int press_gpu_batch_depress(int method, const uint8_t *d_in, const uint64_t *d_in_offsets,
			    uint64_t in_bytes, const uint64_t *d_offsets, uint64_t nreads,
			    uint64_t nsamples, int16_t *d_samples_out, void *d_scratch,
			    size_t scratch_bytes, void *stream)
{
	if (ensure_init())
		return -1;
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream
	const ScratchLayout l = scratch_layout_for(method, nreads, nsamples, scratch_bytes);
	if (scratch_bytes < l.total) {
		set_err("scratch too small: %zu < %zu", scratch_bytes, l.total);
		return -1;
	}
	if ((reinterpret_cast<uintptr_t>(d_samples_out) & 15) || (reinterpret_cast<uintptr_t>(d_in) & 15) ||
	    (reinterpret_cast<uintptr_t>(d_scratch) & 255)) {
		set_err("d_samples_out and d_in must be 16-byte and d_scratch 256-byte aligned");
		return -1;
	}
	if (l.x_pool_cap >= (1ull << 32)) {
		set_err("batch too large for the exception pool (positions are 32-bit): split it");
		return -1;
	}
	if (nsamples / 4096 + 2 * nreads + 2 >= (1ull << 32)) {
		set_err("batch too large: split it (tile ids are 32-bit)");
		return -1;
	}
	uint8_t *scr = static_cast<uint8_t *>(d_scratch);

	if (is_svb(method)) {
		const bool frame = method == PRESS_GPU_VBZ_SVB16_ZD || method == PRESS_GPU_VBZ_SVB_ZD;
		const bool k2 = method == PRESS_GPU_SVB_ZD || method == PRESS_GPU_VBZ_SVB_ZD;
		CK(cudaMemsetAsync(scr, 0, sizeof(Hdr), st));
		SvbDec2Args a;
		a.in = d_in;
		a.in_bytes = in_bytes;
		a.in_offsets = d_in_offsets;
		a.offsets = d_offsets;
		a.nreads = nreads;
		a.samples_out = d_samples_out;
		a.hdr = reinterpret_cast<Hdr *>(scr + l.hdr);
		if (!k2 && !frame) return launch_svb_decode<1, false>(a, st);
		if (!k2 && frame) return launch_svb_decode<1, true>(a, st);
		if (k2 && !frame) return launch_svb_decode<2, false>(a, st);
		return launch_svb_decode<2, true>(a, st);
	}
	if (method == PRESS_GPU_SVB0124_ZD || method == PRESS_GPU_UINT_ZD)
		return mc_batch_depress(method, d_in, d_in_offsets, in_bytes, d_offsets, nreads, d_samples_out, l, scr, st);
	if (method >= PRESS_GPU_VBE21_ZD && method < PRESS_GPU_NMETHODS)
		return ex_batch_depress(method, d_in, d_in_offsets, in_bytes, d_offsets, nreads, nsamples,
					d_samples_out, l, scr, st);
	set_err("batch_depress: method %d not implemented", method);
	return -1;
}

size_t press_gpu_sidecar_bytes(int method, uint64_t nreads, uint64_t nsamples)
{
	return ex_sidecar_bytes(method, nreads, nsamples);
}

int press_gpu_batch_press_sidecar(int method, const int16_t *d_samples, const uint64_t *d_offsets,
				  uint64_t nreads, uint64_t nsamples, uint8_t *d_out, uint64_t out_capacity,
				  uint64_t *d_out_offsets, void *d_scratch, size_t scratch_bytes,
				  uint8_t *d_sidecar, size_t sidecar_bytes, void *stream)
{
	if (d_sidecar && (sidecar_bytes < ex_sidecar_bytes(method, nreads, nsamples) ||
			  (reinterpret_cast<uintptr_t>(d_sidecar) & 15))) {
		set_err("side-car buffer too small or not 16-byte aligned (press_gpu_sidecar_bytes)");
		return -1;
	}
	g_sidecar = ex_sidecar_bytes(method, nreads, nsamples) ? d_sidecar : nullptr;
	const int rc = press_gpu_batch_press(method, d_samples, d_offsets, nreads, nsamples, d_out, out_capacity,
					     d_out_offsets, d_scratch, scratch_bytes, stream);
	g_sidecar = nullptr;
	return rc;
}

int press_gpu_batch_depress_sidecar(int method, const uint8_t *d_in, const uint64_t *d_in_offsets,
				    uint64_t in_bytes, const uint64_t *d_offsets, uint64_t nreads,
				    uint64_t nsamples, int16_t *d_samples_out, void *d_scratch,
				    size_t scratch_bytes, const uint8_t *d_sidecar, size_t sidecar_bytes,
				    void *stream)
{
	if (d_sidecar && (sidecar_bytes < ex_sidecar_bytes(method, nreads, nsamples) ||
			  (reinterpret_cast<uintptr_t>(d_sidecar) & 15))) {
		set_err("side-car buffer too small or not 16-byte aligned (press_gpu_sidecar_bytes)");
		return -1;
	}
	g_sidecar = ex_sidecar_bytes(method, nreads, nsamples) ? const_cast<uint8_t *>(d_sidecar) : nullptr;
	const int rc = press_gpu_batch_depress(method, d_in, d_in_offsets, in_bytes, d_offsets, nreads, nsamples,
					       d_samples_out, d_scratch, scratch_bytes, stream);
	g_sidecar = nullptr;
	return rc;
}

int press_gpu_batch_status(void *d_scratch, void *stream)
{
	if (ensure_init())
		return -1;
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream
	Hdr h;
	if (cudaMemcpyAsync(&h, d_scratch, sizeof h, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
	    cudaStreamSynchronize(st) != cudaSuccess) {
		set_err("batch_status: %s", cudaGetErrorString(cudaGetLastError()));
		return -1;
	}
	return (int) h.error;
}

#if defined(PG_ROLE_TIMING) || defined(PG_DEBUG_AUX)
/* debug builds only (tools/role_timing.py, tools/debug_huff.py): the scratch header's aux words */
int press_gpu_debug_aux(void *d_scratch, void *stream, uint64_t out[29])
{
	Hdr h;
	cudaStream_t st = (cudaStream_t) stream;
	if (cudaMemcpyAsync(&h, d_scratch, sizeof h, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
	    cudaStreamSynchronize(st) != cudaSuccess)
		return -1;
	memcpy(out, h.aux, sizeof h.aux);
	return 0;
}
#endif

void press_gpu_profile_enable(int on) { g_prof = on != 0; }

int press_gpu_profile_collect(double ms[2], uint64_t launches[2])
{
	ms[0] = ms[1] = 0;
	launches[0] = launches[1] = 0;
	for (EvPair &p : g_ev_pending) {
		float t = 0;
		CK(cudaEventSynchronize(p.b));
		CK(cudaEventElapsedTime(&t, p.a, p.b));
		ms[p.kind] += t;
		launches[p.kind]++;
		g_ev_free.push_back(p);
	}
	g_ev_pending.clear();
	return 0;
}

__global__ void global_offsets_kernel(const uint64_t *gathered, int nranks, uint64_t nreads,
				      uint64_t *global)
{
	const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const uint64_t total = (uint64_t) nranks * nreads;
	if (i > total)
		return;
	const int k = i == total ? nranks : (int) (i / nreads);
	uint64_t base = 0;
	for (int j = 0; j < k; j++)
		base += gathered[(uint64_t) j * (nreads + 1) + nreads];
	global[i] = i == total ? base : base + gathered[(uint64_t) k * (nreads + 1) + (i - (uint64_t) k * nreads)];
}

int press_gpu_global_offsets(const uint64_t *d_gathered, int nranks, uint64_t nreads,
			     uint64_t *d_global, void *stream)
{
	if (ensure_init())
		return -1;
	if (nranks < 1 || nreads == 0) {
		set_err("global_offsets: nranks >= 1 and nreads >= 1 required");
		return -1;
	}
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream
	const uint64_t total = (uint64_t) nranks * nreads + 1;
	global_offsets_kernel<<<(unsigned) ((total + 255) / 256), 256, 0, st>>>(d_gathered, nranks, nreads, d_global);
	g_launches++;
	CK(cudaGetLastError());
	return 0;
}

void *press_gpu_host_alloc(size_t bytes)
{
	void *p = nullptr;
	if (ensure_init() || cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
		set_err("host_alloc(%zu) failed", bytes);
		return nullptr;
	}
	return p;
}

void press_gpu_host_free(void *p)
{
	if (p)
		cudaFreeHost(p);
}

}  // extern "C"

#include "host_paths.inc"
#include "comm.inc"

// ---------------------------------------------------------------------------------------------
// svb16-zd compression of a batch with NO allocation anywhere in the call
// (no counterpart in the reference: added
// for callers that cannot call cudaMalloc -- CUDA graph capture, tight
// serving loops, harnesses that forbid it).
//
// press_gpu_batch_press takes its scratch (per-tile status words, the plan tables, the 256-byte
// header with the error word) from the caller.  Here the same kernels run on a file-scope device
// array of a fixed, modest size: every GPU that loads the library holds
// PRESS_GPU_STATIC_SCRATCH_BYTES of it until the process exits.  All calls share the array and it
// keeps the last call's contents: the part that must start at zero (header + status words) is
// cleared on the stream at the start of every call, and two calls that overlap on the GPU (two
// streams, or two host threads without ordering between their streams) RACE on it.
//
// g_static_scratch below is, in host code, a host-side stand-in for the symbol, not a device
// address: the address is looked up once per device with cudaGetSymbolAddress, with that device
// current, and that pointer is what the kernels get.

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[PRESS_GPU_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEV = 64;
std::mutex g_static_mutex;
uint8_t *g_static_ptr[STATIC_MAX_DEV];	// device address of g_static_scratch, per device
int g_static_grid[STATIC_MAX_DEV];		// resident CTAs of the encoder, per device

// once per device, with that device current: the array's device address, the encoder's shared-memory
// attribute and its cooperative grid
int static_setup(int dev)
{
	std::lock_guard<std::mutex> lock(g_static_mutex);
	if (g_static_ptr[dev])
		return 0;
	void *p = nullptr;
	cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
	if (e != cudaSuccess || !p) {
		set_err("static scratch: cudaGetSymbolAddress failed on device %d (%s)", dev, cudaGetErrorString(e));
		return -1;
	}
	if (reinterpret_cast<uintptr_t>(p) & 255) {
		set_err("static scratch: device address is not 256-byte aligned");
		return -1;
	}
	auto kernel = svb_enc_kernel<1, false>;
	const size_t smem = sizeof(EnSmem);
	CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	int per_sm = 0, sms = 0;
	CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, EN_THREADS, smem));
	CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	if (per_sm < 1 || sms < 1) {
		set_err("static scratch: svb_enc_kernel does not fit on an SM (%zu B shared)", smem);
		return -1;
	}
	int grid = per_sm * sms;
	if (grid > EN_MAX_GRID)
		grid = EN_MAX_GRID;
	g_static_grid[dev] = grid;
	g_static_ptr[dev] = static_cast<uint8_t *>(p);
	return 0;
}

}  // namespace

extern "C" int press_gpu_svb16_zd_press_static(const int16_t *d_samples, const uint64_t *d_offsets, uint64_t nreads,
						uint64_t nsamples, uint8_t *d_out, uint64_t out_capacity,
						uint64_t *d_out_offsets, uint32_t *d_status, void *stream)
{
	if (ensure_init())
		return -1;
	if (nreads > PRESS_GPU_STATIC_MAX_READS || nsamples > PRESS_GPU_STATIC_MAX_SAMPLES) {
		set_err("press_gpu_svb16_zd_press_static: %llu reads / %llu samples exceed the static scratch's limits "
			"(%llu reads, %llu samples): split the batch or use press_gpu_batch_press",
			(unsigned long long) nreads, (unsigned long long) nsamples,
			(unsigned long long) PRESS_GPU_STATIC_MAX_READS, (unsigned long long) PRESS_GPU_STATIC_MAX_SAMPLES);
		return -1;
	}
	const ScratchLayout l = scratch_layout(PRESS_GPU_SVB16_ZD, nreads, nsamples, 0.0);
	if (l.total > (size_t) PRESS_GPU_STATIC_SCRATCH_BYTES) {	// (cannot happen within the limits above)
		set_err("press_gpu_svb16_zd_press_static: batch needs %zu B of scratch, the static array holds %llu",
			l.total, (unsigned long long) PRESS_GPU_STATIC_SCRATCH_BYTES);
		return -1;
	}
	if (reinterpret_cast<uintptr_t>(d_samples) & 15) {
		set_err("d_samples must be 16-byte aligned");
		return -1;
	}
	int dev = -1;
	CK(cudaGetDevice(&dev));
	if (dev < 0 || dev >= STATIC_MAX_DEV) {
		set_err("static scratch: device %d out of range", dev);
		return -1;
	}
	if (!g_static_ptr[dev] && static_setup(dev))
		return -1;
	uint8_t *scr = g_static_ptr[dev];
	cudaStream_t st = (cudaStream_t) stream;	// NULL = the CUDA default stream

	CK(cudaMemsetAsync(d_out_offsets, 0xff, (nreads + 1) * sizeof(uint64_t), st));
	// header (tickets, error word) and status words start at zero; the plan tables are rewritten
	CK(cudaMemsetAsync(scr, 0, l.zero_bytes, st));
	PlanArgs pa;
	pa.offsets = d_offsets;
	pa.nreads = nreads;
	pa.tiling = TILING_ALIGNED;
	pa.tile = EN_TS;
	pa.frame = 0;
	pa.keydiv = 8;
	pa.fixed = 0;
	pa.tile_start = reinterpret_cast<uint32_t *>(scr + l.tile_start);
	pa.kx = reinterpret_cast<uint64_t *>(scr + l.kx);
	pa.hdr = reinterpret_cast<Hdr *>(scr + l.hdr);
	plan_scan_kernel<<<1, 1024, 0, st>>>(pa);
	g_launches++;
	if (nreads) {
		const unsigned blocks = (unsigned) ((nreads * 32 + 255) / 256);
		plan_fill_kernel<<<blocks, 256, 0, st>>>(pa.tile_start, nreads, reinterpret_cast<uint32_t *>(scr + l.tile_read));
		g_launches++;
	}
	SvbEncArgs a;
	a.samples = reinterpret_cast<const uint4 *>(d_samples);
	a.nsamples = nsamples;
	a.offsets = d_offsets;
	a.nreads = nreads;
	a.out = d_out;
	a.out_cap = out_capacity;
	a.out_offsets = d_out_offsets;
	a.hdr = reinterpret_cast<Hdr *>(scr + l.hdr);
	a.tile_start = reinterpret_cast<const uint32_t *>(scr + l.tile_start);
	a.tile_read = reinterpret_cast<const uint32_t *>(scr + l.tile_read);
	a.kx = reinterpret_cast<const uint64_t *>(scr + l.kx);
	a.status = reinterpret_cast<uint64_t *>(scr + l.status);
	void *params[] = {&a};
	CK(cudaLaunchCooperativeKernel((const void *) svb_enc_kernel<1, false>, dim3(g_static_grid[dev]), dim3(EN_THREADS),
				       params, sizeof(EnSmem), st));
	g_launches++;
	if (d_status)	// the error word of this call (press_gpu_batch_status's bits), device to device
		CK(cudaMemcpyAsync(d_status, scr + l.hdr + offsetof(Hdr, error), sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
	CK(cudaGetLastError());
	return 0;
}
