/* press_gpu.h -- C ABI of libpress_gpu.so: the B200 (sm_100a) implementation of
 * the per-read zd / svb16 / svb32 / exception / static-Huffman hot path of the
 * `press` benchmark (sashajenner/honours, directory press/).
 *
 * Two layers, both plain C (no CUDA or torch types in any signature):
 *
 *  1. Per-read drop-ins with the reference's own press.h signatures, suffixed
 *     _gpu, so press/test.c can select them with TEST(<method>_gpu, ...)
 *     (press/test.h:10-37).  Host pointers in and out, synchronous.
 *  2. A batch API on DEVICE-resident reads (this is what is measured): all
 *     reads concatenated, one launch sequence per batch, tight container out.
 *
 * There is no CPU fallback: every entry point fails (nonzero / 0 bytes + a
 * line on stderr, press_gpu_last_error()) when no sm_100 device is usable.
 *
 * Reference interface replaced by each entry point is cited as
 * (press/<file>:<line>) relative to /root/reference.
 */
#ifndef PRESS_GPU_H
#define PRESS_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- methods (batch API selector) --------------------------------------- */
enum press_gpu_method {
	PRESS_GPU_SVB16_ZD = 0,		/* svb12_zd_*              press.h:346-352  */
	PRESS_GPU_SVB_ZD = 1,		/* svb_zd_*_16             press.h:322-328  */
	PRESS_GPU_VBZ_SVB16_ZD = 2,	/* zstd_svb12_zd_* minus zstd: [n u32] + svb16-zd, press.c:2025-2045 */
	PRESS_GPU_VBZ_SVB_ZD = 3,	/* zstd_svb_zd_*_16 minus zstd: [n u32] + svb-zd,  press.c:1865-1885 */
	PRESS_GPU_VBE21_ZD = 4,		/* vbe21_zd_*_16           press.h:496-502  */
	PRESS_GPU_VBSSE21_ZD = 5,	/* vbsse21_zd_*_16         press.c:3543-3580 */
	PRESS_GPU_EX_ZD = 6,		/* hasgam_vbsse21_zdq_*_16 press.h:958-964, ex_zd.c */
	PRESS_GPU_SHUFF_VBE21_ZD = 7,	/* shuffman_vbe21_zd_*_16  press.h:632-638  */
	PRESS_GPU_SHUFF_VBSSE21_ZD = 8,	/* shuffman_vbsse21_zd_*_16 press.h:658-663, press.c:4742-4845 */
	PRESS_GPU_VBBE21_ZD = 9,	/* vbbe21_zd_*_16          press.h (press.c:3455-3500, 2775-2970): bit-packed lists */
	PRESS_GPU_SHUFF_VBBE21_ZD = 10,	/* shuffman_vbbe21_zd_*_16 press.c:4505-4614 */
	PRESS_GPU_SVB0124_ZD = 11,	/* svb0124_zd_*_16         press.h:332-336, press.c:1616-1643 */
	PRESS_GPU_UINT_ZD = 12,		/* uint_zd_*_16            press.h:106-112, press.c:639-687 */
	PRESS_GPU_NMETHODS = 13
};

/* ---- context ------------------------------------------------------------- */
/* Bind the calling process to CUDA device `device` (-1: current device) and
 * create the library's stream and staging buffers.  Idempotent.  0 on success. */
int press_gpu_init(int device);
void press_gpu_shutdown(void);
/* Last error text of this process ("" if none). */
const char *press_gpu_last_error(void);
/* Library version / build arch string, e.g. "press_gpu 0.1 sm_100a". */
const char *press_gpu_version(void);
/* Number of kernel launches issued by this library so far (bench evidence). */
uint64_t press_gpu_launch_count(void);
/* 1 if the batch calls accept `method` in this build, else 0. */
int press_gpu_method_implemented(int method);

/* Install the static Huffman code table from a code-table FILE IMAGE (the
 * 1042-byte press/NA12878_zd.huffman; format of huffman.c:440-480).  Replaces
 * read_code_table + build_symbol_encoder (press/test.c:3786-3791,
 * huffman.c:549-694, 353-365).  0 on success. */
int press_gpu_set_huffman_table(const uint8_t *file_image, size_t len);

/* ---- 1. per-read drop-ins (host pointers; press.h signatures) ------------ */
/* svb16-zd: press.h:346-352 (press.c:1678-1694).  depress nin = SAMPLES. */
uint64_t svb12_zd_gpu_bound(uint64_t nin);
void svb12_zd_gpu_press(const int16_t *in, uint64_t nin, uint8_t *out, uint64_t *nout);
void svb12_zd_gpu_depress(const uint8_t *in, uint64_t nin, int16_t *out, uint64_t *nout);

/* scalar twin of the above: press.h:340-344 (press.c:1647-1674).  Same bytes, same kernels. */
uint64_t svb12_zd_nosimd_gpu_bound(uint64_t nin);
void svb12_zd_nosimd_gpu_press(const int16_t *in, uint64_t nin, uint8_t *out, uint64_t *nout);
void svb12_zd_nosimd_gpu_depress(const uint8_t *in, uint64_t nin, int16_t *out, uint64_t *nout);

/* classical svb-zd: press.h:322-328 (press.c:1585-1612).  depress nin = SAMPLES. */
uint64_t svb_zd_gpu_bound_16(uint64_t nin);
void svb_zd_gpu_press_16(const int16_t *in, uint64_t nin, uint8_t *out, uint64_t *nout);
void svb_zd_gpu_depress_16(const uint8_t *in, uint64_t nin, int16_t *out, uint64_t *nout);

/* VBZ, the part before / after zstd (zstd itself stays on the host, outside
 * this library): press.h:378-384, 402-408 (press.c:1860-1910, 2020-2070).
 * press writes tmp = [nin u32 LE] ++ stream, exactly what the reference hands
 * to zstd_press (press.c:1881, 2041); depress takes that tmp.
 * Every depress whose sample count comes from the stream itself (these two, the exception
 * methods) reads *nout on entry as the room in `out`, in samples, the way press/test.c sets
 * it (test.c:2195, 5933), and fails instead of writing past it; 0 = unknown, no check. */
uint64_t vbz_svb_zd_gpu_bound_16(uint32_t nin);
int vbz_svb_zd_gpu_press_16(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
int vbz_svb_zd_gpu_depress_16(const uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);
uint64_t vbz_svb12_zd_gpu_bound(uint32_t nin);
int vbz_svb12_zd_gpu_press(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
int vbz_svb12_zd_gpu_depress(const uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);

/* vbe21-zd: press.h:496-502 (press.c:3411-3448).  depress nin = BYTES. */
uint64_t vbe21_zd_gpu_bound_16(uint32_t nin);
void vbe21_zd_gpu_press_16(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
void vbe21_zd_gpu_depress_16(uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);

/* vbsse21-zd: press.c:3543-3580.  depress nin = BYTES. */
uint64_t vbsse21_zd_gpu_bound_16(uint32_t nin);
void vbsse21_zd_gpu_press_16(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
void vbsse21_zd_gpu_depress_16(uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);

/* ex-zd v0: press.h:958-964 (press.c:8461-8501, ex_zd.c).  0 ok / -1 failure
 * (including the reference's own 2n+1024 scratch-cap failure, ex_zd.c:411). */
uint64_t hasgam_vbsse21_zdq_gpu_bound_16(uint32_t nin);
int hasgam_vbsse21_zdq_gpu_press_16(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
int hasgam_vbsse21_zdq_gpu_depress_16(uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);

/* static-Huffman-zd: press.h:632-638 (press.c:4409-4498).  The first argument is what the
 * reference passes: press takes the caller's `SymbolEncoder *` (huffman.h:50, built by
 * build_symbol_encoder, test.c:3786-3791), depress the caller's `huffman_node *` tree root
 * (read_code_table, huffman.c:549).  The device tables are derived from those structures
 * once per distinct pointer (layout mirrors below); NULL selects the table installed by
 * press_gpu_set_huffman_table.  The code must cover all 256 symbols with codes of <= 24 bits.
 * depress: *nout = expected sample count on entry (press.c:4472).  0 ok / 1 error. */
typedef struct press_gpu_huffman_node {		/* == huffman_node, huffman/huffman.h:22-37 */
	unsigned char isLeaf;
	unsigned long count;
	struct press_gpu_huffman_node *parent;
	union {
		struct {
			struct press_gpu_huffman_node *zero, *one;
		};
		unsigned char symbol;
	};
} press_gpu_huffman_node;
typedef struct press_gpu_huffman_code {		/* == huffman_code, huffman/huffman.h:38-48 */
	unsigned long numbits;
	unsigned char *bits;	/* bit k of the code at bit k % 8 of bits[k / 8] */
} press_gpu_huffman_code;
uint64_t shuffman_vbe21_zd_gpu_bound_16(uint32_t nin);
int shuffman_vbe21_zd_gpu_press_16(const void *table, const int16_t *in, uint32_t nin,
				   uint8_t *out, uint64_t *nout);
int shuffman_vbe21_zd_gpu_depress_16(const void *table, uint8_t *in, uint64_t nin,
				     int16_t *out, uint32_t *nout);

/* static Huffman over the vbsse21 header: press.h:658-663 (press.c:4737-4845).  Same calling
 * convention as above.  The reference keeps the header length in a uint16_t (press.c:4753)
 * and cannot read back a stream whose coded exception lists pass 65535 bytes; press fails
 * (1) for exactly those reads instead of writing such a stream. */
uint64_t shuffman_vbsse21_zd_gpu_bound_16(uint32_t nin);
int shuffman_vbsse21_zd_gpu_press_16(const void *table, const int16_t *in, uint32_t nin,
				     uint8_t *out, uint64_t *nout);
int shuffman_vbsse21_zd_gpu_depress_16(const void *table, uint8_t *in, uint64_t nin,
				       int16_t *out, uint32_t *nout);

/* vbbe21-zd (exception lists bit-packed big-endian at their minimal width, press.c:2775-2970)
 * and static Huffman over it (press.c:4505-4614; thesis' best non-range-coder ratio).  Same
 * conventions as vbsse21_zd_gpu_* / shuffman_vbsse21_zd_gpu_* above, same 16-bit header limit
 * for the Huffman variant (press.c:4521). */
uint64_t vbbe21_zd_gpu_bound_16(uint32_t nin);
void vbbe21_zd_gpu_press_16(const int16_t *in, uint32_t nin, uint8_t *out, uint64_t *nout);
void vbbe21_zd_gpu_depress_16(uint8_t *in, uint64_t nin, int16_t *out, uint32_t *nout);
uint64_t shuffman_vbbe21_zd_gpu_bound_16(uint32_t nin);
int shuffman_vbbe21_zd_gpu_press_16(const void *table, const int16_t *in, uint32_t nin,
				    uint8_t *out, uint64_t *nout);
int shuffman_vbbe21_zd_gpu_depress_16(const void *table, uint8_t *in, uint64_t nin,
				      int16_t *out, uint32_t *nout);

/* svb0124-zd: press.h:332-336 (press.c:1616-1643): streamvbyte with 0 / 1 / 2 / 4-byte codes over
 * zd.  depress nin = SAMPLES. */
uint64_t svb0124_zd_gpu_bound_16(uint64_t nin);
void svb0124_zd_gpu_press_16(const int16_t *in, uint64_t nin, uint8_t *out, uint64_t *nout);
void svb0124_zd_gpu_depress_16(const uint8_t *in, uint64_t nin, int16_t *out, uint64_t *nout);

/* uint-zd: press.h:106-112 (press.c:639-687): first sample, width byte, zd[1..] bit-packed at its
 * minimal width.  The reference splits press in two calls -- uint_zd_get_minbits_16 (which also
 * returns the zd array) and uint_zd_press_16(out_bits, in0, nin, in_zd, ...) -- because its zd
 * lives on the host; here both are one call on the samples (test.c:363-378 collapses to it).
 * depress: press.h:111, nin = SAMPLES.  0 ok / -1 failure. */
uint64_t uint_zd_gpu_bound_16(uint64_t nin);	/* worst case: width 16 */
int uint_zd_gpu_press_16(const int16_t *in, uint64_t nin, uint8_t *out, uint64_t *nout);
int uint_zd_gpu_depress_16(const uint8_t *in, uint64_t nin, int16_t *out, uint64_t *nout);

/* ---- 2. batch API (device pointers) --------------------------------------
 * Layout: d_samples = all reads back to back (int16, base 16-byte aligned),
 * d_offsets[nreads+1] = sample index of each read's first sample (read r is
 * [d_offsets[r], d_offsets[r+1]); any alignment).  The container d_out holds
 * the per-read streams back to back with no gaps, stream r at byte
 * d_out_offsets[r], total bytes in d_out_offsets[nreads]; each stream is byte
 * identical to what the per-read reference function writes.
 * `stream` is a cudaStream_t passed as void* (NULL = the CUDA default stream, as in any CUDA call).
 * All calls are asynchronous on that stream and return 0 when enqueued;
 * failures detected on the device (capacity, malformed stream, ex-zd scratch
 * cap) are reported by press_gpu_batch_status() after synchronisation. */

/* upper bound on container bytes for a batch (sum of the per-read bounds, tightened
 * to what the format can actually produce) */
uint64_t press_gpu_batch_bound(int method, uint64_t nreads, uint64_t nsamples);
/* device scratch the batch calls need for (nreads, nsamples) */
size_t press_gpu_batch_scratch_bytes(int method, uint64_t nreads, uint64_t nsamples);

int press_gpu_batch_press(int method, const int16_t *d_samples, const uint64_t *d_offsets,
			  uint64_t nreads, uint64_t nsamples, uint8_t *d_out,
			  uint64_t out_capacity, uint64_t *d_out_offsets,
			  void *d_scratch, size_t scratch_bytes, void *stream);

int press_gpu_batch_depress(int method, const uint8_t *d_in, const uint64_t *d_in_offsets,
			    uint64_t in_bytes, const uint64_t *d_offsets, uint64_t nreads,
			    uint64_t nsamples, int16_t *d_samples_out, void *d_scratch,
			    size_t scratch_bytes, void *stream);

/* Decode side-car of the static-Huffman methods (7, 8, 10).  The reference's Huffman stream
 * (huffman.c:1199-1209: a symbol count, then codes back to back) has no entry points for a
 * parallel decoder, so press_gpu_batch_depress decodes it self-synchronising, speculatively.
 * When the stream was written HERE, press can also leave, OUTSIDE the container (its bytes stay
 * exactly the reference's), one uint16 per 16 payload symbols: the low 16 bits of that symbol's
 * bit offset in the read's bitstream; entries of read r start at index (offsets[r] >> 4) + r.
 * depress with the side-car decodes every 16-symbol run independently and in one fused kernel
 * (bits -> samples), and reports status bit 1 if the side-car does not match the stream.
 * press_gpu_sidecar_bytes() = bytes to provide (0: the method has no side-car, the pointer is
 * ignored); d_sidecar = NULL selects the plain calls' behaviour. */
size_t press_gpu_sidecar_bytes(int method, uint64_t nreads, uint64_t nsamples);
int press_gpu_batch_press_sidecar(int method, const int16_t *d_samples, const uint64_t *d_offsets,
				  uint64_t nreads, uint64_t nsamples, uint8_t *d_out,
				  uint64_t out_capacity, uint64_t *d_out_offsets, void *d_scratch,
				  size_t scratch_bytes, uint8_t *d_sidecar, size_t sidecar_bytes,
				  void *stream);
int press_gpu_batch_depress_sidecar(int method, const uint8_t *d_in, const uint64_t *d_in_offsets,
				    uint64_t in_bytes, const uint64_t *d_offsets, uint64_t nreads,
				    uint64_t nsamples, int16_t *d_samples_out, void *d_scratch,
				    size_t scratch_bytes, const uint8_t *d_sidecar, size_t sidecar_bytes,
				    void *stream);

/* The exception-coded methods (4-7) collect the values > 255 in a pool inside
 * d_scratch.  Its size is `entries_per_sample` x nsamples + 4096 entries of 8
 * bytes (default 1/64; NA12878 signal has 4.3e-5 exceptions per sample;
 * 1.0 = "every sample may be an exception").  Takes effect in
 * press_gpu_batch_scratch_bytes(); the batch calls use whatever pool fits in the
 * scratch they are given and report status bit 3 when it overflows (nothing is
 * written then: call again with a larger scratch).  The host-buffer and per-read
 * entry points retry with a full-size pool by themselves. */
void press_gpu_set_exception_capacity(double entries_per_sample);

/* Device-side error word of the last batch call that used d_scratch (blocks
 * until `stream` is idle).  0 = ok; bit 0 capacity exceeded, bit 1 malformed
 * input, bit 2 ex-zd scratch cap (the reference would have returned -1; those reads
 * get an empty stream), bit 3 exception pool overflow. */
int press_gpu_batch_status(void *d_scratch, void *stream);

/* Host-buffer batch (end-to-end path): pinned staging, H2D, kernels, D2H, all
 * inside the call.  h_out must hold press_gpu_batch_bound() bytes. */
int press_gpu_batch_press_host(int method, const int16_t *h_samples, const uint64_t *h_offsets,
			       uint64_t nreads, uint8_t *h_out, uint64_t out_capacity,
			       uint64_t *h_out_offsets);
int press_gpu_batch_depress_host(int method, const uint8_t *h_in, const uint64_t *h_in_offsets,
				 const uint64_t *h_offsets, uint64_t nreads, int16_t *h_samples_out);

/* Multi-GPU container index: d_gathered holds every rank's
 * d_out_offsets array ([nranks][nreads + 1], the result of an NCCL allgather
 * of the per-read compressed byte positions); writes the byte offset of every
 * read of every rank inside ONE contiguous container, rank-major:
 * d_global[k * nreads + r], total bytes in d_global[nranks * nreads]. */
int press_gpu_global_offsets(const uint64_t *d_gathered, int nranks, uint64_t nreads,
			     uint64_t *d_global, void *stream);

/* Multi-GPU driver, one process per GPU.  Rank 0 makes an id
 * (press_gpu_comm_unique_id) and hands the 128 bytes to the other ranks by any means (MPI, a
 * file, torch.distributed ...); every rank then calls press_gpu_comm_init(id, rank, nranks)
 * after press_gpu_init.  press_gpu_allgather_offsets enqueues, on `stream`, the path's one
 * exchange -- ncclAllGather of this rank's d_out_offsets[nreads + 1] (every rank passes the same
 * nreads: shards with fewer reads pad d_out_offsets with their total) into
 * d_gathered[nranks][nreads + 1] -- and press_gpu_global_offsets on the result (d_global as
 * above).  NCCL is bound at run time (libnccl.so.2); no call here is needed on one GPU. */
#define PRESS_GPU_COMM_ID_BYTES 128
int press_gpu_comm_unique_id(uint8_t id[PRESS_GPU_COMM_ID_BYTES]);
int press_gpu_comm_init(const uint8_t id[PRESS_GPU_COMM_ID_BYTES], int rank, int nranks);
void press_gpu_comm_destroy(void);
int press_gpu_comm_size(void);
int press_gpu_comm_rank(void);
int press_gpu_allgather_offsets(const uint64_t *d_out_offsets, uint64_t nreads, uint64_t *d_gathered,
				uint64_t *d_global, void *stream);

/* ---- 3. records in, container out -------------------------------------
 * BLOW5 reader: what the reference's benchmark loop does with slow5_open / slow5_get_next
 * (press/test.h:10-37; press/sigtk/slow5lib/src/slow5.c:805-830, 3314, 3925-3960), for this
 * library's batch layout.  press_gpu_blow5_next_batch fills h_samples / h_offsets[0 .. k] with the
 * signals of the next k <= max_reads records (stopping before a read that would pass max_samples;
 * it comes first in the next batch) and returns k (0: end of file, < 0: error).  ids (optional):
 * the read ids as C strings, id_stride bytes apart.  Files whose signals are svb-zd coded
 * (signal-press byte 1) are NOT decoded on the host: the stored [n u32][svb-zd] frames are
 * PRESS_GPU_VBZ_SVB_ZD streams and go through press_gpu_batch_depress_host.  Record compression:
 * none or zlib. */
typedef struct press_gpu_blow5 press_gpu_blow5;
press_gpu_blow5 *press_gpu_blow5_open(const char *path);
void press_gpu_blow5_close(press_gpu_blow5 *f);
int press_gpu_blow5_signal_press(const press_gpu_blow5 *f);	/* 0 raw int16, 1 svb-zd frames */
int64_t press_gpu_blow5_next_batch(press_gpu_blow5 *f, uint64_t max_reads, uint64_t max_samples,
				   int16_t *h_samples, uint64_t *h_offsets, char *ids, size_t id_stride);

/* Container file: [magic "PRSGPU1"][method u64][nreads u64][nsamples u64][nbytes u64]
 * [sample offsets u64 x (nreads + 1)][stream offsets u64 x (nreads + 1)][payload], little endian:
 * the result of press_gpu_batch_press_host on disk.  _info reads the four counts (to size the
 * buffers), _read the tables and the payload. */
int press_gpu_container_write(const char *path, int method, uint64_t nreads, const uint64_t *h_offsets,
			      const uint64_t *h_out_offsets, const uint8_t *h_container);
int press_gpu_container_info(const char *path, int *method, uint64_t *nreads, uint64_t *nsamples,
			     uint64_t *nbytes);
int press_gpu_container_read(const char *path, uint64_t *h_offsets, uint64_t *h_out_offsets,
			     uint8_t *h_container);

/* svb16-zd compression of a batch (method PRESS_GPU_SVB16_ZD; same kernels, same container bytes as
 * press_gpu_batch_press) that ALLOCATES NOTHING: no cudaMalloc / cudaFree or allocating library
 * routine in the call, and no scratch argument.  No counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).
 * The scratch is a file-scope device array of PRESS_GPU_STATIC_SCRATCH_BYTES that every GPU loading
 * the library holds until the process exits; hence the limits: nreads <= PRESS_GPU_STATIC_MAX_READS
 * and nsamples <= PRESS_GPU_STATIC_MAX_SAMPLES, larger batches are rejected (-1, press_gpu_last_error)
 * -- split them or use press_gpu_batch_press.  ALL calls share the array: it keeps the last call's
 * contents (what must start at zero is cleared on `stream` at the start of each call), and calls
 * that overlap on the GPU -- different streams, or host threads without ordering between their
 * streams -- RACE on it: keep them on one stream or order them with events.
 * The first call on a device also does the one-time set-up (library context, the array's device
 * address, kernel attributes): make it once before a graph capture.
 * d_status (device, may be NULL) receives this call's error word (press_gpu_batch_status's bits:
 * bit 0 = out_capacity exceeded) by a device-to-device copy on `stream`. */
#define PRESS_GPU_STATIC_MAX_READS 4096ull
#define PRESS_GPU_STATIC_MAX_SAMPLES 33554432ull	/* 2^25: 64 MiB of raw int16 signal */
#define PRESS_GPU_STATIC_SCRATCH_BYTES 262144ull	/* 256 KiB */
int press_gpu_svb16_zd_press_static(const int16_t *d_samples, const uint64_t *d_offsets, uint64_t nreads,
				    uint64_t nsamples, uint8_t *d_out, uint64_t out_capacity,
				    uint64_t *d_out_offsets, uint32_t *d_status, void *stream);

/* Kernel timing for bench.py's roofline: when enabled every main codec kernel
 * launch is bracketed by CUDA events on its stream.  collect() waits for them
 * and returns, since the last collect, the summed device milliseconds and the
 * launch counts of [0] encode kernels and [1] decode kernels. */
void press_gpu_profile_enable(int on);
int press_gpu_profile_collect(double ms[2], uint64_t launches[2]);

/* pinned host memory for the host-buffer path (cudaHostAlloc / cudaFreeHost) */
void *press_gpu_host_alloc(size_t bytes);
void press_gpu_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* PRESS_GPU_H */
