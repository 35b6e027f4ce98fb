/* oracle/press_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("port") of the reference's per-read hot path, written from
 * the algorithm description, each function citing the reference file:line it
 * follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.  The product (libpress_gpu.so) never
 * links, loads or calls anything in oracle/.
 *
 * Parity status: PINNED.  The restatement is checked byte-for-byte against
 * the reference's own code compiled in place (oracle/_ref/libpress_ref.so,
 * built by oracle/Makefile) on data/three-reads.blow5 and seeded random reads
 * (tests/test_oracle.py), and against fixtures generated from that compiled
 * reference (tests/golden/, script tests/golden/make_golden.py).
 */
#ifndef PRESS_ORACLE_H
#define PRESS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Method ids -- identical to include/press_gpu.h (PRESS_GPU_*). */
enum {
	ORC_SVB16_ZD = 0,	/* press.h svb12_zd_*             (a3/a4) */
	ORC_SVB_ZD = 1,		/* press.h svb_zd_*_16            (a6)    */
	ORC_VBZ_SVB16_ZD = 2,	/* pre-zstd half of zstd_svb12_zd_* (a7) */
	ORC_VBZ_SVB_ZD = 3,	/* pre-zstd half of zstd_svb_zd_*_16 (a7) */
	ORC_VBE21_ZD = 4,	/* press.h vbe21_zd_*_16          (a8)    */
	ORC_VBSSE21_ZD = 5,	/* press.h vbsse21_zd_*_16        (a9)    */
	ORC_EX_ZD = 6,		/* press.h hasgam_vbsse21_zdq_*_16 (a10-a12) */
	ORC_SHUFF_VBE21_ZD = 7,	/* press.h shuffman_vbe21_zd_*_16 (a13-a16) */
	ORC_SHUFF_VBSSE21_ZD = 8,	/* press.h shuffman_vbsse21_zd_*_16 (a17) */
	ORC_VBBE21_ZD = 9,	/* press.h vbbe21_zd_*_16 (a18: bit-packed exception lists) */
	ORC_SHUFF_VBBE21_ZD = 10,	/* press.h shuffman_vbbe21_zd_*_16 (a19) */
	ORC_SVB0124_ZD = 11,	/* press.h svb0124_zd_*_16 */
	ORC_UINT_ZD = 12,	/* press.h uint_zd_*_16 */
	ORC_NMETHODS = 13
};

/* Uniform adapter (the same three calls exist with the ref_ prefix in
 * oracle/ref_shim.c, routed to the reference's own functions).
 *   bound:   upper bound in bytes for a read of n samples
 *   press:   in[n] -> out, *nout = bytes written; returns 0 ok / nonzero error
 *   depress: stream of nbytes, expected n samples -> out[n], *nout = samples
 */
uint64_t oracle_bound(int method, uint64_t n);
int oracle_press(int method, const int16_t *in, uint64_t n, uint8_t *out,
		 uint64_t *nout);
int oracle_depress(int method, const uint8_t *in, uint64_t nbytes, uint64_t n,
		   int16_t *out, uint64_t *nout);
/* parse a code-table file image (format of huffman.c:440-480); 0 on success */
int oracle_set_huffman_table(const uint8_t *file, size_t len);
/* expose the parsed table: 256 x (code bits LSB-first, length); 0 if unset */
int oracle_get_huffman_codes(uint32_t code_le[256], uint8_t len[256]);

/* building blocks, exported for unit tests */
uint16_t oracle_zigzag16(int16_t x);
int16_t oracle_unzigzag16(uint16_t z);
void oracle_zigdelta(const int16_t *in, uint64_t n, uint16_t *zd);
void oracle_unzigdelta(const uint16_t *zd, uint64_t n, int16_t *out);
uint64_t oracle_svb32_encode(const uint32_t *in, uint32_t count, uint8_t *out);
uint64_t oracle_svb32_decode(const uint8_t *in, uint32_t *out, uint32_t count);
uint8_t oracle_find_qts(const int16_t *in, uint64_t n, uint8_t max);

#ifdef __cplusplus
}
#endif
#endif
