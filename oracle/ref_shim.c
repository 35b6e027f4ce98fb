/* oracle/ref_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Thin adapter that gives the reference's OWN functions (compiled in place
 * from /root/reference/press by oracle/Makefile into oracle/_ref/) the same
 * uniform three-call shape as oracle/press_oracle.h, so tests and the CPU
 * baseline can drive either implementation through one code path.  Nothing
 * here restates an algorithm: every case is a direct call into press.h.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <stdbool.h>
#include <zstd.h>
#include "press.h"		/* the reference's header, via -I$(REF) */
#include "ex_zd.h"

enum { M_SVB16_ZD, M_SVB_ZD, M_VBZ_SVB16_ZD, M_VBZ_SVB_ZD, M_VBE21_ZD,
       M_VBSSE21_ZD, M_EX_ZD, M_SHUFF_VBE21_ZD, M_SHUFF_VBSSE21_ZD, M_VBBE21_ZD, M_SHUFF_VBBE21_ZD, M_SVB0124_ZD, M_UINT_ZD };

static huffman_node *g_root;
static SymbolEncoder g_se;
static int g_have_table;

int ref_set_huffman_table(const uint8_t *file, size_t len)
{
	unsigned int data_bytes = 0;
	FILE *fp = fmemopen((void *) file, len, "rb");
	if (!fp)
		return -1;
	huffman_node *root = NULL;
	bool ok = read_code_table(fp, &root, &data_bytes);	/* huffman.c:549 */
	fclose(fp);
	if (!ok || !root)
		return -1;
	memset(g_se, 0, sizeof g_se);
	build_symbol_encoder(root, &g_se);			/* huffman.c:353 */
	g_root = root;
	g_have_table = 1;
	return 0;
}

uint64_t ref_bound(int method, uint64_t n)
{
	switch (method) {
	case M_SVB16_ZD: return svb12_zd_bound(n);
	case M_SVB_ZD: return svb_zd_bound_16(n);
	case M_VBZ_SVB16_ZD: return sizeof(uint32_t) + svb12_zd_bound(n);
	case M_VBZ_SVB_ZD: return sizeof(uint32_t) + svb_zd_bound_16(n);
	case M_VBE21_ZD: return vbe21_zd_bound_16((uint32_t) n);
	case M_VBSSE21_ZD: return vbsse21_zd_bound_16((uint32_t) n);
	case M_EX_ZD: return hasgam_vbsse21_zdq_bound_16((uint32_t) n);
	case M_SHUFF_VBE21_ZD: return shuffman_vbe21_zd_bound_16((uint32_t) n);
	case M_SHUFF_VBSSE21_ZD: return shuffman_vbsse21_zd_bound_16((uint32_t) n);
	case M_VBBE21_ZD: return vbbe21_zd_bound_16((uint32_t) n);
	case M_SHUFF_VBBE21_ZD: return shuffman_vbbe21_zd_bound_16((uint32_t) n);
	case M_SVB0124_ZD: return svb0124_zd_bound_16(n);
	/* (the width is not known before the data is seen; the reference's bound wraps for one sample) */
	case M_UINT_ZD: return n < 2 ? 8 : uint_zd_bound_16(16, n);
	}
	return 0;
}

/* The reference only exposes the VBZ framing fused with zstd
 * (zstd_svb*_zd_press); the pre-zstd stream is recovered by running the real
 * function and undoing the zstd stage with the real libzstd. */
static int vbz_pre_zstd(int method, const int16_t *in, uint64_t n, uint8_t *out,
			uint64_t *nout)
{
	uint64_t zb = method == M_VBZ_SVB16_ZD ? zstd_svb12_zd_bound((uint32_t) n)
					       : zstd_svb_zd_bound_16((uint32_t) n);
	uint8_t *z = malloc(zb);
	uint64_t nz = zb;
	int r = method == M_VBZ_SVB16_ZD ? zstd_svb12_zd_press(in, (uint32_t) n, z, &nz)
					 : zstd_svb_zd_press_16(in, (uint32_t) n, z, &nz);
	if (r == 0) {
		size_t got = ZSTD_decompress(out, ref_bound(method, n), z, nz);
		if (ZSTD_isError(got))
			r = -1;
		else
			*nout = got;
	}
	free(z);
	return r;
}

int ref_press(int method, const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	switch (method) {
	case M_SVB16_ZD: svb12_zd_press(in, n, out, nout); return 0;
	case M_SVB_ZD: svb_zd_press_16(in, n, out, nout); return 0;
	case M_VBZ_SVB16_ZD:
	case M_VBZ_SVB_ZD: return vbz_pre_zstd(method, in, n, out, nout);
	case M_VBE21_ZD: vbe21_zd_press_16(in, (uint32_t) n, out, nout); return 0;
	case M_VBSSE21_ZD: vbsse21_zd_press_16(in, (uint32_t) n, out, nout); return 0;
	case M_EX_ZD: return hasgam_vbsse21_zdq_press_16(in, (uint32_t) n, out, nout);
	case M_SHUFF_VBE21_ZD:
		if (!g_have_table)
			return 1;
		return shuffman_vbe21_zd_press_16(&g_se, in, (uint32_t) n, out, nout);
	case M_SHUFF_VBSSE21_ZD:
		if (!g_have_table)
			return 1;
		return shuffman_vbsse21_zd_press_16(&g_se, in, (uint32_t) n, out, nout);	/* press.c:4742 */
	case M_VBBE21_ZD: vbbe21_zd_press_16(in, (uint32_t) n, out, nout); return 0;	/* press.c:3460 */
	case M_SHUFF_VBBE21_ZD:
		if (!g_have_table)
			return 1;
		return shuffman_vbbe21_zd_press_16(&g_se, in, (uint32_t) n, out, nout);	/* press.c:4510 */
	case M_SVB0124_ZD: svb0124_zd_press_16(in, n, out, nout); return 0;	/* press.c:1621 */
	case M_UINT_ZD: {	/* the two calls of press/test.c:363-378 */
		uint16_t *zd = NULL;
		if (n == 0)
			return -1;
		uint8_t bits = uint_zd_get_minbits_16(in, n, &zd);	/* press.c:639 */
		int r = uint_zd_press_16(bits, in[0], n, zd, out, nout);	/* press.c:659 */
		free(zd);
		return r;
	}
	}
	return -1;
}

int ref_depress(int method, const uint8_t *in, uint64_t nbytes, uint64_t n,
		int16_t *out, uint64_t *nout)
{
	uint32_t n32 = (uint32_t) n;
	int r = 0;
	switch (method) {
	case M_SVB16_ZD: svb12_zd_depress(in, n, out, nout); return 0;
	case M_SVB_ZD: svb_zd_depress_16(in, n, out, nout); return 0;
	case M_VBZ_SVB16_ZD: {	/* second half of zstd_svb12_zd_depress, press.c:2062-2065 */
		uint32_t cnt;
		memcpy(&cnt, in, sizeof cnt);
		svb12_zd_depress(in + sizeof cnt, cnt, out, nout);
		return 0;
	}
	case M_VBZ_SVB_ZD: {	/* press.c:1902-1906 */
		uint32_t cnt;
		memcpy(&cnt, in, sizeof cnt);
		svb_zd_depress_16(in + sizeof cnt, cnt, out, nout);
		return 0;
	}
	case M_VBE21_ZD: vbe21_zd_depress_16((uint8_t *) in, nbytes, out, &n32); break;
	case M_VBSSE21_ZD: vbsse21_zd_depress_16((uint8_t *) in, nbytes, out, &n32); break;
	case M_EX_ZD: r = hasgam_vbsse21_zdq_depress_16((uint8_t *) in, nbytes, out, &n32); break;
	case M_SHUFF_VBE21_ZD:
		if (!g_have_table)
			return 1;
		r = shuffman_vbe21_zd_depress_16(g_root, (uint8_t *) in, nbytes, out, &n32);
		break;
	case M_SHUFF_VBSSE21_ZD:
		if (!g_have_table)
			return 1;
		r = shuffman_vbsse21_zd_depress_16(g_root, (uint8_t *) in, nbytes, out, &n32);	/* press.c:4792 */
		break;
	case M_VBBE21_ZD: vbbe21_zd_depress_16((uint8_t *) in, nbytes, out, &n32); break;	/* press.c:3476 */
	case M_SHUFF_VBBE21_ZD:
		if (!g_have_table)
			return 1;
		r = shuffman_vbbe21_zd_depress_16(g_root, (uint8_t *) in, nbytes, out, &n32);	/* press.c:4560 */
		break;
	case M_SVB0124_ZD: svb0124_zd_depress_16(in, n, out, nout); return 0;	/* press.c:1631 */
	case M_UINT_ZD: *nout = n; return uint_zd_depress_16(in, n, out, nout);	/* press.c:674 */
	default: return -1;
	}
	*nout = n32;
	return r;
}

/* a5 pin: the scalar twin svb12_zd_nosimd_* (press.c:1647-1674 ->
 * streamvbyte_12_encode.c:90-110) called directly, so a committed test can assert that it
 * writes the same bytes as the SIMD svb12_zd_* path and as the GPU. */
uint64_t ref_nosimd_bound(uint64_t n) { return svb12_zd_nosimd_bound(n); }
int ref_nosimd_press(const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	svb12_zd_nosimd_press(in, n, out, nout);
	return 0;
}
int ref_nosimd_depress(const uint8_t *in, uint64_t n, int16_t *out, uint64_t *nout)
{
	svb12_zd_nosimd_depress(in, n, out, nout);
	return 0;
}
