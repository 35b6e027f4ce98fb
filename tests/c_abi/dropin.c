/* tests/c_abi/dropin.c -- the drop-in boundary exercised from plain C (gcc -std=c99).
 *
 * What a maintainer of press/test.c adds to select the GPU methods:
 * test_<method>_gpu functions with the body of the reference's test_<method>
 * (press/test.c:1756-1817 svb12_zd, :3770-3830 shuffman_vbe21_zd, :5900-5950
 * hasgam_vbsse21_zdq) and the _gpu suffix on the three calls.  This file holds those
 * functions, compiled as C against include/press_gpu.h and linked to libpress_gpu.so, and a
 * small main() that feeds them reads from a raw file instead of slow5lib:
 *
 *     dropin READS.bin TABLE.huffman
 *     READS.bin = [nreads u64][offsets u64 x (nreads + 1)][samples i16 x offsets[nreads]]
 *
 * Unlike test.c:1804 the round trip compares ALL samples.  Exit 0 iff every method round
 * trips every read; one TSV line per (method, read) on stdout: method, n, bytes, FNV-1a-64.
 * The Huffman methods get the code table the way the reference passes it: a SymbolEncoder
 * (huffman_code *[256]) for press and a huffman_node tree for depress, built here from the
 * table file (what read_code_table + build_symbol_encoder produce, huffman.c:549-694, 353-365).
 */
#include <inttypes.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "press_gpu.h"

#define ASSERT(x)                                                                  \
	do {                                                                       \
		if (!(x)) {                                                        \
			fprintf(stderr, "dropin: %s:%d: ASSERT(%s) failed: %s\n", \
				__FILE__, __LINE__, #x, press_gpu_last_error());   \
			return -1;                                                 \
		}                                                                  \
	} while (0)

static uint64_t fnv1a64(const uint8_t *p, uint64_t n)
{
	uint64_t h = 0xCBF29CE484222325ull;
	uint64_t i;
	for (i = 0; i < n; i++)
		h = (h ^ p[i]) * 0x100000001B3ull;
	return h;
}

static void report(const char *method, uint64_t n, const uint8_t *st, uint64_t len)
{
	printf("%s\t%" PRIu64 "\t%" PRIu64 "\t%016" PRIx64 "\n", method, n, len, fnv1a64(st, len));
}

/* press/test.c:1756-1817 with svb12_zd -> svb12_zd_gpu */
static int test_svb12_zd_gpu(const int16_t *sigs, uint64_t nr_sigs)
{
	uint64_t pressbound = svb12_zd_gpu_bound(nr_sigs);
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound, depress_len = nr_sigs;
	int16_t *sigs_depress;
	ASSERT(sigs_press);
	svb12_zd_gpu_press(sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(press_len > 0 && press_len <= pressbound);
	sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	ASSERT(sigs_depress);
	svb12_zd_gpu_depress(sigs_press, nr_sigs, sigs_depress, &depress_len);
	ASSERT(depress_len == nr_sigs);
	ASSERT(memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("svb12_zd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* the scalar twin's slot (press.h:340-344) */
static int test_svb12_zd_nosimd_gpu(const int16_t *sigs, uint64_t nr_sigs)
{
	uint64_t pressbound = svb12_zd_nosimd_gpu_bound(nr_sigs);
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound, depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	ASSERT(sigs_press && sigs_depress);
	svb12_zd_nosimd_gpu_press(sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(press_len > 0);
	svb12_zd_nosimd_gpu_depress(sigs_press, nr_sigs, sigs_depress, &depress_len);
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("svb12_zd_nosimd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* press/test.c (vbe21_zd): depress nin = BYTES, *nout = room on entry */
static int test_vbe21_zd_gpu(const int16_t *sigs, uint32_t nr_sigs)
{
	uint64_t pressbound = vbe21_zd_gpu_bound_16(nr_sigs) + 8ull * nr_sigs;	/* harness slack */
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound;
	uint32_t depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	ASSERT(sigs_press && sigs_depress);
	vbe21_zd_gpu_press_16(sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(press_len > 0);
	vbe21_zd_gpu_depress_16(sigs_press, press_len, sigs_depress, &depress_len);
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("vbe21_zd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* press/test.c:5900-5950 with hasgam_vbsse21_zdq -> hasgam_vbsse21_zdq_gpu */
static int test_hasgam_vbsse21_zdq_gpu(const int16_t *sigs, uint32_t nr_sigs)
{
	uint64_t pressbound = hasgam_vbsse21_zdq_gpu_bound_16(nr_sigs);
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound;
	uint32_t depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	int ret;
	ASSERT(sigs_press && sigs_depress);
	ret = hasgam_vbsse21_zdq_gpu_press_16(sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(ret == 0);
	ret = hasgam_vbsse21_zdq_gpu_depress_16(sigs_press, press_len, sigs_depress, &depress_len);
	ASSERT(ret == 0);
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("hasgam_vbsse21_zdq_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* ---- the code table as the reference's structures (huffman.h:22-51) ---------------- */
static press_gpu_huffman_code g_codes[256];
static unsigned char g_bits[256][4];
static press_gpu_huffman_code *g_se[256];		/* == SymbolEncoder */
static press_gpu_huffman_node g_nodes[1024];
static int g_nnodes;

static press_gpu_huffman_node *new_node(void)
{
	press_gpu_huffman_node *n = &g_nodes[g_nnodes++];
	memset(n, 0, sizeof *n);
	return n;
}

/* table file (huffman.c:440-480): [count u32 BE][bytes u32 BE] then [sym][nbits][code bytes] */
static int load_table(const char *path)
{
	FILE *fp = fopen(path, "rb");
	unsigned char hdr[8], e[2];
	uint32_t count, i, k;
	press_gpu_huffman_node *root;
	if (!fp || fread(hdr, 1, 8, fp) != 8)
		return -1;
	count = ((uint32_t) hdr[0] << 24) | ((uint32_t) hdr[1] << 16) | ((uint32_t) hdr[2] << 8) | hdr[3];
	if (count != 256)
		return -1;
	root = new_node();
	for (i = 0; i < count; i++) {
		press_gpu_huffman_node *n = root;
		uint32_t sym, nb;
		if (fread(e, 1, 2, fp) != 2)
			return -1;
		sym = e[0];
		nb = e[1];
		if (nb == 0 || nb > 32 || fread(g_bits[sym], 1, (nb + 7) / 8, fp) != (nb + 7) / 8)
			return -1;
		g_codes[sym].numbits = nb;
		g_codes[sym].bits = g_bits[sym];
		g_se[sym] = &g_codes[sym];
		for (k = 0; k < nb; k++) {	/* bit k of the code: branch k from the root */
			int b = (g_bits[sym][k / 8] >> (k % 8)) & 1;
			press_gpu_huffman_node **c = b ? &n->one : &n->zero;
			if (!*c) {
				*c = new_node();
				(*c)->parent = n;
			}
			n = *c;
		}
		n->isLeaf = 1;
		n->symbol = (unsigned char) sym;
	}
	fclose(fp);
	return 0;
}

/* press/test.c:3770-3830 with shuffman_vbe21_zd -> shuffman_vbe21_zd_gpu; se / root as there */
static int test_shuffman_vbe21_zd_gpu(const int16_t *sigs, uint32_t nr_sigs)
{
	const void *se = g_se, *root = &g_nodes[0];
	uint64_t pressbound = shuffman_vbe21_zd_gpu_bound_16(nr_sigs) + 8ull * nr_sigs;
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound;
	uint32_t depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	int ret;
	ASSERT(sigs_press && sigs_depress);
	ret = shuffman_vbe21_zd_gpu_press_16(se, sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(ret == 0);
	ret = shuffman_vbe21_zd_gpu_depress_16(root, sigs_press, press_len, sigs_depress, &depress_len);
	ASSERT(ret == 0);
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("shuffman_vbe21_zd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* press/test.c (svb0124_zd, press.h:332-336): depress nin = SAMPLES */
static int test_svb0124_zd_gpu(const int16_t *sigs, uint64_t nr_sigs)
{
	uint64_t pressbound = svb0124_zd_gpu_bound_16(nr_sigs);
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound, depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	ASSERT(sigs_press && sigs_depress);
	svb0124_zd_gpu_press_16(sigs, nr_sigs, sigs_press, &press_len);
	ASSERT(press_len > 0 && press_len <= pressbound);
	svb0124_zd_gpu_depress_16(sigs_press, nr_sigs, sigs_depress, &depress_len);
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("svb0124_zd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

/* press/test.c (uint_zd, press.h:106-112: uint_zd_get_minbits_16 + uint_zd_press_16 in one call) */
static int test_uint_zd_gpu(const int16_t *sigs, uint64_t nr_sigs)
{
	uint64_t pressbound = uint_zd_gpu_bound_16(nr_sigs);
	uint8_t *sigs_press = malloc(pressbound);
	uint64_t press_len = pressbound, depress_len = nr_sigs;
	int16_t *sigs_depress = malloc(nr_sigs * sizeof *sigs_depress);
	ASSERT(sigs_press && sigs_depress);
	ASSERT(uint_zd_gpu_press_16(sigs, nr_sigs, sigs_press, &press_len) == 0);
	ASSERT(press_len > 0 && press_len <= pressbound);
	ASSERT(uint_zd_gpu_depress_16(sigs_press, nr_sigs, sigs_depress, &depress_len) == 0);	/* nin = samples, press.h:111 */
	ASSERT(depress_len == nr_sigs && memcmp(sigs, sigs_depress, nr_sigs * sizeof *sigs) == 0);
	report("uint_zd_gpu", nr_sigs, sigs_press, press_len);
	free(sigs_press);
	free(sigs_depress);
	return 0;
}

int main(int argc, char **argv)
{
	FILE *fp;
	uint64_t nreads, r, *offsets;
	int16_t *samples;
	int bad = 0;
	if (argc < 3) {
		fprintf(stderr, "usage: %s READS.bin TABLE.huffman\n", argv[0]);
		return 2;
	}
	if (press_gpu_init(-1) != 0) {	/* no CPU fallback: fails loudly without a B200 */
		fprintf(stderr, "dropin: %s\n", press_gpu_last_error());
		return 3;
	}
	if (load_table(argv[2]) != 0) {
		fprintf(stderr, "dropin: cannot read the code table %s\n", argv[2]);
		return 2;
	}
	fp = fopen(argv[1], "rb");
	if (!fp || fread(&nreads, 8, 1, fp) != 1)
		return 2;
	offsets = malloc((nreads + 1) * 8);
	if (fread(offsets, 8, nreads + 1, fp) != nreads + 1)
		return 2;
	samples = malloc(offsets[nreads] * 2 + 2);
	if (fread(samples, 2, offsets[nreads], fp) != offsets[nreads])
		return 2;
	fclose(fp);
	/* TEST(method, &res, fp) loop of press/test.h:10-37: every read through every method */
	for (r = 0; r < nreads; r++) {
		const int16_t *s = samples + offsets[r];
		const uint64_t n = offsets[r + 1] - offsets[r];
		bad |= test_svb12_zd_gpu(s, n);
		bad |= test_svb12_zd_nosimd_gpu(s, n);
		bad |= test_vbe21_zd_gpu(s, (uint32_t) n);
		bad |= test_hasgam_vbsse21_zdq_gpu(s, (uint32_t) n);
		bad |= test_shuffman_vbe21_zd_gpu(s, (uint32_t) n);
		bad |= test_svb0124_zd_gpu(s, n);
		bad |= test_uint_zd_gpu(s, n);
	}
	press_gpu_shutdown();
	free(offsets);
	free(samples);
	if (bad) {
		fprintf(stderr, "dropin: FAILED\n");
		return 1;
	}
	fprintf(stderr, "dropin ok: %" PRIu64 " reads x 7 methods\n", nreads);
	return 0;
}
