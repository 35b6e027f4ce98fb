/* oracle/press_oracle.c -- TEST INFRASTRUCTURE ONLY (see press_oracle.h).
 *
 * Plain-C restatement of the reference's per-read zd / svb16 / svb32 / vbe21 /
 * vbsse21 / ex-zd / static-Huffman path.  Written from the algorithm, not
 * copied: every routine is a single straightforward loop over the read with
 * explicit 16-bit wrap-around, and cites the reference lines it restates
 * (paths relative to /root/reference/press).
 *
 * Parity: PINNED against the compiled reference (oracle/_ref) and against
 * tests/golden/ (see press_oracle.h).
 */
#include "press_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ---------- little helpers ------------------------------------------------ */

static void put_u16(uint8_t *p, uint16_t v) { p[0] = v & 0xff; p[1] = v >> 8; }
static void put_u32(uint8_t *p, uint32_t v)
{
	p[0] = v & 0xff; p[1] = (v >> 8) & 0xff; p[2] = (v >> 16) & 0xff; p[3] = v >> 24;
}
static void put_u64(uint8_t *p, uint64_t v)
{
	put_u32(p, (uint32_t) v);
	put_u32(p + 4, (uint32_t) (v >> 32));
}
static uint16_t get_u16(const uint8_t *p) { return (uint16_t) (p[0] | (p[1] << 8)); }
static uint32_t get_u32(const uint8_t *p)
{
	return (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) |
	       ((uint32_t) p[3] << 24);
}
static uint64_t get_u64(const uint8_t *p)
{
	return (uint64_t) get_u32(p) | ((uint64_t) get_u32(p + 4) << 32);
}
static void put_u32_be(uint8_t *p, uint32_t v)
{
	p[0] = v >> 24; p[1] = (v >> 16) & 0xff; p[2] = (v >> 8) & 0xff; p[3] = v & 0xff;
}
static uint32_t get_u32_be(const uint8_t *p)
{
	return ((uint32_t) p[0] << 24) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 8) | p[3];
}

/* ---------- a1/a2: zigzag, zigzag-delta ----------------------------------- */

/* trans.c:75-78 (zigzag_one_16); svb16/encode_scalar.hpp:9-12 */
uint16_t oracle_zigzag16(int16_t x)
{
	uint16_t u = (uint16_t) x;
	uint16_t sign = (u & 0x8000u) ? 0xffffu : 0u;	/* arithmetic x >> 15 */
	return (uint16_t) ((uint16_t) (u << 1) ^ sign);
}

/* trans.c:80-83 (unzigzag_one_16); svb16/decode_scalar.hpp:10-13 */
int16_t oracle_unzigzag16(uint16_t z)
{
	uint16_t u = (uint16_t) ((z >> 1) ^ (uint16_t) (0u - (z & 1u)));
	return (int16_t) u;
}

/* trans.c:215-231 (zigdelta_16_u16): prev starts at 0, all arithmetic mod 2^16 */
void oracle_zigdelta(const int16_t *in, uint64_t n, uint16_t *zd)
{
	uint16_t prev = 0;
	for (uint64_t i = 0; i < n; i++) {
		uint16_t cur = (uint16_t) in[i];
		zd[i] = oracle_zigzag16((int16_t) (uint16_t) (cur - prev));
		prev = cur;
	}
}

/* trans.c:260-270 (unzigdelta_u16_16): running sum from 0, mod 2^16 */
void oracle_unzigdelta(const uint16_t *zd, uint64_t n, int16_t *out)
{
	uint16_t acc = 0;
	for (uint64_t i = 0; i < n; i++) {
		acc = (uint16_t) (acc + (uint16_t) oracle_unzigzag16(zd[i]));
		out[i] = (int16_t) acc;
	}
}

/* ---------- svb16: 1 key bit per value, 1 or 2 data bytes ------------------ */

/* svb16/svb16.h:11-15 */
static uint64_t svb16_keylen(uint64_t n) { return (n + 7) / 8; }

/* svb16/encode_scalar.hpp:14-62 on already-transformed u16 values.
 * key bit (i & 7) of byte (i >> 3) is 1 iff v[i] >= 256; data is 1 byte or 2
 * bytes little-endian; unused bits of the last key byte stay 0. */
static uint64_t svb16_pack(const uint16_t *v, uint64_t n, uint8_t *out)
{
	uint64_t nk = svb16_keylen(n);
	uint8_t *data = out + nk;
	memset(out, 0, nk);
	for (uint64_t i = 0; i < n; i++) {
		if (v[i] < 256) {
			*data++ = (uint8_t) v[i];
		} else {
			put_u16(data, v[i]);
			data += 2;
			out[i >> 3] |= (uint8_t) (1u << (i & 7));
		}
	}
	return (uint64_t) (data - out);
}

/* svb16/decode_scalar.hpp:31-71 */
static uint64_t svb16_unpack(const uint8_t *in, uint64_t n, uint16_t *v)
{
	const uint8_t *data = in + svb16_keylen(n);
	for (uint64_t i = 0; i < n; i++) {
		if ((in[i >> 3] >> (i & 7)) & 1) {
			v[i] = get_u16(data);
			data += 2;
		} else {
			v[i] = *data++;
		}
	}
	return (uint64_t) (data - in);
}

/* ---------- svb32: classical streamvbyte, 2-bit keys ----------------------- */

/* streamvbyte/src/streamvbyte_encode.c:11-59, 70-94: key code = bytes-1 at bit
 * 2*(i&3) of key byte i>>2; keys first ((count+3)/4 bytes), then data LE. */
uint64_t oracle_svb32_encode(const uint32_t *in, uint32_t count, uint8_t *out)
{
	uint64_t nk = ((uint64_t) count + 3) / 4;
	uint8_t *data = out + nk;
	memset(out, 0, nk);
	for (uint32_t i = 0; i < count; i++) {
		uint32_t v = in[i];
		unsigned nb = v < (1u << 8) ? 1 : v < (1u << 16) ? 2 : v < (1u << 24) ? 3 : 4;
		for (unsigned b = 0; b < nb; b++)
			*data++ = (uint8_t) (v >> (8 * b));
		out[i >> 2] |= (uint8_t) ((nb - 1) << (2 * (i & 3)));
	}
	return (uint64_t) (data - out);
}

/* streamvbyte/src/streamvbyte_decode.c:16-86; returns bytes consumed */
uint64_t oracle_svb32_decode(const uint8_t *in, uint32_t *out, uint32_t count)
{
	if (count == 0)
		return 0;
	const uint8_t *data = in + ((uint64_t) count + 3) / 4;
	for (uint32_t i = 0; i < count; i++) {
		unsigned nb = ((in[i >> 2] >> (2 * (i & 3))) & 3u) + 1;
		uint32_t v = 0;
		for (unsigned b = 0; b < nb; b++)
			v |= (uint32_t) (*data++) << (8 * b);
		out[i] = v;
	}
	return (uint64_t) (data - in);
}

/* streamvbyte/include/streamvbyte.h:35-41 and 42-49 (STREAMVBYTE_PADDING = 16) */
static uint64_t svb32_bound(uint64_t n) { return (n + 3) / 4 + 4 * n + 16; }
static uint64_t svb16_bound(uint64_t n) { return svb16_keylen(n) + 4 * n + 16; }

/* ---------- a3/a4: svb16-zd (press.c:1678-1694) ---------------------------- */

static int svb16_zd_press(const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	uint16_t *zd = malloc((n ? n : 1) * sizeof *zd);
	if (!zd)
		return -1;
	oracle_zigdelta(in, n, zd);
	*nout = svb16_pack(zd, n, out);
	free(zd);
	return 0;
}

static int svb16_zd_depress(const uint8_t *in, uint64_t n, int16_t *out, uint64_t *nout)
{
	uint16_t *zd = malloc((n ? n : 1) * sizeof *zd);
	if (!zd)
		return -1;
	(void) svb16_unpack(in, n, zd);
	oracle_unzigdelta(zd, n, out);
	free(zd);
	*nout = n;
	return 0;
}

/* ---------- a6: classical svb-zd (press.c:1585-1612) ----------------------- */

static int svb32_zd_press(const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	uint32_t *w = malloc((n ? n : 1) * sizeof *w);
	uint16_t *zd = malloc((n ? n : 1) * sizeof *zd);
	if (!w || !zd) {
		free(w); free(zd);
		return -1;
	}
	oracle_zigdelta(in, n, zd);		/* trans.c:233-249 widens the same values */
	for (uint64_t i = 0; i < n; i++)
		w[i] = zd[i];
	*nout = oracle_svb32_encode(w, (uint32_t) n, out);
	free(w); free(zd);
	return 0;
}

static int svb32_zd_depress(const uint8_t *in, uint64_t n, int16_t *out, uint64_t *nout)
{
	uint32_t *w = malloc((n ? n : 1) * sizeof *w);
	uint16_t *zd = malloc((n ? n : 1) * sizeof *zd);
	if (!w || !zd) {
		free(w); free(zd);
		return -1;
	}
	(void) oracle_svb32_decode(in, w, (uint32_t) n);
	for (uint64_t i = 0; i < n; i++)
		zd[i] = (uint16_t) w[i];	/* trans.c:272-282 truncates via unzigzag_one_16 */
	oracle_unzigdelta(zd, n, out);
	free(w); free(zd);
	*nout = n;
	return 0;
}

/* ---------- exception lists shared by a8/a9/a11 ---------------------------- */

struct exlist {
	uint32_t nex;
	uint32_t *pos;		/* index into v[] (= sample index - 1) */
	uint32_t *val;		/* v[pos] (raw, >= 256) */
};

/* first pass of press.c:2692-2705 / 3213-3227 / ex_zd.c:45-74 */
static int collect_exceptions(const uint16_t *v, uint32_t m, struct exlist *e)
{
	uint32_t cnt = 0;
	for (uint32_t i = 0; i < m; i++)
		cnt += v[i] > 255;
	e->nex = cnt;
	e->pos = malloc((cnt ? cnt : 1) * sizeof *e->pos);
	e->val = malloc((cnt ? cnt : 1) * sizeof *e->val);
	if (!e->pos || !e->val)
		return -1;
	cnt = 0;
	for (uint32_t i = 0; i < m; i++) {
		if (v[i] > 255) {
			e->pos[cnt] = i;
			e->val[cnt] = v[i];
			cnt++;
		}
	}
	return 0;
}

static void free_exceptions(struct exlist *e) { free(e->pos); free(e->val); }

/* second pass of press.c:2717-2725 / 3277-3285 / ex_zd.c:156-165 */
static uint8_t *emit_low_bytes(const uint16_t *v, uint32_t m, uint8_t *dst)
{
	for (uint32_t i = 0; i < m; i++)
		if (v[i] <= 255)
			*dst++ = (uint8_t) v[i];
	return dst;
}

/* merge loop of press.c:2750-2761 / 3342-3353 / ex_zd.c:274-287.  The
 * exceptions have already been scattered into v[]; walk until the input is
 * used up and every exception has been stepped over.  Returns the count. */
static uint32_t merge_low_bytes(const uint8_t *src, uint64_t nsrc, const uint32_t *pos,
				uint32_t nex, uint16_t *v)
{
	uint64_t off = 0;
	uint32_t i = 0, j = 0;
	while (off < nsrc || j < nex) {
		if (j < nex && i == pos[j])
			j++;
		else
			v[i] = src[off++];
		i++;
	}
	return i;
}

/* trans.c:129-149 */
static void pos_to_gaps(const uint32_t *pos, uint32_t nex, uint32_t *gap)
{
	for (uint32_t k = 0; k < nex; k++)
		gap[k] = k ? pos[k] - pos[k - 1] - 1 : pos[0];
}

/* trans.c:186-197 */
static void gaps_to_pos(uint32_t *p, uint32_t nex)
{
	for (uint32_t k = 1; k < nex; k++)
		p[k] += p[k - 1] + 1;
}

/* press.c:2575-2583 evaluated the way C does: 1 + n*0.2*6 + n*0.8 in double,
 * truncated on return.  (sizeof(uint8_t) + nin*0.2*(sizeof(uint32_t)+2) + nin*0.8) */
static uint64_t vb1e2_bound(uint32_t m)
{
	return (uint64_t) (1 + m * 0.2 * 6 + m * 0.8);
}

/* ---------- a8: vbe21-zd (press.c:2679-2765, 3411-3448) --------------------- */

static int vbe21_zd_press(const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (n == 0)
		return -1;	/* reference underflows nin - 1 (press.c:3426); undefined */
	uint16_t *zd = malloc(n * sizeof *zd);
	struct exlist e;
	if (!zd)
		return -1;
	oracle_zigdelta(in, n, zd);
	if (collect_exceptions(zd + 1, n - 1, &e))
		return -1;
	uint8_t *p = out;
	put_u16(p, zd[0]); p += 2;
	put_u32(p, e.nex); p += 4;
	for (uint32_t k = 0; k < e.nex; k++, p += 4)
		put_u32(p, e.pos[k]);
	for (uint32_t k = 0; k < e.nex; k++, p += 2)
		put_u16(p, (uint16_t) e.val[k]);
	p = emit_low_bytes(zd + 1, n - 1, p);
	*nout = (uint64_t) (p - out);
	free_exceptions(&e);
	free(zd);
	return 0;
}

static int vbe21_zd_depress(const uint8_t *in, uint64_t nbytes, uint64_t ncap,
			    int16_t *out, uint64_t *nout)
{
	/* the reference sizes its scratch by nbytes (press.c:3438); we take the
	 * larger of that and the caller's expected count */
	uint64_t cap = (nbytes > ncap ? nbytes : ncap) + 1;
	uint16_t *zd = malloc(cap * sizeof *zd);
	if (!zd || nbytes < 6)
		return -1;
	zd[0] = get_u16(in);
	uint32_t nex = get_u32(in + 2);
	uint64_t off = 6;
	uint32_t *pos = malloc((nex ? nex : 1) * sizeof *pos);
	for (uint32_t k = 0; k < nex; k++, off += 4)
		pos[k] = get_u32(in + off);
	for (uint32_t k = 0; k < nex; k++, off += 2)
		zd[1 + pos[k]] = get_u16(in + off);
	uint32_t m = merge_low_bytes(in + off, nbytes - off, pos, nex, zd + 1);
	oracle_unzigdelta(zd, (uint64_t) m + 1, out);
	*nout = (uint64_t) m + 1;
	free(pos);
	free(zd);
	return 0;
}

/* ---------- a9: vbsse21-zd (press.c:3191-3360, 3543-3580) ------------------- */

static int vbsse21_zd_press(const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (n == 0)
		return -1;
	uint16_t *zd = malloc(n * sizeof *zd);
	struct exlist e;
	if (!zd)
		return -1;
	oracle_zigdelta(in, n, zd);
	if (collect_exceptions(zd + 1, n - 1, &e))
		return -1;
	uint8_t *p = out;
	put_u16(p, zd[0]); p += 2;
	put_u32(p, e.nex); p += 4;
	if (e.nex > 1) {
		uint32_t *gap = malloc(e.nex * sizeof *gap);
		uint16_t *exv = malloc(e.nex * sizeof *exv);
		pos_to_gaps(e.pos, e.nex, gap);
		uint32_t len = (uint32_t) oracle_svb32_encode(gap, e.nex, p + 4);
		put_u32(p, len); p += 4 + len;
		/* values: (v - 256) as int16, svb16 WITHOUT zd (press.c:3255 -> 1573) */
		for (uint32_t k = 0; k < e.nex; k++)
			exv[k] = (uint16_t) (e.val[k] - 256);
		len = (uint32_t) svb16_pack(exv, e.nex, p + 4);
		put_u32(p, len); p += 4 + len;
		free(gap); free(exv);
	} else if (e.nex == 1) {
		put_u32(p, e.pos[0]); p += 4;
		put_u16(p, (uint16_t) (e.val[0] - 256)); p += 2;	/* int16: 2 bytes, press.c:3272 */
	}
	p = emit_low_bytes(zd + 1, n - 1, p);
	*nout = (uint64_t) (p - out);
	free_exceptions(&e);
	free(zd);
	return 0;
}

static int vbsse21_zd_depress(const uint8_t *in, uint64_t nbytes, uint64_t ncap,
			      int16_t *out, uint64_t *nout)
{
	uint64_t cap = (nbytes > ncap ? nbytes : ncap) + 1;
	uint16_t *zd = malloc(cap * sizeof *zd);
	if (!zd || nbytes < 6)
		return -1;
	zd[0] = get_u16(in);
	uint32_t nex = get_u32(in + 2);
	uint64_t off = 6;
	uint32_t *pos = malloc((nex ? nex : 1) * sizeof *pos);
	uint16_t *exv = malloc((nex ? nex : 1) * sizeof *exv);
	if (nex > 1) {
		uint32_t len = get_u32(in + off); off += 4;
		(void) oracle_svb32_decode(in + off, pos, nex); off += len;
		gaps_to_pos(pos, nex);
		len = get_u32(in + off); off += 4;
		(void) svb16_unpack(in + off, nex, exv); off += len;
	} else if (nex == 1) {
		pos[0] = get_u32(in + off); off += 4;
		exv[0] = get_u16(in + off); off += 2;
	}
	for (uint32_t k = 0; k < nex; k++)	/* int16 + 256 stored into uint16: wraps */
		zd[1 + pos[k]] = (uint16_t) (exv[k] + 256);
	uint32_t m = merge_low_bytes(in + off, nbytes - off, pos, nex, zd + 1);
	oracle_unzigdelta(zd, (uint64_t) m + 1, out);
	*nout = (uint64_t) m + 1;
	free(pos); free(exv); free(zd);
	return 0;
}

/* ---------- a10-a12: ex-zd v0 (ex_zd.c whole file, press.c:8461-8501) ------- */

/* ex_zd.c:358-381: largest q <= max such that every sample has q low zero bits
 * (a zero sample never lowers q) */
uint8_t oracle_find_qts(const int16_t *in, uint64_t n, uint8_t max)
{
	uint8_t q = max;
	for (uint64_t i = 0; i < n && q; i++) {
		uint16_t s = (uint16_t) in[i];
		while (q && (s & ((1u << q) - 1u)))
			q--;
	}
	return q;
}

static int ex_zd_press(const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (n == 0)
		return -1;
	uint64_t cap = 2 * (uint64_t) n + 1024;		/* ex_zd.c:411 scratch cap */
	/* worst case stream: 16 + 2 lists of (4 + svb32 bound) + payload */
	uint8_t *tmp = malloc(64 + 12 * (uint64_t) n);
	int16_t *s = malloc(n * sizeof *s);
	uint16_t *zd = malloc(n * sizeof *zd);
	struct exlist e;
	if (!tmp || !s || !zd)
		return -1;

	uint8_t q = oracle_find_qts(in, n, 5);
	for (uint32_t i = 0; i < n; i++)		/* ex_zd.c:383-394: arithmetic >> on int16 */
		s[i] = (int16_t) (in[i] >> q);
	oracle_zigdelta(s, n, zd);
	if (collect_exceptions(zd + 1, n - 1, &e))
		return -1;

	uint8_t *p = tmp;
	*p++ = 0;				/* version, ex_zd.c:408,420-423 */
	put_u64(p, n); p += 8;			/* ex_zd.c:425-428 */
	*p++ = q;				/* ex_zd.c:436-439 */
	put_u16(p, zd[0]); p += 2;		/* ex_zd.c:311-314 */
	put_u32(p, e.nex); p += 4;		/* ex_zd.c:82-84 */
	if (e.nex > 1) {			/* ex_zd.c:86-144: both lists svb32 */
		uint32_t *w = malloc(e.nex * sizeof *w);
		pos_to_gaps(e.pos, e.nex, w);
		uint32_t len = (uint32_t) oracle_svb32_encode(w, e.nex, p + 4);
		put_u32(p, len); p += 4 + len;
		for (uint32_t k = 0; k < e.nex; k++)
			w[k] = e.val[k] - 256;
		len = (uint32_t) oracle_svb32_encode(w, e.nex, p + 4);
		put_u32(p, len); p += 4 + len;
		free(w);
	} else if (e.nex == 1) {		/* ex_zd.c:146-153: raw u32, raw u32 */
		put_u32(p, e.pos[0]); p += 4;
		put_u32(p, e.val[0] - 256); p += 4;
	}
	p = emit_low_bytes(zd + 1, n - 1, p);
	uint64_t len = (uint64_t) (p - tmp);
	int ret = 0;
	if (len > cap) {			/* some ASSERT(cap_out - offset >= ..) fired */
		ret = -1;
	} else {
		memcpy(out, tmp, len);		/* press.c:8475 */
		*nout = len;
	}
	free_exceptions(&e);
	free(tmp); free(s); free(zd);
	return ret;
}

static int ex_zd_depress(const uint8_t *in, uint64_t nbytes, uint64_t ncap,
			 int16_t *out, uint64_t *nout)
{
	if (nbytes < 16 || in[0] != 0)		/* ex_zd.c:501-512 version switch */
		return -1;
	uint64_t nhdr = get_u64(in + 1);	/* ex_zd.c:461-465 */
	uint8_t q = in[9];
	if (q > 5)				/* ex_zd.c:477 */
		return -1;
	uint64_t cap = nbytes;
	if (ncap > cap) cap = ncap;
	if (nhdr > cap) cap = nhdr;
	uint16_t *zd = malloc((cap + 1) * sizeof *zd);
	if (!zd)
		return -1;
	zd[0] = get_u16(in + 10);
	uint32_t nex = get_u32(in + 12);
	uint64_t off = 16;
	uint32_t *pos = malloc((nex ? nex : 1) * sizeof *pos);
	uint32_t *exv = malloc((nex ? nex : 1) * sizeof *exv);
	int ret = 0;
	if (nex > 1) {
		uint32_t len = get_u32(in + off); off += 4;
		if (oracle_svb32_decode(in + off, pos, nex) != len)	/* ex_zd.c:222-232 */
			ret = -1;
		off += len;
		gaps_to_pos(pos, nex);
		len = get_u32(in + off); off += 4;
		if (oracle_svb32_decode(in + off, exv, nex) != len)	/* ex_zd.c:252-262 */
			ret = -1;
		off += len;
	} else if (nex == 1) {
		pos[0] = get_u32(in + off); off += 4;
		exv[0] = get_u32(in + off); off += 4;
	}
	if (ret == 0) {
		for (uint32_t k = 0; k < nex; k++)
			zd[1 + pos[k]] = (uint16_t) (exv[k] + 256);
		uint32_t m = merge_low_bytes(in + off, nbytes - off, pos, nex, zd + 1);
		oracle_unzigdelta(zd, (uint64_t) m + 1, out);
		for (uint64_t i = 0; i < (uint64_t) m + 1; i++)	/* ex_zd.c:396-401 */
			out[i] = (int16_t) (uint16_t) ((uint16_t) out[i] << q);
		*nout = (uint64_t) m + 1;
	}
	free(pos); free(exv); free(zd);
	return ret;
}

/* ---------- a13: static Huffman table -------------------------------------- */

/* Code table parsed from the file image written by huffman.c:440-480:
 * [count u32 BE][dataBytes u32 BE] then count x [sym u8][numbits u8][code
 * bytes], code bit k stored at bit k%8 of byte k/8, bit 0 = root-most branch
 * (huffman.c:27-30 get_bit, :59-101 new_code + :32-53 reverse_bits).
 * The decode tree is the one read_code_table builds (huffman.c:549-694). */
static struct {
	int set;
	uint32_t code[256];	/* bit k = k-th emitted bit */
	uint8_t len[256];
	/* binary tree in arrays: child[node][bit]; >= 0 internal node id, < 0 leaf ~symbol */
	int32_t child[1024][2];
	int32_t nnodes;
} H;

int oracle_set_huffman_table(const uint8_t *file, size_t flen)
{
	if (flen < 8)
		return -1;
	uint32_t count = get_u32_be(file);
	if (count == 0 || count > 256)
		return -1;
	memset(&H, 0, sizeof H);
	H.nnodes = 1;
	H.child[0][0] = H.child[0][1] = 0;	/* 0 = absent (root is never a child) */
	size_t off = 8;
	for (uint32_t e = 0; e < count; e++) {
		if (off + 2 > flen)
			return -1;
		uint8_t sym = file[off], nb = file[off + 1];
		size_t nbytes = (nb + 7u) / 8u;
		off += 2;
		if (nb == 0 || nb > 32 || off + nbytes > flen)
			return -1;
		uint32_t code = 0;
		for (unsigned k = 0; k < nb; k++)
			code |= (uint32_t) ((file[off + k / 8] >> (k % 8)) & 1u) << k;
		off += nbytes;
		H.code[sym] = code;
		H.len[sym] = nb;
		int32_t node = 0;
		for (unsigned k = 0; k < nb; k++) {
			unsigned bit = (code >> k) & 1u;
			if (k == nb - 1u) {
				if (H.child[node][bit] != 0)
					return -1;	/* not a prefix code */
				H.child[node][bit] = ~(int32_t) sym;
			} else {
				int32_t nx = H.child[node][bit];
				if (nx < 0)
					return -1;
				if (nx == 0) {
					if (H.nnodes >= 1024)
						return -1;
					nx = H.nnodes++;
					H.child[nx][0] = H.child[nx][1] = 0;
					H.child[node][bit] = nx;
				}
				node = nx;
			}
		}
	}
	H.set = 1;
	return 0;
}

int oracle_get_huffman_codes(uint32_t code_le[256], uint8_t len[256])
{
	if (!H.set)
		return -1;
	memcpy(code_le, H.code, sizeof H.code);
	memcpy(len, H.len, sizeof H.len);
	return 0;
}

/* ---------- a14: shuffman_encode_memory (huffman.c:1184-1217, 848-887) ------ */
/* [count u32 big-endian] then the codes back to back, first bit of the stream
 * in bit 0 of byte 0; the last byte is zero padded. */
static uint64_t huff_encode(const uint8_t *sym, uint32_t m, uint8_t *out)
{
	put_u32_be(out, m);
	uint8_t *p = out + 4;
	uint64_t acc = 0;
	unsigned nacc = 0;
	for (uint32_t i = 0; i < m; i++) {
		acc |= (uint64_t) H.code[sym[i]] << nacc;
		nacc += H.len[sym[i]];
		while (nacc >= 8) {
			*p++ = (uint8_t) acc;
			acc >>= 8;
			nacc -= 8;
		}
	}
	if (nacc)
		*p++ = (uint8_t) acc;
	return (uint64_t) (p - out);
}

/* ---------- a15: shuffman_decode_memory (huffman.c:1219-1265, 696-711) ------ */
static int huff_decode(const uint8_t *in, uint64_t nbytes, uint8_t *sym, uint32_t *m)
{
	if (4 >= nbytes)			/* memread: readlen + index >= buflen fails */
		return 1;
	uint32_t want = get_u32_be(in), got = 0;
	int32_t node = 0;
	for (uint64_t i = 4; i < nbytes && got < want; i++) {
		for (unsigned b = 0; b < 8 && got < want; b++) {
			int32_t nx = H.child[node][(in[i] >> b) & 1u];
			if (nx < 0) {
				sym[got++] = (uint8_t) ~nx;
				node = 0;
			} else {
				node = nx;	/* nx == 0 (absent) cannot occur for a full tree */
			}
		}
	}
	*m = got;
	return 0;
}

/* ---------- a16: shuffman_vbe21_zd (press.c:4409-4498) ---------------------- */

static int shuff_vbe21_zd_press(const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (!H.set || n == 0)
		return 1;
	uint64_t cap = 8 + 8 * (uint64_t) n;
	uint8_t *vb = malloc(cap);
	uint64_t nvb = 0;
	if (!vb || vbe21_zd_press(in, n, vb, &nvb))
		return 1;
	uint32_t nex = get_u32(vb + 2);
	uint64_t hdr = 2 + 4 + 6 * (uint64_t) nex;	/* zd0, nex, pos[], val[] verbatim */
	memcpy(out, vb, hdr);
	*nout = hdr + huff_encode(vb + hdr, (uint32_t) (nvb - hdr), out + hdr);
	free(vb);
	return 0;
}

static int shuff_vbe21_zd_depress(const uint8_t *in, uint64_t nbytes, uint64_t n,
				  int16_t *out, uint64_t *nout)
{
	if (!H.set || nbytes < 6)
		return 1;
	uint32_t nex = get_u32(in + 2);
	uint64_t hdr = 2 + 4 + 6 * (uint64_t) nex;
	if (hdr > nbytes)
		return 1;
	uint32_t want = nbytes - hdr > 4 ? get_u32_be(in + hdr) : 0;
	uint8_t *vb = malloc(hdr + (uint64_t) want + 16);
	uint32_t m = 0;
	if (!vb)
		return 1;
	memcpy(vb, in, hdr);
	int ret = huff_decode(in + hdr, nbytes - hdr, vb + hdr, &m);
	if (ret == 0)
		ret = vbe21_zd_depress(vb, hdr + m, n, out, nout) ? 1 : 0;
	free(vb);
	return ret;
}

/* ---------- a18: vbbe21-zd (press.c:2775-2970, 3455-3500) ------------------- */

/* uint_get_minbits (press.c:461-471): smallest i with max < 2^i */
static uint8_t minbits_of(uint32_t max)
{
	uint8_t i = 0;
	while (i < 32 && ((uint64_t) max >> i) != 0)
		i++;
	return i;
}

/* uint_press_16 / uint_press_32 (press.c:489-507, 546-564 -> uintx_press_core 285-397): one byte
 * holding the width, then the values as `bits`-wide big-endian fields back to back, most
 * significant bit first, the last byte padded with zero bits; width 0 writes nothing more. */
static uint64_t bitpack_be(const uint32_t *v, uint32_t count, uint8_t bits, uint8_t *out)
{
	out[0] = bits;
	uint64_t nbytes = ((uint64_t) count * bits + 7) / 8;
	memset(out + 1, 0, nbytes);
	uint64_t bp = 0;
	for (uint32_t k = 0; k < count; k++)
		for (int b = bits - 1; b >= 0; b--, bp++)
			if ((v[k] >> b) & 1u)
				out[1 + bp / 8] |= (uint8_t) (0x80u >> (bp % 8));
	return 1 + nbytes;
}

static void bitunpack_be(const uint8_t *in, uint32_t count, uint32_t *v)
{
	uint8_t bits = in[0];
	uint64_t bp = 0;
	for (uint32_t k = 0; k < count; k++) {
		uint32_t x = 0;
		for (int b = 0; b < bits; b++, bp++)
			x = (x << 1) | ((in[1 + bp / 8] >> (7 - bp % 8)) & 1u);
		v[k] = x;
	}
}

static uint32_t max_u32(const uint32_t *v, uint32_t n)
{
	uint32_t m = 0;
	for (uint32_t k = 0; k < n; k++)
		m = v[k] > m ? v[k] : m;
	return m;
}

static int vbbe21_zd_press(const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (n == 0)
		return -1;
	uint16_t *zd = malloc(n * sizeof *zd);
	struct exlist e;
	if (!zd)
		return -1;
	oracle_zigdelta(in, n, zd);
	if (collect_exceptions(zd + 1, n - 1, &e))
		return -1;
	uint8_t *p = out;
	put_u16(p, zd[0]); p += 2;
	put_u32(p, e.nex); p += 4;
	if (e.nex > 1) {	/* press.c:2822-2866 */
		uint32_t *tmp = malloc(e.nex * sizeof *tmp);
		pos_to_gaps(e.pos, e.nex, tmp);
		uint32_t len = (uint32_t) bitpack_be(tmp, e.nex, minbits_of(max_u32(tmp, e.nex)), p + 4);
		put_u32(p, len); p += 4 + len;
		for (uint32_t k = 0; k < e.nex; k++)
			tmp[k] = (uint16_t) (e.val[k] - 256);
		len = (uint32_t) bitpack_be(tmp, e.nex, minbits_of(max_u32(tmp, e.nex)), p + 4);
		put_u32(p, len); p += 4 + len;
		free(tmp);
	} else if (e.nex == 1) {	/* press.c:2867-2872 */
		put_u32(p, e.pos[0]); p += 4;
		put_u16(p, (uint16_t) (e.val[0] - 256)); p += 2;
	}
	p = emit_low_bytes(zd + 1, n - 1, p);
	*nout = (uint64_t) (p - out);
	free_exceptions(&e);
	free(zd);
	return 0;
}

static int vbbe21_zd_depress(const uint8_t *in, uint64_t nbytes, uint64_t ncap,
			     int16_t *out, uint64_t *nout)
{
	uint64_t cap = (nbytes > ncap ? nbytes : ncap) + 1;
	uint16_t *zd = malloc(cap * sizeof *zd);
	if (!zd || nbytes < 6)
		return -1;
	zd[0] = get_u16(in);
	uint32_t nex = get_u32(in + 2);
	uint64_t off = 6;
	uint32_t *pos = malloc((nex ? nex : 1) * sizeof *pos);
	uint32_t *exv = malloc((nex ? nex : 1) * sizeof *exv);
	if (nex > 1) {		/* press.c:2912-2938 */
		uint32_t len = get_u32(in + off); off += 4;
		bitunpack_be(in + off, nex, pos); off += len;
		gaps_to_pos(pos, nex);
		len = get_u32(in + off); off += 4;
		bitunpack_be(in + off, nex, exv); off += len;
	} else if (nex == 1) {
		pos[0] = get_u32(in + off); off += 4;
		exv[0] = get_u16(in + off); off += 2;
	}
	for (uint32_t k = 0; k < nex; k++)
		zd[1 + pos[k]] = (uint16_t) (exv[k] + 256);
	uint32_t m = merge_low_bytes(in + off, nbytes - off, pos, nex, zd + 1);
	oracle_unzigdelta(zd, (uint64_t) m + 1, out);
	*nout = (uint64_t) m + 1;
	free(pos); free(exv); free(zd);
	return 0;
}

/* ---------- a17: shuffman_vbsse21_zd (press.c:4737-4845) -------------------- */

/* length of [nex][coded lists] at the head of a vbsse21 body (press.c:4766-4775, 4808-4817);
 * the reference keeps it in a uint16_t: once the coded lists pass 65535 bytes its press copies
 * a wrapped length and its depress looks for the second list length in the wrong place.  Those
 * reads are reported as failures here (return 1) instead of restating the wrap. */
static int vbsse21_exlen(const uint8_t *body, uint64_t nbody, uint64_t *exlen)
{
	if (nbody < 4)
		return 1;
	uint32_t nex = get_u32(body);
	uint64_t e = 4;
	if (nex > 1) {
		if (e + 4 > nbody)
			return 1;
		e += 4 + (uint64_t) get_u32(body + e);
		if (e + 4 > nbody)
			return 1;
		e += 4 + (uint64_t) get_u32(body + e);
	} else if (nex == 1) {
		e += 6;
	}
	if (e > nbody || e > 0xffff)
		return 1;
	*exlen = e;
	return 0;
}

/* a19: shuffman_vbbe21_zd (press.c:4505-4614) has the same frame over the vbbe21 body */
static int shuff_lists_zd_press(int bitpacked, const int16_t *in, uint32_t n, uint8_t *out, uint64_t *nout)
{
	if (!H.set || n == 0)
		return 1;
	uint64_t cap = 64 + 8 * (uint64_t) n;
	uint8_t *vb = malloc(cap);
	uint64_t nvb = 0, exlen = 0;
	if (!vb || (bitpacked ? vbbe21_zd_press(in, n, vb, &nvb) : vbsse21_zd_press(in, n, vb, &nvb)) ||
	    vbsse21_exlen(vb + 2, nvb - 2, &exlen)) {
		free(vb);
		return 1;
	}
	uint64_t hdr = 2 + exlen;	/* zd0, then the exception lists verbatim (press.c:4757, 4776) */
	memcpy(out, vb, hdr);
	*nout = hdr + huff_encode(vb + hdr, (uint32_t) (nvb - hdr), out + hdr);
	free(vb);
	return 0;
}

static int shuff_lists_zd_depress(int bitpacked, const uint8_t *in, uint64_t nbytes, uint64_t n,
				  int16_t *out, uint64_t *nout)
{
	uint64_t exlen = 0;
	if (!H.set || nbytes < 6 || vbsse21_exlen(in + 2, nbytes - 2, &exlen))
		return 1;
	uint64_t hdr = 2 + exlen;
	uint32_t want = nbytes - hdr > 4 ? get_u32_be(in + hdr) : 0;
	uint8_t *vb = malloc(hdr + (uint64_t) want + 16);
	uint32_t m = 0;
	if (!vb)
		return 1;
	memcpy(vb, in, hdr);
	int ret = huff_decode(in + hdr, nbytes - hdr, vb + hdr, &m);	/* press.c:4826-4829 */
	if (ret == 0)
		ret = (bitpacked ? vbbe21_zd_depress(vb, hdr + m, n, out, nout)
				 : vbsse21_zd_depress(vb, hdr + m, n, out, nout)) ? 1 : 0;
	free(vb);
	return ret;
}

/* ---------- svb0124-zd (press.c:1616-1643; streamvbyte_0124_encode.c:8-59) ------------------ */

/* keys[(n + 3) / 4] then data: value i has a 2-bit code at bit 2 (i & 3) of key byte i >> 2:
 * 0 = no byte (value 0), 1 = one byte, 2 = two bytes LE, 3 = four bytes LE; the values are zd of
 * all n samples widened to u32 (zigdelta_16_u32, trans.c:233-249: the first against 0). */
static int svb0124_zd_press(const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	uint64_t nk = (n + 3) / 4;
	uint8_t *data = out + nk;
	memset(out, 0, nk);
	int16_t prev = 0;
	for (uint64_t i = 0; i < n; i++) {
		uint32_t v = oracle_zigzag16((int16_t) (in[i] - prev));
		prev = in[i];
		uint8_t code;
		if (v == 0) {
			code = 0;
		} else if (v < 256) {
			*data++ = (uint8_t) v;
			code = 1;
		} else {	/* zd < 65536: the four-byte code never occurs */
			put_u16(data, (uint16_t) v);
			data += 2;
			code = 2;
		}
		out[i >> 2] |= (uint8_t) (code << (2 * (i & 3)));
	}
	*nout = (uint64_t) (data - out);
	return 0;
}

static int svb0124_zd_depress(const uint8_t *in, uint64_t n, int16_t *out, uint64_t *nout)
{
	const uint8_t *data = in + (n + 3) / 4;
	uint16_t acc = 0;
	for (uint64_t i = 0; i < n; i++) {	/* streamvbyte_0124_decode.c scalar form, then trans.c:272-282 */
		uint8_t code = (in[i >> 2] >> (2 * (i & 3))) & 3;
		uint32_t v = 0;
		if (code == 1) {
			v = *data++;
		} else if (code == 2) {
			v = get_u16(data);
			data += 2;
		} else if (code == 3) {
			v = get_u32(data);
			data += 4;
		}
		/* trans.c:272-282 hands the uint32 value to unzigzag_one_16(uint16_t), trans.c:80-83: a four-byte
		 * value counts by its low half (a 16-bit signal never makes the encoder write one) */
		const uint16_t z = (uint16_t) v;
		acc = (uint16_t) (acc + (uint16_t) ((z >> 1) ^ (uint16_t) (0u - (z & 1u))));
		out[i] = (int16_t) acc;
	}
	*nout = n;
	return 0;
}

/* ---------- uint-zd (press.c:639-687) ----------------------------------------------------- */

/* [first sample i16 LE] then uint_press_16 of zd[1 .. n - 1] (zigdelta_16, trans.c:199-213) at the
 * width of their maximum: [width u8][fields] (bitpack_be above). */
static int uint_zd_press(const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	if (n == 0)
		return -1;
	uint32_t m = (uint32_t) (n - 1);
	uint32_t *v = malloc((m ? m : 1) * sizeof *v);
	for (uint32_t i = 0; i < m; i++)
		v[i] = oracle_zigzag16((int16_t) (in[i + 1] - in[i]));
	put_u16(out, (uint16_t) in[0]);
	*nout = 2 + bitpack_be(v, m, minbits_of(max_u32(v, m)), out + 2);
	free(v);
	return 0;
}

static int uint_zd_depress(const uint8_t *in, uint64_t n, int16_t *out, uint64_t *nout)
{
	if (n == 0 || in[2] > 16)
		return -1;
	uint32_t m = (uint32_t) (n - 1);
	uint32_t *v = malloc((m ? m : 1) * sizeof *v);
	bitunpack_be(in + 2, m, v);
	uint16_t acc = get_u16(in);
	out[0] = (int16_t) acc;
	for (uint32_t i = 0; i < m; i++) {	/* trans.c:251-258 */
		acc = (uint16_t) (acc + (uint16_t) ((v[i] >> 1) ^ (0u - (v[i] & 1u))));
		out[i + 1] = (int16_t) acc;
	}
	free(v);
	*nout = n;
	return 0;
}

/* ---------- uniform adapter ------------------------------------------------ */

uint64_t oracle_bound(int method, uint64_t n)
{
	switch (method) {
	case ORC_SVB16_ZD: return svb16_bound(n);			/* press.c:1678-1681 */
	case ORC_SVB_ZD: return svb32_bound(n);				/* press.c:1585-1588 */
	case ORC_VBZ_SVB16_ZD: return 4 + svb16_bound(n);		/* press.c:2036 scratch */
	case ORC_VBZ_SVB_ZD: return 4 + svb32_bound(n);			/* press.c:1876 scratch */
	case ORC_VBE21_ZD:						/* press.c:3411-3414 */
	case ORC_VBSSE21_ZD:						/* press.c:3543-3546 */
	case ORC_SHUFF_VBE21_ZD:					/* press.c:4409-4412 -> 3965 */
	case ORC_SHUFF_VBSSE21_ZD:					/* press.c:4737-4740 -> 4293 */
	case ORC_VBBE21_ZD:						/* press.c:3455-3458 */
	case ORC_SHUFF_VBBE21_ZD:					/* press.c:4505-4508 */
		return 2 + vb1e2_bound((uint32_t) (n - 1));
	case ORC_EX_ZD:							/* press.c:8461-8464 */
		return svb32_bound(2 + vb1e2_bound((uint32_t) (n - 1)));
	case ORC_SVB0124_ZD: return svb32_bound(n);			/* press.c:1616-1619, 1531-1535 */
	case ORC_UINT_ZD: return 3 + 2 * n;				/* press.c:654-657 at width 16 */
	}
	return 0;
}

int oracle_press(int method, const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout)
{
	switch (method) {
	case ORC_SVB16_ZD: return svb16_zd_press(in, n, out, nout);
	case ORC_SVB_ZD: return svb32_zd_press(in, n, out, nout);
	case ORC_VBZ_SVB16_ZD:	/* press.c:2035-2039: [n u32 LE] then the svb16-zd stream */
	case ORC_VBZ_SVB_ZD: {	/* press.c:1875-1879 */
		put_u32(out, (uint32_t) n);
		int r = method == ORC_VBZ_SVB16_ZD ? svb16_zd_press(in, n, out + 4, nout)
						   : svb32_zd_press(in, n, out + 4, nout);
		*nout += 4;
		return r;
	}
	case ORC_VBE21_ZD: return vbe21_zd_press(in, (uint32_t) n, out, nout);
	case ORC_VBSSE21_ZD: return vbsse21_zd_press(in, (uint32_t) n, out, nout);
	case ORC_EX_ZD: return ex_zd_press(in, (uint32_t) n, out, nout);
	case ORC_SHUFF_VBE21_ZD: return shuff_vbe21_zd_press(in, (uint32_t) n, out, nout);
	case ORC_SHUFF_VBSSE21_ZD: return shuff_lists_zd_press(0, in, (uint32_t) n, out, nout);
	case ORC_VBBE21_ZD: return vbbe21_zd_press(in, (uint32_t) n, out, nout);
	case ORC_SHUFF_VBBE21_ZD: return shuff_lists_zd_press(1, in, (uint32_t) n, out, nout);
	case ORC_SVB0124_ZD: return svb0124_zd_press(in, n, out, nout);
	case ORC_UINT_ZD: return uint_zd_press(in, n, out, nout);
	}
	return -1;
}

int oracle_depress(int method, const uint8_t *in, uint64_t nbytes, uint64_t n,
		   int16_t *out, uint64_t *nout)
{
	switch (method) {
	case ORC_SVB16_ZD: return svb16_zd_depress(in, n, out, nout);
	case ORC_SVB_ZD: return svb32_zd_depress(in, n, out, nout);
	case ORC_VBZ_SVB16_ZD:	/* press.c:2062-2065: count comes from the prefix */
		return svb16_zd_depress(in + 4, get_u32(in), out, nout);
	case ORC_VBZ_SVB_ZD:	/* press.c:1902-1906 */
		return svb32_zd_depress(in + 4, get_u32(in), out, nout);
	case ORC_VBE21_ZD: return vbe21_zd_depress(in, nbytes, n, out, nout);
	case ORC_VBSSE21_ZD: return vbsse21_zd_depress(in, nbytes, n, out, nout);
	case ORC_EX_ZD: return ex_zd_depress(in, nbytes, n, out, nout);
	case ORC_SHUFF_VBE21_ZD: return shuff_vbe21_zd_depress(in, nbytes, n, out, nout);
	case ORC_SHUFF_VBSSE21_ZD: return shuff_lists_zd_depress(0, in, nbytes, n, out, nout);
	case ORC_VBBE21_ZD: return vbbe21_zd_depress(in, nbytes, n, out, nout);
	case ORC_SHUFF_VBBE21_ZD: return shuff_lists_zd_depress(1, in, nbytes, n, out, nout);
	case ORC_SVB0124_ZD: return svb0124_zd_depress(in, n, out, nout);
	case ORC_UINT_ZD: return uint_zd_depress(in, n, out, nout);
	}
	return -1;
}
