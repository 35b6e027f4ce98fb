/* oracle/cpu_driver.c -- TEST INFRASTRUCTURE ONLY.
 *
 * The CPU baseline loop, shared by the restatement (liboracle.so, prefix
 * "oracle") and the compiled reference (_ref/libpress_ref.so, prefix "ref").
 * Per read it does what press/test.c:1756-1817 does for one method:
 * bound -> malloc -> press -> malloc -> depress -> compare -> free, except
 * that ALL samples are compared (test.c:1804 only checks the first half) and
 * that the wall clock (CLOCK_MONOTONIC) is used next to per-thread CPU time.
 * The reference has no threading (results.tex:33-36); reads are independent,
 * so "all host cores" = T threads over disjoint, sample-balanced read ranges.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(DRIVER_PREFIX, name)

uint64_t FN(bound)(int method, uint64_t n);
int FN(press)(int method, const int16_t *in, uint64_t n, uint8_t *out, uint64_t *nout);
int FN(depress)(int method, const uint8_t *in, uint64_t nbytes, uint64_t n,
		int16_t *out, uint64_t *nout);

struct job {
	int method;
	const int16_t *samples;
	const uint64_t *offsets;
	uint64_t r0, r1;
	pthread_barrier_t *bar;
	double t_press, t_depress;	/* this thread's wall time in each phase */
	uint64_t bytes, mismatches, errors;
};

static double now(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static void *worker(void *arg)
{
	struct job *j = arg;
	uint64_t nr = j->r1 - j->r0;
	uint8_t **streams = calloc(nr ? nr : 1, sizeof *streams);
	uint64_t *lens = calloc(nr ? nr : 1, sizeof *lens);

	pthread_barrier_wait(j->bar);
	double t0 = now();
	for (uint64_t r = j->r0; r < j->r1; r++) {
		uint64_t n = j->offsets[r + 1] - j->offsets[r];
		/* generous: the reference bounds are "practical" (press.c:2578) and a
		 * few methods write past them on exception-heavy reads */
		uint64_t cap = FN(bound)(j->method, n) + 8 * n + 64;
		uint8_t *out = malloc(cap);
		uint64_t nout = cap;
		if (!out || FN(press)(j->method, j->samples + j->offsets[r], n, out, &nout)) {
			j->errors++;
			nout = 0;
		}
		streams[r - j->r0] = out;
		lens[r - j->r0] = nout;
		j->bytes += nout;
	}
	j->t_press = now() - t0;

	pthread_barrier_wait(j->bar);
	t0 = now();
	for (uint64_t r = j->r0; r < j->r1; r++) {
		uint64_t n = j->offsets[r + 1] - j->offsets[r];
		int16_t *back = malloc((n + 64) * sizeof *back);
		uint64_t nback = n;
		if (!lens[r - j->r0])
			goto next;
		if (FN(depress)(j->method, streams[r - j->r0], lens[r - j->r0], n, back, &nback))
			j->errors++;
		else if (nback != n || memcmp(back, j->samples + j->offsets[r], n * sizeof *back))
			j->mismatches++;
next:
		free(back);
	}
	j->t_depress = now() - t0;
	pthread_barrier_wait(j->bar);

	for (uint64_t r = 0; r < nr; r++)
		free(streams[r]);
	free(streams);
	free(lens);
	return NULL;
}

/* res[0] = press wall seconds (slowest thread), res[1] = depress wall seconds,
 * res[2] = compressed bytes, res[3] = round-trip mismatches, res[4] = errors.
 * Returns 0 when every read round-tripped. */
int FN(run_batch)(int method, const int16_t *samples, const uint64_t *offsets,
		  uint64_t nreads, int nthreads, double res[5])
{
	if (nthreads < 1)
		nthreads = 1;
	if ((uint64_t) nthreads > nreads && nreads)
		nthreads = (int) nreads;
	pthread_t *tid = calloc(nthreads, sizeof *tid);
	struct job *jobs = calloc(nthreads, sizeof *jobs);
	pthread_barrier_t bar;
	pthread_barrier_init(&bar, NULL, nthreads);

	/* contiguous ranges balanced by sample count */
	uint64_t total = offsets[nreads] - offsets[0], r = 0;
	for (int t = 0; t < nthreads; t++) {
		uint64_t goal = offsets[0] + total / nthreads * (t + 1);
		jobs[t].r0 = r;
		if (t == nthreads - 1)
			r = nreads;
		else
			while (r < nreads && offsets[r + 1] <= goal)
				r++;
		jobs[t].r1 = r;
		jobs[t].method = method;
		jobs[t].samples = samples;
		jobs[t].offsets = offsets;
		jobs[t].bar = &bar;
		pthread_create(&tid[t], NULL, worker, &jobs[t]);
	}
	memset(res, 0, 5 * sizeof *res);
	for (int t = 0; t < nthreads; t++) {
		pthread_join(tid[t], NULL);
		if (jobs[t].t_press > res[0]) res[0] = jobs[t].t_press;
		if (jobs[t].t_depress > res[1]) res[1] = jobs[t].t_depress;
		res[2] += (double) jobs[t].bytes;
		res[3] += (double) jobs[t].mismatches;
		res[4] += (double) jobs[t].errors;
	}
	pthread_barrier_destroy(&bar);
	free(tid);
	free(jobs);
	return res[3] == 0 && res[4] == 0 ? 0 : -1;
}

/* ---- per-read fingerprints (FNV-1a-64, the fingerprint of SURVEY.md 8(c)) ----------------
 * hash_batch: every read of the batch pressed by this implementation, hashes[r] = FNV-1a-64
 * of its stream, sizes[r] = its length (0 and hash of nothing when press fails).
 * hash_slices: the same fingerprint of slices [oo[r], oo[r + 1]) of a container somebody
 * else wrote (the GPU).  Both split the reads over T threads by sample / byte count, so a
 * whole bench-scale chunk is compared byte for byte in seconds. */
static uint64_t fnv1a64(const uint8_t *p, uint64_t n)
{
	uint64_t h = 0xCBF29CE484222325ull;
	for (uint64_t i = 0; i < n; i++)
		h = (h ^ p[i]) * 0x100000001B3ull;
	return h;
}

struct hjob {
	int method;
	const int16_t *samples;
	const uint8_t *buf;
	const uint64_t *offsets;
	uint64_t r0, r1;
	uint64_t *hashes, *sizes;
};

static void *hash_press_worker(void *arg)
{
	struct hjob *j = arg;
	for (uint64_t r = j->r0; r < j->r1; r++) {
		uint64_t n = j->offsets[r + 1] - j->offsets[r];
		uint64_t cap = FN(bound)(j->method, n) + 8 * n + 64;
		uint8_t *out = malloc(cap);
		uint64_t nout = cap;
		if (!out || n == 0 || FN(press)(j->method, j->samples + j->offsets[r], n, out, &nout))
			nout = 0;
		j->hashes[r] = fnv1a64(out, nout);
		j->sizes[r] = nout;
		free(out);
	}
	return NULL;
}

static void *hash_slice_worker(void *arg)
{
	struct hjob *j = arg;
	for (uint64_t r = j->r0; r < j->r1; r++)
		j->hashes[r] = fnv1a64(j->buf + j->offsets[r], j->offsets[r + 1] - j->offsets[r]);
	return NULL;
}

static int hash_run(struct hjob proto, uint64_t nreads, int nthreads, void *(*fn)(void *))
{
	if (nthreads < 1)
		nthreads = 1;
	if ((uint64_t) nthreads > nreads && nreads)
		nthreads = (int) nreads;
	pthread_t *tid = calloc(nthreads, sizeof *tid);
	struct hjob *jobs = calloc(nthreads, sizeof *jobs);
	const uint64_t *o = proto.offsets;
	uint64_t total = o[nreads] - o[0], r = 0;
	for (int t = 0; t < nthreads; t++) {
		uint64_t goal = o[0] + total / nthreads * (t + 1);
		jobs[t] = proto;
		jobs[t].r0 = r;
		if (t == nthreads - 1)
			r = nreads;
		else
			while (r < nreads && o[r + 1] <= goal)
				r++;
		jobs[t].r1 = r;
		pthread_create(&tid[t], NULL, fn, &jobs[t]);
	}
	for (int t = 0; t < nthreads; t++)
		pthread_join(tid[t], NULL);
	free(tid);
	free(jobs);
	return 0;
}

int FN(hash_batch)(int method, const int16_t *samples, const uint64_t *offsets, uint64_t nreads,
		   int nthreads, uint64_t *hashes, uint64_t *sizes)
{
	struct hjob p = {method, samples, NULL, offsets, 0, 0, hashes, sizes};
	return hash_run(p, nreads, nthreads, hash_press_worker);
}

int FN(hash_slices)(const uint8_t *buf, const uint64_t *oo, uint64_t nreads, int nthreads,
		    uint64_t *hashes)
{
	struct hjob p = {0, NULL, buf, oo, 0, 0, hashes, NULL};
	return hash_run(p, nreads, nthreads, hash_slice_worker);
}
