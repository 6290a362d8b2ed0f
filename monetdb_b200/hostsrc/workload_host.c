/* workload_host.c — threaded host generator for the synthetic TPC-H-shaped columns of SURVEY.md §8(d).
 *
 * Same definition as monetdb_b200/workload.py (the numpy form the tests use as the golden one):
 * splitmix64, one stream per column, seed = 0x5EED0000 + column index, value = f(z of row i), so any row
 * range can be produced independently.  bench.py needs SF100 (600 M rows x 7 columns) on the host for the
 * end-to-end leg and for the reference arm; numpy takes minutes for that, this takes seconds.
 * Plain C, no CUDA, no GDK: built by monetdb_b200/build.py into libworkload.so and loaded by workload.py.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

enum { C_SHIPDATE, C_DISCOUNT, C_QUANTITY, C_EXTPRICE, C_TAX, C_RETFLAG };

static inline uint64_t smix(uint64_t seed, uint64_t row)
{
	uint64_t z = seed + (row + 1) * 0x9E3779B97F4A7C15ULL;
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
	return z ^ (z >> 31);
}
#define SEED(c) (0x5EED0000ULL + (c))

struct job {
	uint64_t lo, hi, base;
	int nils;
	const int32_t *dates;
	int ndates;
	int32_t cut_ls, cut_band;
	int32_t *shipdate;
	int64_t *discount, *quantity, *extprice, *tax;
	int8_t *returnflag, *linestatus;
};

static void *run(void *arg)
{
	struct job *j = (struct job *) arg;
	const int nils = j->nils;
	for (uint64_t r = j->lo; r < j->hi; r++) {
		const uint64_t o = r - j->base;
		if (j->shipdate || j->returnflag || j->linestatus) {
			const uint64_t z = smix(SEED(C_SHIPDATE), r);
			const int32_t ship = j->dates[z % (uint64_t) j->ndates];
			if (j->shipdate)
				j->shipdate[o] = (nils && z % 100 == 0) ? INT32_MIN : ship;
			if (j->returnflag || j->linestatus) {
				const int8_t ls = ship > j->cut_ls;
				int8_t rf;
				if (ls)
					rf = 1;
				else if (ship > j->cut_band)
					rf = 1;
				else
					rf = (smix(SEED(C_RETFLAG), r) % 2 == 1) ? 0 : 2;
				if (j->returnflag)
					j->returnflag[o] = rf;
				if (j->linestatus)
					j->linestatus[o] = ls;
			}
		}
		if (j->discount) {
			const uint64_t z = smix(SEED(C_DISCOUNT), r);
			j->discount[o] = (nils && z % 100 == 0) ? INT64_MIN : (int64_t) (z % 11);
		}
		if (j->quantity) {
			const uint64_t z = smix(SEED(C_QUANTITY), r);
			j->quantity[o] = (nils && z % 100 == 0) ? INT64_MIN : (int64_t) ((z % 50 + 1) * 100);
		}
		if (j->extprice) {
			const uint64_t z = smix(SEED(C_EXTPRICE), r);
			j->extprice[o] = (nils && z % 100 == 0) ? INT64_MIN : (int64_t) (z % 10400001ULL + 90000);
		}
		if (j->tax) {
			const uint64_t z = smix(SEED(C_TAX), r);
			j->tax[o] = (nils && z % 100 == 0) ? INT64_MIN : (int64_t) (z % 9);
		}
	}
	return NULL;
}

/* rows [lo,hi) of lineitem; any output pointer may be NULL (column not wanted).  nils != 0: the 1 %-nil
 * variant (a value is nil where its own z % 100 == 0; the two bte key columns never are). */
int wl_lineitem(uint64_t lo, uint64_t hi, int nthreads, int nils, const int32_t *dates, int ndates,
		int32_t cut_ls, int32_t cut_band, int32_t *shipdate, int64_t *discount, int64_t *quantity,
		int64_t *extprice, int64_t *tax, int8_t *returnflag, int8_t *linestatus)
{
	if (nthreads < 1)
		nthreads = 1;
	if (nthreads > 256)
		nthreads = 256;
	pthread_t th[256];
	struct job jobs[256];
	const uint64_t n = hi - lo;
	int started = 0;
	for (int t = 0; t < nthreads; t++) {
		struct job *j = &jobs[t];
		j->lo = lo + n * (uint64_t) t / (uint64_t) nthreads;
		j->hi = lo + n * (uint64_t) (t + 1) / (uint64_t) nthreads;
		j->base = lo;
		j->nils = nils;
		j->dates = dates;
		j->ndates = ndates;
		j->cut_ls = cut_ls;
		j->cut_band = cut_band;
		j->shipdate = shipdate;
		j->discount = discount;
		j->quantity = quantity;
		j->extprice = extprice;
		j->tax = tax;
		j->returnflag = returnflag;
		j->linestatus = linestatus;
		if (t == nthreads - 1 || pthread_create(&th[t], NULL, run, j) != 0) {
			/* the last share (or any share a thread could not be made for) runs here */
			if (t != nthreads - 1) {
				j->hi = hi;
				run(j);
				break;
			}
			run(j);
		} else {
			started++;
		}
	}
	for (int t = 0; t < started; t++)
		pthread_join(th[t], NULL);
	return 0;
}

/* ---- config 5: the window-analytics table (workload.py: window()).  Rows come in blocks of sum(lens) rows;
 * block k holds one partition of each length in the order of permutation z(k) % 6, so blocks are independent
 * and the threads take whole blocks.  pkey ascends, ord ascends inside a partition with duplicates
 * (0 at the partition start, then + z % 3), val = z % 10000 - 5000. */
enum { C_WPERM = 10, C_WORD = 11, C_WVAL = 12 };
static const int wperms[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};

struct wjob {
	uint64_t b0, b1, lo, hi, blk;
	uint64_t lens[3];
	int64_t *pkey, *ord, *val;
};

static void *wrun(void *arg)
{
	struct wjob *j = (struct wjob *) arg;
	for (uint64_t k = j->b0; k < j->b1; k++) {
		const int *perm = wperms[smix(SEED(C_WPERM), k) % 6];
		uint64_t row = k * j->blk;
		for (int p = 0; p < 3; p++) {
			const uint64_t len = j->lens[perm[p]];
			int64_t o = 0;
			for (uint64_t t = 0; t < len; t++, row++) {
				if (row < j->lo || row >= j->hi)
					continue;
				const uint64_t i = row - j->lo;
				if (t > 0)
					o += (int64_t) (smix(SEED(C_WORD), row) % 3);
				j->pkey[i] = (int64_t) (3 * k + (uint64_t) p);
				j->ord[i] = o;
				j->val[i] = (int64_t) (smix(SEED(C_WVAL), row) % 10000) - 5000;
			}
		}
	}
	return NULL;
}

int wl_window(uint64_t lo, uint64_t hi, int nthreads, uint64_t len0, uint64_t len1, uint64_t len2, int64_t *pkey, int64_t *ord,
	      int64_t *val)
{
	const uint64_t blk = len0 + len1 + len2;
	if (blk == 0 || lo % blk != 0)
		return -1;
	const uint64_t b0 = lo / blk, b1 = (hi + blk - 1) / blk, nb = b1 - b0;
	if (nthreads < 1)
		nthreads = 1;
	if (nthreads > 256)
		nthreads = 256;
	if ((uint64_t) nthreads > nb)
		nthreads = (int) (nb ? nb : 1);
	pthread_t th[256];
	struct wjob jobs[256];
	int started = 0;
	for (int t = 0; t < nthreads; t++) {
		struct wjob *j = &jobs[t];
		j->b0 = b0 + nb * (uint64_t) t / (uint64_t) nthreads;
		j->b1 = b0 + nb * (uint64_t) (t + 1) / (uint64_t) nthreads;
		j->lo = lo;
		j->hi = hi;
		j->blk = blk;
		j->lens[0] = len0;
		j->lens[1] = len1;
		j->lens[2] = len2;
		j->pkey = pkey;
		j->ord = ord;
		j->val = val;
		if (t == nthreads - 1 || pthread_create(&th[t], NULL, wrun, j) != 0) {
			if (t != nthreads - 1)
				j->b1 = b1;
			wrun(j);
			if (t != nthreads - 1)
				break;
		} else {
			started++;
		}
	}
	for (int t = 0; t < started; t++)
		pthread_join(th[t], NULL);
	return 0;
}
