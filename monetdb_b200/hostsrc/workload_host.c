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
