/*
 * ref_harness.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin helper layer compiled against the reference's own headers where they lie
 * under /root/reference and linked with the reference's GDK built by
 * oracle/Makefile into oracle/_ref/libbat_ref.so.  It lets the Python tests and
 * bench.py's CPU-baseline leg create in-memory BATs from raw arrays, inspect
 * result BATs, and drive the reference's real entry points (BATselect,
 * BATproject, BATgroup, BATgroupsum, BATjoin, GDKanalytical*, ...) which are
 * called directly through ctypes with the BAT* handles made here.
 *
 * Initialisation follows the embedded in-memory start-up of the reference
 * (tools/monetdbe/monetdbe.c:646,742; gdk_utils.c:933).
 */
#include "monetdb_config.h"
#include "gdk.h"
#include "gdk_analytic.h"
#include "monet_options.h"
#include <pthread.h>

static int refh_ready = 0;

int
refh_init(int nthreads)
{
	if (refh_ready)
		return 0;
	opt *set = NULL;
	int setlen = mo_builtin_settings(&set);
	if (setlen == 0)
		return -1;
	if (nthreads > 0) {
		char buf[32];
		snprintf(buf, sizeof(buf), "%d", nthreads);
		setlen = mo_add_option(&set, setlen, opt_cmdline, "gdk_nr_threads", buf);
	}
	if (BBPaddfarm(NULL, (1U << PERSISTENT) | (1U << TRANSIENT), false) != GDK_SUCCEED)
		return -2;
	if (GDKinit(set, setlen, true, NULL) != GDK_SUCCEED)
		return -3;
	/* GDKerror only records its text when the thread has an error buffer (gdk_tracer.c:586-597) */
	if (GDKerrbuf == NULL)
		GDKsetbuf(calloc(1, GDKMAXERRLEN));
	refh_ready = 1;
	return 0;
}

/* a worker thread that wants to call GDK must be known to it (for GDKerrbuf etc.) */
int
refh_thread_attach(const char *name)
{
	if (!MT_thread_register())
		return -1;
	if (GDKerrbuf == NULL)
		GDKsetbuf(calloc(1, GDKMAXERRLEN));
	return 0;
}

void
refh_thread_detach(void)
{
	MT_thread_deregister();
}

const char *
refh_errbuf(void)
{
	char *e = GDKerrbuf;
	return e ? e : "";
}

void
refh_clrerr(void)
{
	GDKclrerr();
}

BAT *
refh_bat_new(int type, uint64_t count, uint64_t hseq, const void *data,
	     int sorted, int revsorted, int key, int nonil, int nil)
{
	BAT *b = COLnew((oid) hseq, type, (BUN) count, TRANSIENT);
	if (b == NULL)
		return NULL;
	if (count > 0 && data != NULL)
		memcpy(Tloc(b, 0), data, (size_t) count << b->tshift);
	BATsetcount(b, (BUN) count);
	b->tsorted = sorted != 0;
	b->trevsorted = revsorted != 0;
	b->tkey = key != 0;
	b->tnonil = nonil != 0;
	b->tnil = nil != 0;
	b->tseqbase = oid_nil;
	if (count <= 1) {
		/* BATsetcount has already set the trivial properties */
		b->tsorted = b->trevsorted = true;
		b->tkey = true;
	}
	return b;
}

/* a str BAT: `n` NUL-terminated strings packed back to back in `packed` ("\200" = nil), appended one by one so
 * the heap, its duplicate elimination and the offset width are whatever GDK makes of them */
BAT *
refh_str_new(uint64_t hseq, uint64_t n, const char *packed)
{
	BAT *b = COLnew((oid) hseq, TYPE_str, (BUN) n, TRANSIENT);
	if (b == NULL)
		return NULL;
	for (uint64_t i = 0; i < n; i++) {
		if (BUNappend(b, packed, false) != GDK_SUCCEED) {
			BBPreclaim(b);
			return NULL;
		}
		packed += strlen(packed) + 1;
	}
	return b;
}

/* how GDK stores it: offset width, used bytes of the string heap, whether the heap is duplicate-free */
void
refh_str_parts(BAT *b, int *width, uint64_t *heapfree, int *elim)
{
	*width = b->twidth;
	*heapfree = b->tvheap->free;
	*elim = GDK_ELIMDOUBLES(b->tvheap);
}

void
refh_str_heap(BAT *b, void *dst)
{
	memcpy(dst, b->tvheap->base, b->tvheap->free);
}

/* the decoders of binary COPY INTO (sql/backends/monet5/sql_bincopyconvert.c:168-280) are loops around the GDK
 * constructors below; sql/ itself cannot be compiled here (it needs the MAL / SQL headers), so the harness runs the
 * same loops over the real date_create / daytime_create / timestamp_create of libbat */
#include "gdk_time.h"
void
refh_bincopy_date(int32_t *dst, const uint8_t *src, uint64_t n)
{
	for (uint64_t i = 0; i < n; i++, src += 4) {
		int16_t year;
		memcpy(&year, src + 2, 2);
		dst[i] = date_create(year, src[1], src[0]);
	}
}

void
refh_bincopy_time(int64_t *dst, const uint8_t *src, uint64_t n)
{
	for (uint64_t i = 0; i < n; i++, src += 8) {
		uint32_t ms;
		memcpy(&ms, src, 4);
		dst[i] = daytime_create(src[6], src[5], src[4], ms);
	}
}

void
refh_bincopy_timestamp(int64_t *dst, const uint8_t *src, uint64_t n)
{
	for (uint64_t i = 0; i < n; i++, src += 12) {
		uint32_t ms;
		int16_t year;
		memcpy(&ms, src, 4);
		memcpy(&year, src + 10, 2);
		dst[i] = timestamp_create(date_create(year, src[9], src[8]), daytime_create(src[6], src[5], src[4], ms));
	}
}

/* encode_date / encode_time / encode_timestamp (:186-330) with the real accessors */
void
refh_bincopy_encode(int kind, uint8_t *dst, const void *src, uint64_t n)
{
	for (uint64_t i = 0; i < n; i++) {
		date dt = date_nil;
		daytime tm = daytime_nil;
		bool nil;
		if (kind == 0) {
			dt = ((const date *) src)[i];
			nil = is_date_nil(dt);
		} else if (kind == 1) {
			tm = ((const daytime *) src)[i];
			nil = is_daytime_nil(tm);
		} else {
			timestamp ts = ((const timestamp *) src)[i];
			nil = is_timestamp_nil(ts);
			if (!nil) {
				dt = timestamp_date(ts);
				tm = timestamp_daytime(ts);
			}
		}
		const int rs = kind == 0 ? 4 : kind == 1 ? 8 : 12;
		uint8_t *r = dst + i * rs;
		if (nil) {
			memset(r, 0xff, rs);
			continue;
		}
		if (kind != 0) {
			uint32_t ms = daytime_usec(tm);
			memcpy(r, &ms, 4);
			r[4] = daytime_sec(tm);
			r[5] = daytime_min(tm);
			r[6] = daytime_hour(tm);
			r[7] = 0;
		}
		if (kind != 1) {
			uint8_t *d = r + (kind == 0 ? 0 : 8);
			int16_t year = date_year(dt);
			d[0] = date_day(dt);
			d[1] = date_month(dt);
			memcpy(d + 2, &year, 2);
		}
	}
}

/* a msk BAT from uint32 words holding `nbits` bits */
BAT *
refh_msk_new(uint64_t hseq, uint64_t nbits, const uint32_t *words)
{
	BAT *b = COLnew((oid) hseq, TYPE_msk, (BUN) nbits, TRANSIENT);
	if (b == NULL)
		return NULL;
	memcpy(Tloc(b, 0), words, ((nbits + 31) / 32) * 4);
	BATsetcount(b, (BUN) nbits);
	return b;
}

BAT *
refh_dense(uint64_t hseq, uint64_t tseq, uint64_t cnt)
{
	return BATdense((oid) hseq, (oid) tseq, (BUN) cnt);
}

void *
refh_tloc(BAT *b)
{
	return b->ttype == TYPE_void ? NULL : Tloc(b, 0);
}

struct refh_info {
	int32_t i_type;
	int32_t i_ccand;		/* 0 none, 1 negative oids, 2 mask */
	uint64_t i_hseq;
	uint64_t i_count;
	uint64_t i_tseq;
	uint8_t p_sorted, p_revsorted, p_key, p_nonil, p_nil, pad[3];
};

void
refh_info(BAT *b, struct refh_info *out)
{
	out->i_type = b->ttype;
	out->i_ccand = 0;
	if (b->ttype == TYPE_void && b->tvheap != NULL)
		out->i_ccand = ((ccand_t *) b->tvheap->base)->type == CAND_NEGOID ? 1 : 2;
	out->i_hseq = b->hseqbase;
	out->i_count = BATcount(b);
	out->i_tseq = b->tseqbase;
	out->p_sorted = b->tsorted;
	out->p_revsorted = b->trevsorted;
	out->p_key = b->tkey;
	out->p_nonil = b->tnonil;
	out->p_nil = b->tnil;
}

int
refh_width(BAT *b)
{
	return b->ttype == TYPE_void ? 8 : b->twidth;
}

/* logical tail: void / candidate lists are materialised as oids */
int
refh_tail(BAT *b, void *dst)
{
	BUN n = BATcount(b);
	if (b->ttype == TYPE_void) {
		oid *o = dst;
		if (b->tvheap != NULL || !is_oid_nil(b->tseqbase)) {
			if (b->tvheap != NULL) {
				struct canditer ci;
				canditer_init(&ci, NULL, b);
				for (BUN i = 0; i < ci.ncand; i++)
					o[i] = canditer_next(&ci);
			} else {
				for (BUN i = 0; i < n; i++)
					o[i] = b->tseqbase + i;
			}
		} else {
			for (BUN i = 0; i < n; i++)
				o[i] = oid_nil;
		}
		return 0;
	}
	if (b->ttype == TYPE_msk) {
		memcpy(dst, Tloc(b, 0), ((n + 31) / 32) * 4);
		return 0;
	}
	memcpy(dst, Tloc(b, 0), (size_t) n << b->tshift);
	return 0;
}

void
refh_free(BAT *b)
{
	BBPreclaim(b);
}

/* run the reference's own property checker on a result (gdk_bat.c:2699) */
void
refh_assert_props(BAT *b)
{
	BATassertProps(b);
}

/* --- bounded multi-core CPU baseline (mitosis-style slicing, opt_mitosis.c:188-214):
 * the caller supplies a per-slice worker; threads are registered with GDK. --- */
typedef void (*refh_slice_fn)(void *arg, int slice);
struct refh_job { refh_slice_fn fn; void *arg; int next, nslices; pthread_mutex_t mu; };

static void *
refh_worker(void *p)
{
	struct refh_job *j = p;
	MT_thread_register();
	for (;;) {
		pthread_mutex_lock(&j->mu);
		int s = j->next < j->nslices ? j->next++ : -1;
		pthread_mutex_unlock(&j->mu);
		if (s < 0)
			break;
		j->fn(j->arg, s);
	}
	MT_thread_deregister();
	return NULL;
}

int
refh_parallel(refh_slice_fn fn, void *arg, int nslices, int nthreads)
{
	struct refh_job j = { fn, arg, 0, nslices, PTHREAD_MUTEX_INITIALIZER };
	pthread_t th[256];
	if (nthreads > 256)
		nthreads = 256;
	for (int i = 0; i < nthreads; i++)
		if (pthread_create(&th[i], NULL, refh_worker, &j) != 0)
			return -1;
	for (int i = 0; i < nthreads; i++)
		pthread_join(th[i], NULL);
	return 0;
}

/* --- constant-operand calc: the Python side passes (pointer, type id) and we build
 * the ValRecord the reference wants (VALinit, gdk_value.c). op: 0 add 1 sub 2 mul 3 div 4 mod;
 * side: 0 = BAT op cst, 1 = cst op BAT --- */
BAT *
refh_calc_cst(int op, int side, BAT *b, const void *v, int vtype, BAT *s, int tp)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	switch (op * 2 + side) {
	case 0: bn = BATcalcaddcst(b, &val, s, tp); break;
	case 1: bn = BATcalccstadd(&val, b, s, tp); break;
	case 2: bn = BATcalcsubcst(b, &val, s, tp); break;
	case 3: bn = BATcalccstsub(&val, b, s, tp); break;
	case 4: bn = BATcalcmulcst(b, &val, s, tp); break;
	case 5: bn = BATcalccstmul(&val, b, s, tp); break;
	case 6: bn = BATcalcdivcst(b, &val, s, tp); break;
	case 7: bn = BATcalccstdiv(&val, b, s, tp); break;
	case 8: bn = BATcalcmodcst(b, &val, s, tp); break;
	case 9: bn = BATcalccstmod(&val, b, s, tp); break;
	}
	VALclear(&val);
	return bn;
}

/* op: 0 and 1 or 2 xor (BATcalc<op>cst, gdk_calc.c:2967,2736,2506) */
BAT *
refh_logic_cst(int op, BAT *b, const void *v, int vtype, BAT *s)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	switch (op) {
	case 0: bn = BATcalcandcst(b, &val, s); break;
	case 1: bn = BATcalcorcst(b, &val, s); break;
	case 2: bn = BATcalcxorcst(b, &val, s); break;
	}
	VALclear(&val);
	return bn;
}

/* op: 0 lsh 1 rsh; side 0: b OP v, side 1: v OP b (gdk_calc.c:3305,3356,3655,3706) */
BAT *
refh_shift_cst(int op, int side, BAT *b, const void *v, int vtype, BAT *s)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	if (op == 0)
		bn = side == 0 ? BATcalclshcst(b, &val, s) : BATcalccstlsh(&val, b, s);
	else
		bn = side == 0 ? BATcalcrshcst(b, &val, s) : BATcalccstrsh(&val, b, s);
	VALclear(&val);
	return bn;
}

/* which: 0 min 1 min_no_nil 2 max 3 max_no_nil (gdk_calc.c:1403,1563,2086,2219) */
BAT *
refh_minmax_cst(int which, BAT *b, const void *v, int vtype, BAT *s)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	switch (which) {
	case 0: bn = BATcalcmincst(b, &val, s); break;
	case 1: bn = BATcalcmincst_no_nil(b, &val, s); break;
	case 2: bn = BATcalcmaxcst(b, &val, s); break;
	case 3: bn = BATcalcmaxcst_no_nil(b, &val, s); break;
	}
	VALclear(&val);
	return bn;
}

/* cmp: 0 lt 1 le 2 gt 3 ge 4 eq 5 ne */
BAT *
refh_cmp_cst(int cmp, BAT *b, const void *v, int vtype, BAT *s, int nil_matches)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	switch (cmp) {
	case 0: bn = BATcalcltcst(b, &val, s); break;
	case 1: bn = BATcalclecst(b, &val, s); break;
	case 2: bn = BATcalcgtcst(b, &val, s); break;
	case 3: bn = BATcalcgecst(b, &val, s); break;
	case 4: bn = BATcalceqcst(b, &val, s, nil_matches); break;
	case 5: bn = BATcalcnecst(b, &val, s, nil_matches); break;
	}
	VALclear(&val);
	return bn;
}

/* constant on the left: BATcalccst<op>(v, b, s) */
BAT *
refh_cst_cmp(int cmp, const void *v, int vtype, BAT *b, BAT *s, int nil_matches)
{
	ValRecord val;
	if (VALinit(&val, vtype, v) == NULL)
		return NULL;
	BAT *bn = NULL;
	switch (cmp) {
	case 0: bn = BATcalccstlt(&val, b, s); break;
	case 1: bn = BATcalccstle(&val, b, s); break;
	case 2: bn = BATcalccstgt(&val, b, s); break;
	case 3: bn = BATcalccstge(&val, b, s); break;
	case 4: bn = BATcalccsteq(&val, b, s, nil_matches); break;
	case 5: bn = BATcalccstne(&val, b, s, nil_matches); break;
	}
	VALclear(&val);
	return bn;
}

/* hashjoin is what BATjoin picks for unsorted inputs; expose the algorithm string of the
 * last call so tests can assert which implementation the oracle exercised */
const char *
refh_algorithm(void)
{
	const char *a = MT_thread_getalgorithm();
	return a ? a : "";
}

/* timing helper for the CPU baseline: microseconds, monotonic */
int64_t
refh_usec(void)
{
	return (int64_t) GDKusec();
}
