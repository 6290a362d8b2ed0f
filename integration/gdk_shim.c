/*
 * gdk_shim.c — the reference-side binding of the B200 GDK path (form B).
 *
 * A shared library that EXPORTS THE REFERENCE'S OWN ENTRY POINTS with their exact prototypes — this
 * file includes the reference headers (gdk.h, gdk_calc.h, gdk_analytic.h) where they lie, so the compiler
 * checks every definition below against the reference declaration (gdk.h:2245-2274, gdk.h:1447,
 * gdk_calc.h:36-147, gdk_analytic.h:21-42).  Loaded before libbat (LD_PRELOAD, or first in the link
 * order), it interposes those symbols: MAL modules, the interpreter and SQL call BATselect & co unchanged.
 *
 *   eligible call   -> BAT* arguments are converted to b200_bat descriptors (no copy of GDK structures,
 *                      only exported helpers: COLnew, BATdense, bat_iterator, BBP_desc, GDKtracer_log),
 *                      the operator runs in libgdk_b200.so (CUDA), the result becomes a new TRANSIENT
 *                      BAT owned by the caller;
 *   anything else   -> forwarded to the original (dlsym(RTLD_NEXT, ...)): variable-sized atoms, complex
 *                      candidate encodings, dense-candidate BATproject (a view in the reference, no
 *                      work), inputs below B200_SHIM_MIN_ROWS.
 *
 * The device heap tier as seen from GDK (row a23).  A result BAT's tail heap is page-locked host memory
 * owned by this library and attached with storage = STORE_NOWN (gdk.h:534-543; the zero-copy pattern of
 * the reference's own pyapi3, convert_loops.h:76-122), so every GDK reader can dereference Tloc() as
 * before.  Its DEVICE TWIN — the b200_bat the operator produced — stays resident and is registered under
 * the heap's base address: when a later call receives that BAT (or a GDK view of it) as input, the twin is
 * used and nothing crosses PCIe.  Because the base address belongs to this library until it releases it,
 * a hit can never be stale.  GDK gives a preloaded library no callback when a BAT dies, so the library
 * holds one reference of its own on the Heap record (HEAPincref, gdk.h:897) and sweeps: once that is the
 * only reference left, the BAT and every GDK view of it are gone — the reference is dropped (GDK frees the
 * record, not the memory: STORE_NOWN), the page-locked block and the device twin are released.
 * Inputs that were not produced here (base columns) are staged by the library on every call; with
 * B200_SHIM_TRUST_BASE=1 staged copies of BATs that are not writable (batRestricted != BAT_WRITE) are kept
 * until b200shim_forget() — the mode a patched libbat with a HEAPfree hook (form A) would make default.
 *
 * Errors: a failing device operator leaves the reference's message text in b200_errbuf(); it is recorded
 * with GDKtracer_log under the entry point's name and NULL / GDK_FAIL is returned (gdk.h:1710-1717).
 * Threading: every thread that calls in gets its own CUDA stream inside libgdk_b200; the twin table is
 * guarded by one mutex and entries in use by a running call are pinned.
 */
#include "monetdb_config.h"
/* gdk.h turns BAT field names into macros (#define ttype T.type, tkey T.key, ...) that would rewrite the
 * members of b200_bat: the C-ABI header and the accessors for its fields come first */
#include "gdk_b200.h"
static inline int d_type(const b200_bat *d) { return d->ttype; }
static inline uint64_t d_count(const b200_bat *d) { return d->count; }
static inline uint64_t d_hseq(const b200_bat *d) { return d->hseqbase; }
static inline uint64_t d_tseq(const b200_bat *d) { return d->tseqbase; }
static inline void *d_base(const b200_bat *d) { return d->base; }
static inline bool d_owns(const b200_bat *d) { return d->owns != 0; }
static inline bool d_sorted(const b200_bat *d) { return d->tsorted != 0; }
static inline bool d_revsorted(const b200_bat *d) { return d->trevsorted != 0; }
static inline bool d_key(const b200_bat *d) { return d->tkey != 0; }
static inline bool d_nonil(const b200_bat *d) { return d->tnonil != 0; }
static inline bool d_nil(const b200_bat *d) { return d->tnil != 0; }
static inline void d_oidrange(b200_bat *d, uint64_t first, uint64_t last)
{
	d->oidrange_known = 1;
	d->first_oid = first;
	d->last_oid = last;
}
#include "gdk.h"
#include "gdk_calc.h"
#include "gdk_analytic.h"
#include <dlfcn.h>
#include <pthread.h>

/* ------------------------------------------------------------------ originals */
#define REAL(name) real_##name
#define LOAD_REAL(name)								\
	do {									\
		if (REAL(name) == NULL)						\
			*(void **) &REAL(name) = dlsym(RTLD_NEXT, #name);	\
	} while (0)

static size_t min_rows = 0;
static int trust_base = 0, disabled = 0, verbose = 0;
static pthread_once_t once = PTHREAD_ONCE_INIT;
static _Atomic unsigned long n_device = 0, n_forward = 0, n_twin_hits = 0, n_staged = 0;

static void
shim_init(void)
{
	const char *e;
	if ((e = getenv("B200_SHIM_MIN_ROWS")) != NULL)
		min_rows = (size_t) strtoull(e, NULL, 10);
	if ((e = getenv("B200_SHIM_TRUST_BASE")) != NULL)
		trust_base = atoi(e);
	if ((e = getenv("B200_SHIM_DISABLE")) != NULL)
		disabled = atoi(e);
	if ((e = getenv("B200_SHIM_VERBOSE")) != NULL)
		verbose = atoi(e);
	if (!disabled && b200_device_count() < 1)
		disabled = 1;		/* no GPU: everything is forwarded to the reference */
	if (trust_base)
		b200_stage_cache(1);
}

/* counters for the tests / for tuning: calls run on the device, forwarded, inputs served by a twin, staged */
void
b200shim_stats(unsigned long out[4])
{
	out[0] = n_device;
	out[1] = n_forward;
	out[2] = n_twin_hits;
	out[3] = n_staged;
}

/* ------------------------------------------------------------------ pinned heap blocks (cached by size) */
typedef struct pblock {
	struct pblock *next;
	char *p;
	size_t size;
} pblock;
static pblock *pfree = NULL;
static pthread_mutex_t plock = PTHREAD_MUTEX_INITIALIZER;

static char *
pin_alloc(size_t bytes, size_t *got)
{
	size_t size = (bytes + 4095) & ~(size_t) 4095;
	if (size == 0)
		size = 4096;
	pthread_mutex_lock(&plock);
	for (pblock **pp = &pfree; *pp; pp = &(*pp)->next) {
		if ((*pp)->size >= size && (*pp)->size <= size + size / 4) {
			pblock *b = *pp;
			*pp = b->next;
			pthread_mutex_unlock(&plock);
			char *p = b->p;
			*got = b->size;
			free(b);
			return p;
		}
	}
	pthread_mutex_unlock(&plock);
	*got = size;
	return (char *) b200_pinned_alloc(size);
}

static void
pin_free(char *p, size_t size)
{
	pblock *b = malloc(sizeof(pblock));
	if (b == NULL) {
		b200_pinned_free(p);
		return;
	}
	b->p = p;
	b->size = size;
	pthread_mutex_lock(&plock);
	b->next = pfree;
	pfree = b;
	pthread_mutex_unlock(&plock);
}

/* ------------------------------------------------------------------ device twins, keyed by heap base */
typedef struct twin {
	struct twin *next;
	Heap *heap;		/* the Heap record; we hold one reference on it */
	char *hbase;		/* the NOWN heap (pinned host memory of ours) */
	size_t hsize;
	b200_bat *dev;		/* the device-resident column */
	int inuse;
} twin;
#define TWIN_BUCKETS 1024
static twin *twins[TWIN_BUCKETS];
static pthread_mutex_t tlock = PTHREAD_MUTEX_INITIALIZER;
static size_t ntwins = 0;

static inline unsigned
twin_bucket(const void *p)
{
	return (unsigned) (((uintptr_t) p >> 12) * 2654435761u) % TWIN_BUCKETS;
}

/* the top two bits of Heap.refs are flags (gdk_private.h:433-435) */
#define SHIM_HEAPREFS (((ATOMIC_BASE_TYPE) 1 << (sizeof(ATOMIC_BASE_TYPE) * 8 - 2)) - 1)

static bool
twin_alive(const twin *t)
{
	return (ATOMIC_GET(&t->heap->refs) & SHIM_HEAPREFS) > 1;	/* someone besides us still refers to the heap */
}

/* release the twins whose BAT is gone (GDK has no death callback for a preloaded library) */
static void
twin_sweep(void)
{
	twin *dead = NULL;
	pthread_mutex_lock(&tlock);
	for (unsigned i = 0; i < TWIN_BUCKETS; i++) {
		for (twin **pp = &twins[i]; *pp;) {
			twin *t = *pp;
			if (t->inuse == 0 && !twin_alive(t)) {
				*pp = t->next;
				t->next = dead;
				dead = t;
				ntwins--;
			} else {
				pp = &t->next;
			}
		}
	}
	pthread_mutex_unlock(&tlock);
	while (dead) {
		twin *t = dead;
		dead = t->next;
		HEAPdecref(t->heap, false);	/* the last reference: GDK frees the record, the memory is ours */
		if (t->dev)
			b200_BATfree(t->dev);
		pin_free(t->hbase, t->hsize);
		free(t);
	}
}

static twin *
twin_find(const char *hbase)
{
	pthread_mutex_lock(&tlock);
	twin *t = twins[twin_bucket(hbase)];
	while (t && t->hbase != hbase)
		t = t->next;
	if (t)
		t->inuse++;
	pthread_mutex_unlock(&tlock);
	return t;
}

static void
twin_release(twin *t)
{
	if (t) {
		pthread_mutex_lock(&tlock);
		t->inuse--;
		pthread_mutex_unlock(&tlock);
	}
}

/* the caller (or a test) says: drop every twin and every staged copy now */
void
b200shim_forget(void)
{
	twin_sweep();
	b200_stage_cache_clear();
}

/* ------------------------------------------------------------------ BAT* -> b200_bat */
typedef struct inarg {
	b200_bat *d;		/* descriptor handed to the device operator (NULL for a NULL BAT) */
	twin *t;		/* pinned twin, if the input was served by one */
	BATiter bi;
	bool iter;
} inarg;

static bool
type_ok(int tt)
{
	switch (tt) {
	case TYPE_void: case TYPE_bit: case TYPE_bte: case TYPE_sht: case TYPE_int: case TYPE_oid:
	case TYPE_flt: case TYPE_dbl: case TYPE_lng:
#ifdef HAVE_HGE
	case TYPE_hge:
#endif
	case TYPE_date: case TYPE_daytime: case TYPE_timestamp:
		return true;
	default:
		return false;
	}
}

/* may this BAT travel to the device at all?  (NULL is fine: optional arguments) */
static bool
bat_ok(BAT *b)
{
	if (b == NULL)
		return true;
	if (!type_ok(b->ttype) || b->tvheap != NULL)
		return false;		/* var-sized atoms; negative / mask candidate encodings live in tvheap */
	return true;
}

static bool
in_open(inarg *a, BAT *b)
{
	memset(a, 0, sizeof(*a));
	if (b == NULL)
		return true;
	if (b->ttype == TYPE_void || BATtdense(b)) {
		a->d = b200_BATdense(b->hseqbase, b->tseqbase, BATcount(b));
		return a->d != NULL;
	}
	a->bi = bat_iterator(b);	/* pins the heap for the duration of the call (gdk.h:1142-1210) */
	a->iter = true;
	const size_t cnt = a->bi.count;
	const char *tail = (const char *) a->bi.base;	/* = theap->base + (tbaseoff << tshift) */
	const char *hbase = a->bi.h ? a->bi.h->base : NULL;
	twin *t = hbase ? twin_find(hbase) : NULL;
	if (t != NULL && t->dev != NULL && tail >= t->hbase && (size_t) (tail - t->hbase) + (cnt << a->bi.shift) <= t->hsize &&
	    d_base(t->dev) != NULL) {
		/* the BAT (or a GDK view of it) was produced here: use the device-resident twin */
		a->t = t;
		a->d = b200_BATwrap((char *) d_base(t->dev) + (tail - t->hbase), a->bi.type, cnt, b->hseqbase);
		n_twin_hits++;
	} else {
		twin_release(t);
		a->d = b200_BAThostwrap(tail, a->bi.type, cnt, b->hseqbase);
		n_staged++;
	}
	if (a->d == NULL)
		return false;
	b200_BATsetprops(a->d, a->bi.sorted, a->bi.revsorted, a->bi.key, a->bi.nonil, a->bi.nil);
	if (a->bi.type == TYPE_oid && a->bi.sorted && cnt > 0) {
		/* a materialised candidate list: the library wants its oid range (it keeps it for its own lists) */
		const oid *o = (const oid *) tail;
		d_oidrange(a->d, o[0], o[cnt - 1]);
	}
	return true;
}

/* dictionary strings: a str tail travels as GDK stores it — offsets + string heap.  Only
 * b of BATselect / BATthetaselect (equality forms) and BATgroup (duplicate-free heap) may be one. */
static bool
str_ok(BAT *b)
{
	return b != NULL && b->ttype == TYPE_str && b->tvheap != NULL && b->tvheap->base != NULL;
}

static bool
in_open_str(inarg *a, BAT *b)
{
	memset(a, 0, sizeof(*a));
	a->bi = bat_iterator(b);
	a->iter = true;
	a->d = b200_BATstrfromhost(a->bi.base, a->bi.width, a->bi.count, b->hseqbase, a->bi.vh->base, a->bi.vhfree,
				   a->bi.vhfree < GDK_ELIMLIMIT);
	n_staged++;
	return a->d != NULL;
}

static inline bool
in_open_any(inarg *a, BAT *b)
{
	return str_ok(b) ? in_open_str(a, b) : in_open(a, b);
}

static void
in_close(inarg *a)
{
	if (a->d)
		b200_BATfree(a->d);	/* descriptors only: neither form owns memory */
	twin_release(a->t);
	if (a->iter)
		bat_iterator_end(&a->bi);
	memset(a, 0, sizeof(*a));
}

/* ------------------------------------------------------------------ b200_bat -> BAT* */
static void
set_props(BAT *bn, const b200_bat *r)
{
	bn->tsorted = d_sorted(r) || d_count(r) <= 1;
	bn->trevsorted = d_revsorted(r) || d_count(r) <= 1;
	bn->tkey = d_key(r) || d_count(r) <= 1;
	bn->tnonil = d_nonil(r);
	bn->tnil = d_nil(r);
}

/* takes r over (it becomes the twin, or is freed) */
static BAT *
out_bat(b200_bat *r)
{
	if (r == NULL)
		return NULL;
	if (d_type(r) == B200_void) {
		BAT *bn = BATdense(d_hseq(r), d_tseq(r), d_count(r));
		b200_BATfree(r);
		return bn;
	}
	const int tt = d_type(r);
	const size_t w = ATOMsize(tt), bytes = (size_t) d_count(r) * w;
	size_t hsize = 0;
	char *h = pin_alloc(bytes, &hsize);
	BAT *bn = h ? COLnew(d_hseq(r), TYPE_void, 0, TRANSIENT) : NULL;
	if (bn == NULL || (d_count(r) && b200_BATtohost(r, h) != 0)) {
		if (h)
			pin_free(h, hsize);
		if (bn)
			BBPreclaim(bn);
		else
			GDKerror("b200 shim: cannot create the result BAT\n");
		b200_BATfree(r);
		return NULL;
	}
	/* attach our page-locked heap, not owned by GDK (cf. pyapi3 CREATE_BAT_ZEROCOPY) */
	bn->ttype = tt;
	bn->twidth = (unsigned short) w;
	bn->tshift = ATOMelmshift(w);
	bn->tseqbase = oid_nil;
	assert(bn->theap->base == NULL);
	bn->theap->base = h;
	bn->theap->free = bytes;
	bn->theap->size = hsize;
	bn->theap->storage = STORE_NOWN;
	bn->theap->newstorage = STORE_MEM;
	bn->theap->dirty = true;
	bn->batCapacity = (BUN) (hsize / w);
	bn->batCount = (BUN) d_count(r);
	bn->batCopiedtodisk = false;
	set_props(bn, r);
	twin *t = malloc(sizeof(twin));
	if (t != NULL && (!d_owns(r) || d_base(r) == NULL)) {
		/* the device result is a view of someone else's memory: keep only the host copy as the twin's
		 * payload-less marker so the page-locked block is still released with the BAT */
		b200_BATfree(r);
		r = NULL;
	}
	if (t == NULL) {
		/* out of memory for the registry entry: the host copy is complete; the block is leaked to GDK's
		 * lifetime (never freed) rather than risking a dangling heap */
		b200_BATfree(r);
		return bn;
	}
	HEAPincref(bn->theap);
	t->heap = bn->theap;
	t->hbase = h;
	t->hsize = hsize;
	t->dev = r;
	t->inuse = 0;
	pthread_mutex_lock(&tlock);
	unsigned k = twin_bucket(h);
	t->next = twins[k];
	twins[k] = t;
	ntwins++;
	pthread_mutex_unlock(&tlock);
	return bn;
}

static void *
fail(const char *func)
{
	const char *m = b200_errbuf();
	GDKtracer_log(__FILE__, func, __LINE__, M_ERROR, GDK, NULL, "%s", (m && *m) ? m : "b200: device operator failed\n");
	return NULL;
}

static bool
enter(size_t rows)
{
	pthread_once(&once, shim_init);
	if (disabled || rows < min_rows)
		return false;
	static _Atomic unsigned calls = 0;
	if ((++calls & 15) == 0 || ntwins > 256)
		twin_sweep();
	if (!trust_base)
		b200_stage_cache_clear();
	return true;
}

#define FORWARD(name, ...)				\
	do {						\
		LOAD_REAL(name);			\
		n_forward++;				\
		return REAL(name)(__VA_ARGS__);		\
	} while (0)

/* ------------------------------------------------------------------ selection: gdk.h:2245-2246 */
static BAT *(*real_BATselect)(BAT *, BAT *, const void *, const void *, bool, bool, bool, bool);
static BAT *(*real_BATthetaselect)(BAT *, BAT *, const void *, const char *);

BAT *
BATselect(BAT *b, BAT *s, const void *tl, const void *th, bool li, bool hi, bool anti, bool nil_matches)
{
	const bool strsel = str_ok(b) && tl != NULL && li && hi && (th == NULL || strcmp((const char *) tl, (const char *) th) == 0);
	if (b == NULL || !(strsel || bat_ok(b)) || !bat_ok(s) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATselect, b, s, tl, th, li, hi, anti, nil_matches);
	inarg ab = {0}, as = {0};
	BAT *bn = NULL;
	if (in_open_any(&ab, b) && in_open(&as, s)) {
		b200_bat *r = b200_BATselect(ab.d, as.d, tl, th, li, hi, anti, nil_matches);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return bn;
}

BAT *
BATthetaselect(BAT *b, BAT *s, const void *val, const char *op)
{
	const bool strsel = str_ok(b) && val != NULL && op != NULL &&
		(strcmp(op, "=") == 0 || strcmp(op, "==") == 0 || strcmp(op, "!=") == 0 || strcmp(op, "<>") == 0 ||
		 strcmp(op, "eq") == 0 || strcmp(op, "ne") == 0);
	if (b == NULL || !(strsel || bat_ok(b)) || !bat_ok(s) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATthetaselect, b, s, val, op);
	inarg ab = {0}, as = {0};
	BAT *bn = NULL;
	if (in_open_any(&ab, b) && in_open(&as, s)) {
		b200_bat *r = b200_BATthetaselect(ab.d, as.d, val, op);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return bn;
}

/* ------------------------------------------------------------------ projection: gdk.h:2273-2274 */
static BAT *(*real_BATproject)(BAT *restrict, BAT *restrict);
static BAT *(*real_BATproject2)(BAT *restrict, BAT *restrict, BAT *restrict);

BAT *
BATproject2(BAT *restrict l, BAT *restrict r1, BAT *restrict r2)
{
	/* dense candidates: the reference returns a view of r (no data moves) — nothing to accelerate */
	if (l == NULL || r1 == NULL || !bat_ok(l) || !bat_ok(r1) || !bat_ok(r2) || BATtdense(l) || l->ttype == TYPE_void ||
	    r1->ttype == TYPE_void || (r2 && r2->ttype == TYPE_void) || !enter(BATcount(l)))
		FORWARD(BATproject2, l, r1, r2);
	inarg al = {0}, a1 = {0}, a2 = {0};
	BAT *bn = NULL;
	if (in_open(&al, l) && in_open(&a1, r1) && in_open(&a2, r2)) {
		b200_bat *r = b200_BATproject2(al.d, a1.d, a2.d);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&a2);
	in_close(&a1);
	in_close(&al);
	return bn;
}

BAT *
BATproject(BAT *restrict l, BAT *restrict r)
{
	return BATproject2(l, r, NULL);
}

/* ------------------------------------------------------------------ grouping: gdk.h:1447 */
static gdk_return (*real_BATgroup)(BAT **, BAT **, BAT **, BAT *, BAT *, BAT *, BAT *, BAT *);

gdk_return
BATgroup(BAT **groups, BAT **extents, BAT **histo, BAT *b, BAT *s, BAT *g, BAT *e, BAT *h)
{
	const bool strgrp = str_ok(b) && b->tvheap->free < GDK_ELIMLIMIT;	/* grouped by offset, gdk_group.c:897-919 */
	if (b == NULL || groups == NULL || !(strgrp || bat_ok(b)) || !bat_ok(s) || !bat_ok(g) || !bat_ok(e) || !bat_ok(h) ||
	    b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATgroup, groups, extents, histo, b, s, g, e, h);
	inarg ab = {0}, as = {0}, ag = {0}, ae = {0}, ah = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *rg = NULL, *re = NULL, *rh = NULL;
	*groups = NULL;
	if (extents)
		*extents = NULL;
	if (histo)
		*histo = NULL;
	if (in_open_any(&ab, b) && in_open(&as, s) && in_open(&ag, g) && in_open(&ae, e) && in_open(&ah, h)) {
		n_device++;
		if (b200_BATgroup(&rg, extents ? &re : NULL, histo ? &rh : NULL, ab.d, as.d, ag.d, ae.d, ah.d) == B200_SUCCEED) {
			*groups = out_bat(rg);
			if (extents)
				*extents = out_bat(re);
			if (histo)
				*histo = out_bat(rh);
			if (*groups && (!extents || *extents) && (!histo || *histo)) {
				rc = GDK_SUCCEED;
			} else {
				BBPreclaim(*groups);
				*groups = NULL;
				if (extents) {
					BBPreclaim(*extents);
					*extents = NULL;
				}
				if (histo) {
					BBPreclaim(*histo);
					*histo = NULL;
				}
			}
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ah);
	in_close(&ae);
	in_close(&ag);
	in_close(&as);
	in_close(&ab);
	return rc;
}

/* ------------------------------------------------------------------ aggregates: gdk_calc.h:127-147 */
#define GROUPAGGR(NAME)											\
static BAT *(*real_##NAME)(BAT *, BAT *, BAT *, BAT *, int, bool);					\
BAT *													\
NAME(BAT *b, BAT *g, BAT *e, BAT *s, int tp, bool skip_nils)						\
{													\
	if (b == NULL || g == NULL || !bat_ok(b) || !bat_ok(g) || !bat_ok(e) || !bat_ok(s) || !type_ok(tp) ||	\
	    b->ttype == TYPE_void || !enter(BATcount(b)))							\
		FORWARD(NAME, b, g, e, s, tp, skip_nils);						\
	inarg ab = {0}, ag = {0}, ae = {0}, as = {0};										\
	BAT *bn = NULL;											\
	if (in_open(&ab, b) && in_open(&ag, g) && in_open(&ae, e) && in_open(&as, s)) {			\
		b200_bat *r = b200_##NAME(ab.d, ag.d, ae.d, as.d, tp, skip_nils);			\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&as);											\
	in_close(&ae);											\
	in_close(&ag);											\
	in_close(&ab);											\
	return bn;											\
}
GROUPAGGR(BATgroupsum)
GROUPAGGR(BATgroupprod)
GROUPAGGR(BATgroupcount)
GROUPAGGR(BATgroupmin)
GROUPAGGR(BATgroupmax)

static gdk_return (*real_BATgroupavg)(BAT **, BAT **, BAT *, BAT *, BAT *, BAT *, int, bool, int);

gdk_return
BATgroupavg(BAT **bnp, BAT **cntsp, BAT *b, BAT *g, BAT *e, BAT *s, int tp, bool skip_nils, int scale)
{
	if (b == NULL || g == NULL || bnp == NULL || !bat_ok(b) || !bat_ok(g) || !bat_ok(e) || !bat_ok(s) || tp != TYPE_dbl ||
	    b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATgroupavg, bnp, cntsp, b, g, e, s, tp, skip_nils, scale);
	inarg ab = {0}, ag = {0}, ae = {0}, as = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *ra = NULL, *rc_ = NULL;
	*bnp = NULL;
	if (cntsp)
		*cntsp = NULL;
	if (in_open(&ab, b) && in_open(&ag, g) && in_open(&ae, e) && in_open(&as, s)) {
		n_device++;
		if (b200_BATgroupavg(&ra, cntsp ? &rc_ : NULL, ab.d, ag.d, ae.d, as.d, tp, skip_nils, scale) == B200_SUCCEED) {
			*bnp = out_bat(ra);
			if (cntsp)
				*cntsp = out_bat(rc_);
			if (*bnp && (!cntsp || *cntsp))
				rc = GDK_SUCCEED;
			else {
				BBPreclaim(*bnp);
				*bnp = NULL;
				if (cntsp) {
					BBPreclaim(*cntsp);
					*cntsp = NULL;
				}
			}
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ae);
	in_close(&ag);
	in_close(&ab);
	return rc;
}

static gdk_return (*real_BATsum)(void *, int, BAT *, BAT *, bool, bool);

gdk_return
BATsum(void *res, int tp, BAT *b, BAT *s, bool skip_nils, bool nil_if_empty)
{
	if (b == NULL || res == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(tp) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATsum, res, tp, b, s, skip_nils, nil_if_empty);
	inarg ab = {0}, as = {0};
	gdk_return rc = GDK_FAIL;
	if (in_open(&ab, b) && in_open(&as, s)) {
		n_device++;
		if (b200_BATsum(res, tp, ab.d, as.d, skip_nils, nil_if_empty) == B200_SUCCEED)
			rc = GDK_SUCCEED;
		else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return rc;
}

/* ------------------------------------------------------------------ calc: gdk_calc.h:36-47, compares :66-83 */
#define CALC_BINARY(OP)											\
static BAT *(*real_BATcalc##OP)(BAT *, BAT *, BAT *, BAT *, int);					\
BAT *													\
BATcalc##OP(BAT *b1, BAT *b2, BAT *s1, BAT *s2, int tp)							\
{													\
	if (b1 == NULL || b2 == NULL || !bat_ok(b1) || !bat_ok(b2) || !bat_ok(s1) || !bat_ok(s2) || !type_ok(tp) ||	\
	    b1->ttype == TYPE_void || b2->ttype == TYPE_void || !enter(BATcount(b1)))				\
		FORWARD(BATcalc##OP, b1, b2, s1, s2, tp);						\
	inarg a1 = {0}, a2 = {0}, c1 = {0}, c2 = {0};										\
	BAT *bn = NULL;											\
	if (in_open(&a1, b1) && in_open(&a2, b2) && in_open(&c1, s1) && in_open(&c2, s2)) {		\
		b200_bat *r = b200_BATcalc##OP(a1.d, a2.d, c1.d, c2.d, tp);				\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&c2);											\
	in_close(&c1);											\
	in_close(&a2);											\
	in_close(&a1);											\
	return bn;											\
}													\
static BAT *(*real_BATcalc##OP##cst)(BAT *, const ValRecord *, BAT *, int);				\
BAT *													\
BATcalc##OP##cst(BAT *b, const ValRecord *v, BAT *s, int tp)						\
{													\
	if (b == NULL || v == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(tp) || !type_ok(v->vtype) ||	\
	    v->vtype == TYPE_void || b->ttype == TYPE_void || !enter(BATcount(b)))				\
		FORWARD(BATcalc##OP##cst, b, v, s, tp);							\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;											\
	if (in_open(&ab, b) && in_open(&as, s)) {							\
		b200_bat *r = b200_BATcalc##OP##cst(ab.d, VALptr(v), v->vtype, as.d, tp);		\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&as);											\
	in_close(&ab);											\
	return bn;											\
}													\
static BAT *(*real_BATcalccst##OP)(const ValRecord *, BAT *, BAT *, int);				\
BAT *													\
BATcalccst##OP(const ValRecord *v, BAT *b, BAT *s, int tp)						\
{													\
	if (b == NULL || v == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(tp) || !type_ok(v->vtype) ||	\
	    v->vtype == TYPE_void || b->ttype == TYPE_void || !enter(BATcount(b)))				\
		FORWARD(BATcalccst##OP, v, b, s, tp);							\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;											\
	if (in_open(&ab, b) && in_open(&as, s)) {							\
		b200_bat *r = b200_BATcalccst##OP(VALptr(v), v->vtype, ab.d, as.d, tp);			\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&as);											\
	in_close(&ab);											\
	return bn;											\
}
CALC_BINARY(add)
CALC_BINARY(sub)
CALC_BINARY(mul)
CALC_BINARY(div)

#define CMP_BINARY(OP, EXTRA_DECL, EXTRA_ARG)								\
static BAT *(*real_BATcalc##OP)(BAT *, BAT *, BAT *, BAT * EXTRA_DECL);					\
BAT *													\
BATcalc##OP(BAT *b1, BAT *b2, BAT *s1, BAT *s2 EXTRA_DECL)						\
{													\
	if (b1 == NULL || b2 == NULL || !bat_ok(b1) || !bat_ok(b2) || !bat_ok(s1) || !bat_ok(s2) ||	\
	    b1->ttype == TYPE_void || b2->ttype == TYPE_void || !enter(BATcount(b1)))				\
		FORWARD(BATcalc##OP, b1, b2, s1, s2 EXTRA_ARG);						\
	inarg a1 = {0}, a2 = {0}, c1 = {0}, c2 = {0};										\
	BAT *bn = NULL;											\
	if (in_open(&a1, b1) && in_open(&a2, b2) && in_open(&c1, s1) && in_open(&c2, s2)) {		\
		b200_bat *r = b200_BATcalc##OP(a1.d, a2.d, c1.d, c2.d EXTRA_ARG);			\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&c2);											\
	in_close(&c1);											\
	in_close(&a2);											\
	in_close(&a1);											\
	return bn;											\
}													\
static BAT *(*real_BATcalc##OP##cst)(BAT *, const ValRecord *, BAT * EXTRA_DECL);			\
BAT *													\
BATcalc##OP##cst(BAT *b, const ValRecord *v, BAT *s EXTRA_DECL)						\
{													\
	if (b == NULL || v == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(v->vtype) || v->vtype == TYPE_void ||	\
	    b->ttype == TYPE_void || !enter(BATcount(b)))							\
		FORWARD(BATcalc##OP##cst, b, v, s EXTRA_ARG);						\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;											\
	if (in_open(&ab, b) && in_open(&as, s)) {							\
		b200_bat *r = b200_BATcalc##OP##cst(ab.d, VALptr(v), v->vtype, as.d EXTRA_ARG);		\
		bn = r ? out_bat(r) : fail(__func__);							\
		n_device++;										\
	} else												\
		fail(__func__);										\
	in_close(&as);											\
	in_close(&ab);											\
	return bn;											\
}
#define NOTHING
#define COMMA_NILM_DECL , bool nil_matches
#define COMMA_NILM_ARG , nil_matches
CMP_BINARY(lt, NOTHING, NOTHING)
CMP_BINARY(le, NOTHING, NOTHING)
CMP_BINARY(gt, NOTHING, NOTHING)
CMP_BINARY(ge, NOTHING, NOTHING)
CMP_BINARY(eq, COMMA_NILM_DECL, COMMA_NILM_ARG)
CMP_BINARY(ne, COMMA_NILM_DECL, COMMA_NILM_ARG)

CALC_BINARY(mod)

/* ------------------------------------------------------------------ the rest of the calc surface: gdk_calc.h:12-95 */
#define CALC_UNARY(OP)												\
static BAT *(*real_BATcalc##OP)(BAT *, BAT *);									\
BAT *														\
BATcalc##OP(BAT *b, BAT *s)											\
{														\
	if (b == NULL || !bat_ok(b) || !bat_ok(s) || b->ttype == TYPE_void || !enter(BATcount(b)))			\
		FORWARD(BATcalc##OP, b, s);										\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;													\
	if (in_open(&ab, b) && in_open(&as, s)) {									\
		b200_bat *r = b200_BATcalc##OP(ab.d, as.d);								\
		bn = r ? out_bat(r) : fail(__func__);									\
		n_device++;												\
	} else														\
		fail(__func__);												\
	in_close(&as);													\
	in_close(&ab);													\
	return bn;													\
}
CALC_UNARY(not)
CALC_UNARY(negate)
CALC_UNARY(absolute)
CALC_UNARY(iszero)
CALC_UNARY(sign)
CALC_UNARY(isnil)
CALC_UNARY(isnotnil)

/* binary operators without a result-type argument; BB / BC / CB = the names of the bat-bat, bat-cst, cst-bat forms */
#define CALC_NOTP(BB, BC, CB)											\
static BAT *(*real_##BB)(BAT *, BAT *, BAT *, BAT *);								\
BAT *														\
BB(BAT *b1, BAT *b2, BAT *s1, BAT *s2)										\
{														\
	if (b1 == NULL || b2 == NULL || !bat_ok(b1) || !bat_ok(b2) || !bat_ok(s1) || !bat_ok(s2) ||			\
	    b1->ttype == TYPE_void || b2->ttype == TYPE_void || !enter(BATcount(b1)))					\
		FORWARD(BB, b1, b2, s1, s2);										\
	inarg a1 = {0}, a2 = {0}, c1 = {0}, c2 = {0};									\
	BAT *bn = NULL;													\
	if (in_open(&a1, b1) && in_open(&a2, b2) && in_open(&c1, s1) && in_open(&c2, s2)) {				\
		b200_bat *r = b200_##BB(a1.d, a2.d, c1.d, c2.d);							\
		bn = r ? out_bat(r) : fail(__func__);									\
		n_device++;												\
	} else														\
		fail(__func__);												\
	in_close(&c2);													\
	in_close(&c1);													\
	in_close(&a2);													\
	in_close(&a1);													\
	return bn;													\
}														\
static BAT *(*real_##BC)(BAT *, const ValRecord *, BAT *);							\
BAT *														\
BC(BAT *b, const ValRecord *v, BAT *s)										\
{														\
	if (b == NULL || v == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(v->vtype) || v->vtype == TYPE_void ||	\
	    b->ttype == TYPE_void || !enter(BATcount(b)))								\
		FORWARD(BC, b, v, s);											\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;													\
	if (in_open(&ab, b) && in_open(&as, s)) {									\
		b200_bat *r = b200_##BC(ab.d, VALptr(v), v->vtype, as.d);						\
		bn = r ? out_bat(r) : fail(__func__);									\
		n_device++;												\
	} else														\
		fail(__func__);												\
	in_close(&as);													\
	in_close(&ab);													\
	return bn;													\
}														\
static BAT *(*real_##CB)(const ValRecord *, BAT *, BAT *);							\
BAT *														\
CB(const ValRecord *v, BAT *b, BAT *s)										\
{														\
	if (b == NULL || v == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(v->vtype) || v->vtype == TYPE_void ||	\
	    b->ttype == TYPE_void || !enter(BATcount(b)))								\
		FORWARD(CB, v, b, s);											\
	inarg ab = {0}, as = {0};											\
	BAT *bn = NULL;													\
	if (in_open(&ab, b) && in_open(&as, s)) {									\
		b200_bat *r = b200_##CB(VALptr(v), v->vtype, ab.d, as.d);						\
		bn = r ? out_bat(r) : fail(__func__);									\
		n_device++;												\
	} else														\
		fail(__func__);												\
	in_close(&as);													\
	in_close(&ab);													\
	return bn;													\
}
CALC_NOTP(BATcalcand, BATcalcandcst, BATcalccstand)
CALC_NOTP(BATcalcor, BATcalcorcst, BATcalccstor)
CALC_NOTP(BATcalcxor, BATcalcxorcst, BATcalccstxor)
CALC_NOTP(BATcalclsh, BATcalclshcst, BATcalccstlsh)
CALC_NOTP(BATcalcrsh, BATcalcrshcst, BATcalccstrsh)
/* min / max: the cst-first forms are wrappers of the cst-second ones in the reference too (gdk_calc.c:1530,1708,2213,2365) */
#define b200_BATcalccstmin(v, t, b, s) b200_BATcalcmincst(b, v, t, s)
#define b200_BATcalccstmin_no_nil(v, t, b, s) b200_BATcalcmincst_no_nil(b, v, t, s)
#define b200_BATcalccstmax(v, t, b, s) b200_BATcalcmaxcst(b, v, t, s)
#define b200_BATcalccstmax_no_nil(v, t, b, s) b200_BATcalcmaxcst_no_nil(b, v, t, s)
CALC_NOTP(BATcalcmin, BATcalcmincst, BATcalccstmin)
CALC_NOTP(BATcalcmin_no_nil, BATcalcmincst_no_nil, BATcalccstmin_no_nil)
CALC_NOTP(BATcalcmax, BATcalcmaxcst, BATcalccstmax)
CALC_NOTP(BATcalcmax_no_nil, BATcalcmaxcst_no_nil, BATcalccstmax_no_nil)

/* between / ifthenelse (gdk_calc.h:87-95) */
static BAT *(*real_BATcalcbetween)(BAT *, BAT *, BAT *, BAT *, BAT *, BAT *, bool, bool, bool, bool, bool);
BAT *
BATcalcbetween(BAT *b, BAT *lo, BAT *hi, BAT *s, BAT *slo, BAT *shi, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	if (b == NULL || lo == NULL || hi == NULL || !bat_ok(b) || !bat_ok(lo) || !bat_ok(hi) || !bat_ok(s) || !bat_ok(slo) || !bat_ok(shi) ||
	    b->ttype == TYPE_void || lo->ttype == TYPE_void || hi->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcbetween, b, lo, hi, s, slo, shi, symmetric, linc, hinc, nils_false, anti);
	inarg ab = {0}, al = {0}, ah = {0}, as = {0}, asl = {0}, ash = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&al, lo) && in_open(&ah, hi) && in_open(&as, s) && in_open(&asl, slo) && in_open(&ash, shi)) {
		b200_bat *r = b200_BATcalcbetween(ab.d, al.d, ah.d, as.d, asl.d, ash.d, symmetric, linc, hinc, nils_false, anti);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&ash);
	in_close(&asl);
	in_close(&as);
	in_close(&ah);
	in_close(&al);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATcalcbetweencstcst)(BAT *, const ValRecord *, const ValRecord *, BAT *, bool, bool, bool, bool, bool);
BAT *
BATcalcbetweencstcst(BAT *b, const ValRecord *lo, const ValRecord *hi, BAT *s, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	if (b == NULL || lo == NULL || hi == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(lo->vtype) || !type_ok(hi->vtype) ||
	    lo->vtype == TYPE_void || hi->vtype == TYPE_void || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcbetweencstcst, b, lo, hi, s, symmetric, linc, hinc, nils_false, anti);
	inarg ab = {0}, as = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&as, s)) {
		b200_bat *r = b200_BATcalcbetweencstcst(ab.d, VALptr(lo), lo->vtype, VALptr(hi), hi->vtype, as.d, symmetric, linc, hinc, nils_false, anti);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATcalcifthenelse)(BAT *, BAT *, BAT *);
BAT *
BATcalcifthenelse(BAT *b, BAT *b1, BAT *b2)
{
	if (b == NULL || b1 == NULL || b2 == NULL || !bat_ok(b) || !bat_ok(b1) || !bat_ok(b2) || b->ttype == TYPE_void ||
	    b1->ttype == TYPE_void || b2->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcifthenelse, b, b1, b2);
	inarg ab = {0}, a1 = {0}, a2 = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&a1, b1) && in_open(&a2, b2)) {
		b200_bat *r = b200_BATcalcifthenelse(ab.d, a1.d, a2.d);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&a2);
	in_close(&a1);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATcalcifthenelsecst)(BAT *, BAT *, const ValRecord *);
BAT *
BATcalcifthenelsecst(BAT *b, BAT *b1, const ValRecord *c2)
{
	if (b == NULL || b1 == NULL || c2 == NULL || !bat_ok(b) || !bat_ok(b1) || !type_ok(c2->vtype) || c2->vtype == TYPE_void ||
	    b->ttype == TYPE_void || b1->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcifthenelsecst, b, b1, c2);
	inarg ab = {0}, a1 = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&a1, b1)) {
		b200_bat *r = b200_BATcalcifthenelsecst(ab.d, a1.d, VALptr(c2), c2->vtype);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&a1);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATcalcifthencstelse)(BAT *, const ValRecord *, BAT *);
BAT *
BATcalcifthencstelse(BAT *b, const ValRecord *c1, BAT *b2)
{
	if (b == NULL || b2 == NULL || c1 == NULL || !bat_ok(b) || !bat_ok(b2) || !type_ok(c1->vtype) || c1->vtype == TYPE_void ||
	    b->ttype == TYPE_void || b2->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcifthencstelse, b, c1, b2);
	inarg ab = {0}, a2 = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&a2, b2)) {
		b200_bat *r = b200_BATcalcifthencstelse(ab.d, VALptr(c1), c1->vtype, a2.d);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&a2);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATcalcifthencstelsecst)(BAT *, const ValRecord *, const ValRecord *);
BAT *
BATcalcifthencstelsecst(BAT *b, const ValRecord *c1, const ValRecord *c2)
{
	if (b == NULL || c1 == NULL || c2 == NULL || !bat_ok(b) || !type_ok(c1->vtype) || !type_ok(c2->vtype) || c1->vtype == TYPE_void ||
	    c2->vtype == TYPE_void || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATcalcifthencstelsecst, b, c1, c2);
	inarg ab = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b)) {
		b200_bat *r = b200_BATcalcifthencstelsecst(ab.d, VALptr(c1), c1->vtype, VALptr(c2), c2->vtype);
		bn = r ? out_bat(r) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&ab);
	return bn;
}

/* ------------------------------------------------------------------ join: gdk.h:2266 */
static gdk_return (*real_BATjoin)(BAT **, BAT **, BAT *, BAT *, BAT *, BAT *, bool, BUN);

gdk_return
BATjoin(BAT **r1p, BAT **r2p, BAT *l, BAT *r, BAT *sl, BAT *sr, bool nil_matches, BUN estimate)
{
	if (l == NULL || r == NULL || r1p == NULL || !bat_ok(l) || !bat_ok(r) || !bat_ok(sl) || !bat_ok(sr) ||
	    !enter(BATcount(l) + BATcount(r)))
		FORWARD(BATjoin, r1p, r2p, l, r, sl, sr, nil_matches, estimate);
	inarg al = {0}, ar = {0}, asl = {0}, asr = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *o1 = NULL, *o2 = NULL;
	*r1p = NULL;
	if (r2p)
		*r2p = NULL;
	if (in_open(&al, l) && in_open(&ar, r) && in_open(&asl, sl) && in_open(&asr, sr)) {
		n_device++;
		if (b200_BATjoin(&o1, r2p ? &o2 : NULL, al.d, ar.d, asl.d, asr.d, nil_matches, (uint64_t) estimate) == B200_SUCCEED) {
			*r1p = out_bat(o1);
			if (r2p)
				*r2p = out_bat(o2);
			if (*r1p && (!r2p || *r2p))
				rc = GDK_SUCCEED;
			else {
				BBPreclaim(*r1p);
				*r1p = NULL;
				if (r2p) {
					BBPreclaim(*r2p);
					*r2p = NULL;
				}
			}
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&asr);
	in_close(&asl);
	in_close(&ar);
	in_close(&al);
	return rc;
}

/* ------------------------------------------------------------------ the leftjoin family: gdk.h:2254-2265 */
#define JOIN2(NAME, XDECL, XARG)											\
static gdk_return (*real_##NAME)(BAT **, BAT **, BAT *, BAT *, BAT *, BAT *, bool XDECL, BUN);				\
gdk_return													\
NAME(BAT **r1p, BAT **r2p, BAT *l, BAT *r, BAT *sl, BAT *sr, bool nil_matches XDECL, BUN estimate)			\
{														\
	if (l == NULL || r == NULL || r1p == NULL || !bat_ok(l) || !bat_ok(r) || !bat_ok(sl) || !bat_ok(sr) ||		\
	    !enter(BATcount(l) + BATcount(r)))										\
		FORWARD(NAME, r1p, r2p, l, r, sl, sr, nil_matches XARG, estimate);					\
	inarg al = {0}, ar = {0}, asl = {0}, asr = {0};									\
	gdk_return rc = GDK_FAIL;											\
	b200_bat *o1 = NULL, *o2 = NULL;										\
	*r1p = NULL;													\
	if (r2p)													\
		*r2p = NULL;												\
	if (in_open(&al, l) && in_open(&ar, r) && in_open(&asl, sl) && in_open(&asr, sr)) {				\
		n_device++;												\
		if (b200_##NAME(&o1, r2p ? &o2 : NULL, al.d, ar.d, asl.d, asr.d, nil_matches XARG, (uint64_t) estimate) == B200_SUCCEED) { \
			*r1p = out_bat(o1);										\
			if (r2p)											\
				*r2p = out_bat(o2);									\
			if (*r1p && (!r2p || *r2p))									\
				rc = GDK_SUCCEED;									\
			else {												\
				BBPreclaim(*r1p);									\
				*r1p = NULL;										\
				if (r2p) {										\
					BBPreclaim(*r2p);								\
					*r2p = NULL;									\
				}											\
			}												\
		} else													\
			fail(__func__);											\
	} else														\
		fail(__func__);												\
	in_close(&asr);													\
	in_close(&asl);													\
	in_close(&ar);													\
	in_close(&al);													\
	return rc;													\
}
#define COMMA_FLAG_DECL(F) , bool F
#define COMMA_FLAG_ARG(F) , F
JOIN2(BATleftjoin, NOTHING, NOTHING)
JOIN2(BATouterjoin, COMMA_FLAG_DECL(match_one), COMMA_FLAG_ARG(match_one))
JOIN2(BATsemijoin, COMMA_FLAG_DECL(max_one), COMMA_FLAG_ARG(max_one))

#define JOIN1(NAME, FLAG)												\
static BAT *(*real_##NAME)(BAT *, BAT *, BAT *, BAT *, bool, bool, BUN);						\
BAT *														\
NAME(BAT *l, BAT *r, BAT *sl, BAT *sr, bool nil_matches, bool FLAG, BUN estimate)					\
{														\
	if (l == NULL || r == NULL || !bat_ok(l) || !bat_ok(r) || !bat_ok(sl) || !bat_ok(sr) ||			\
	    !enter(BATcount(l) + BATcount(r)))										\
		FORWARD(NAME, l, r, sl, sr, nil_matches, FLAG, estimate);						\
	inarg al = {0}, ar = {0}, asl = {0}, asr = {0};									\
	BAT *bn = NULL;													\
	if (in_open(&al, l) && in_open(&ar, r) && in_open(&asl, sl) && in_open(&asr, sr)) {				\
		b200_bat *o = b200_##NAME(al.d, ar.d, asl.d, asr.d, nil_matches, FLAG, (uint64_t) estimate);		\
		bn = o ? out_bat(o) : fail(__func__);									\
		n_device++;												\
	} else														\
		fail(__func__);												\
	in_close(&asr);													\
	in_close(&asl);													\
	in_close(&ar);													\
	in_close(&al);													\
	return bn;													\
}
JOIN1(BATintersect, max_one)
JOIN1(BATdiff, not_in)

/* ------------------------------------------------------------------ BATunique gdk.h:2279, BATconvert gdk_calc.h:123 */
static BAT *(*real_BATunique)(BAT *, BAT *);
BAT *
BATunique(BAT *b, BAT *s)
{
	if (b == NULL || !bat_ok(b) || !bat_ok(s) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATunique, b, s);
	inarg ab = {0}, as = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&as, s)) {
		b200_bat *o = b200_BATunique(ab.d, as.d);
		bn = o ? out_bat(o) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return bn;
}

static BAT *(*real_BATconvert)(BAT *, BAT *, int, uint8_t, uint8_t, uint8_t);
BAT *
BATconvert(BAT *b, BAT *s, int tp, uint8_t scale1, uint8_t scale2, uint8_t precision)
{
	if (b == NULL || !bat_ok(b) || !bat_ok(s) || !type_ok(tp) || tp == TYPE_void || b->ttype == TYPE_void || tp == TYPE_oid ||
	    b->ttype == TYPE_oid || !enter(BATcount(b)))
		FORWARD(BATconvert, b, s, tp, scale1, scale2, precision);
	inarg ab = {0}, as = {0};
	BAT *bn = NULL;
	if (in_open(&ab, b) && in_open(&as, s)) {
		b200_bat *o = b200_BATconvert(ab.d, as.d, tp, scale1, scale2, precision);
		bn = o ? out_bat(o) : fail(__func__);
		n_device++;
	} else
		fail(__func__);
	in_close(&as);
	in_close(&ab);
	return bn;
}

/* ------------------------------------------------------------------ BATsort gdk.h:1526, BATfirstn gdk.h:2281 */
static gdk_return (*real_BATsort)(BAT **, BAT **, BAT **, BAT *, BAT *, BAT *, bool, bool, bool);
gdk_return
BATsort(BAT **sorted, BAT **order, BAT **groups, BAT *b, BAT *o, BAT *g, bool reverse, bool nilslast, bool stable)
{
	if (b == NULL || !bat_ok(b) || !bat_ok(o) || !bat_ok(g) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATsort, sorted, order, groups, b, o, g, reverse, nilslast, stable);
	inarg ab = {0}, ao = {0}, ag = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *r[3] = {NULL, NULL, NULL};
	BAT **outs[3] = {sorted, order, groups};
	for (int i = 0; i < 3; i++)
		if (outs[i])
			*outs[i] = NULL;
	if (in_open(&ab, b) && in_open(&ao, o) && in_open(&ag, g)) {
		n_device++;
		if (b200_BATsort(sorted ? &r[0] : NULL, order ? &r[1] : NULL, groups ? &r[2] : NULL, ab.d, ao.d, ag.d, reverse, nilslast,
				 stable) == B200_SUCCEED) {
			rc = GDK_SUCCEED;
			for (int i = 0; i < 3; i++)
				if (outs[i] && (*outs[i] = out_bat(r[i])) == NULL)
					rc = GDK_FAIL;
			if (rc != GDK_SUCCEED)
				for (int i = 0; i < 3; i++)
					if (outs[i]) {
						BBPreclaim(*outs[i]);
						*outs[i] = NULL;
					}
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ag);
	in_close(&ao);
	in_close(&ab);
	return rc;
}

static gdk_return (*real_BATfirstn)(BAT **, BAT **, BAT *, BAT *, BAT *, BUN, bool, bool, bool);
gdk_return
BATfirstn(BAT **topn, BAT **gids, BAT *b, BAT *s, BAT *g, BUN n, bool asc, bool nilslast, bool distinct)
{
	if (b == NULL || topn == NULL || !bat_ok(b) || !bat_ok(s) || !bat_ok(g) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(BATfirstn, topn, gids, b, s, g, n, asc, nilslast, distinct);
	inarg ab = {0}, as = {0}, ag = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *o1 = NULL, *o2 = NULL;
	*topn = NULL;
	if (gids)
		*gids = NULL;
	if (in_open(&ab, b) && in_open(&as, s) && in_open(&ag, g)) {
		n_device++;
		if (b200_BATfirstn(&o1, gids ? &o2 : NULL, ab.d, as.d, ag.d, (uint64_t) n, asc, nilslast, distinct) == B200_SUCCEED) {
			*topn = out_bat(o1);
			if (gids)
				*gids = out_bat(o2);
			if (*topn && (!gids || *gids))
				rc = GDK_SUCCEED;
			else {
				BBPreclaim(*topn);
				*topn = NULL;
				if (gids) {
					BBPreclaim(*gids);
					*gids = NULL;
				}
			}
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ag);
	in_close(&as);
	in_close(&ab);
	return rc;
}

/* ------------------------------------------------------------------ window functions: gdk_analytic.h:21-42.
 * The caller pre-allocates r with capacity BATcount(b) (sql_rank.c:161,1393); the device result is copied
 * into Tloc(r, 0) and count / nil properties are set as the reference's functions do. */
static bool
analytic_finish(BAT *r, b200_bat *dr, const char *func)
{
	if (BATcapacity(r) < d_count(dr)) {
		GDKtracer_log(__FILE__, func, __LINE__, M_ERROR, GDK, NULL, "result BAT too small\n");
		return false;
	}
	if (d_count(dr) && b200_BATtohost(dr, Tloc(r, 0)) != 0) {
		fail(func);
		return false;
	}
	BATsetcount(r, (BUN) d_count(dr));
	r->tnonil = d_nonil(dr);
	r->tnil = d_nil(dr);
	return true;
}

static gdk_return (*real_GDKanalyticaldiff)(BAT *, BAT *, BAT *, const bit *restrict, int);

gdk_return
GDKanalyticaldiff(BAT *r, BAT *b, BAT *p, const bit *restrict npbit, int tpe)
{
	if (r == NULL || b == NULL || !bat_ok(b) || !bat_ok(p) || !type_ok(tpe) || b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(GDKanalyticaldiff, r, b, p, npbit, tpe);
	inarg ab = {0}, ap = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *dr = b200_COLnew(r->hseqbase, TYPE_bit, BATcount(b));
	if (dr && in_open(&ab, b) && in_open(&ap, p)) {
		n_device++;
		if (b200_GDKanalyticaldiff(dr, ab.d, ap.d, (const int8_t *) npbit, tpe) == B200_SUCCEED) {
			if (analytic_finish(r, dr, __func__))
				rc = GDK_SUCCEED;
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ap);
	in_close(&ab);
	b200_BATfree(dr);
	return rc;
}

static gdk_return (*real_GDKanalyticalwindowbounds)(BAT *, BAT *, BAT *, BAT *, const void *restrict, int, int, int, bool, oid);

gdk_return
GDKanalyticalwindowbounds(BAT *r, BAT *b, BAT *p, BAT *l, const void *restrict bound, int tp1, int tp2, int unit,
			  bool preceding, oid first_half)
{
	BAT *any = b ? b : p ? p : l;
	if (r == NULL || any == NULL || !bat_ok(b) || !bat_ok(p) || !bat_ok(l) || !enter(BATcount(any)))
		FORWARD(GDKanalyticalwindowbounds, r, b, p, l, bound, tp1, tp2, unit, preceding, first_half);
	inarg ab = {0}, ap = {0}, al = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *dr = b200_COLnew(r->hseqbase, TYPE_oid, BATcount(any));
	if (dr && in_open(&ab, b) && in_open(&ap, p) && in_open(&al, l)) {
		n_device++;
		if (b200_GDKanalyticalwindowbounds(dr, ab.d, ap.d, al.d, bound, tp1, tp2, unit, preceding, first_half) == B200_SUCCEED) {
			if (analytic_finish(r, dr, __func__))
				rc = GDK_SUCCEED;
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&al);
	in_close(&ap);
	in_close(&ab);
	b200_BATfree(dr);
	return rc;
}

static gdk_return (*real_GDKanalyticalsum)(BAT *, BAT *, BAT *, BAT *, BAT *, BAT *, int, int, int);

gdk_return
GDKanalyticalsum(BAT *r, BAT *p, BAT *o, BAT *b, BAT *s, BAT *e, int tp1, int tp2, int frame_type)
{
	if (r == NULL || b == NULL || !bat_ok(b) || !bat_ok(p) || !bat_ok(o) || !bat_ok(s) || !bat_ok(e) || !type_ok(tp2) ||
	    b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(GDKanalyticalsum, r, p, o, b, s, e, tp1, tp2, frame_type);
	inarg ab = {0}, ap = {0}, ao = {0}, as = {0}, ae = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *dr = b200_COLnew(r->hseqbase, tp2, BATcount(b));
	if (dr && in_open(&ap, p) && in_open(&ao, o) && in_open(&ab, b) && in_open(&as, s) && in_open(&ae, e)) {
		n_device++;
		if (b200_GDKanalyticalsum(dr, ap.d, ao.d, ab.d, as.d, ae.d, tp1, tp2, frame_type) == B200_SUCCEED) {
			if (analytic_finish(r, dr, __func__))
				rc = GDK_SUCCEED;
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ae);
	in_close(&as);
	in_close(&ab);
	in_close(&ao);
	in_close(&ap);
	b200_BATfree(dr);
	return rc;
}

#define ANALYTIC_AGGR(NAME, RTYPE)									\
static gdk_return (*real_##NAME)(BAT *, BAT *, BAT *, BAT *, BAT *, BAT *, int, int);			\
gdk_return												\
NAME(BAT *r, BAT *p, BAT *o, BAT *b, BAT *s, BAT *e, int tpe, int frame_type)				\
{													\
	if (r == NULL || b == NULL || !bat_ok(b) || !bat_ok(p) || !bat_ok(o) || !bat_ok(s) || !bat_ok(e) ||	\
	    !type_ok(tpe) || b->ttype == TYPE_void || !enter(BATcount(b)))					\
		FORWARD(NAME, r, p, o, b, s, e, tpe, frame_type);					\
	inarg ab = {0}, ap = {0}, ao = {0}, as = {0}, ae = {0};									\
	gdk_return rc = GDK_FAIL;									\
	b200_bat *dr = b200_COLnew(r->hseqbase, (RTYPE), BATcount(b));					\
	if (dr && in_open(&ap, p) && in_open(&ao, o) && in_open(&ab, b) && in_open(&as, s) && in_open(&ae, e)) {	\
		n_device++;										\
		if (b200_##NAME(dr, ap.d, ao.d, ab.d, as.d, ae.d, tpe, frame_type) == B200_SUCCEED) {	\
			if (analytic_finish(r, dr, __func__))						\
				rc = GDK_SUCCEED;							\
		} else											\
			fail(__func__);									\
	} else												\
		fail(__func__);										\
	in_close(&ae);											\
	in_close(&as);											\
	in_close(&ab);											\
	in_close(&ao);											\
	in_close(&ap);											\
	b200_BATfree(dr);										\
	return rc;											\
}
ANALYTIC_AGGR(GDKanalyticalavg, TYPE_dbl)
ANALYTIC_AGGR(GDKanalyticalavginteger, tpe)
ANALYTIC_AGGR(GDKanalyticalmin, tpe)
ANALYTIC_AGGR(GDKanalyticalmax, tpe)

static gdk_return (*real_GDKanalyticalcount)(BAT *, BAT *, BAT *, BAT *, BAT *, BAT *, bit, int, int);

gdk_return
GDKanalyticalcount(BAT *r, BAT *p, BAT *o, BAT *b, BAT *s, BAT *e, bit ignore_nils, int tpe, int frame_type)
{
	if (r == NULL || b == NULL || !bat_ok(b) || !bat_ok(p) || !bat_ok(o) || !bat_ok(s) || !bat_ok(e) || !type_ok(tpe) ||
	    b->ttype == TYPE_void || !enter(BATcount(b)))
		FORWARD(GDKanalyticalcount, r, p, o, b, s, e, ignore_nils, tpe, frame_type);
	inarg ab = {0}, ap = {0}, ao = {0}, as = {0}, ae = {0};
	gdk_return rc = GDK_FAIL;
	b200_bat *dr = b200_COLnew(r->hseqbase, TYPE_lng, BATcount(b));
	if (dr && in_open(&ap, p) && in_open(&ao, o) && in_open(&ab, b) && in_open(&as, s) && in_open(&ae, e)) {
		n_device++;
		if (b200_GDKanalyticalcount(dr, ap.d, ao.d, ab.d, as.d, ae.d, ignore_nils, tpe, frame_type) == B200_SUCCEED) {
			if (analytic_finish(r, dr, __func__))
				rc = GDK_SUCCEED;
		} else
			fail(__func__);
	} else
		fail(__func__);
	in_close(&ae);
	in_close(&as);
	in_close(&ab);
	in_close(&ao);
	in_close(&ap);
	b200_BATfree(dr);
	return rc;
}
