/*
 * gdk_b200.h — C-ABI of the B200-native GDK column-at-a-time hot path.
 *
 * One entry point per reference GDK operator on the path, with the same name
 * (prefixed b200_), the same argument order, meaning and error behaviour.  The
 * only change is that the reference's `BAT *` (gdk.h:778-804) is replaced by a
 * plain descriptor `b200_bat` holding exactly the BAT/COLrec fields the path
 * reads: no reference struct layout, no torch types, plain pointers and sizes.
 * The reference-side binding (a patched libbat entry point that converts
 * BAT* <-> b200_bat and tail-calls these functions) is shown in INTEGRATION.md
 *.
 *
 * Conventions (all cited files are relative to the MonetDB reference tree):
 *  - oid / BUN are 64-bit unsigned (gdk.h:460-461,507); oid_nil = 1<<63.
 *  - fixed-width atoms only; nil is in-band (gdk_atoms.h:149-218): integer nil
 *    = type minimum, flt/dbl nil = NaN.  date is stored as int, timestamp and
 *    daytime as lng (ATOMbasetype, gdk_atoms.h:285-289).
 *  - BAT-returning functions return NULL after recording a message retrievable
 *    with b200_errbuf() (GDKerror/GDKerrbuf, gdk.h:1710-1717); gdk_return
 *    functions return B200_FAIL likewise.  Messages carry the reference's
 *    SQLSTATE-prefixed texts.
 *  - inputs are borrowed; every returned BAT is new, device-resident and owned
 *    by the caller (release with b200_BATfree = BBPreclaim).
 *  - there is NO CPU fallback: if no CUDA device is usable every entry point
 *    fails loudly.
 */
#ifndef GDK_B200_H
#define GDK_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint64_t b200_oid;
#define B200_OID_NIL ((b200_oid) 1 << 63)
#define B200_BUN_NONE ((uint64_t) INT64_MAX)

/* GDK type ids, gdk.h:429-451 (HAVE_HGE build) */
enum {
	B200_void = 0, B200_msk = 1, B200_bit = 2, B200_bte = 3, B200_sht = 4,
	B200_int = 5, B200_oid = 6, B200_ptr = 7, B200_flt = 8, B200_dbl = 9,
	B200_lng = 10, B200_hge = 11, B200_date = 12, B200_daytime = 13,
	B200_timestamp = 14,
	B200_str = 16		/* dictionary-encoded: fixed-width offsets (width in b200_bat.pad) + string heap in vheap */
};

typedef enum { B200_FAIL = 0, B200_SUCCEED = 1 } b200_return;	/* gdk.h:348-350 */

/* candidate-list encodings beyond dense/oid[] (gdk_cand.h:16-45) */
enum { B200_CAND_NONE = 0, B200_CAND_NEGOID = 1, B200_CAND_MSK = 2 };

/* The slice of BAT + COLrec (gdk.h:712-804) the hot path uses. */
typedef struct b200_bat {
	int32_t  ttype;		/* tail type id; B200_void => virtual dense oids from tseqbase */
	int32_t  device;	/* CUDA ordinal owning base/vheap; -1 = host memory */
	uint64_t hseqbase;	/* oid of row 0 */
	uint64_t count;		/* batCount */
	uint64_t tseqbase;	/* void/dense-oid tail start, else oid_nil */
	void    *base;		/* Tloc(b,0); NULL for void */
	uint64_t capacity;	/* allocated items behind base */
	int32_t  ccand;		/* B200_CAND_*: complex candidate list in vheap */
	uint32_t firstbit;	/* CAND_MSK: first valid bit in word 0 */
	void    *vheap;		/* NEGOID: sorted exception oids; MSK: uint32 words */
	uint64_t nvheap;	/* number of exceptions / mask words */
	uint8_t  tsorted, trevsorted, tkey, tnonil, tnil;
	uint8_t  owns;		/* descriptor owns base/vheap (freed by b200_BATfree) */
	uint8_t  oidrange_known;	/* first_oid/last_oid valid (sorted oid tails we produced) */
	uint8_t  pad;
	uint64_t first_oid, last_oid;	/* smallest / largest oid of a sorted oid tail (cf. tminpos/tmaxpos) */
	void    *parent;	/* views (owns == 0 slices of a device BAT of ours): the tail heap they share.  The heap
				   carries a share count like the reference's BBP share on a view's parent
				   (theap->parentid, gdk_bat.c VIEWcreate): it is released when the owner AND every
				   view have been freed, in either order */
	uint32_t producer;	/* id of the thread context (stream) that produced this BAT, 0 = none: a consumer or
				   b200_BATfree on another thread orders its own stream behind that stream first */
	uint32_t pad2;
} b200_bat;

/* ---- runtime / device heap tier (the tier added next to gdk_heap.c:141-225) ---- */
int  b200_init(int device);			/* bind calling thread to a device; 0 on success */
int  b200_device_count(void);
const char *b200_errbuf(void);			/* GDKerrbuf analogue (thread-local) */
void b200_clrerr(void);
const char *b200_algorithm(void);		/* MT_thread_setalgorithm tag of the last call */
uint64_t b200_kernel_launches(void);		/* kernels launched by this library so far */
void b200_sync(void);				/* wait for the calling thread's stream */
void *b200_stream(void);			/* cudaStream_t of the calling thread */

/* per-kernel timing for bench.py: CUDA events on the launching stream around the dominant
 * kernels.  b200_timing_read syncs and writes lines "name launches total_ms algorithmic_bytes\n". */
void b200_timing_enable(int on);
int  b200_timing_read(char *buf, size_t bufsize);

/* freed device heaps are cached by size and reused without a driver call (cf. HEAPcache, gdk_heap.c:60-139);
 * the cache is bounded by B200_HEAP_CACHE_GB (default 60 % of the device) and emptied on allocation failure */
void b200_heap_cache_trim(void);		/* give every cached heap back to the driver */
uint64_t b200_heap_cache_bytes(void);

void *b200_pinned_alloc(size_t bytes);		/* page-locked host staging memory */
void  b200_pinned_free(void *p);

b200_bat *b200_COLnew(uint64_t hseq, int tt, uint64_t cap);			/* gdk_bat.c:292 */
b200_bat *b200_BATdense(uint64_t hseq, uint64_t tseq, uint64_t cnt);		/* gdk_bat.c:298 */
b200_bat *b200_BATfromhost(const void *src, int tt, uint64_t cnt, uint64_t hseq);	/* upload a tail heap */
b200_bat *b200_BATwrap(void *devptr, int tt, uint64_t cnt, uint64_t hseq);	/* borrow device memory */
b200_bat *b200_BAThostwrap(const void *hostptr, int tt, uint64_t cnt, uint64_t hseq);	/* borrow HOST memory (device = -1):
							   every operator stages it to the device on use */
void b200_stage_cache(int on);		/* keep staged copies of HOST heaps (device = -1) until cleared: a host column used
					   by several operators of one query is uploaded once */
void b200_stage_cache_clear(void);	/* drop every staged copy (call when a host heap is written or freed) */
b200_bat *b200_BATnegcands_fromhost(uint64_t seq, uint64_t cnt, const b200_oid *exceptions, uint64_t nexc); /* gdk_cand.c:1296 */
b200_bat *b200_BATmaskedcands_fromhost(uint64_t hseq, uint64_t nrows, const uint32_t *words, bool selected); /* gdk_cand.c:1366 */
b200_bat *b200_BATunmask(const b200_bat *b);	/* gdk_cand.c:1464: mask / negative candidates -> dense or oid[] */
b200_return b200_BATextend(b200_bat *b, uint64_t newcap);	/* gdk_bat.c:529 / HEAPextend gdk_heap.c:246 */
void b200_BATsetprops(b200_bat *b, bool sorted, bool revsorted, bool key, bool nonil, bool nil);
int  b200_BATtohost(const b200_bat *b, void *dst);	/* logical tail -> host (void/cands materialised as oids) */
void b200_BATfree(b200_bat *b);				/* BBPreclaim */
/* the owner of a WRAPPED heap (b200_BATwrap / b200_BAThostwrap) rewrote it in place: forget what the library
 * cached about the old contents (BAThash index, staged device copy).  Heaps the library owns are immutable. */
void b200_BATinvalidate(const b200_bat *b);

/* ---- selection: gdk_select.c:1342 (BATselect), :2102 (BATthetaselect) ---- */
b200_bat *b200_BATselect(const b200_bat *b, const b200_bat *s, const void *tl, const void *th,
			 bool li, bool hi, bool anti, bool nil_matches);
b200_bat *b200_BATthetaselect(const b200_bat *b, const b200_bat *s, const void *val, const char *op);

/* ---- candidate set algebra: gdk_cand.c:46,184,259 ---- */
b200_bat *b200_BATmergecand(const b200_bat *a, const b200_bat *b);
b200_bat *b200_BATintersectcand(const b200_bat *a, const b200_bat *b);
b200_bat *b200_BATdiffcand(const b200_bat *a, const b200_bat *b);

/* ---- projection: gdk_project.c:857 (BATproject), :590 (BATproject2) ---- */
b200_bat *b200_BATproject(const b200_bat *l, const b200_bat *r);
b200_bat *b200_BATproject2(const b200_bat *l, const b200_bat *r1, const b200_bat *r2);

/* ---- calc: gdk_calc_addsub.c:1480-1664,3166-3280; gdk_calc_mul.c:2085-2113;
 *      gdk_calc_div.c:1945; gdk_calc_mod.c:1179,1186,1236; compares gdk_calc_compare_*.c; BATconvert gdk_calc_convert.c:1415.
 *      A ValRecord constant (gdk.h:567-600) is passed as (pointer, type id). ---- */
b200_bat *b200_BATcalcadd(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, int tp);
b200_bat *b200_BATcalcsub(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, int tp);
b200_bat *b200_BATcalcmul(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, int tp);
b200_bat *b200_BATcalcdiv(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, int tp);
b200_bat *b200_BATcalcmod(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, int tp);
b200_bat *b200_BATcalcaddcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, int tp);
b200_bat *b200_BATcalcsubcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, int tp);
b200_bat *b200_BATcalcmulcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, int tp);
b200_bat *b200_BATcalcdivcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, int tp);
b200_bat *b200_BATcalcmodcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, int tp);
b200_bat *b200_BATcalccstadd(const void *v, int vtype, const b200_bat *b, const b200_bat *s, int tp);
b200_bat *b200_BATcalccstsub(const void *v, int vtype, const b200_bat *b, const b200_bat *s, int tp);
b200_bat *b200_BATcalccstmul(const void *v, int vtype, const b200_bat *b, const b200_bat *s, int tp);
b200_bat *b200_BATcalccstdiv(const void *v, int vtype, const b200_bat *b, const b200_bat *s, int tp);
b200_bat *b200_BATcalccstmod(const void *v, int vtype, const b200_bat *b, const b200_bat *s, int tp);
/* min / max against a constant: gdk_calc.c:1403 (mincst), 1563 (mincst_no_nil), 2086 (maxcst), 2219 (maxcst_no_nil);
 * the reference's BATcalccstmin/max[_no_nil] forward to these (:1530,1708) */
b200_bat *b200_BATcalcmincst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcmincst_no_nil(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcmaxcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcmaxcst_no_nil(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
/* and / or / xor: gdk_calc.c:2901,2967,3030 / 2670,2736,2799 / 2439,2506,2563 (three-valued on bit, bitwise on integers) */
b200_bat *b200_BATcalcand(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcor(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcxor(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcandcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcorcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcxorcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalccstand(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstor(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstxor(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
/* shifts: gdk_calc.c:3246 (BATcalclsh), 3305 (lshcst), 3356 (cstlsh), 3596-3770 (rsh forms); result type = left operand */
b200_bat *b200_BATcalclsh(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcrsh(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalclshcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcrshcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalccstlsh(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstrsh(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalclt(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcle(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcgt(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcge(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalceq(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, bool nil_matches);
b200_bat *b200_BATcalcne(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, bool nil_matches);
b200_bat *b200_BATcalcltcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalclecst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcgtcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalcgecst(const b200_bat *b, const void *v, int vtype, const b200_bat *s);
b200_bat *b200_BATcalceqcst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, bool nil_matches);
b200_bat *b200_BATcalcnecst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, bool nil_matches);
/* constant on the left: BATcalccstop, gdk_calc_compare.h:928 */
b200_bat *b200_BATcalccstlt(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstle(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstgt(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccstge(const void *v, int vtype, const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalccsteq(const void *v, int vtype, const b200_bat *b, const b200_bat *s, bool nil_matches);
b200_bat *b200_BATcalccstne(const void *v, int vtype, const b200_bat *b, const b200_bat *s, bool nil_matches);
b200_bat *b200_BATcalcisnil(const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalcisnotnil(const b200_bat *b, const b200_bat *s);	/* gdk_calc.c:916 */
b200_bat *b200_BATconvert(const b200_bat *b, const b200_bat *s, int tp, uint8_t scale1, uint8_t scale2, uint8_t precision);

/* ---- grouping: gdk_group.c:1347 ---- */
b200_return b200_BATgroup(b200_bat **groups, b200_bat **extents, b200_bat **histo,
			  const b200_bat *b, const b200_bat *s, const b200_bat *g,
			  const b200_bat *e, const b200_bat *h);

/* ---- grouped / scalar aggregates: gdk_aggr.c:900,1018,1801,3069,3561,3720 ---- */
b200_bat *b200_BATgroupsum(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils);
b200_bat *b200_BATgroupcount(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils);
/* gdk_aggr.c:1575 — products in row order with the reference's overflow verdict ("22003!overflow in product aggregate.");
 * integer results by an exact order-independent form, flt/dbl results by an ordered replay per group */
b200_bat *b200_BATgroupprod(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils);
/* gdk_aggr.c:1650 — the scalar form; *res is an atom of type tp in host memory; a flt / dbl product of more than 2^24 rows is
 * refused (order-dependent, one sequential pass) */
b200_return b200_BATprod(void *res, int tp, const b200_bat *b, const b200_bat *s, bool skip_nils, bool nil_if_empty);
b200_bat *b200_BATgroupmin(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils);
b200_bat *b200_BATgroupmax(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils);
b200_return b200_BATgroupavg(b200_bat **bnp, b200_bat **cntsp, const b200_bat *b, const b200_bat *g,
			     const b200_bat *e, const b200_bat *s, int tp, bool skip_nils, int scale);

/* exact integer average as (floor-rounded avg, rem, cnt) triples and their recombination —
 * mergetable's phase 1 / phase 2 (gdk_aggr.c:1996, 2634); scalar average (gdk_aggr.c:2987) */
b200_return b200_BATgroupavg3(b200_bat **avgp, b200_bat **remp, b200_bat **cntp, const b200_bat *b, const b200_bat *g,
			      const b200_bat *e, const b200_bat *s, bool skip_nils);
b200_bat *b200_BATgroupavg3combine(const b200_bat *avg, const b200_bat *rem, const b200_bat *cnt, const b200_bat *g,
				   const b200_bat *e, bool skip_nils);
b200_return b200_BATcalcavg(const b200_bat *b, const b200_bat *s, double *avg, uint64_t *vals, int scale);
/* phase 2 of the two-phase average (mergetable splits avg into per-slice sum + count, opt_mergetable.c:1204-1215):
 * dbl averages from merged lng/hge sums and lng counts with BATgroupavg's own closing formula
 * floor(S/n) + (dbl) rem / n, then / 10^scale (gdk_aggr.c:1741-1748,1952-1957) — bit-identical to one
 * BATgroupavg over all the rows */
b200_bat *b200_BATcalcavgsum(const b200_bat *sum, const b200_bat *cnt, int scale);
b200_return b200_BATsum(void *res, int tp, const b200_bat *b, const b200_bat *s, bool skip_nils, bool nil_if_empty);

/* ---- min/max, between, ifthenelse: gdk_calc.h:24-33,87-95 (gdk_calc.c:976-2300, 3968-4290, 4376-4800).
 *      ValRecord arguments travel as (pointer to the atom, type id), as in the arithmetic cst forms. ---- */
b200_bat *b200_BATcalcmin(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcmin_no_nil(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcmax(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcmax_no_nil(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2);
b200_bat *b200_BATcalcbetween(const b200_bat *b, const b200_bat *lo, const b200_bat *hi, const b200_bat *s,
			      const b200_bat *slo, const b200_bat *shi, bool symmetric, bool linc, bool hinc,
			      bool nils_false, bool anti);
b200_bat *b200_BATcalcbetweencstcst(const b200_bat *b, const void *lo, int lotype, const void *hi, int hitype,
				    const b200_bat *s, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti);
b200_bat *b200_BATcalcbetweenbatcst(const b200_bat *b, const b200_bat *lo, const void *hi, int hitype, const b200_bat *s,
				    const b200_bat *slo, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti);
b200_bat *b200_BATcalcbetweencstbat(const b200_bat *b, const void *lo, int lotype, const b200_bat *hi, const b200_bat *s,
				    const b200_bat *shi, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti);
b200_bat *b200_BATcalcifthenelse(const b200_bat *b, const b200_bat *b1, const b200_bat *b2);
b200_bat *b200_BATcalcifthenelsecst(const b200_bat *b, const b200_bat *b1, const void *c2, int c2type);
b200_bat *b200_BATcalcifthencstelse(const b200_bat *b, const void *c1, int c1type, const b200_bat *b2);
b200_bat *b200_BATcalcifthencstelsecst(const b200_bat *b, const void *c1, int c1type, const void *c2, int c2type);

/* unary operators: gdk_calc.h:15-23 (gdk_calc.c:41 not, :233 negate, :371 absolute, :513 iszero -> bit, :652 sign -> bte) */
b200_bat *b200_BATcalcnot(const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalcnegate(const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalcabsolute(const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalciszero(const b200_bat *b, const b200_bat *s);
b200_bat *b200_BATcalcsign(const b200_bat *b, const b200_bat *s);

/* BATunique (gdk.h:2279, gdk_unique.c:30): candidate list of the first row of each distinct value of b within s */
b200_bat *b200_BATunique(const b200_bat *b, const b200_bat *s);

/* ---- sort / top-N: gdk_batop.c:2342 (BATsort: sorted copy, order oids, group ids; o / g chain
 *      the calls of a multi-column ORDER BY), gdk_firstn.c:1280 (BATfirstn, plain form: g == NULL, gids == NULL,
 *      !distinct).  One stable LSD radix sort serves stable and !stable alike. ---- */
b200_return b200_BATsort(b200_bat **sorted, b200_bat **order, b200_bat **groups, const b200_bat *b, const b200_bat *o,
			 const b200_bat *g, bool reverse, bool nilslast, bool stable);
/* order index, gdk_orderidx.c:74 (BATcheckorderidx), :184 (BATorderidx): the stable ascending order of a device column, kept
 * beside its heap; BATsort (no o / g, ascending, nils first) and selective range selects on large columns use it */
b200_return b200_BATorderidx(const b200_bat *b, bool stable);
int b200_BATcheckorderidx(const b200_bat *b);
b200_return b200_BATfirstn(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, const b200_bat *g,
			   uint64_t n, bool asc, bool nilslast, bool distinct);

/* ---- dictionary strings: a str tail as GDK stores it — offsets of `width` bytes (1/2: minus
 * GDK_VAROFFSET, gdk_atoms.h:370-377) into a heap of NUL-terminated strings, nil = "\200"; elim_doubles = the heap is
 * duplicate-free (GDK_ELIMDOUBLES, gdk.h).  On such columns BATgroup (gdk_group.c:897-919), the (anti-)equality forms of
 * BATselect / BATthetaselect (gdk_select.c:608-656) and BATproject (gdk_project.c:681-695) run on the device. ---- */
b200_bat *b200_BATstrfromhost(const void *offsets, int width, uint64_t cnt, uint64_t hseq, const void *heap,
			      uint64_t heapbytes, bool elim_doubles);
int b200_BATstrheap_tohost(const b200_bat *b, void *dst);	/* nvheap bytes; the offsets come with b200_BATtohost */
int b200_BATstrwidth(const b200_bat *b);

/* ---- data-format edges of the device tier.
 * Binary COPY INTO records (sql/backends/monet5/sql_bincopyconvert.c:590-616; layouts common/utils/copybinary.h:15-33):
 * `src` (host) holds `count` records of `type` — the atom itself for bte/bit/sht/int/lng/hge/flt/dbl, the 4 / 8 / 12-byte
 * copy_binary_date / _time / _timestamp for the temporal types; byteswap = the file has the other endianness;
 * width > 0 = DECIMAL(width) range validation (VALIDATE_DECIMAL :45-58); filename only appears in error texts
 * ("22003!invalid boolean byte value %d in %s", "22003!decimal out of range in %s").  Impossible dates / times decode to
 * nil as date_create / daytime_create / timestamp_create do (gdk_time.c:62,281,401).  encode is the way back
 * (dump_binary_column): count * record-size bytes into `dst` (host). */
b200_bat *b200_bincopy_decode(int type, const void *src, uint64_t count, bool byteswap, int width, const char *filename,
			      uint64_t hseq);
int b200_bincopy_encode(const b200_bat *b, void *dst, bool byteswap);
/* gdk_heap.c:729 (HEAPload) / :860 (HEAPsave -> GDKsave): a tail heap file — the raw array of values — read straight into
 * / written straight out of device memory through page-locked staging; properties are the BBP directory's business
 * (b200_BATsetprops).  A file shorter than count * width is refused. */
b200_bat *b200_HEAPload(const char *path, int type, uint64_t count, uint64_t hseq);
int b200_HEAPsave(const b200_bat *b, const char *path, bool dosync);

/* ---- allocation-free range select.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Selects the rows with lo <= vals[i] <= hi (nil never matches; nil bounds are refused) of a DEVICE-resident lng
 * column and writes their oids (hseqbase + i, ascending) to oids_out (device, room for n) and their number to
 * *count_out (device); same result as b200_BATselect(b, NULL, &lo, &hi, true, true, false, false).
 * The scratch (mask bits, per-chunk counts / offsets) is a file-scope static __device__ array of about 2.1 MiB that
 * every GPU loading the library holds until exit: n <= B200_NOALLOC_MAX_ROWS, larger inputs are refused (-2).
 * Calls no cudaMalloc / cudaFree / pool routine and creates no thread context: three launches on `stream`
 * (a cudaStream_t, NULL = the legacy default stream), no synchronisation — read the outputs in stream order.
 * The first call on a device looks the array's address up (cudaGetSymbolAddress) and must happen outside stream
 * capture.  All calls share the array and it keeps the last call's contents (each call overwrites what it reads, so
 * nothing needs clearing): calls that OVERLAP on the GPU race — serialise them on one stream or with events.
 * Returns 0, or -1 bad arguments, -2 too many rows, -3 CUDA error; text in b200_errbuf() (stderr if the calling
 * thread has no context). ---- */
#define B200_NOALLOC_MAX_ROWS ((uint64_t) 1 << 24)
int b200_select_lng_noalloc(const int64_t *vals, uint64_t n, uint64_t hseqbase, int64_t lo, int64_t hi, uint64_t *oids_out,
			    uint64_t *count_out, void *stream);

/* ---- hash + join: gdk_hash.c:989 (BAThash), gdk_join.c:4451 (BATjoin) ---- */
b200_return b200_BAThash(const b200_bat *b);
b200_return b200_BATjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
			 const b200_bat *sl, const b200_bat *sr, bool nil_matches, uint64_t estimate);
/* the leftjoin family (gdk.h:2254-2265; gdk_join.c:4049 leftjoin and its flag wrappers :4320-4405).  r1 is
 * always in l's candidate order.  semijoin/intersect/diff return r1 as a candidate list (dense when contiguous);
 * outerjoin writes oid_nil into r2 for left rows without a match; match_one / max_one violations fail with
 * "more than one match" / "not enough matches" like the reference's hashjoin (:2759-2762,2869-2871). */
b200_return b200_BATleftjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
			     const b200_bat *sl, const b200_bat *sr, bool nil_matches, uint64_t estimate);
b200_return b200_BATouterjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
			      const b200_bat *sl, const b200_bat *sr, bool nil_matches, bool match_one, uint64_t estimate);
b200_return b200_BATsemijoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
			     const b200_bat *sl, const b200_bat *sr, bool nil_matches, bool max_one, uint64_t estimate);
b200_bat *b200_BATintersect(const b200_bat *l, const b200_bat *r, const b200_bat *sl, const b200_bat *sr,
			    bool nil_matches, bool max_one, uint64_t estimate);
b200_bat *b200_BATdiff(const b200_bat *l, const b200_bat *r, const b200_bat *sl, const b200_bat *sr,
		       bool nil_matches, bool not_in, uint64_t estimate);

/* ---- the one exchange step of the multi-GPU path (no reference counterpart: mitosis only slices by row
 *      range, opt_mitosis.c:188-214): split (key, oid) tuples by hash(key) mod nparts into nparts
 *      contiguous send runs for an NCCL all-to-all.  counts[d] = tuples for destination d; pass
 *      keys_out = oids_out = NULL to size the buffers first. ---- */
b200_return b200_hash_partition(const b200_bat *b, const b200_bat *s, int nparts, bool keep_nil, void *keys_out,
				b200_oid *oids_out, uint64_t *counts);

b200_return b200_BATpartition(b200_bat **keysp, b200_bat **oidsp, const b200_bat *b, const b200_bat *s, int nparts,
			      bool keep_nil, uint64_t *counts);	/* one-call form: the send runs as two new BATs */

/* ---- multi-GPU inside the library (one process per GPU, NCCL on the calling thread's stream).
 *      rank 0 makes the 128-byte id (ncclGetUniqueId) and the launcher hands it to every rank (bench.py: one
 *      torch.distributed broadcast).  Collective entry points must be called by every rank, in the same order,
 *      from the thread that called b200_comm_init.  world == 1 needs no NCCL and no id.
 *  b200_comm_pack     = mat.pack across the ranks (mergetable's packing step, opt_mergetable.c:1204-1215):
 *      out[i] = rank-order concatenation of every rank's in[i] (k <= 16 small BATs of at most max_rows rows each,
 *      the same k / types / max_rows on every rank).  One ncclAllGather, one sync.  Phase 2 of a two-phase
 *      aggregate is then ordinary GDK calls on the packed BATs.
 *  b200_comm_exchange = the all-to-all of the repartitioning plans: k parallel device columns, each laid out as
 *      `world` contiguous runs of send_counts[d] rows (b200_BATpartition), are exchanged with one grouped
 *      ncclSend/ncclRecv; out[i] = run `rank` of every rank in rank order; recv_counts[r] (may be NULL) = rows
 *      received from rank r.  The outputs are ready in stream order (no sync after the data movement). ---- */
int  b200_comm_unique_id(void *out128);
int  b200_comm_init(int rank, int world, const void *id128);
int  b200_comm_rank(void);
int  b200_comm_world(void);
int  b200_comm_nccl_version(void);
void b200_comm_destroy(void);
b200_return b200_comm_pack(int k, const b200_bat *const *in, b200_bat **out, uint64_t max_rows);
b200_return b200_comm_exchange(int k, const b200_bat *const *in, const uint64_t *send_counts, b200_bat **out,
			       uint64_t *recv_counts);

/* ---- window frames and analytic aggregates: gdk_analytic.h:21-42.
 *      As in the reference the CALLER allocates r (b200_COLnew of the result type,
 *      capacity >= BATcount(b)); the callee fills it and sets count/tnil/tnonil. ---- */
b200_return b200_GDKanalyticaldiff(b200_bat *r, const b200_bat *b, const b200_bat *p, const int8_t *npbit, int tpe);
b200_return b200_GDKanalyticalwindowbounds(b200_bat *r, const b200_bat *b, const b200_bat *p, const b200_bat *l,
					   const void *bound, int tp1, int tp2, int unit,
					   bool preceding, b200_oid first_half);
b200_return b200_GDKanalyticalsum(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
				  const b200_bat *s, const b200_bat *e, int tp1, int tp2, int frame_type);
b200_return b200_GDKanalyticalavg(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
				  const b200_bat *s, const b200_bat *e, int tpe, int frame_type);
b200_return b200_GDKanalyticalavginteger(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					 const b200_bat *s, const b200_bat *e, int tpe, int frame_type);	/* gdk_analytic_statistics.c:631 */
b200_return b200_GDKanalyticalcount(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
				    const b200_bat *s, const b200_bat *e, int8_t ignore_nils, int tpe, int frame_type);
b200_return b200_GDKanalyticalmin(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
				  const b200_bat *s, const b200_bat *e, int tpe, int frame_type);
b200_return b200_GDKanalyticalmax(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
				  const b200_bat *s, const b200_bat *e, int tpe, int frame_type);

#ifdef __cplusplus
}
#endif
#endif /* GDK_B200_H */
