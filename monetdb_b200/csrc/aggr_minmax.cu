// aggr_minmax.cu — BATgroupmin / BATgroupmax on the device
// (gdk_aggr.c:3207-3240 AGGR_CMP, :3247/:3369 do_groupmin/max, :3487-3565,3720).
//
// Result per group: the oid of the FIRST candidate row holding the smallest
// (largest) non-nil value (strict comparison: the earliest row wins ties);
// with skip_nils == false the first nil row of a group wins; a group without a
// qualifying row yields oid_nil.
//
// Parallel formulation: every value is mapped to an order-preserving unsigned
// key.  Pass 1 reduces the key per group (64 most significant bits), pass 2 (hge
// only) the low 64 bits among the rows that tie on the high part, and the last
// pass takes the smallest candidate position among the rows equal to the
// group's extreme.  A plain read of the running extreme filters out almost every
// atomic because rows are visited in ascending order.
#include "aggr_init.cuh"

typedef unsigned long long u64;
typedef long long i64;

#define MM_THREADS 256
#define MM_NONE (~0ULL)

// order-preserving map to unsigned, split in (hi, lo) 64-bit halves; lo is only used for hge
template <typename T> struct Ord {
	__device__ static inline void get(T v, u64 &hi, u64 &lo)
	{
		hi = (u64) ((i64) v) ^ (1ULL << 63);
		lo = 0;
	}
};
template <> struct Ord<hge> {
	__device__ static inline void get(hge v, u64 &hi, u64 &lo)
	{
		hi = (u64) (v >> 64) ^ (1ULL << 63);
		lo = (u64) v;
	}
};
template <> struct Ord<dbl> {
	__device__ static inline void get(dbl v, u64 &hi, u64 &lo)
	{
		if (v == 0)
			v = 0;		/* -0 == +0 */
		u64 b = (u64) __double_as_longlong(v);
		hi = (b >> 63) ? ~b : (b | (1ULL << 63));
		lo = 0;
	}
};
template <> struct Ord<flt> {
	__device__ static inline void get(flt v, u64 &hi, u64 &lo) { Ord<dbl>::get((dbl) v, hi, lo); }
};

template <typename T> struct MMIn {
	const T *vals;		/* dense: vals[r]; else vals[cand[r] - hseq] */
	const oid *cand;
	oid hseq, seq;
	const oid *gids;	/* NULL: gdense / single group */
	oid gmin, gmax;
	bool gconst;		/* gids == NULL: every row in group gmin (else gid = gmin + r) */
	uint64_t n;
	bool ismax, skip_nils;
	__device__ __forceinline__ bool fetch(uint64_t r, T &x, u64 &gi) const
	{
		x = cand ? vals[__ldcs(cand + r) - hseq] : ld_stream(vals + r);
		oid g = gids ? __ldcs(gids + r) : (gconst ? gmin : gmin + r);
		gi = g - gmin;
		return g >= gmin && g <= gmax;
	}
	__device__ __forceinline__ void key(T x, u64 &hi, u64 &lo) const
	{
		Ord<T>::get(x, hi, lo);
		if (ismax) {
			hi = ~hi;
			lo = ~lo;
		}
	}
};

__device__ __forceinline__ void min_update(u64 *p, u64 v)
{
	if (v < __ldcg(p))
		atomicMin(p, v);
}

// PHASE 0: best_hi (+ first nil row);  1: best_lo among hi ties (hge);  2: first row equal to the extreme
template <typename T, int PHASE>
__global__ void __launch_bounds__(MM_THREADS)
minmax_kernel(MMIn<T> in, u64 *__restrict__ best_hi, u64 *__restrict__ best_lo, u64 *__restrict__ firstnil,
	      u64 *__restrict__ row)
{
	const uint64_t stride = (uint64_t) gridDim.x * MM_THREADS;
	for (uint64_t r = (uint64_t) blockIdx.x * MM_THREADS + threadIdx.x; r < in.n; r += stride) {
		T x;
		u64 gi;
		if (!in.fetch(r, x, gi))
			continue;
		if (Atom<T>::is_nil(x)) {
			if (PHASE == 0 && !in.skip_nils)
				min_update(firstnil + gi, r);
			continue;
		}
		u64 hi, lo;
		in.key(x, hi, lo);
		if (PHASE == 0) {
			min_update(best_hi + gi, hi);
		} else if (PHASE == 1) {
			if (hi == __ldg(best_hi + gi))
				min_update(best_lo + gi, lo);
		} else {
			if (hi == __ldg(best_hi + gi) && (sizeof(T) < 16 || lo == __ldg(best_lo + gi)))
				min_update(row + gi, r);
		}
	}
}

template <typename T>
__global__ void minmax_finish_kernel(MMIn<T> in, const u64 *__restrict__ firstnil, const u64 *__restrict__ row, u64 ngrp,
				     oid *__restrict__ out, unsigned long long *__restrict__ nils)
{
	unsigned long long mynils = 0;
	for (u64 g = (u64) blockIdx.x * blockDim.x + threadIdx.x; g < ngrp; g += (u64) gridDim.x * blockDim.x) {
		u64 r = firstnil && firstnil[g] != MM_NONE ? firstnil[g] : row[g];
		oid o = OID_NIL;
		if (r != MM_NONE)
			o = in.cand ? in.cand[r] : in.seq + r;
		else
			mynils++;
		out[g] = o;
	}
	if (mynils)
		atomicAdd(nils, mynils);
}

template <typename T>
static bool minmax_typed(const AggrSetup &a, bool ismax, bool skip_nils, oid *out, uint64_t &nils)
{
	Ctx *c = ctx();
	const u64 ngrp = a.ngrp;
	MMIn<T> in;
	in.hseq = a.ci.hseq;
	in.seq = a.ci.seq;
	in.n = a.ci.ncand;
	if (a.ci.dense) {
		in.vals = (const T *) a.b->base + (a.ci.seq - a.ci.hseq);
		in.cand = nullptr;
	} else {
		in.vals = (const T *) a.b->base;
		in.cand = a.ci.oids;
	}
	in.gmin = a.min;
	in.gmax = a.max;
	in.gids = nullptr;
	in.gconst = true;
	if (a.g != nullptr) {
		if (bat_is_dense(a.g)) {
			in.gconst = false;
			in.gmin = a.min;	/* gid of row r = g->tseqbase + r */
			if (a.g->tseqbase != a.min) {
				set_err("BATgroupminmax: dense g not aligned with e\n");
				return false;
			}
		} else {
			in.gids = (const oid *) a.g->base;
		}
	}
	in.ismax = ismax;
	in.skip_nils = skip_nils;
	const bool niltrack = !skip_nils && !a.b->tnonil;
	const bool wide = sizeof(T) == 16;
	/* [best_hi][best_lo][firstnil][row] ngrp entries each, all initialised to MM_NONE, + nil counter */
	u64 *buf = (u64 *) dev_alloc((4 * ngrp + 1) * 8);
	if (buf == nullptr)
		return false;
	bool ok = false;
	u64 *best_hi = buf, *best_lo = buf + ngrp, *firstnil = buf + 2 * ngrp, *row = buf + 3 * ngrp;
	unsigned long long *nilcnt = (unsigned long long *) (buf + 4 * ngrp);
	const unsigned grid = grid_for(in.n, MM_THREADS, NSM * 8);
	if (!cuda_ok(cudaMemsetAsync(buf, 0xff, 4 * ngrp * 8, c->stream), "memset minmax") ||
	    !cuda_ok(cudaMemsetAsync(nilcnt, 0, 8, c->stream), "memset nils"))
		goto done;
	{
		KTimer kt("groupminmax", in.n * (sizeof(T) + (in.gids ? 8 : 0)) * (wide ? 3 : 2));
		minmax_kernel<T, 0><<<grid, MM_THREADS, 0, c->stream>>>(in, best_hi, best_lo, firstnil, row);
		COUNT_LAUNCH();
		if (wide) {
			minmax_kernel<T, 1><<<grid, MM_THREADS, 0, c->stream>>>(in, best_hi, best_lo, firstnil, row);
			COUNT_LAUNCH();
		}
		minmax_kernel<T, 2><<<grid, MM_THREADS, 0, c->stream>>>(in, best_hi, best_lo, firstnil, row);
		COUNT_LAUNCH();
	}
	minmax_finish_kernel<T><<<grid_for(ngrp, 256, NSM * 4), 256, 0, c->stream>>>(in, niltrack ? firstnil : nullptr, row, ngrp, out, nilcnt);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "minmax launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, nilcnt, 8, cudaMemcpyDeviceToHost, c->stream), "read nils") || !stream_sync())
		goto done;
	nils = c->slot_h->v[0];
	ok = true;
done:
	dev_free(buf);
	return ok;
}

static b200_bat *groupminmax(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp,
			     bool skip_nils, bool ismax, const char *name)
{
	(void) tp;	/* always TYPE_oid (gdk_aggr.c:3506-3507) */
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s)) {
		/* the reference prefixes BATgroupaggrinit's message with the function name (:3515-3518) */
		char msg[512];
		snprintf(msg, sizeof(msg), "%s: %s", name, b200_errbuf());
		set_err("%s", msg);
		return nullptr;
	}
	if (a.ci.ncand == 0 || a.ngrp == 0) {
		const oid nil = OID_NIL;
		return bat_constant(a.ngrp == 0 ? 0 : a.min, B200_oid, &nil, a.ngrp);
	}
	if (a.b->ttype == B200_void) {
		set_err("%s: void columns are not on the device path\n", name);
		return nullptr;
	}
	b200_bat *bn = bat_new(a.min, B200_oid, a.ngrp);
	if (bn == nullptr)
		return nullptr;
	uint64_t nils = 0;
	bool ok = false;
	ctx()->algo = ismax ? "groupmax: cuda key reduction" : "groupmin: cuda key reduction";
	switch (basetype(a.b->ttype)) {
	case B200_bte: ok = minmax_typed<bte>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_sht: ok = minmax_typed<sht>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_int: ok = minmax_typed<int>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_lng: ok = minmax_typed<lng>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_hge: ok = minmax_typed<hge>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_flt: ok = minmax_typed<flt>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	case B200_dbl: ok = minmax_typed<dbl>(a, ismax, skip_nils, (oid *) bn->base, nils); break;
	default:
		set_err("%s: cannot determine minimum on non-linear type %s\n", name, atomname(a.b->ttype));
	}
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = a.ngrp;
	bn->tkey = bn->tsorted = bn->trevsorted = a.ngrp <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	bn->tseqbase = OID_NIL;
	return bn;
}

extern "C" b200_bat *b200_BATgroupmin(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils)
{
	b200_clrerr();
	return groupminmax(b, g, e, s, tp, skip_nils, false, "BATgroupmin");
}

extern "C" b200_bat *b200_BATgroupmax(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils)
{
	b200_clrerr();
	return groupminmax(b, g, e, s, tp, skip_nils, true, "BATgroupmax");
}
