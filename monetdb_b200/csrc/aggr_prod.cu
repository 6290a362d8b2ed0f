// aggr_prod.cu — BATgroupprod on the device (gdk_aggr.c:1188-1346 AGGR_PROD*, :1349-1571 doprod, :1575-1647).
//
// The reference multiplies every group's values IN ROW ORDER with a checked multiplication and fails the whole
// call ("22003!overflow in product aggregate.") as soon as one running product leaves [-max, max].  Its verdict
// and result depend on the order only through three per-group facts, which is what makes an exact parallel form
// possible for the integer result types:
//   * while no zero has been met every factor has magnitude >= 1, so the running magnitude never shrinks: some
//     prefix overflows  <=>  the product of |v| over all rows BEFORE the group's first zero (or first nil that
//     ends the product) exceeds max.  Magnitudes are multiplied with a saturating, order-independent CAS update;
//     rows with |v| == 1 never touch it and a group can absorb at most 63 (127) larger factors before it
//     saturates, so the atomics are naturally rare;
//   * after the first zero the product is 0 and nothing can overflow any more;
//   * nil handling (skip_nils == false).  hge results: any nil makes the group nil (AGGR_PROD_HGE tests `seen`
//     first).  Narrower results (AGGR_PROD): a nil sets the product to nil WITHOUT marking the group seen, so
//     the first non-nil value that follows resets the product to 1 — nils in front of a group's first value are
//     forgotten, a nil behind it is final.  Reproduced: pass A finds the first non-nil row, pass B the first
//     zero and the first nil that counts.
// Floating-point results (AGGR_PROD_FLOAT) multiply rounded values and can shrink again, so both the rounding
// and the overflow verdict depend on the full order: the rows are stably sorted by group id (sort.cu) and one
// thread per group replays the reference's loop — bit-identical, parallel over groups only.
#include "aggr_init.cuh"
#include <type_traits>

typedef unsigned long long u64;

#define PR_THREADS 256
#define PR_NONE (~0ULL)

template <typename T> struct PIn {
	const T *vals;		/* dense: vals[r]; else vals[cand[r] - hseq] */
	const oid *cand;
	oid hseq;
	const oid *gids;	/* NULL: dense g, gid = gmin + r */
	oid gmin, gmax;
	uint64_t n;
	bool skip_nils;
	bool gconst;		/* BATprod: every row in the one group gmin */
	bool nil_if_empty;	/* false (BATprod only): the product starts at 1 and ANY nil met is final (doprod without `seen`) */
	__device__ __forceinline__ T value(uint64_t r) const { return cand ? vals[__ldcs(cand + r) - hseq] : ld_stream(vals + r); }
	__device__ __forceinline__ bool group(uint64_t r, u64 &gi) const
	{
		const oid g = gids ? __ldcs(gids + r) : (gconst ? gmin : gmin + r);
		gi = g - gmin;
		return g >= gmin && g <= gmax;
	}
};

__device__ __forceinline__ void pr_min(u64 *p, u64 v)
{
	if (v < __ldcg(p))
		atomicMin(p, v);
}

// pass A: first non-nil row per group
template <typename T>
__global__ void __launch_bounds__(PR_THREADS)
prod_first_kernel(PIn<T> in, u64 *__restrict__ fnn)
{
	for (uint64_t r = (uint64_t) blockIdx.x * PR_THREADS + threadIdx.x; r < in.n; r += (uint64_t) gridDim.x * PR_THREADS) {
		u64 gi;
		if (in.group(r, gi) && !Atom<T>::is_nil(in.value(r)))
			pr_min(fnn + gi, r);
	}
}

// pass B: first zero, first nil that ends the product, seen flag; counts nil rows (the reference's `nils`)
template <typename T, bool HGE_RESULT>
__global__ void __launch_bounds__(PR_THREADS)
prod_cut_kernel(PIn<T> in, const u64 *__restrict__ fnn, u64 *__restrict__ fzero, u64 *__restrict__ fnil, unsigned *__restrict__ seen,
		unsigned long long *__restrict__ nilrows)
{
	unsigned mynils = 0;
	for (uint64_t r = (uint64_t) blockIdx.x * PR_THREADS + threadIdx.x; r < in.n; r += (uint64_t) gridDim.x * PR_THREADS) {
		u64 gi;
		if (!in.group(r, gi))
			continue;
		const T v = in.value(r);
		if (Atom<T>::is_nil(v)) {
			if (!in.skip_nils) {
				mynils++;
				if (HGE_RESULT || !in.nil_if_empty || (fnn != nullptr && r > __ldg(fnn + gi)))
					pr_min(fnil + gi, r);
			}
			if (HGE_RESULT && !(__ldcg(seen + (gi >> 5)) & (1u << (gi & 31))))
				atomicOr(seen + (gi >> 5), 1u << (gi & 31));	/* AGGR_PROD_HGE marks the group before the nil test */
			continue;
		}
		if (!(__ldcg(seen + (gi >> 5)) & (1u << (gi & 31))))
			atomicOr(seen + (gi >> 5), 1u << (gi & 31));
		if (v == 0)
			pr_min(fzero + gi, r);
	}
	for (int o = 16; o > 0; o >>= 1)
		mynils += __shfl_down_sync(0xffffffffu, mynils, o);
	if ((threadIdx.x & 31) == 0 && mynils)
		atomicAdd(nilrows, (unsigned long long) mynils);
}

// ---------------------------------------------------------------- saturating magnitude product
template <typename M> struct MagOps;
template <> struct MagOps<u64> {
	__device__ static __forceinline__ u64 load(const u64 *p) { return __ldcg(p); }
	__device__ static __forceinline__ u64 cas(u64 *p, u64 cmp, u64 val) { return atomicCAS(p, cmp, val); }
};
template <> struct MagOps<uhge> {
	__device__ static __forceinline__ uhge load(const uhge *p)
	{
		const ulonglong2 q = __ldcg(reinterpret_cast<const ulonglong2 *>(p));
		return ((uhge) q.y << 64) | q.x;
	}
	__device__ static __forceinline__ uhge cas(uhge *p, uhge cmp, uhge val)
	{
		u64 clo = (u64) cmp, chi = (u64) (cmp >> 64), vlo = (u64) val, vhi = (u64) (val >> 64), olo, ohi;
		asm volatile("{\n\t.reg .b128 c, v, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 v, {%4, %5};\n\t"
			     "atom.global.cas.b128 o, [%6], c, v;\n\tmov.b128 {%0, %1}, o;\n\t}"
			     : "=l"(olo), "=l"(ohi)
			     : "l"(clo), "l"(chi), "l"(vlo), "l"(vhi), "l"(p)
			     : "memory");
		return ((uhge) ohi << 64) | olo;
	}
};

// pass C: multiply the magnitudes of the rows in front of the group's cut, count the negative ones.
// M = u64 for results up to lng, uhge for hge; `limit` = the result type's maximum; ~0 = saturated (overflow).
template <typename T, typename M>
__global__ void __launch_bounds__(PR_THREADS)
prod_mul_kernel(PIn<T> in, const u64 *__restrict__ fzero, const u64 *__restrict__ fnil, M *__restrict__ mag, unsigned *__restrict__ neg,
		M limit)
{
	const M SAT = ~(M) 0;
	unsigned myneg = 0;	/* one group: counted per thread, one atomic per warp at the end */
	for (uint64_t r = (uint64_t) blockIdx.x * PR_THREADS + threadIdx.x; r < in.n; r += (uint64_t) gridDim.x * PR_THREADS) {
		u64 gi;
		if (!in.group(r, gi))
			continue;
		const T v = in.value(r);
		if (Atom<T>::is_nil(v) || v == 0)
			continue;
		const u64 z = __ldg(fzero + gi), nl = fnil ? __ldg(fnil + gi) : PR_NONE;
		if (r > (z < nl ? z : nl))
			continue;	/* behind the first zero / final nil: the running product is 0 or nil there */
		if (v < 0) {
			if (in.gconst)
				myneg++;
			else
				atomicAdd(neg + gi, 1u);
		}
		const M f = v < 0 ? (M) 0 - (M) v : (M) v;
		if (f == 1)
			continue;
		M cur = MagOps<M>::load(mag + gi);
		while (cur != SAT) {
			const M nxt = cur > limit / f ? SAT : cur * f;
			const M old = MagOps<M>::cas(mag + gi, cur, nxt);
			if (old == cur)
				break;
			cur = old;
		}
	}
	if (in.gconst) {
		for (int o = 16; o > 0; o >>= 1)
			myneg += __shfl_down_sync(0xffffffffu, myneg, o);
		if ((threadIdx.x & 31) == 0 && (myneg & 1u))
			atomicAdd(neg, 1u);	/* only the parity matters */
	}
}

// finish: signed results, overflow flag, nil count
template <typename TR, typename M>
__global__ void __launch_bounds__(256)
prod_finish_kernel(u64 ngrp, const u64 *__restrict__ fzero, const u64 *__restrict__ fnil, const unsigned *__restrict__ seen,
		   const M *__restrict__ mag, const unsigned *__restrict__ neg, TR *__restrict__ out, unsigned long long *__restrict__ state,
		   bool nil_if_empty = true)
{
	unsigned long long nils = 0, unseen = 0;
	bool ovf = false;
	for (u64 g = (u64) blockIdx.x * blockDim.x + threadIdx.x; g < ngrp; g += (u64) gridDim.x * blockDim.x) {
		const bool s = (seen[g >> 5] >> (g & 31)) & 1u;
		const M m = mag[g];
		TR res;
		if (m == ~(M) 0)
			ovf = true;
		if (fnil && fnil[g] != PR_NONE) {
			res = Atom<TR>::nil();		/* a final nil always belongs to a seen group (see prod_cut_kernel) */
		} else if (!s) {
			res = nil_if_empty ? Atom<TR>::nil() : (TR) 1;
			unseen++;
		} else if (fzero[g] != PR_NONE) {
			res = 0;
		} else {
			res = (neg[g] & 1u) ? (TR) ((M) 0 - m) : (TR) m;
		}
		if (Atom<TR>::is_nil(res))
			nils++;
		out[g] = res;
	}
	if (ovf)
		atomicOr(&state[0], 1ULL);
	if (nils)
		atomicAdd(&state[3], nils);	/* nil RESULTS; state[1] holds the nil ROWS pass B counted (the reference's `nils`) */
	if (unseen)
		atomicAdd(&state[2], unseen);
}

template <typename M> __global__ void prod_fill_kernel(M *__restrict__ mag, u64 ngrp)
{
	for (u64 g = (u64) blockIdx.x * blockDim.x + threadIdx.x; g < ngrp; g += (u64) gridDim.x * blockDim.x)
		mag[g] = 1;
}

// ---------------------------------------------------------------- floating-point results: ordered replay
// sg = the group ids in ascending order, ord = the candidate positions (+ ghseq) in that order, stable: inside a
// group the rows keep their order.  One thread per group runs AGGR_PROD_FLOAT's loop (gdk_aggr.c:1308-1346).
template <typename T1, typename T2>
__global__ void __launch_bounds__(128)
prod_float_kernel(PIn<T1> in, const oid *__restrict__ sg, const oid *__restrict__ ord, oid ghseq, u64 ngrp, T2 *__restrict__ out,
		  unsigned long long *__restrict__ state)
{
	unsigned long long nilrows = 0, unseen = 0;
	bool ovf = false;
	for (u64 g = (u64) blockIdx.x * blockDim.x + threadIdx.x; g < ngrp; g += (u64) gridDim.x * blockDim.x) {
		const oid want = in.gmin + g;
		uint64_t lo = 0, hi = sg == nullptr ? 0 : in.n;	/* sg == NULL (BATprod): one group, the rows in candidate order */
		while (lo < hi) {
			const uint64_t mid = (lo + hi) >> 1;
			if (sg[mid] == OID_NIL || sg[mid] < want)	/* BATsort puts nil group ids first */
				lo = mid + 1;
			else
				hi = mid;
		}
		T2 prod = in.nil_if_empty ? Atom<T2>::nil() : (T2) 1;
		bool seen = !in.nil_if_empty;
		for (uint64_t i = lo; i < in.n && (sg == nullptr || sg[i] == want) && !ovf; i++) {
			const T1 v = in.value(sg == nullptr ? i : ord[i] - ghseq);
			if (Atom<T1>::is_nil(v)) {
				if (!in.skip_nils) {
					prod = Atom<T2>::nil();
					nilrows++;
				}
				continue;
			}
			if (!seen) {
				seen = true;
				prod = 1;
			}
			if (!Atom<T2>::is_nil(prod)) {
				/* ABSOLUTE(vals[i]) > 1 && GDK_<T2>_max / ABSOLUTE(vals[i]) < ABSOLUTE(prods[gid]) with C's usual
				 * arithmetic conversions: an integer or flt operand is converted to the type of the other side */
				const T1 av = v < 0 ? -v : v;
				const T2 ap = prod < 0 ? -prod : prod;
				typedef typename std::conditional<Atom<T1>::is_float && (sizeof(T1) > sizeof(T2)), T1, T2>::type TC;
				if (av > 1 && (TC) Atom<T2>::maxval() / (TC) av < (TC) ap)
					ovf = true;
				else
					prod = (T2) ((TC) prod * (TC) v);
			}
		}
		if (!seen)
			unseen++;
		out[g] = prod;
	}
	if (ovf)
		atomicOr(&state[0], 1ULL);
	if (nilrows)
		atomicAdd(&state[1], nilrows);
	if (unseen)
		atomicAdd(&state[2], unseen);
}

// ---------------------------------------------------------------- host side
static thread_local bool tl_prod_nil_if_empty = true;	/* false only inside b200_BATprod(..., nil_if_empty = false) */
#define PROD_SCALAR_FLOAT_MAX ((uint64_t) 16 << 20)	/* rows one thread may replay for an order-dependent scalar product */

template <typename T> static PIn<T> prod_input(const AggrSetup &a, bool skip_nils)
{
	PIn<T> in;
	in.hseq = a.ci.hseq;
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
	in.gconst = a.g == nullptr;
	in.gids = (a.g == nullptr || bat_is_dense(a.g)) ? nullptr : (const oid *) a.g->base;
	in.skip_nils = skip_nils;
	in.nil_if_empty = tl_prod_nil_if_empty;
	return in;
}

// st: [0] overflow, [1] nil rows (reference's `nils`), [2] groups never seen
template <typename T, typename TR>
static bool prod_int(const AggrSetup &a, bool skip_nils, void *out, unsigned long long st[3])
{
	typedef typename std::conditional<sizeof(TR) == 16, uhge, u64>::type M;
	constexpr bool HGE_RESULT = sizeof(TR) == 16;
	Ctx *c = ctx();
	const u64 ngrp = a.ngrp;
	const PIn<T> in = prod_input<T>(a, skip_nils);
	const bool nilpath = !skip_nils && !a.b->tnonil;
	const bool need_first = nilpath && !HGE_RESULT && in.nil_if_empty;
	const size_t words = (ngrp + 31) / 32;
	/* [mag: ngrp x M][fzero][fnil][fnn: ngrp x u64 each][neg: ngrp x u32][seen: words x u32][state 4 x u64] */
	const size_t off_fz = ngrp * sizeof(M), off_fnil = off_fz + ngrp * 8, off_fnn = off_fnil + ngrp * 8, off_neg = off_fnn + ngrp * 8,
		     off_seen = off_neg + ((ngrp * 4 + 7) & ~(size_t) 7), off_state = off_seen + ((words * 4 + 7) & ~(size_t) 7);
	char *buf = (char *) dev_alloc(off_state + 32);
	if (buf == nullptr)
		return false;
	M *mag = (M *) buf;
	u64 *fzero = (u64 *) (buf + off_fz), *fnil = (u64 *) (buf + off_fnil), *fnn = (u64 *) (buf + off_fnn);
	unsigned *neg = (unsigned *) (buf + off_neg), *seen = (unsigned *) (buf + off_seen);
	unsigned long long *state = (unsigned long long *) (buf + off_state);
	bool ok = false;
	const unsigned grid = grid_for(in.n, PR_THREADS, NSM * 8);
	if (!cuda_ok(cudaMemsetAsync(buf + off_fz, 0xff, 3 * ngrp * 8, c->stream), "memset prod") ||
	    !cuda_ok(cudaMemsetAsync(buf + off_neg, 0, off_state + 32 - off_neg, c->stream), "memset prod2"))
		goto done;
	prod_fill_kernel<M><<<grid_for(ngrp, 256, NSM * 4), 256, 0, c->stream>>>(mag, ngrp);
	COUNT_LAUNCH();
	{
		KTimer kt("groupprod", in.n * (sizeof(T) + (in.gids ? 8 : 0)) * (need_first ? 3 : 2));
		if (need_first) {
			prod_first_kernel<T><<<grid, PR_THREADS, 0, c->stream>>>(in, fnn);
			COUNT_LAUNCH();
		}
		prod_cut_kernel<T, HGE_RESULT><<<grid, PR_THREADS, 0, c->stream>>>(in, need_first ? fnn : nullptr, fzero, fnil, seen, state + 1);
		COUNT_LAUNCH();
		prod_mul_kernel<T, M><<<grid, PR_THREADS, 0, c->stream>>>(in, fzero, nilpath ? fnil : nullptr, mag, neg, (M) Atom<TR>::maxval());
		COUNT_LAUNCH();
	}
	prod_finish_kernel<TR, M><<<grid_for(ngrp, 256, NSM * 4), 256, 0, c->stream>>>(ngrp, fzero, nilpath ? fnil : nullptr, seen, mag, neg,
										    (TR *) out, state, in.nil_if_empty);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "prod launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 32, cudaMemcpyDeviceToHost, c->stream), "read prod state") || !stream_sync())
		goto done;
	st[0] = c->slot_h->v[0];
	st[1] = c->slot_h->v[1];
	st[2] = c->slot_h->v[2];
	ok = true;
done:
	dev_free(buf);
	return ok;
}

template <typename T1, typename T2>
static bool prod_float(const AggrSetup &a, bool skip_nils, void *out, unsigned long long st[3])
{
	Ctx *c = ctx();
	const PIn<T1> in = prod_input<T1>(a, skip_nils);
	b200_bat *sg = nullptr, *ord = nullptr;
	if (a.g == nullptr) {
		if (in.n > PROD_SCALAR_FLOAT_MAX) {
			set_err("BATprod: a floating-point product is order-dependent (one sequential pass): more than %llu rows are not on the device path\n",
				(unsigned long long) PROD_SCALAR_FLOAT_MAX);
			return false;
		}
	} else if (b200_BATsort(&sg, &ord, nullptr, a.g, nullptr, nullptr, false, false, true) != B200_SUCCEED)
		return false;
	bool ok = false;
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	if (sg != nullptr && (sg->ttype == B200_void || ord->ttype == B200_void)) {
		set_err("BATgroupprod: unexpected dense sort result\n");
		goto done;
	}
	if (state == nullptr || !cuda_ok(cudaMemsetAsync(state, 0, 32, c->stream), "memset prod state"))
		goto done;
	{
		KTimer kt("groupprod_float", in.n * (sizeof(T1) + 16));
		prod_float_kernel<T1, T2><<<grid_for(a.ngrp, 128, NSM * 16), 128, 0, c->stream>>>(in, sg ? (const oid *) sg->base : nullptr,
												    ord ? (const oid *) ord->base : nullptr,
												    a.g ? a.g->hseqbase : 0, a.ngrp, (T2 *) out, state);
		COUNT_LAUNCH();
	}
	if (!cuda_ok(cudaGetLastError(), "prod launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 32, cudaMemcpyDeviceToHost, c->stream), "read prod state") || !stream_sync())
		goto done;
	st[0] = c->slot_h->v[0];
	st[1] = c->slot_h->v[1];
	st[2] = c->slot_h->v[2];
	ok = true;
done:
	b200_BATfree(sg);
	b200_BATfree(ord);
	return ok;
}

// the reference's type table, gdk_aggr.c:1366-1535: -1 unsupported, 0 failed, 1 ok
static int prod_dispatch(const AggrSetup &a, int t1, int t2, bool skip_nils, void *out, unsigned long long st[3])
{
#define PI(T, TR) return prod_int<T, TR>(a, skip_nils, out, st) ? 1 : 0
#define PF(T, TR) return prod_float<T, TR>(a, skip_nils, out, st) ? 1 : 0
	switch (t2) {
	case B200_bte:
		if (t1 == B200_bte) PI(bte, bte);
		break;
	case B200_sht:
		if (t1 == B200_bte) PI(bte, sht);
		if (t1 == B200_sht) PI(sht, sht);
		break;
	case B200_int:
		if (t1 == B200_bte) PI(bte, int);
		if (t1 == B200_sht) PI(sht, int);
		if (t1 == B200_int) PI(int, int);
		break;
	case B200_lng:
		if (t1 == B200_bte) PI(bte, lng);
		if (t1 == B200_sht) PI(sht, lng);
		if (t1 == B200_int) PI(int, lng);
		if (t1 == B200_lng) PI(lng, lng);
		break;
	case B200_hge:
		if (t1 == B200_bte) PI(bte, hge);
		if (t1 == B200_sht) PI(sht, hge);
		if (t1 == B200_int) PI(int, hge);
		if (t1 == B200_lng) PI(lng, hge);
		if (t1 == B200_hge) PI(hge, hge);
		break;
	case B200_flt:
		if (t1 == B200_bte) PF(bte, flt);
		if (t1 == B200_sht) PF(sht, flt);
		if (t1 == B200_int) PF(int, flt);
		if (t1 == B200_lng) PF(lng, flt);
		if (t1 == B200_hge) PF(hge, flt);
		if (t1 == B200_flt) PF(flt, flt);
		break;
	case B200_dbl:
		if (t1 == B200_bte) PF(bte, dbl);
		if (t1 == B200_sht) PF(sht, dbl);
		if (t1 == B200_int) PF(int, dbl);
		if (t1 == B200_lng) PF(lng, dbl);
		if (t1 == B200_hge) PF(hge, dbl);
		if (t1 == B200_flt) PF(flt, dbl);
		if (t1 == B200_dbl) PF(dbl, dbl);
		break;
	}
#undef PI
#undef PF
	return -1;
}

extern "C" b200_bat *b200_BATgroupprod(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils)
{
	b200_clrerr();
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s))
		return nullptr;		/* the reference reports BATgroupaggrinit's text as is (:1590-1593) */
	if (a.g == nullptr) {
		set_err("b and g must be aligned\n");
		return nullptr;
	}
	const int t2 = basetype(tp);
	unsigned char nilb[16];
	if (atomsize(t2) <= 0) {
		set_err("BATgroupprod: type combination (mul(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
		return nullptr;
	}
	atom_nil_bits(t2, nilb);
	if (a.ci.ncand == 0 || a.ngrp == 0)
		return bat_constant(a.ngrp == 0 ? 0 : a.min, tp, nilb, a.ngrp);
	const bool gdense = bat_is_dense(a.g);
	if ((a.e == nullptr || (a.e->count == a.ci.ncand && a.e->hseqbase == a.ci.rhseq)) && (gdense || (a.g->tkey && a.g->tnonil))) {
		/* singleton groups: the inputs in the result type (:1602-1608) */
		return b200_BATconvert(b, s, tp, 0, 0, 0);
	}
	if (a.b->ttype == B200_void || gdense) {
		set_err("BATgroupprod: void columns / dense group ids with unaligned extents are not on the device path\n");
		return nullptr;
	}
	b200_bat *bn = bat_new(a.min, tp, a.ngrp);
	if (bn == nullptr)
		return nullptr;
	unsigned long long st[3] = {0, 0, 0};
	ctx()->algo = (t2 == B200_flt || t2 == B200_dbl) ? "groupprod: cuda ordered replay per group" : "groupprod: cuda saturating magnitude product";
	const int rc = prod_dispatch(a, basetype(a.b->ttype), t2, skip_nils, bn->base, st);
	if (rc <= 0 || (st[0] & 1)) {
		b200_BATfree(bn);
		if (rc < 0)
			set_err("BATgroupprod: type combination (mul(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
		else if (rc > 0)
			set_err("22003!overflow in product aggregate.\n");
		return nullptr;
	}
	/* gdk_aggr.c:1539-1552, 1627-1633: nils = nil rows met (skip_nils == false), else 1 if a group stayed empty */
	const bool nils = st[1] != 0 || st[2] != 0;
	bn->count = a.ngrp;
	bn->tkey = bn->tsorted = bn->trevsorted = a.ngrp <= 1;
	bn->tnil = nils;
	bn->tnonil = !nils;
	return bn;
}

// BATprod (gdk_aggr.c:1650): the product of all candidates as one value in *res (host memory, atom of type tp)
extern "C" b200_return b200_BATprod(void *res, int tp, const b200_bat *b, const b200_bat *s, bool skip_nils, bool nil_if_empty)
{
	b200_clrerr();
	if (res == nullptr) {
		set_err("BATprod: result pointer required\n");
		return B200_FAIL;
	}
	AggrSetup a;
	if (!aggr_init(a, b, nullptr, nullptr, s))
		return B200_FAIL;
	const int t2 = basetype(tp), w = atomsize(t2);
	if (w <= 0 || t2 == B200_void) {
		set_err("type combination (prod(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
		return B200_FAIL;
	}
	/* *res = nil_if_empty ? nil : 1 (:1666-1694) */
	unsigned char init[16];
	atom_nil_bits(t2, init);
	if (!nil_if_empty) {
		memset(init, 0, sizeof(init));
		if (t2 == B200_flt) {
			const flt one = 1;
			memcpy(init, &one, 4);
		} else if (t2 == B200_dbl) {
			const dbl one = 1;
			memcpy(init, &one, 8);
		} else
			init[0] = 1;
	}
	memcpy(res, init, w);
	if (a.ci.ncand == 0)
		return B200_SUCCEED;
	if (a.b->ttype == B200_void) {
		set_err("BATprod: void columns are not on the device path\n");
		return B200_FAIL;
	}
	Ctx *c = ctx();
	void *out = dev_alloc(16);
	if (out == nullptr)
		return B200_FAIL;
	unsigned long long st[3] = {0, 0, 0};
	tl_prod_nil_if_empty = nil_if_empty;
	c->algo = (t2 == B200_flt || t2 == B200_dbl) ? "prod: cuda ordered replay" : "prod: cuda saturating magnitude product";
	const int rc = prod_dispatch(a, basetype(a.b->ttype), t2, skip_nils, out, st);
	tl_prod_nil_if_empty = true;
	bool ok = rc > 0 && !(st[0] & 1);
	if (rc < 0)
		set_err("BATprod: type combination (mul(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
	else if (rc > 0 && (st[0] & 1))
		set_err("22003!overflow in product aggregate.\n");
	if (ok)
		ok = cuda_ok(cudaMemcpyAsync(&c->slot_h->v[8], out, 16, cudaMemcpyDeviceToHost, c->stream), "read product") && stream_sync();
	dev_free(out);
	if (!ok)
		return B200_FAIL;
	memcpy(res, &c->slot_h->v[8], w);
	return B200_SUCCEED;
}
