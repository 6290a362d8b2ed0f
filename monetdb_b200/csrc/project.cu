// project.cu — BATproject / BATproject2 (positional fetch-join) on the device.
// bn[i] = r[l[i] - r.hseqbase]  (gdk_project.c:590-854; kernels project1_TYPE
// :27-67, project_TYPE :81-161, project_oid :177-290).  A dense l yields a view
// on r's heap (BATslice, :623-645); a nil oid yields the type's nil; an oid
// outside r (and r2) fails with "does not match always".
// Device layout: each thread produces one 128-bit store worth of output values
// (16/w rows), reading its oids with 128-bit loads; the value reads are the
// gather.
#include "common.cuh"
#include <utility>
#include "compact.cuh"

template <int W> struct WordOf;
template <> struct WordOf<1> { typedef uint8_t type; };
template <> struct WordOf<2> { typedef uint16_t type; };
template <> struct WordOf<4> { typedef uint32_t type; };
template <> struct WordOf<8> { typedef uint64_t type; };
template <> struct WordOf<16> { typedef uint4 type; };

struct ProjSrc {
	const void *r1;		/* NULL when r1 is a dense void column */
	const void *r2;
	oid r1seq, r1end;	/* oid range of r1 */
	oid r2end;		/* end of r2's oid range (== r1end when no r2) */
	oid r1t, r2t;		/* tseqbase when r1/r2 are void */
	bool r1void, r2void;
};

template <int W, bool ALIGNED>
__global__ void __launch_bounds__(256)
project_kernel(const oid *__restrict__ l, uint64_t n, ProjSrc src, typename WordOf<W>::type nilword,
	       typename WordOf<W>::type *__restrict__ dst, unsigned int *errflag)
{
	typedef typename WordOf<W>::type word;
	constexpr int K = 16 / W;	/* rows per thread per step */
	const uint64_t nchunks = (n + K - 1) / K;
	bool bad = false;
	for (uint64_t c = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += (uint64_t) gridDim.x * blockDim.x) {
		const uint64_t i0 = c * K;
		oid o[K];
		if (i0 + K <= n) {
			if (ALIGNED && K >= 2) {
#pragma unroll
				for (int k = 0; k < K; k += 2) {
					ulonglong2 q = __ldcs(reinterpret_cast<const ulonglong2 *>(l + i0 + k));
					o[k] = q.x;
					o[k + 1 < K ? k + 1 : k] = q.y;
				}
			} else {
#pragma unroll
				for (int k = 0; k < K; k++)
					o[k] = __ldcs(l + i0 + k);
			}
			union { uint4 q; word v[K]; } u;
#pragma unroll
			for (int k = 0; k < K; k++) {
				word v = nilword;
				const oid x = o[k];
				if (x != OID_NIL) {
					if (x >= src.r1seq && x < src.r1end)
						v = static_cast<const word *>(src.r1)[x - src.r1seq];
					else if (x >= src.r1end && x < src.r2end)
						v = static_cast<const word *>(src.r2)[x - src.r1end];
					else
						bad = true;
				}
				u.v[k] = v;
			}
			*reinterpret_cast<uint4 *>(dst + i0) = u.q;
		} else {
			for (uint64_t i = i0; i < n; i++) {
				word v = nilword;
				const oid x = l[i];
				if (x != OID_NIL) {
					if (x >= src.r1seq && x < src.r1end)
						v = static_cast<const word *>(src.r1)[x - src.r1seq];
					else if (x >= src.r1end && x < src.r2end)
						v = static_cast<const word *>(src.r2)[x - src.r1end];
					else
						bad = true;
				}
				dst[i] = v;
			}
		}
	}
	if (bad)
		atomicOr(errflag, 1u);
}

// right side is a dense (void) column: the fetched "value" is computed
__global__ void __launch_bounds__(256)
project_void_kernel(const oid *__restrict__ l, uint64_t n, ProjSrc src, oid *__restrict__ dst, unsigned int *errflag)
{
	bool bad = false;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const oid x = l[i];
		oid v = OID_NIL;
		if (x != OID_NIL) {
			if (x >= src.r1seq && x < src.r1end)
				v = src.r1void ? (src.r1t == OID_NIL ? OID_NIL : src.r1t + (x - src.r1seq))
					       : static_cast<const oid *>(src.r1)[x - src.r1seq];
			else if (x >= src.r1end && x < src.r2end)
				v = src.r2void ? (src.r2t == OID_NIL ? OID_NIL : src.r2t + (x - src.r1end))
					       : static_cast<const oid *>(src.r2)[x - src.r1end];
			else
				bad = true;
		}
		dst[i] = v;
	}
	if (bad)
		atomicOr(errflag, 1u);
}

template <int W>
static void launch_project(const oid *l, uint64_t n, const ProjSrc &src, const void *nilbits, void *dst, unsigned int *errflag)
{
	typedef typename WordOf<W>::type word;
	word nilword;
	memcpy(&nilword, nilbits, W);
	constexpr int K = 16 / W;
	uint64_t nchunks = (n + K - 1) / K;
	unsigned grid = grid_for(nchunks, 256, NSM * 32);
	KTimer kt("project", n * (8 + 2 * W));
	if (((uintptr_t) l & 15) == 0)
		project_kernel<W, true><<<grid, 256, 0, ctx()->stream>>>(l, n, src, nilword, (word *) dst, errflag);
	else
		project_kernel<W, false><<<grid, 256, 0, ctx()->stream>>>(l, n, src, nilword, (word *) dst, errflag);
	COUNT_LAUNCH();
}

// BATproject through a select result that still has its scan bitmask (SHADOW_SELMASK)
template <int W>
static void launch_project_mask(const Shadow &sh, const void *r1base, int64_t rowoff, void *dst, uint64_t n)
{
	typedef typename WordOf<W>::type word;
	const word *src = (const word *) r1base + rowoff;	/* row of bit 0; only set bits are dereferenced */
	KTimer kt("project", n * (8 + 2 * W));			/* algorithmic: the API's oid list + gather */
	if (sh.a2 == CP_CHUNK)
		compact_gather_kernel<CP_CHUNK / 32, word><<<(unsigned) sh.a1, CP_CHUNK / 32, 0, ctx()->stream>>>(
			(const uint32_t *) sh.p, (const unsigned long long *) ((const char *) sh.p + sh.a3), src, (word *) dst);
	else
		compact_gather_kernel<CP_CHUNK_SMALL / 32, word><<<(unsigned) sh.a1, CP_CHUNK_SMALL / 32, 0, ctx()->stream>>>(
			(const uint32_t *) sh.p, (const unsigned long long *) ((const char *) sh.p + sh.a3), src, (word *) dst);
	COUNT_LAUNCH();
}

static b200_bat *constant_nil(uint64_t hseq, int tt, uint64_t n)
{
	return bat_constant(hseq, tt, nullptr, n);
}

// BATslice (gdk_batop.c:1825) of a column for BATproject with dense candidates.  `r` is the staged
// form of the caller's BAT and `st` says how it was staged:
//  * device-resident column of ours: a view that shares the heap and holds a share on it
//    (heap_share, the BBP share the reference's VIEWcreate takes on the parent), so owner and views
//    can be freed in any order;
//  * a wrapped device heap (b200_BATwrap): a plain alias, the wrapper's owner controls the lifetime
//    exactly as it does for the wrapped BAT itself;
//  * HOST column staged for this call: the staged device copy is a temporary, so the result takes
//    the copy over when it is the whole column and uncached, and is copied out of it otherwise.
static b200_bat *slice_view(const b200_bat *r, const Staged &st, uint64_t lo, uint64_t hi, uint64_t hseq)
{
	if (r->ttype == B200_void) {
		b200_bat *bn = bat_dense(hseq, r->tseqbase == OID_NIL ? OID_NIL : r->tseqbase + lo, hi - lo);
		return bn;
	}
	const int w = atomsize(r->ttype);
	const bool from_host = st.tmp != nullptr;
	b200_bat *bn = nullptr;
	if (from_host) {
		if (lo == 0 && st.tmp->owns) {	/* owns: a private staged copy (not the stage cache's, not mapped host memory) */
			/* take the staged heap over */
			bn = bat_alloc_desc();
			if (bn == nullptr)
				return nullptr;
			bn->base = st.tmp->base;
			bn->capacity = st.tmp->capacity;
			bn->owns = 1;
			st.tmp->base = nullptr;
		} else {
			bn = bat_new(hseq, r->ttype, hi - lo);
			if (bn == nullptr)
				return nullptr;
			if (hi > lo && !cuda_ok(cudaMemcpyAsync(bn->base, (const char *) r->base + lo * w, (hi - lo) * w,
								 cudaMemcpyDefault, ctx()->stream), "project: slice copy")) {
				b200_BATfree(bn);
				return nullptr;
			}
		}
	} else {
		bn = bat_alloc_desc();
		if (bn == nullptr)
			return nullptr;
		bn->base = (char *) r->base + lo * w;
		bn->capacity = hi - lo;
		bn->owns = 0;
		bn->producer = r->producer;	/* a view is as ready as the heap it shares */
		const void *root = r->parent ? r->parent : (r->owns ? r->base : nullptr);
		if (root) {
			heap_share(root);
			bn->parent = (void *) root;
		}
	}
	bn->ttype = r->ttype;
	bn->device = r->device;
	bn->hseqbase = hseq;
	bn->count = hi - lo;
	bn->tsorted = r->tsorted || hi - lo <= 1;
	bn->trevsorted = r->trevsorted || hi - lo <= 1;
	bn->tkey = r->tkey || hi - lo <= 1;
	bn->tnonil = r->tnonil;
	bn->tnil = 0;
	bn->oidrange_known = 0;
	if (r->oidrange_known && r->tsorted && lo == 0 && hi == r->count) {
		bn->oidrange_known = 1;
		bn->first_oid = r->first_oid;
		bn->last_oid = r->last_oid;
	}
	return bn;
}

extern "C" b200_bat *b200_BATproject2(const b200_bat *l, const b200_bat *r1, const b200_bat *r2)
{
	b200_clrerr();
	if (l == nullptr || r1 == nullptr) {
		set_err("BATproject: BAT required\n");
		return nullptr;
	}
	if (r1->ttype == B200_STR_TYPE) {
		/* gdk_project.c:681-695: gather the offsets, the result refers to the same strings */
		if (r2 != nullptr || r1->device < 0) {
			set_err("BATproject: str columns must be device-resident and single (no r2) on the device path\n");
			return nullptr;
		}
		if (l->ttype != B200_void && l->ccand == B200_CAND_NONE && !l->tnonil) {
			set_err("BATproject: nil oids into a str column are not on the device path\n");
			return nullptr;
		}
		b200_bat view;
		if (!str_as_offsets(r1, view, false, "BATproject"))
			return nullptr;
		b200_bat *bn = b200_BATproject2(l, &view, nullptr);
		if (bn == nullptr)
			return nullptr;
		if (bn->ttype == B200_void || !str_adopt_heap(bn, r1)) {
			b200_BATfree(bn);
			if (b200_errbuf()[0] == 0)
				set_err("BATproject: unexpected result shape for a str column\n");
			return nullptr;
		}
		bn->tnonil = r1->tnonil;
		bn->tnil = 0;
		bn->tsorted = bn->trevsorted = bn->tkey = bn->count <= 1;
		return bn;
	}
	Staged sl, s1, s2;
	if (!stage(l, sl) || !stage_sparse(r1, s1, r2 == nullptr && l != nullptr ? l->count : UINT64_MAX) || !stage(r2, s2))
		return nullptr;
	l = sl.b;
	r1 = s1.b;
	r2 = s2.b;
	if (r2 && r1->count == 0) {
		r1 = r2;
		r2 = nullptr;
		std::swap(s1.b, s2.b);
		std::swap(s1.tmp, s2.tmp);
		std::swap(s1.cached, s2.cached);
	}
	if (r1->ccand != B200_CAND_NONE || (r2 && r2->ccand != B200_CAND_NONE) || r1->ttype == B200_msk) {
		set_err("BATproject: complex candidate lists / msk on the right-hand side are not on the device path\n");
		return nullptr;
	}
	if (basetype(l->ttype) != B200_lng && l->ttype != B200_void) {
		set_err("BATproject: left input must be an oid/void BAT\n");
		return nullptr;
	}
	int tpe = r1->ttype == B200_void ? B200_oid : r1->ttype;
	uint64_t lcount = l->count;
	if (bat_is_dense(l) && lcount > 0) {
		oid lo = l->tseqbase, hi = l->tseqbase + lcount;
		if (lo >= r1->hseqbase && hi <= r1->hseqbase + r1->count) {
			ctx()->algo = "project: slice view";
			return slice_view(r1, s1, lo - r1->hseqbase, hi - r1->hseqbase, l->hseqbase);
		}
		if (lo < r1->hseqbase || r2 == nullptr || hi > r2->hseqbase + r2->count) {
			set_err("does not match always\n");
			return nullptr;
		}
		if (lo >= r2->hseqbase) {
			ctx()->algo = "project: slice2 view";
			return slice_view(r2, s2, lo - r2->hseqbase, hi - r2->hseqbase, l->hseqbase);
		}
	}
	/* normalise l to an oid array */
	Cand lc;
	const oid *loids = nullptr;
	bool l_sorted_known = false;
	oid lfirst = 0, llast = 0;
	if (l->ttype == B200_void || l->ccand != B200_CAND_NONE) {
		if (l->ccand == B200_CAND_NONE && l->tseqbase == OID_NIL) {
			ctx()->algo = "project: constant nil";
			return constant_nil(l->hseqbase, tpe, lcount);
		}
		if (!cand_init(lc, nullptr, l))
			return nullptr;
		lcount = lc.ncand;
	} else {
		loids = (const oid *) l->base;
		if (l->oidrange_known && l->tsorted) {
			l_sorted_known = true;
			lfirst = l->first_oid;
			llast = l->last_oid;
		}
	}
	if (lcount == 0 ||
	    (r1->ttype == B200_void && r1->tseqbase == OID_NIL && (r2 == nullptr || (r2->ttype == B200_void && r2->tseqbase == OID_NIL)))) {
		ctx()->algo = "project: constant nil";
		return constant_nil(l->hseqbase, tpe, lcount);
	}
	oid *ltmp = nullptr;
	if (loids == nullptr) {
		if (lc.dense) {
			/* dense l straddling r1 and r2: materialise the oids */
			ltmp = (oid *) dev_alloc(lcount * sizeof(oid));
			if (ltmp == nullptr)
				return nullptr;
			launch_iota(ltmp, lcount, lc.seq);
			loids = ltmp;
			l_sorted_known = true;
			lfirst = lc.seq;
			llast = lc.seq + lcount - 1;
		} else {
			loids = lc.oids;
		}
	}
	ProjSrc src;
	src.r1 = r1->base;
	src.r2 = r2 ? r2->base : nullptr;
	src.r1seq = r1->hseqbase;
	src.r1end = r1->hseqbase + r1->count;
	src.r2end = r2 ? src.r1end + r2->count : src.r1end;
	src.r1void = r1->ttype == B200_void;
	src.r2void = r2 && r2->ttype == B200_void;
	src.r1t = r1->tseqbase;
	src.r2t = r2 ? r2->tseqbase : OID_NIL;
	if (l_sorted_known && (lfirst < src.r1seq || llast >= src.r2end)) {
		if (ltmp) dev_free(ltmp);
		set_err("does not match always\n");
		return nullptr;
	}
	b200_bat *bn = bat_new(l->hseqbase, tpe, lcount);
	if (bn == nullptr) {
		if (ltmp) dev_free(ltmp);
		return nullptr;
	}
	Ctx *c = ctx();
	unsigned int *errflag = (unsigned int *) scratch_get(64);
	bool need_check = !l_sorted_known;
	if (errflag == nullptr ||
	    (need_check && !cuda_ok(cudaMemsetAsync(errflag, 0, 4, c->stream), "memset"))) {
		if (ltmp) dev_free(ltmp);
		b200_BATfree(bn);
		return nullptr;
	}
	unsigned char nilb[16];
	atom_nil_bits(tpe, nilb);
	ctx()->algo = "project: cuda gather";
	Shadow msh;
	const int w = atomsize(tpe);
	/* dense enough for streaming the whole column to beat the gather (8-byte oid + 32-byte sector per row) */
	if (r2 == nullptr && !src.r1void && l_sorted_known && loids == (const oid *) l->base &&
	    shadow_get(l->base, SHADOW_SELMASK, lcount, msh) &&
	    lcount * 4 >= msh.a1 * msh.a2) {
		const int64_t rowoff = (int64_t) msh.a0 - (int64_t) r1->hseqbase;
		switch (w) {
		case 1: launch_project_mask<1>(msh, r1->base, rowoff, bn->base, lcount); break;
		case 2: launch_project_mask<2>(msh, r1->base, rowoff, bn->base, lcount); break;
		case 4: launch_project_mask<4>(msh, r1->base, rowoff, bn->base, lcount); break;
		case 8: launch_project_mask<8>(msh, r1->base, rowoff, bn->base, lcount); break;
		case 16: launch_project_mask<16>(msh, r1->base, rowoff, bn->base, lcount); break;
		}
		ctx()->algo = "project: cuda bitmask expand";
	} else if (src.r1void || src.r2void) {
		project_void_kernel<<<grid_for(lcount, 256, NSM * 32), 256, 0, c->stream>>>(loids, lcount, src, (oid *) bn->base, errflag);
		COUNT_LAUNCH();
	} else {
		switch (atomsize(tpe)) {
		case 1: launch_project<1>(loids, lcount, src, nilb, bn->base, errflag); break;
		case 2: launch_project<2>(loids, lcount, src, nilb, bn->base, errflag); break;
		case 4: launch_project<4>(loids, lcount, src, nilb, bn->base, errflag); break;
		case 8: launch_project<8>(loids, lcount, src, nilb, bn->base, errflag); break;
		case 16: launch_project<16>(loids, lcount, src, nilb, bn->base, errflag); break;
		}
	}
	if (ltmp) dev_free(ltmp);
	if (!cuda_ok(cudaGetLastError(), "project launch")) {
		b200_BATfree(bn);
		return nullptr;
	}
	if (need_check) {
		if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[8], errflag, 4, cudaMemcpyDeviceToHost, c->stream), "flag copy") || !stream_sync()) {
			b200_BATfree(bn);
			return nullptr;
		}
		if ((unsigned) c->slot_h->v[8] & 1u) {
			b200_BATfree(bn);
			set_err("does not match always\n");
			return nullptr;
		}
	}
	bn->count = lcount;
	/* result properties, gdk_project.c:738-755 */
	bool lnonil = l->tnonil || l->ttype == B200_void || l->ccand != B200_CAND_NONE;
	bool lsorted = l->tsorted || l->ttype == B200_void, lrev = l->trevsorted, lkey = l->tkey || l->ttype == B200_void;
	bn->tnil = 0;
	if (r2) {
		bn->tnonil = lnonil && r1->tnonil && r2->tnonil;
		bn->tsorted = bn->trevsorted = bn->tkey = lcount <= 1;
	} else {
		bn->tnonil = lnonil && r1->tnonil;
		bn->tsorted = lcount <= 1 || (lsorted && r1->tsorted) || (lrev && r1->trevsorted) || r1->count <= 1;
		bn->trevsorted = lcount <= 1 || (lsorted && r1->trevsorted) || (lrev && r1->tsorted) || r1->count <= 1;
		bn->tkey = lcount <= 1 || (lkey && r1->tkey);
	}
	return bn;
}

extern "C" b200_bat *b200_BATproject(const b200_bat *l, const b200_bat *r)
{
	return b200_BATproject2(l, r, nullptr);
}

__global__ void iota_kernel(oid *dst, uint64_t n, oid start)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x)
		dst[i] = start + i;
}
void launch_iota(oid *dst, uint64_t n, oid start)
{
	if (n == 0)
		return;
	iota_kernel<<<grid_for(n, 256), 256, 0, ctx()->stream>>>(dst, n, start);
	COUNT_LAUNCH();
}
