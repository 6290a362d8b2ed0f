// join_void.cu — joins with a DENSE side: mergejoin_void (gdk_join.c:563-1010), the foreign-key fetch join of
// SURVEY.md §8f rank 1.  When r is a dense (void) column and its candidate list, if any, is dense too, nothing has
// to be searched: a value v of l matches iff lo <= v < hi, where [lo, hi) are the tail values of r the candidates
// leave, and its partner is the row at oid v - r.tseqbase + r.hseqbase.  The reference does this with BATselect on
// l plus index arithmetic (or, with nil_on_miss, with three hand-written loops); here it is one ballot-mask
// compaction over l's candidates (select + arithmetic fused) or, for nil_on_miss, one element-wise kernel.
//   inner / semi        (o, v - tseq + hseq) for every matching candidate o of l, in candidate order
//   only_misses         the candidates that do not match — nil values of l included (:618-636)
//   nil_on_miss         every candidate, beside its partner or nil
// r has no nils and is a key, so semi / max_one / min_one / nil_matches change nothing (the reference does not even
// pass them on).  A dense side that cannot take this path (non-dense candidates on r, dense l against a real
// column) is materialised as an oid column and goes through the hash join like any other column.
#include "common.cuh"
#include "compact.cuh"

enum { FV_NIL_ON_MISS = 2, FV_ONLY_MISSES = 8 };	/* = JM_NIL_ON_MISS / JM_ONLY_MISSES of join.cu */

// value of l at candidate idx: a real oid column, or the virtual tail of a dense one
struct FetchL {
	const oid *vals;	/* NULL: dense l, value = tseq + (o - hseq) */
	oid tseq, hseq;
	const oid *cand;	/* NULL: dense candidates from seq */
	oid seq;
	__device__ __forceinline__ oid at(uint64_t idx) const { return cand ? cand[idx] : seq + idx; }
	__device__ __forceinline__ oid value(oid o) const { return vals ? vals[o - hseq] : tseq + (o - hseq); }
};

struct FetchPred {
	FetchL l;
	oid lo, hi;
	bool misses;
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &o) const
	{
		o = l.cand ? __ldcs(l.cand + idx) : l.seq + idx;
		const oid v = l.value(o);
		const bool hit = v != OID_NIL && v >= lo && v < hi;
		return hit != misses;
	}
	__device__ __forceinline__ oid at(uint64_t idx) const { return l.at(idx); }
};

// r2 of the matches: partner oid of every oid in r1 (dense r1: first + i)
__global__ void __launch_bounds__(256)
fetch_partner_kernel(const oid *__restrict__ r1, oid first, uint64_t n, FetchL l, long long delta, oid *__restrict__ r2)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const oid o = r1 ? r1[i] : first + i;
		r2[i] = (oid) ((long long) l.value(o) + delta);
	}
}

// nil_on_miss: every candidate gets its partner or nil; state[0] = number of nils
__global__ void __launch_bounds__(256)
fetch_outer_kernel(FetchL l, uint64_t n, oid lo, oid hi, long long delta, oid *__restrict__ r2, unsigned long long *__restrict__ state)
{
	unsigned nils = 0;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const oid v = l.value(l.at(i));
		const bool hit = v != OID_NIL && v >= lo && v < hi;
		r2[i] = hit ? (oid) ((long long) v + delta) : OID_NIL;
		nils += !hit;
	}
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[0], (unsigned long long) nils);
}

// lb: oid column or dense void; rb: dense void with dense candidates (the caller checked).  mode: FV_* bits.
bool join_dense_right(b200_bat **r1p, b200_bat **r2p, const b200_bat *lb, const b200_bat *rb, const Cand &lci, const Cand &rci,
		      unsigned mode)
{
	Ctx *c = ctx();
	c->algo = "join: cuda mergejoin_void (dense right-hand side)";
	/* the tail values of r that can match: [lo, hi) (gdk_join.c:597-604) */
	const oid lo = rb->tseqbase + (rci.seq - rb->hseqbase), hi = lo + rci.ncand;
	const long long delta = (long long) rb->hseqbase - (long long) rb->tseqbase;
	FetchL fl;
	fl.vals = lb->ttype == B200_void ? nullptr : (const oid *) lb->base;
	fl.tseq = lb->tseqbase;
	fl.hseq = lci.hseq;
	fl.cand = lci.dense ? nullptr : lci.oids;
	fl.seq = lci.seq;
	const uint64_t n = lci.ncand;
	b200_bat *r1 = nullptr, *r2 = nullptr;
	if (mode & FV_NIL_ON_MISS) {
		r1 = cand_to_bat(lci);
		if (r1 == nullptr)
			return false;
		if (r2p) {
			r2 = bat_new(0, B200_oid, n);
			unsigned long long *state = (unsigned long long *) scratch_get(64);
			bool ok = r2 != nullptr && state != nullptr && cuda_ok(cudaMemsetAsync(state, 0, 8, c->stream), "memset");
			if (ok) {
				KTimer kt("join_fetch", n * (8 + (fl.vals ? 8 : 0) + (fl.cand ? 8 : 0)));
				fetch_outer_kernel<<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(fl, n, lo, hi, delta, (oid *) r2->base, state);
				COUNT_LAUNCH();
			}
			ok = ok && cuda_ok(cudaGetLastError(), "fetch launch") &&
			     cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 8, cudaMemcpyDeviceToHost, c->stream), "read") && stream_sync();
			if (!ok) {
				b200_BATfree(r1);
				if (r2)
					b200_BATfree(r2);
				return false;
			}
			const uint64_t nils = c->slot_h->v[0];
			r2->count = n;
			r2->tnil = nils != 0;
			r2->tnonil = nils == 0;
			r2->tsorted = r2->trevsorted = r2->tkey = n <= 1;
			r2->tseqbase = OID_NIL;
		}
		*r1p = r1;
		if (r2p)
			*r2p = r2;
		return true;
	}
	/* matches (or misses): select + compaction in one pass */
	oid *dst = (oid *) dev_alloc(n * sizeof(oid));
	if (dst == nullptr)
		return false;
	CompactResult res;
	FetchPred pred{fl, lo, hi, (mode & FV_ONLY_MISSES) != 0};
	if (!run_compact<1, FetchPred>(n, pred, dst, res)) {
		dev_free(dst);
		return false;
	}
	r1 = oidlist_to_bat(dst, n, res);	/* takes dst over; dense results become void */
	if (r1 == nullptr)
		return false;
	if (r2p && !(mode & FV_ONLY_MISSES)) {
		const uint64_t m = r1->count;
		if (m == 0) {
			r2 = bat_dense(0, 0, 0);
		} else if (r1->ttype == B200_void && lb->ttype == B200_void) {
			/* dense matches of a dense l: the partners are dense too (gdk_join.c:643-649) */
			r2 = bat_dense(0, (oid) ((long long) (lb->tseqbase + (r1->tseqbase - lb->hseqbase)) + delta), m);
		} else {
			r2 = bat_new(0, B200_oid, m);
			if (r2 != nullptr) {
				KTimer kt("join_fetch", m * (16 + (fl.vals ? 8 : 0)));
				fetch_partner_kernel<<<grid_for(m, 256, NSM * 32), 256, 0, c->stream>>>(
					r1->ttype == B200_void ? nullptr : (const oid *) r1->base, r1->tseqbase, m, fl, delta, (oid *) r2->base);
				COUNT_LAUNCH();
				if (!cuda_ok(cudaGetLastError(), "fetch launch")) {
					b200_BATfree(r2);
					r2 = nullptr;
				} else {
					r2->count = m;
					r2->tnonil = 1;
					r2->tnil = 0;
					/* the partners follow l's values: as ordered / unique as l is (gdk_join.c:677-679) */
					const bool lv = lb->ttype == B200_void;
					r2->tsorted = m <= 1 || lv || lb->tsorted;
					r2->trevsorted = m <= 1 || (!lv && lb->trevsorted);
					r2->tkey = m <= 1 || lv || lb->tkey;
					r2->tseqbase = OID_NIL;
				}
			}
		}
		if (r2 == nullptr) {
			b200_BATfree(r1);
			return false;
		}
	}
	*r1p = r1;
	if (r2p)
		*r2p = r2;
	return true;
}

// a dense (void) column as a real oid column: the general fallback for joins with a dense side
b200_bat *join_materialise_dense(const b200_bat *b)
{
	b200_bat *bn = bat_new(b->hseqbase, B200_oid, b->count);
	if (bn == nullptr)
		return nullptr;
	launch_iota((oid *) bn->base, b->count, b->tseqbase);
	bn->count = b->count;
	bn->tsorted = bn->tkey = bn->tnonil = 1;
	bn->trevsorted = b->count <= 1;
	bn->tnil = 0;
	bn->tseqbase = OID_NIL;
	return bn;
}
