// operand.cuh — one side of an element-wise operator: a column read densely, a column
// gathered through an oid candidate list, or a constant (the three shapes the reference's
// BATcalc drivers hand to their typeswitch loops as (pointer, incr, canditer)).
#pragma once
#include "common.cuh"

template <typename T> struct CmpOperand {
	const T *vals;		/* dense: vals[k]; gather: vals[cand[k] - hseq]; constant: vals[0] (device copy) */
	const oid *cand;
	oid hseq;
	bool cst;
	T cval;
	__device__ __forceinline__ T get(uint64_t k) const
	{
		if (cst)
			return cval;
		return cand ? vals[__ldcs(cand + k) - hseq] : ld_stream(vals + k);
	}
};

struct CmpSide {
	const b200_bat *b;	/* NULL: constant */
	const Cand *ci;
	const void *cst;
	int type;		/* storage type */
};

template <typename T> static CmpOperand<T> make_operand(const CmpSide &s)
{
	CmpOperand<T> o;
	o.cand = nullptr;
	o.hseq = 0;
	o.cst = s.b == nullptr;
	o.vals = nullptr;
	o.cval = T();
	if (o.cst) {
		o.cval = *(const T *) s.cst;
	} else if (s.ci->dense) {
		o.vals = (const T *) s.b->base + (s.ci->seq - s.ci->hseq);
	} else {
		o.vals = (const T *) s.b->base;
		o.cand = s.ci->oids;
		o.hseq = s.ci->hseq;
	}
	return o;
}

static inline bool cst_is_nil(int t, const void *v)
{
	switch (t) {
	case B200_bte: return Atom<bte>::is_nil(*(const bte *) v);
	case B200_sht: return Atom<sht>::is_nil(*(const sht *) v);
	case B200_int: return Atom<int>::is_nil(*(const int *) v);
	case B200_lng: return Atom<lng>::is_nil(*(const lng *) v);
	case B200_hge: return Atom<hge>::is_nil(*(const hge *) v);
	case B200_flt: return Atom<flt>::is_nil(*(const flt *) v);
	case B200_dbl: return Atom<dbl>::is_nil(*(const dbl *) v);
	}
	return false;
}

