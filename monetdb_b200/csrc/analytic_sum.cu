// analytic_sum.cu — GDKanalyticalsum (gdk_analytic_func.c:1959-2022) and the argument
// handling shared by the analytic aggregates.
#include "analytic_aggr.cuh"
#include "analytic_prefix.cuh"

bool an_frame_prepare(const char *fn, b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b, const b200_bat *s,
		      const b200_bat *e, int rtype, int frame_type, FrameStage &fs, FrameArgs &fa)
{
	b200_clrerr();
	if (r == nullptr || b == nullptr) {
		set_err("%s: BAT required\n", fn);
		return false;
	}
	if (!stage(p, fs.p) || !stage(o, fs.o) || !stage(b, fs.b) || !stage(s, fs.s) || !stage(e, fs.e))
		return false;
	fa.p = fs.p.b;
	fa.o = fs.o.b;
	fa.b = fs.b.b;
	fa.s = fs.s.b;
	fa.e = fs.e.b;
	fa.frame_type = frame_type;
	fa.n = fa.b->count;
	const uint64_t n = fa.n;
	if (r->device >= 0)
		shadow_drop(r->base);
	if (r->device < 0 || atomsize(r->ttype) != atomsize(rtype) || r->capacity < n || (r->base == nullptr && n > 0)) {
		set_err("%s: result must be a device %s BAT with capacity >= BATcount(b)\n", fn, atomname(rtype));
		return false;
	}
	if (n > 0 && (fa.b->base == nullptr || atomsize(fa.b->ttype) <= 0)) {
		set_err("%s: input must be a materialised fixed-width column\n", fn);
		return false;
	}
	if (!an_check_bit(fa.p, "p", n) || !an_check_bit(fa.o, "o", n))
		return false;
	if (frame_type < 3 || frame_type > 6) {
		/* explicit frames: start and end oid columns from GDKanalyticalwindowbounds */
		if (fa.s == nullptr || fa.e == nullptr || fa.s->count < n || fa.e->count < n ||
		    (n > 0 && (fa.s->base == nullptr || fa.e->base == nullptr)) ||
		    basetype(fa.s->ttype) != B200_lng || basetype(fa.e->ttype) != B200_lng) {
			set_err("%s: frame types 0-2 need materialised start and end oid columns aligned with b\n", fn);
			return false;
		}
	}
	return true;
}

b200_return an_frame_done(b200_bat *r, uint64_t n, bool has_nils)
{
	r->count = n;
	r->tnonil = !has_nils;
	r->tnil = has_nils;
	r->tsorted = r->trevsorted = r->tkey = n <= 1;
	return B200_SUCCEED;
}

// integer -> hge: prefix sums (analytic_prefix.cuh)
template <typename T1>
static bool sum_prefix_run(const FrameArgs &fa, hge *rb, unsigned long long st[2])
{
	Ctx *c = ctx();
	const uint64_t n = fa.n;
	const bool peers = fa.frame_type == 3 || fa.frame_type == 4;
	const bool framed = fa.frame_type < 3 || fa.frame_type > 6;
	SegList S, O;
	if (fa.frame_type != 6 && !an_partitions(fa.p, nullptr, n, S))
		return false;
	if (peers && !an_partitions(fa.p, fa.o, n, O))
		return false;
	PrefixArrays pa;
	if (!an_prefix_get<T1>(fa.b, n, pa))
		return false;
	unsigned long long *state = an_state_init();
	if (state == nullptr)
		return false;
	{
		KTimer kt("an_sum", n * ((framed ? 16 : 0) + 32 + 16));
		const unsigned grid = (unsigned) ((n + AN_TILE - 1) / AN_TILE);
		const oid *fs = framed ? (const oid *) fa.s->base : nullptr, *fe = framed ? (const oid *) fa.e->base : nullptr;
		if (peers)
			an_frame_prefix_kernel<true><<<grid, AN_THREADS, 0, c->stream>>>(S.S, S.cnt, O.S, O.cnt, n, fs, fe, fa.frame_type, pa.P, pa.C, rb, state);
		else
			an_frame_prefix_kernel<false><<<grid, AN_THREADS, 0, c->stream>>>(S.S, S.cnt, nullptr, 0, n, fs, fe, fa.frame_type, pa.P, pa.C, rb, state);
		COUNT_LAUNCH();
		ctx()->algo = "analyticalsum: cuda prefix sums";
	}
	return an_read_state(state, st);
}

template <typename T1, typename T2>
static bool sum_run(const FrameArgs &fa, void *rb, unsigned long long st[2])
{
	if constexpr (sizeof(T2) == 16 && sizeof(T1) <= 8 && !Atom<T1>::is_float)
		return sum_prefix_run<T1>(fa, (hge *) rb, st);
	ctx()->algo = "analyticalsum: cuda fan-out-16 segment tree";
	return an_run<SumOp<T1, T2>>(fa, (T2 *) rb, st, "an_sum");
}

extern "C" b200_return b200_GDKanalyticalsum(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					     const b200_bat *s, const b200_bat *e, int tp1, int tp2, int frame_type)
{
	FrameStage fs;
	FrameArgs fa;
	if (!an_frame_prepare("GDKanalyticalsum", r, p, o, b, s, e, tp2, frame_type, fs, fa))
		return B200_FAIL;
	const int t1 = basetype(tp1), t2 = basetype(tp2);
	if (fa.n == 0)
		return an_frame_done(r, 0, false);
	if (atomsize(fa.b->ttype) != atomsize(t1)) {
		set_err("GDKanalyticalsum: tp1 does not match the input column\n");
		return B200_FAIL;
	}
	unsigned long long st[2] = {0, 0};
	int rc = -1;	/* -1 unsupported, 0 failed, 1 ok */
	void *rb = r->base;
#define SUM(T1, T2) rc = sum_run<T1, T2>(fa, rb, st) ? 1 : 0
	switch (t2) {
	case B200_bte:
		if (t1 == B200_bte) SUM(bte, bte);
		break;
	case B200_sht:
		if (t1 == B200_bte) SUM(bte, sht);
		else if (t1 == B200_sht) SUM(sht, sht);
		break;
	case B200_int:
		if (t1 == B200_bte) SUM(bte, int);
		else if (t1 == B200_sht) SUM(sht, int);
		else if (t1 == B200_int) SUM(int, int);
		break;
	case B200_lng:
		if (t1 == B200_bte) SUM(bte, lng);
		else if (t1 == B200_sht) SUM(sht, lng);
		else if (t1 == B200_int) SUM(int, lng);
		else if (t1 == B200_lng) SUM(lng, lng);
		break;
	case B200_hge:
		if (t1 == B200_bte) SUM(bte, hge);
		else if (t1 == B200_sht) SUM(sht, hge);
		else if (t1 == B200_int) SUM(int, hge);
		else if (t1 == B200_lng) SUM(lng, hge);
		else if (t1 == B200_hge) SUM(hge, hge);
		break;
	case B200_flt:
		if (t1 == B200_flt) SUM(flt, flt);
		break;
	case B200_dbl:
		if (t1 == B200_flt) SUM(flt, dbl);
		else if (t1 == B200_dbl) SUM(dbl, dbl);
		break;
	}
#undef SUM
	if (rc < 0) {
		set_err("42000!type combination (sum(%s)->%s) not supported.\n", atomname(tp1), atomname(tp2));
		return B200_FAIL;
	}
	if (rc == 0)
		return B200_FAIL;
	if (st[0] & AN_ERR_OVERFLOW) {
		set_err("22003!overflow in calculation.\n");
		return B200_FAIL;
	}
	return an_frame_done(r, fa.n, st[1] != 0);
}
