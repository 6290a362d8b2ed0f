// analytic_stat.cu — GDKanalyticalavg (gdk_analytic_statistics.c:364-437),
// GDKanalyticalcount (gdk_analytic_func.c:1626-1681), GDKanalyticalmin/max (:1263-1323).
#include "analytic_aggr.cuh"
#include "analytic_prefix.cuh"

// ---------------------------------------------------------------- avg
template <typename T> static bool avg_run(const FrameArgs &fa, dbl *rb, unsigned long long st[2])
{
	const bool framed = fa.frame_type < 3 || fa.frame_type > 6;
	if (!framed) {
		ctx()->algo = "analyticalavg: cuda exact sum/count";
		return an_run<AvgExactOp<T>>(fa, rb, st, "an_avg");
	}
	ctx()->algo = "analyticalavg: cuda fan-out-16 segment tree";
	if constexpr (Atom<T>::is_float)
		return an_run<AvgTreeFltOp<T>>(fa, rb, st, "an_avg");
	else {
		/* level 0 of the tree walk from the column's running totals (shared with GDKanalyticalsum) */
		PrefixArrays pa;
		if (sizeof(T) <= 8 && (fa.frame_type < 3 || fa.frame_type > 6) && an_prefix_get<T>(fa.b, fa.n, pa)) {
			ctx()->algo = "analyticalavg: cuda prefix sums (leaves) + fan-out-16 segment tree";
			return an_run<AvgTreeIntOp<T>>(fa, rb, st, "an_avg", pa.P, pa.C);
		}
		b200_clrerr();
		return an_run<AvgTreeIntOp<T>>(fa, rb, st, "an_avg");
	}
}

extern "C" b200_return b200_GDKanalyticalavg(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					     const b200_bat *s, const b200_bat *e, int tpe, int frame_type)
{
	FrameStage fs;
	FrameArgs fa;
	if (!an_frame_prepare("GDKanalyticalavg", r, p, o, b, s, e, B200_dbl, frame_type, fs, fa))
		return B200_FAIL;
	if (fa.n == 0)
		return an_frame_done(r, 0, false);
	const int t = basetype(tpe);
	if (atomsize(fa.b->ttype) != atomsize(t)) {
		set_err("GDKanalyticalavg: tpe does not match the input column\n");
		return B200_FAIL;
	}
	unsigned long long st[2] = {0, 0};
	bool ok;
	dbl *rb = (dbl *) r->base;
	switch (t) {
	case B200_bte: ok = avg_run<bte>(fa, rb, st); break;
	case B200_sht: ok = avg_run<sht>(fa, rb, st); break;
	case B200_int: ok = avg_run<int>(fa, rb, st); break;
	case B200_lng: ok = avg_run<lng>(fa, rb, st); break;
	case B200_flt: ok = avg_run<flt>(fa, rb, st); break;
	case B200_dbl: ok = avg_run<dbl>(fa, rb, st); break;
	case B200_hge:
		set_err("GDKanalyticalavg: hge input (sum may exceed 128 bits) is not on the device path\n");
		return B200_FAIL;
	default:
		set_err("42000!average of type %s to dbl unsupported.\n", atomname(tpe));
		return B200_FAIL;
	}
	if (!ok)
		return B200_FAIL;
	return an_frame_done(r, fa.n, st[1] != 0);
}

// ---------------------------------------------------------------- avg with an integer result
template <typename T> static bool avgint_run(const FrameArgs &fa, T *rb, unsigned long long st[2])
{
	const bool framed = fa.frame_type < 3 || fa.frame_type > 6;
	ctx()->algo = framed ? "analyticalavginteger: cuda fan-out-16 segment tree" : "analyticalavginteger: cuda exact sum/count";
	if (framed) {
		if constexpr (sizeof(T) <= 8) {
			PrefixArrays pa;
			if (an_prefix_get<T>(fa.b, fa.n, pa))
				return an_run<AvgIntegerOp<T, false>>(fa, rb, st, "an_avg", pa.P, pa.C);
			b200_clrerr();
		}
		return an_run<AvgIntegerOp<T, false>>(fa, rb, st, "an_avg");
	}
	return an_run<AvgIntegerOp<T, true>>(fa, rb, st, "an_avg");
}

extern "C" b200_return b200_GDKanalyticalavginteger(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
						    const b200_bat *s, const b200_bat *e, int tpe, int frame_type)
{
	FrameStage fs;
	FrameArgs fa;
	if (!an_frame_prepare("GDKanalyticalavginteger", r, p, o, b, s, e, tpe, frame_type, fs, fa))
		return B200_FAIL;
	if (fa.n == 0)
		return an_frame_done(r, 0, false);
	if (atomsize(fa.b->ttype) != atomsize(tpe)) {
		set_err("GDKanalyticalavginteger: tpe does not match the input column\n");
		return B200_FAIL;
	}
	unsigned long long st[2] = {0, 0};
	bool ok;
	switch (tpe) {		/* the reference switches on tpe itself, not its base type (:674-690) */
	case B200_bte: ok = avgint_run<bte>(fa, (bte *) r->base, st); break;
	case B200_sht: ok = avgint_run<sht>(fa, (sht *) r->base, st); break;
	case B200_int: ok = avgint_run<int>(fa, (int *) r->base, st); break;
	case B200_lng: ok = avgint_run<lng>(fa, (lng *) r->base, st); break;
	case B200_hge:
		set_err("GDKanalyticalavginteger: hge input (sum may exceed 128 bits) is not on the device path\n");
		return B200_FAIL;
	default:
		set_err("42000!average of type %s to %s unsupported.\n", atomname(tpe), atomname(tpe));
		return B200_FAIL;
	}
	if (!ok)
		return B200_FAIL;
	return an_frame_done(r, fa.n, st[1] != 0);
}

// ---------------------------------------------------------------- count
extern "C" b200_return b200_GDKanalyticalcount(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					       const b200_bat *s, const b200_bat *e, int8_t ignore_nils, int tpe, int frame_type)
{
	FrameStage fs;
	FrameArgs fa;
	if (!an_frame_prepare("GDKanalyticalcount", r, p, o, b, s, e, B200_lng, frame_type, fs, fa))
		return B200_FAIL;
	if (fa.n == 0)
		return an_frame_done(r, 0, false);
	const bool count_all = !ignore_nils || fa.b->tnonil;
	unsigned long long st[2] = {0, 0};
	bool ok;
	lng *rb = (lng *) r->base;
	if (count_all) {
		ctx()->algo = "analyticalcount: cuda frame length";
		ok = an_run<CountAllOp>(fa, rb, st, "an_count");
	} else {
		ctx()->algo = "analyticalcount: cuda fan-out-16 segment tree";
		switch (basetype(tpe)) {
		case B200_bte: ok = an_run<CountOp<bte>>(fa, rb, st, "an_count"); break;
		case B200_sht: ok = an_run<CountOp<sht>>(fa, rb, st, "an_count"); break;
		case B200_int: ok = an_run<CountOp<int>>(fa, rb, st, "an_count"); break;
		case B200_lng: ok = an_run<CountOp<lng>>(fa, rb, st, "an_count"); break;
		case B200_hge: ok = an_run<CountOp<hge>>(fa, rb, st, "an_count"); break;
		case B200_flt: ok = an_run<CountOp<flt>>(fa, rb, st, "an_count"); break;
		case B200_dbl: ok = an_run<CountOp<dbl>>(fa, rb, st, "an_count"); break;
		default:
			set_err("GDKanalyticalcount: type %s is not on the device path (fixed-width atoms only)\n", atomname(tpe));
			return B200_FAIL;
		}
	}
	if (!ok)
		return B200_FAIL;
	return an_frame_done(r, fa.n, false);
}

// ---------------------------------------------------------------- min / max
template <bool ISMIN>
static b200_return minmax(const char *fn, b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
			  const b200_bat *s, const b200_bat *e, int tpe, int frame_type)
{
	FrameStage fs;
	FrameArgs fa;
	if (!an_frame_prepare(fn, r, p, o, b, s, e, tpe, frame_type, fs, fa))
		return B200_FAIL;
	if (fa.n == 0)
		return an_frame_done(r, 0, false);
	unsigned long long st[2] = {0, 0};
	bool ok;
	void *rb = r->base;
	ctx()->algo = "analyticalmin/max: cuda fan-out-16 segment tree";
	switch (basetype(tpe)) {
	case B200_bte: ok = an_run<MinMaxOp<bte, ISMIN>>(fa, (bte *) rb, st, "an_minmax"); break;
	case B200_sht: ok = an_run<MinMaxOp<sht, ISMIN>>(fa, (sht *) rb, st, "an_minmax"); break;
	case B200_int: ok = an_run<MinMaxOp<int, ISMIN>>(fa, (int *) rb, st, "an_minmax"); break;
	case B200_lng: ok = an_run<MinMaxOp<lng, ISMIN>>(fa, (lng *) rb, st, "an_minmax"); break;
	case B200_hge: ok = an_run<MinMaxOp<hge, ISMIN>>(fa, (hge *) rb, st, "an_minmax"); break;
	case B200_flt: ok = an_run<MinMaxOp<flt, ISMIN>>(fa, (flt *) rb, st, "an_minmax"); break;
	case B200_dbl: ok = an_run<MinMaxOp<dbl, ISMIN>>(fa, (dbl *) rb, st, "an_minmax"); break;
	default:
		set_err("%s: type %s is not on the device path (fixed-width atoms only)\n", fn, atomname(tpe));
		return B200_FAIL;
	}
	if (!ok)
		return B200_FAIL;
	return an_frame_done(r, fa.n, st[1] != 0);
}

extern "C" b200_return b200_GDKanalyticalmin(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					     const b200_bat *s, const b200_bat *e, int tpe, int frame_type)
{
	return minmax<true>("GDKanalyticalmin", r, p, o, b, s, e, tpe, frame_type);
}

extern "C" b200_return b200_GDKanalyticalmax(b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b,
					     const b200_bat *s, const b200_bat *e, int tpe, int frame_type)
{
	return minmax<false>("GDKanalyticalmax", r, p, o, b, s, e, tpe, frame_type);
}
