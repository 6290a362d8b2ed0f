// firstn.cu — the grouped / distinct / gids forms of BATfirstn (gdk_firstn.c:1022-1277), i.e. the refinement steps of
// a multi-column ORDER BY ... LIMIT:
//      BATfirstn(&s1, &g1, b1, NULL, NULL, n, asc, nilslast, distinct);
//      BATfirstn(&s2, &g2, b2, s1, g1, n, ...);
//      BATfirstn(&s3, NULL, b3, s2, g2, n, ...);
// The reference builds these forms out of its own operators around two heap-based cores (BATfirstn_unique[_with_groups]);
// the cores are replaced by the device radix sort (the first n of the order by (group ascending, value)), the rest is
// the SAME composition over the device operators — BATunique, BATintersect, BATselect, BATmergecand, BATproject,
// BATsort, BATgroup, BATleftjoin — each of which is bit-identical to the reference's.  Where the reference's heap
// leaves the choice among equal boundary rows open (forms without gids and without distinct) the earliest rows win
// here; with gids or distinct every tie is included, so those results are fully determined and compared as such.
#include "common.cuh"

namespace {
struct Own {
	b200_bat *p = nullptr;
	Own() {}
	explicit Own(b200_bat *q) : p(q) {}
	Own(const Own &) = delete;
	Own &operator=(const Own &) = delete;
	~Own() { if (p) b200_BATfree(p); }
	void reset(b200_bat *q)
	{
		if (p)
			b200_BATfree(p);
		p = q;
	}
	b200_bat *release()
	{
		b200_bat *q = p;
		p = nullptr;
		return q;
	}
	operator const b200_bat *() const { return p; }
	b200_bat **out()
	{
		reset(nullptr);
		return &p;
	}
};

// one atom of a column (by oid) on the host: BUNtail(bi, oid - hseq)
bool fetch_atom(const b200_bat *b, oid o, void *dst16)
{
	Own one(bat_dense(0, o, 1));
	if (one.p == nullptr)
		return false;
	Own v(b200_BATproject(one, b));
	if (v.p == nullptr)
		return false;
	memset(dst16, 0, 16);
	return b200_BATtohost(v, dst16) == 0;
}

// BATfirstn_unique / BATfirstn_unique_with_groups (gdk_firstn.c:212, :717): the first n candidates in the order
// (g ascending,) value ascending / descending with the nils where nilslast puts them, as an ascending candidate list;
// *last = oid of the n-th element of that order (oid_nil when every candidate is returned and g == NULL), *lastg = its
// group id.
bool firstn_core(const b200_bat *b, const b200_bat *s, const b200_bat *g, uint64_t n, bool asc, bool nilslast, b200_bat **topn,
		 oid *last, oid *lastg)
{
	*topn = nullptr;
	if (last)
		*last = OID_NIL;
	if (lastg)
		*lastg = 0;
	Own cand(s ? nullptr : bat_dense(0, b->hseqbase, b->count));
	const b200_bat *sc = s ? s : (const b200_bat *) cand;
	if (sc == nullptr)
		return false;
	Own vals(b200_BATproject(sc, b));		/* value of candidate i at position i */
	if (vals.p == nullptr)
		return false;
	const uint64_t cnt = vals.p->count;
	const bool all = n >= cnt;
	if (n > cnt)
		n = cnt;
	if (n == 0) {
		*topn = bat_dense(0, 0, 0);
		if (last)
			*last = g ? 0 : OID_NIL;
		return *topn != nullptr;
	}
	if (all && g == nullptr) {
		/* trivial: every candidate (:229-241) */
		Cand ci;
		if (!cand_init(ci, b, s))
			return false;
		*topn = cand_to_bat(ci);
		return *topn != nullptr;
	}
	Own ord;
	if (g) {
		if (g->count != cnt) {
			set_err("BATfirstn: g must be aligned with s\n");
			return false;
		}
		Own o1, g1;
		b200_bat galigned = *g;		/* same rows, seen with the hseqbase of the projected values */
		galigned.owns = 0;
		galigned.parent = nullptr;
		galigned.hseqbase = vals.p->hseqbase;
		if (b200_BATsort(nullptr, o1.out(), g1.out(), &galigned, nullptr, nullptr, false, false, true) != B200_SUCCEED ||
		    b200_BATsort(nullptr, ord.out(), nullptr, vals, o1, g1, !asc, nilslast, false) != B200_SUCCEED)
			return false;
	} else if (b200_BATsort(nullptr, ord.out(), nullptr, vals, nullptr, nullptr, !asc, nilslast, false) != B200_SUCCEED) {
		return false;
	}
	Own head(bat_dense(0, ord.p->hseqbase, n));
	if (head.p == nullptr)
		return false;
	Own picked(b200_BATproject(head, ord));		/* positions (oids of vals) of the first n */
	if (picked.p == nullptr)
		return false;
	if (last || lastg) {
		oid pos[2] = {0, 0};
		if (!fetch_atom(ord, ord.p->hseqbase + n - 1, pos))
			return false;
		oid tmp[2];
		if (last) {
			b200_bat scal = *sc;	/* the candidate list as a column indexed like vals */
			scal.owns = 0;
			scal.parent = nullptr;
			scal.hseqbase = vals.p->hseqbase;
			if (!fetch_atom(&scal, pos[0], tmp))
				return false;
			*last = tmp[0];
		}
		if (lastg && g) {
			b200_bat galigned = *g;
			galigned.owns = 0;
			galigned.parent = nullptr;
			galigned.hseqbase = vals.p->hseqbase;
			if (!fetch_atom(&galigned, pos[0], tmp))
				return false;
			*lastg = tmp[0];
		}
	}
	b200_bat scal = *sc;
	scal.owns = 0;
	scal.parent = nullptr;
	scal.hseqbase = vals.p->hseqbase;
	Own oids(b200_BATproject(picked, &scal));	/* their oids in b ... */
	if (oids.p == nullptr)
		return false;
	Own cl;
	if (b200_BATsort(cl.out(), nullptr, nullptr, oids, nullptr, nullptr, false, false, false) != B200_SUCCEED)	/* ... ascending */
		return false;
	cl.p->hseqbase = 0;
	cl.p->tsorted = cl.p->tkey = cl.p->tnonil = 1;
	cl.p->trevsorted = n <= 1;
	*topn = cl.release();
	return true;
}

// group ids of the result rows: dense ranks of (g,) value in the sort order the reference uses (:1090-1116, :1233-1268)
bool firstn_gids(const b200_bat *bn, const b200_bat *b, const b200_bat *s, const b200_bat *g, bool asc, b200_bat **gids)
{
	Own vals(b200_BATproject(bn, b));
	if (vals.p == nullptr)
		return false;
	Own ord, grp, inv;
	if (g) {
		Own pos(b200_BATintersect(s, bn, nullptr, nullptr, false, false, B200_BUN_NONE));	/* where in s the result rows sit */
		if (pos.p == nullptr)
			return false;
		Own gv(b200_BATproject(pos, g));
		if (gv.p == nullptr)
			return false;
		Own o1, g1;
		if (b200_BATsort(nullptr, o1.out(), g1.out(), gv, nullptr, nullptr, false, false, false) != B200_SUCCEED ||
		    b200_BATsort(nullptr, ord.out(), grp.out(), vals, o1, g1, !asc, !asc, false) != B200_SUCCEED)
			return false;
	} else if (b200_BATsort(nullptr, ord.out(), grp.out(), vals, nullptr, nullptr, !asc, !asc, false) != B200_SUCCEED) {
		return false;
	}
	if (b200_BATsort(nullptr, inv.out(), nullptr, ord, nullptr, nullptr, false, false, false) != B200_SUCCEED)
		return false;
	*gids = b200_BATproject(inv, grp);
	return *gids != nullptr;
}
}	// namespace

// BATfirstn_grouped (gdk_firstn.c:1022-1127): g == NULL, gids and / or distinct asked for
static b200_return firstn_grouped(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, uint64_t n, bool asc,
				  bool nilslast, bool distinct)
{
	ctx()->algo = "firstn: cuda grouped (radix sort core + tie extension)";
	const b200_bat *su = s;
	Own uniq;
	if (distinct && !b->tkey) {
		uniq.reset(b200_BATunique(b, s));
		if (uniq.p == nullptr)
			return B200_FAIL;
		s = uniq;
	}
	Own bn;
	oid last = OID_NIL;
	if (!firstn_core(b, s, nullptr, n, asc, nilslast, bn.out(), &last, nullptr))
		return B200_FAIL;
	if (bn.p->count == 0) {
		if (gids && (*gids = bat_dense(0, 0, 0)) == nullptr)
			return B200_FAIL;
		*topn = bn.release();
		return B200_SUCCEED;
	}
	if (!b->tkey) {
		if (distinct) {
			/* every row (of the original candidates) whose value is one of the n chosen ones */
			Own all(b200_BATintersect(b, b, su, bn, true, false, B200_BUN_NONE));
			if (all.p == nullptr)
				return B200_FAIL;
			bn.reset(all.release());
		} else if (last != OID_NIL) {
			/* every row equal to the last value that made it */
			unsigned char val[16];
			if (!fetch_atom(b, last, val))
				return B200_FAIL;
			Own ties(b200_BATselect(b, s, val, nullptr, true, false, false, false));
			if (ties.p == nullptr)
				return B200_FAIL;
			Own mg(b200_BATmergecand(bn, ties));
			if (mg.p == nullptr)
				return B200_FAIL;
			bn.reset(mg.release());
		}
	}
	if (gids && !firstn_gids(bn, b, nullptr, nullptr, asc, gids))
		return B200_FAIL;
	*topn = bn.release();
	return B200_SUCCEED;
}

// BATfirstn_grouped_with_groups (gdk_firstn.c:1129-1277)
static b200_return firstn_grouped_with_groups(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, const b200_bat *g,
					      uint64_t n, bool asc, bool nilslast, bool distinct)
{
	ctx()->algo = "firstn: cuda grouped with groups (radix sort core + tie extension)";
	Own bn;
	oid last = OID_NIL, lastg = 0;
	if (distinct) {
		Own grp, ext;
		if (b200_BATgroup(grp.out(), ext.out(), nullptr, b, s, g, nullptr, nullptr) != B200_SUCCEED)
			return B200_FAIL;
		Own gvals(b200_BATproject(ext, b));		/* one value per (g, value) combination */
		if (gvals.p == nullptr)
			return B200_FAIL;
		Own pos(b200_BATintersect(s, ext, nullptr, nullptr, false, false, B200_BUN_NONE));
		if (pos.p == nullptr)
			return B200_FAIL;
		Own ggrp(b200_BATproject(pos, g));		/* ... and its group id */
		if (ggrp.p == nullptr)
			return B200_FAIL;
		Own chosen;
		if (!firstn_core(gvals, nullptr, ggrp, n, asc, nilslast, chosen.out(), nullptr, nullptr))
			return B200_FAIL;
		Own rows;
		if (b200_BATleftjoin(rows.out(), nullptr, grp, chosen, nullptr, nullptr, false, B200_BUN_NONE) != B200_SUCCEED)
			return B200_FAIL;
		bn.reset(b200_BATproject(rows, s));
		if (bn.p == nullptr)
			return B200_FAIL;
	} else if (!firstn_core(b, s, g, n, asc, nilslast, bn.out(), &last, &lastg)) {
		return B200_FAIL;
	}
	if (bn.p->count == 0) {
		if (gids && (*gids = bat_dense(0, 0, 0)) == nullptr)
			return B200_FAIL;
		*topn = bn.release();
		return B200_SUCCEED;
	}
	if (!distinct && !b->tkey) {
		/* the rows of the last group that made it which hold the last value that made it */
		Own ingrp(b200_BATselect(g, nullptr, &lastg, nullptr, true, false, false, false));
		if (ingrp.p == nullptr)
			return B200_FAIL;
		Own theirs(b200_BATproject(ingrp, s));
		if (theirs.p == nullptr)
			return B200_FAIL;
		unsigned char val[16];
		if (!fetch_atom(b, last, val))
			return B200_FAIL;
		Own ties(b200_BATselect(b, theirs, val, nullptr, true, false, false, false));
		if (ties.p == nullptr)
			return B200_FAIL;
		Own mg(b200_BATmergecand(bn, ties));
		if (mg.p == nullptr)
			return B200_FAIL;
		bn.reset(mg.release());
	}
	if (gids && !firstn_gids(bn, b, s, g, asc, gids))
		return B200_FAIL;
	*topn = bn.release();
	return B200_SUCCEED;
}

// called by b200_BATfirstn (sort.cu) for every form but the plain one
b200_return firstn_forms(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, const b200_bat *g, uint64_t n,
			 bool asc, bool nilslast, bool distinct)
{
	if (g != nullptr && s == nullptr) {
		set_err("BATfirstn: g requires s\n");	/* assert(g == NULL || s != NULL), gdk_firstn.c:1298 */
		return B200_FAIL;
	}
	Staged sb, ss, sg;
	if (!stage(b, sb) || !stage(s, ss) || !stage(g, sg))
		return B200_FAIL;
	b = sb.b;
	s = ss.b;
	g = sg.b;
	if (n == 0 || b->count == 0 || (s != nullptr && s->count == 0)) {
		*topn = bat_dense(0, 0, 0);
		if (*topn && gids)
			*gids = bat_dense(0, 0, 0);
		if (*topn == nullptr || (gids && *gids == nullptr)) {
			if (*topn)
				b200_BATfree(*topn);
			*topn = nullptr;
			return B200_FAIL;
		}
		return B200_SUCCEED;
	}
	b200_return rc;
	if (g == nullptr)
		rc = firstn_grouped(topn, gids, b, s, n, asc, nilslast, distinct);
	else if (gids == nullptr && !distinct) {
		ctx()->algo = "firstn: cuda radix sort by (group, value) + head";
		rc = firstn_core(b, s, g, n, asc, nilslast, topn, nullptr, nullptr) ? B200_SUCCEED : B200_FAIL;
	} else
		rc = firstn_grouped_with_groups(topn, gids, b, s, g, n, asc, nilslast, distinct);
	if (rc != B200_SUCCEED && gids && *gids) {
		b200_BATfree(*gids);
		*gids = nullptr;
	}
	return rc;
}
