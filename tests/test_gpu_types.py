"""date / daytime / timestamp / oid as VALUE columns (only Q6/Q1's shipdate select used a temporal
atom): the atoms are stored as int / lng (ATOMbasetype, gdk_atoms.h:285-289) and every operator of the path
must treat them like their base type, with the type id carried through to the result."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import assert_same, both, same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE
ATOMS = [T.date, T.daytime, T.timestamp, T.oid]


def col(rng, tt, n, nilfrac=0.05, distinct=200):
    dt = G.NP_OF[tt]
    if tt == T.oid:
        a = rng.integers(0, distinct, n).astype(np.uint64)
        nil = np.uint64(1 << 63)
    else:
        base = {T.date: 4_300_000, T.daytime: 0, T.timestamp: 600_000_000_000_000_000}[tt]
        a = (base + rng.integers(0, distinct, n)).astype(dt)
        nil = np.iinfo(dt).min
    if nilfrac:
        a[rng.random(n) < nilfrac] = nil
    return a


def pairs(r1, r2):
    a, b = r1.numpy(), r2.numpy()
    o = np.lexsort((b, a))
    return a[o], b[o]


@pytest.mark.parametrize("tt", ATOMS)
def test_select_project_compare(gdk, ref, tt):
    rng = np.random.default_rng(tt)
    for n in (1, 333, 70_001):
        a = col(rng, tt, n)
        b, rb = both(gdk, ref, a, tt, hseq=10)
        nonnil = a[a != (np.uint64(1 << 63) if tt == T.oid else np.iinfo(a.dtype).min)]
        lo = int(nonnil.min()) + 20 if len(nonnil) else 0
        hi = lo + 90
        for args in ((lo, hi, True, True, False), (lo, hi, False, True, True), (lo, None, True, True, False), (lo, lo, True, True, False)):
            assert_same(gdk.BATselect(b, None, *args), ref.BATselect(rb, None, *args))
        for op in ("<", "<=", "==", "!=", ">=", ">"):
            s, rs = gdk.BATthetaselect(b, None, lo, op), ref.BATthetaselect(rb, None, lo, op)
            assert_same(s, rs)
        s2, rs2 = gdk.BATselect(b, s, lo, hi, True, False, False), ref.BATselect(rb, rs, lo, hi, True, False, False)
        assert_same(s2, rs2)
        pr, rpr = gdk.BATproject(s, b), ref.BATproject(rs, rb)
        assert_same(pr, rpr)
        assert pr.props()["type"] == rpr.props()["type"] or tt == T.oid
        a2 = col(rng, tt, n)
        c, rc = both(gdk, ref, a2, tt, hseq=10)
        for op in ("lt", "le", "eq", "ne", "gt", "ge"):
            assert_same(gdk.BATcalccmp(op, b, c), ref.BATcalccmp(op, rb, rc))
            assert_same(gdk.BATcalccmpcst(op, b, lo, tt), ref.BATcalccmpcst(op, rb, lo, tt))
        assert_same(gdk.BATcalcisnil(b), ref.BATcalcisnil(rb))
        assert_same(gdk.BATcalcminmax("min", b, c), ref.BATcalcminmax("min", rb, rc))
        assert_same(gdk.BATcalcminmax("max_no_nil", b, c), ref.BATcalcminmax("max_no_nil", rb, rc))


@pytest.mark.parametrize("tt", ATOMS)
def test_group_aggr_join_unique(gdk, ref, tt):
    rng = np.random.default_rng(100 + tt)
    for n in (1, 500, 60_007):
        a = col(rng, tt, n, distinct=150)
        b, rb = both(gdk, ref, a, tt)
        g, e, h = gdk.BATgroup(b, None, None, None, None)
        rg, re_, rh = ref.BATgroup(rb, None, None, None, None)
        assert_same(g, rg)
        assert_same(e, re_)
        assert_same(h, rh)
        v = rng.integers(-1000, 1000, n).astype(np.int64)
        bv, rbv = both(gdk, ref, v, T.lng)
        assert_same(gdk.BATgroupsum(bv, g, e, None, T.hge, True), ref.BATgroupsum(rbv, rg, re_, None, T.hge, True))
        # min / max of the temporal column itself per group of another key
        k = rng.integers(0, 17, n).astype(np.int32)
        bk, rbk = both(gdk, ref, k, T.int)
        g2, e2, _ = gdk.BATgroup(bk, None, None, None, None)
        rg2, re2, _ = ref.BATgroup(rbk, None, None, None, None)
        for fn in ("BATgroupmin", "BATgroupmax"):
            assert_same(getattr(gdk, fn)(b, g2, e2, None), getattr(ref, fn)(rb, rg2, re2, None))
        assert_same(gdk.BATgroupcount(b, g2, e2, None, T.lng, True), ref.BATgroupcount(rb, rg2, re2, None, T.lng, True))
        assert_same(gdk.BATunique(b, None), ref.BATunique(rb, None))
        # join on the temporal / oid key
        a2 = col(rng, tt, max(1, n // 3), distinct=150)
        r, rr = both(gdk, ref, a2, tt, hseq=7)
        x1, x2 = pairs(*gdk.BATjoin(b, r, None, None, False))
        y1, y2 = pairs(*ref.BATjoin(rb, rr, None, None, False))
        assert np.array_equal(x1, y1) and np.array_equal(x2, y2)
        x1, x2 = pairs(*gdk.BATjoin(b, r, None, None, True))          # nil_matches
        y1, y2 = pairs(*ref.BATjoin(rb, rr, None, None, True))
        assert np.array_equal(x1, y1) and np.array_equal(x2, y2)
