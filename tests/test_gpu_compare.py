"""BATcalc{lt,le,gt,ge,eq,ne} (+cst) and BATcalcisnil: CUDA path vs the real reference GDK, bit-exact."""
import numpy as np
import pytest

from tests.util import ALL_TYPES, assert_same, both, random_col
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE
OPS = ["lt", "le", "gt", "ge", "eq", "ne"]


@pytest.mark.parametrize("op", OPS)
def test_compare_matrix(gdk, ref, op):
    rng = np.random.default_rng(3)
    n = 4101
    for t1 in ALL_TYPES:
        for t2 in ALL_TYPES:
            a1 = random_col(rng, t1, n, nilfrac=0.05, lo=-9, hi=9)
            a2 = random_col(rng, t2, n, nilfrac=0.05, lo=-9, hi=9)
            if t1 in (T.flt, T.dbl):
                a1 = np.round(a1).astype(a1.dtype)
            if t2 in (T.flt, T.dbl):
                a2 = np.round(a2).astype(a2.dtype)
            b1, r1 = both(gdk, ref, a1, t1)
            b2, r2 = both(gdk, ref, a2, t2)
            assert_same(gdk.BATcalccmp(op, b1, b2), ref.BATcalccmp(op, r1, r2))
            if op in ("eq", "ne"):
                assert_same(gdk.BATcalccmp(op, b1, b2, nil_matches=True), ref.BATcalccmp(op, r1, r2, nil_matches=True))


@pytest.mark.parametrize("op", OPS)
def test_compare_cst_candidates(gdk, ref, op):
    rng = np.random.default_rng(4)
    n = 30011
    a = random_col(rng, T.lng, n, nilfrac=0.02, lo=-100, hi=100)
    b, rb = both(gdk, ref, a, T.lng, hseq=9)
    o = np.sort(rng.choice(n, n // 4, replace=False)).astype(np.uint64) + np.uint64(9)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    for cand, rcand in ((None, None), (s, rs), (gdk.BATdense(0, 500, 1003), ref.dense(0, 500, 1003))):
        for v, vt in ((7, T.lng), (7, T.int), (None, T.lng), (2.5, T.dbl)):
            vv = G.nil_of(vt) if v is None else v
            assert_same(gdk.BATcalccmpcst(op, b, vv, vt, cand), ref.BATcalccmpcst(op, rb, vv, vt, rcand))
            if op in ("eq", "ne"):
                assert_same(gdk.BATcalccmpcst(op, b, vv, vt, cand, True), ref.BATcalccmpcst(op, rb, vv, vt, rcand, True))
            # constant on the left: BATcalccst<op>(v, b, s)
            assert_same(gdk.BATcalccmpcst(op, b, vv, vt, cand, cst_first=True),
                        ref.BATcalccmpcst(op, rb, vv, vt, rcand, cst_first=True))
            if op in ("eq", "ne"):
                assert_same(gdk.BATcalccmpcst(op, b, vv, vt, cand, True, cst_first=True),
                            ref.BATcalccmpcst(op, rb, vv, vt, rcand, True, cst_first=True))


def test_compare_size_mismatch(gdk, ref):
    b1, r1 = both(gdk, ref, np.arange(10, dtype=np.int32), T.int)
    b2, r2 = both(gdk, ref, np.arange(11, dtype=np.int32), T.int)
    with pytest.raises(ref.RefError) as er:
        ref.BATcalccmp("lt", r1, r2)
    with pytest.raises(G.GDKError) as eo:
        gdk.BATcalccmp("lt", b1, b2)
    assert str(eo.value).strip() == str(er.value).strip()


@pytest.mark.parametrize("tt", ALL_TYPES)
def test_isnil(gdk, ref, tt):
    rng = np.random.default_rng(6)
    for nilfrac in (0.0, 0.2):
        a = random_col(rng, tt, 5003, nilfrac=nilfrac)
        b, rb = both(gdk, ref, a, tt, hseq=2)
        assert_same(gdk.BATcalcisnil(b), ref.BATcalcisnil(rb))
        assert_same(gdk.BATcalcisnil(b, gdk.BATdense(0, 100, 50)), ref.BATcalcisnil(rb, ref.dense(0, 100, 50)))
        assert_same(gdk.BATcalcisnotnil(b), ref.BATcalcisnotnil(rb))
        assert_same(gdk.BATcalcisnotnil(b, gdk.BATdense(0, 100, 50)), ref.BATcalcisnotnil(rb, ref.dense(0, 100, 50)))
        if tt not in (T.hge, T.flt, T.dbl):      # integer nils are the type minimum: np.sort puts them first, as GDK orders them
            sb, rsb = both(gdk, ref, np.sort(a), tt, sorted=True)
            for f in ("BATcalcisnil", "BATcalcisnotnil"):
                got, want = getattr(gdk, f)(sb), getattr(ref, f)(rsb)
                assert_same(got, want)
                assert (got.props()["sorted"], got.props()["revsorted"]) == (want.props()["sorted"], want.props()["revsorted"])
