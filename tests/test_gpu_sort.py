"""BATsort / BATfirstn against the reference (gdk_batop.c:2342, gdk_firstn.c:1280):
sorted copies and group ids bit-exact; order columns identical for stable sorts and valid permutations (the
order of equal values is open) otherwise; multi-column ORDER BY through the (order, groups) chain."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import both, keyview, random_col, same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE
MODES = [(False, False, False), (True, True, False), (False, True, False), (True, False, False), (False, False, True), (True, True, True)]


def check_sort(gdk, ref, b, rb, a, o=None, ro=None, g=None, rg=None, mode=(False, False, False), hseq=0):
    rev, nl, st = mode
    s, od, gr = gdk.BATsort(b, o, g, rev, nl, st)
    rs, rod, rgr = ref.BATsort(rb, ro, rg, rev, nl, st)
    sa, ra = s.numpy(), rs.numpy()
    if not st and sa.dtype.names is None and np.issubdtype(sa.dtype, np.floating):
        # -0 == +0: an unstable sort may leave them in any order; compare with the zeros folded
        assert same_bits(sa + 0.0, ra + 0.0), (mode, sa[:8], ra[:8])
    else:
        assert same_bits(sa, ra), (mode, sa[:8], ra[:8])
    assert same_bits(gr.numpy(), rgr.numpy()), mode
    oa, roa = od.numpy(), rod.numpy()
    # a permutation of the input rows that reproduces the sorted column
    assert np.array_equal(np.sort(oa), np.sort(roa))
    assert same_bits(a[(oa - hseq).astype(np.int64)], sa)
    if st:
        assert np.array_equal(oa, roa), mode
    for x, y in ((s, rs), (od, rod), (gr, rgr)):
        px, py = x.props(), y.props()
        assert (px["count"], px["hseq"]) == (py["count"], py["hseq"])
        # order flags may be conservative but never falsely true
        if px["sorted"] and len(sa) > 1 and x is s and a.dtype.names is None and np.issubdtype(a.dtype, np.signedinteger):
            v = sa.astype(object)
            assert all(v[i] <= v[i + 1] for i in range(min(len(v) - 1, 2000)))
    return (s, od, gr), (rs, rod, rgr)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl, T.date, T.timestamp, T.oid])
@pytest.mark.parametrize("mode", MODES)
def test_single_column(gdk, ref, tt, mode):
    rng = np.random.default_rng(tt * 13 + sum(mode))
    for n, distinct, nilfrac in ((1, 1, 0), (2, 2, 0.5), (33, 5, 0.2), (2049, 50, 0.1), (70_001, 1000, 0.05), (70_001, 10**9, 0.0)):
        if tt == T.oid:
            a = rng.integers(0, min(distinct, 10**9), n).astype(np.uint64)
            if nilfrac:
                a[rng.random(n) < nilfrac] = np.uint64(1 << 63)
        elif tt in (T.flt, T.dbl):
            a = np.round(rng.uniform(-distinct, distinct, n)).astype(G.NP_OF[tt])
            a[rng.random(n) < 0.05] = -0.0
            if nilfrac:
                a[rng.random(n) < nilfrac] = np.nan
        else:
            hi = min(distinct, {T.bte: 120, T.sht: 30000}.get(tt, 10**9))
            a = random_col(rng, tt, n, nilfrac=nilfrac, lo=-hi, hi=hi)
        b, rb = both(gdk, ref, a, tt, hseq=7)
        check_sort(gdk, ref, b, rb, a, mode=mode, hseq=7)


@pytest.mark.parametrize("stable", [False, True])
def test_multi_column_order_by(gdk, ref, stable):
    """ORDER BY c1, c2 DESC, c3: each call refines the (order, groups) of the previous one"""
    rng = np.random.default_rng(5 + stable)
    n = 50_003
    c1 = rng.integers(0, 20, n).astype(np.int32)
    c2 = random_col(rng, T.lng, n, nilfrac=0.05, lo=-30, hi=30)
    c3 = np.round(rng.uniform(-5, 5, n)).astype(np.float64) + 0.0        # no -0: its place among zeros is open when unstable
    b1, r1 = both(gdk, ref, c1, T.int)
    b2, r2 = both(gdk, ref, c2, T.lng)
    b3, r3 = both(gdk, ref, c3, T.dbl)
    (s1, o1, g1), (rs1, ro1, rg1) = check_sort(gdk, ref, b1, r1, c1, mode=(False, False, stable))
    # the second key is sorted inside the groups of the first, descending
    m2 = (True, True, stable)
    s2, o2, g2 = gdk.BATsort(b2, o1, g1, *m2)
    rs2, ro2, rg2 = ref.BATsort(r2, ro1 if stable else ref.bat(o1.numpy(), ref.T.oid), rg1, *m2)
    assert same_bits(s2.numpy(), rs2.numpy())
    assert same_bits(g2.numpy(), rg2.numpy())
    assert same_bits(c2[o2.numpy().astype(np.int64)], s2.numpy())
    if stable:
        assert np.array_equal(o2.numpy(), ro2.numpy())
    s3, o3, g3 = gdk.BATsort(b3, o2, g2, False, False, stable)
    rs3, ro3, rg3 = ref.BATsort(r3, ro2 if stable else ref.bat(o2.numpy(), ref.T.oid), rg2, False, False, stable)
    assert same_bits(s3.numpy(), rs3.numpy())
    assert same_bits(g3.numpy(), rg3.numpy())
    # the final order is the lexicographic one
    fo = o3.numpy().astype(np.int64)
    k2 = np.where(c2 == np.iinfo(np.int64).min, np.iinfo(np.int64).min, c2)
    ref_order = np.lexsort((c3, -k2.astype(np.float64), c1))
    assert np.array_equal(c1[fo], c1[ref_order]) and same_bits(c2[fo], c2[ref_order]) and same_bits(c3[fo], c3[ref_order])
    if stable:
        assert np.array_equal(o3.numpy(), ro3.numpy())


def test_sort_only_some_outputs_and_errors(gdk, ref):
    a = np.array([5, 3, 9, 3, 1], dtype=np.int64)
    b, rb = both(gdk, ref, a, T.lng)
    s, o, g = gdk.BATsort(b, None, None, False, False, False, want=(True, False, False))
    assert o is None and g is None and s.numpy().tolist() == [1, 3, 3, 5, 9]
    s, o, g = gdk.BATsort(b, None, None, False, False, False, want=(False, True, False))
    assert s is None and a[o.numpy().astype(np.int64)].tolist() == [1, 3, 3, 5, 9]
    with pytest.raises(G.GDKError) as e1:
        gdk.BATsort(b, None, None, True, False, True)
    with pytest.raises(Exception) as e2:
        ref.BATsort(rb, None, None, True, False, True)
    assert str(e1.value).strip() == str(e2.value).strip() == "stable sort cannot have reverse != nilslast"


@pytest.mark.parametrize("tt", [T.int, T.lng, T.dbl])
@pytest.mark.parametrize("asc", [True, False])
@pytest.mark.parametrize("nilslast", [True, False])
def test_firstn(gdk, ref, tt, asc, nilslast):
    rng = np.random.default_rng(tt + asc * 2 + nilslast)
    for n, k in ((10, 3), (10, 20), (5000, 100), (60_000, 1000), (60_000, 1)):
        if tt == T.dbl:
            a = np.round(rng.uniform(-500, 500, n)).astype(np.float64)
            a[rng.random(n) < 0.01] = np.nan
        else:
            a = random_col(rng, tt, n, nilfrac=0.01, lo=-500, hi=500)
        b, rb = both(gdk, ref, a, tt, hseq=3)
        got, want = gdk.BATfirstn(b, None, k, asc, nilslast), ref.BATfirstn(rb, None, k, asc, nilslast)
        ga, wa = got.numpy(), want.numpy()
        assert len(ga) == len(wa) == min(k, n)
        assert np.all(np.diff(ga.astype(np.int64)) > 0)                # a candidate list
        # the same VALUES are selected (which of several equal boundary rows is open)
        gv, wv = np.sort(a[(ga - 3).astype(np.int64)]), np.sort(a[(wa - 3).astype(np.int64)])
        assert same_bits(gv, wv)
        # with a candidate list
        cand = np.sort(rng.choice(n, max(1, n // 2), replace=False)).astype(np.uint64) + 3
        s, rs = gdk.BATfromhost(cand, T.oid).setprops(sorted=True, key=True, nonil=True), ref.bat(cand, ref.T.oid, sorted=True, key=True, nonil=True)
        ga, wa = gdk.BATfirstn(b, s, k, asc, nilslast).numpy(), ref.BATfirstn(rb, rs, k, asc, nilslast).numpy()
        assert set(ga.tolist()) <= set(cand.tolist())
        assert same_bits(np.sort(a[(ga - 3).astype(np.int64)]), np.sort(a[(wa - 3).astype(np.int64)]))


# ------------------------------------------------------------------ BATfirstn: grouped / distinct / gids forms
def _firstn_both(gdk, ref, b, rb, s, rs, g, rg, n, asc, nilslast, distinct, want_gids):
    """run both sides; a reference failure must be a failure here too (same composition of operators)"""
    try:
        want = ref.BATfirstn(rb, rs, n, asc, nilslast, rg, distinct, want_gids)
    except ref.RefError as ex:
        try:
            gdk.BATfirstn(b, s, n, asc, nilslast, g, distinct, want_gids)
        except G.GDKError as e2:
            assert str(e2).strip() == str(ex).strip()
        return None, None
    got = gdk.BATfirstn(b, s, n, asc, nilslast, g, distinct, want_gids)
    return got, want


@pytest.mark.parametrize("tt", [T.int, T.lng, T.dbl, T.date])
@pytest.mark.parametrize("distinct", [False, True])
@pytest.mark.parametrize("asc", [True, False])
def test_firstn_cascade(gdk, ref, tt, distinct, asc):
    """ORDER BY c1, c2, c3 LIMIT n as the cascade of three BATfirstn calls (gdk_firstn.c:52-60)"""
    rng = np.random.default_rng(tt * 5 + distinct * 2 + asc)
    m = 30_011
    c1 = random_col(rng, tt, m, nilfrac=0.03, lo=0, hi=40) if tt not in (T.flt, T.dbl) else np.round(random_col(rng, tt, m, nilfrac=0.03, lo=0, hi=40))
    c2 = random_col(rng, T.int, m, nilfrac=0.02, lo=0, hi=6)
    c3 = random_col(rng, T.lng, m, nilfrac=0.0, lo=0, hi=500)
    b1, r1 = both(gdk, ref, c1, tt)
    b2, r2 = both(gdk, ref, c2, T.int)
    b3, r3 = both(gdk, ref, c3, T.lng)
    for n in (1, 40, 700):
        for nilslast in (False, True):
            got, want = _firstn_both(gdk, ref, b1, r1, None, None, None, None, n, asc, nilslast, distinct, True)
            if got is None:
                continue
            (s1, g1), (rs1, rg1) = got, want
            assert np.array_equal(s1.numpy(), rs1.numpy()), (n, nilslast, len(s1), len(rs1))
            assert np.array_equal(g1.numpy(), rg1.numpy())
            assert len(s1) >= min(n, m)
            got, want = _firstn_both(gdk, ref, b2, r2, s1, rs1, g1, rg1, n, asc, nilslast, distinct, True)
            if got is None:
                continue
            (s2, g2), (rs2, rg2) = got, want
            assert np.array_equal(s2.numpy(), rs2.numpy()), (n, nilslast, len(s2), len(rs2))
            assert np.array_equal(g2.numpy(), rg2.numpy())
            got, want = _firstn_both(gdk, ref, b3, r3, s2, rs2, g2, rg2, n, asc, nilslast, distinct, False)
            if got is None:
                continue
            x, y = got.numpy().astype(np.int64), want.numpy().astype(np.int64)
            assert np.all(x[:-1] < x[1:])
            if distinct:
                assert np.array_equal(x, y)
            else:
                # which of several equal boundary rows make it is open: the chosen VALUE tuples are not
                assert len(x) == len(y) == min(n, len(s2))
                k1 = c1.view(np.int64) if tt == T.dbl else c1     # bit patterns: nan != nan would spoil the comparison
                kx = sorted(zip(keyview(tt, k1[x]).tolist(), c2[x].tolist(), c3[x].tolist()))
                ky = sorted(zip(keyview(tt, k1[y]).tolist(), c2[y].tolist(), c3[y].tolist()))
                assert kx == ky


@pytest.mark.parametrize("tt", [T.bte, T.int, T.hge, T.flt])
def test_firstn_single_column_forms(gdk, ref, tt):
    rng = np.random.default_rng(tt + 31)
    m = 20_003
    a = random_col(rng, tt, m, nilfrac=0.05, lo=-50, hi=50)
    if tt == T.flt:
        a = np.round(a).astype(a.dtype)
    b, rb = both(gdk, ref, a, tt, hseq=6)
    c = np.sort(rng.choice(m, m // 2, replace=False)).astype(np.uint64) + 6
    s, rs = gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True), ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)
    for cs, rcs in ((None, None), (s, rs)):
        for n in (1, 7, 300, m + 5):
            for asc in (True, False):
                for nilslast in (False, True):
                    for distinct in (False, True):
                        got, want = _firstn_both(gdk, ref, b, rb, cs, rcs, None, None, n, asc, nilslast, distinct, True)
                        if got is None:
                            continue
                        assert np.array_equal(got[0].numpy(), want[0].numpy()), (n, asc, nilslast, distinct)
                        assert np.array_equal(got[1].numpy(), want[1].numpy()), (n, asc, nilslast, distinct)
                    # distinct without gids: n complete groups of values
                    got, want = _firstn_both(gdk, ref, b, rb, cs, rcs, None, None, n, asc, nilslast, True, False)
                    if got is not None:
                        assert np.array_equal(got.numpy(), want.numpy()), (n, asc, nilslast)


# ------------------------------------------------------------------ order index (gdk_orderidx.c)
import os  # noqa: E402

os.environ.setdefault("B200_OIDX_SELECT_MIN_BYTES", "0")     # let small test columns take the index path of BATselect


@pytest.mark.parametrize("tt", [T.bte, T.int, T.lng, T.hge, T.flt, T.dbl, T.date])
def test_order_index(gdk, ref, tt):
    """BATorderidx + its two users: BATsort (order = the index) and selective range selects (two binary searches in
    the index); results must be those of the plain paths and of the reference"""
    rng = np.random.default_rng(tt + 900)
    n = 120_011
    a = random_col(rng, tt, n, nilfrac=0.03, lo=-100 if tt == T.bte else -50_000, hi=100 if tt == T.bte else 50_000)
    # hseqbase 0: the reference's own order-index select returns nothing for columns with another hseqbase (it filters
    # the index oids against a window computed without it, gdk_select.c:1740-1760,1895) — BASELINE.md §5; a second
    # column with hseqbase 21 is checked against the definition below
    b, rb = both(gdk, ref, a, tt, hseq=0)
    plain = gdk.BATsort(b, None, None, False, False, True)
    assert not gdk.BATcheckorderidx(b)
    gdk.BATorderidx(b, True)
    assert gdk.BATcheckorderidx(b)
    gdk.BATorderidx(b, True)                      # idempotent
    via = gdk.BATsort(b, None, None, False, False, True)
    assert gdk.lib.b200_algorithm() == b"sort: order index"
    rs, ro, rg = ref.BATsort(rb, None, None, False, False, True)
    for x, y, z in zip(plain, via, (rs, ro, rg)):
        assert same_bits(x.numpy(), y.numpy()) and same_bits(y.numpy(), z.numpy())
    # other orders do not use it
    check_sort(gdk, ref, b, rb, a, mode=(True, True, False), hseq=0)
    # selective ranges (<= n/64 rows) through the index, wide ones through the scan; both identical to the reference
    if tt in (T.flt, T.dbl):
        ranges = [(10.0, 400.0), (-3.5, -3.5), (49_000.0, 60_000.0), (-40_000.0, 40_000.0)]
    elif tt == T.bte:
        ranges = [(5, 5), (99, 120), (-100, 100)]
    else:
        ranges = [(10, 400), (-7, -7), (49_500, 60_000), (-40_000, 40_000), (60_001, 70_000)]
    used = 0
    for lo, hi in ranges:
        for li, hi_incl in ((True, True), (False, True), (True, False)):
            got = gdk.BATselect(b, None, lo, hi, li, hi_incl, False, False)
            used += gdk.lib.b200_algorithm() == b"select: orderidx"
            want = ref.BATselect(rb, None, lo, hi, li, hi_incl, False, False)
            assert same_bits(got.numpy(), want.numpy()), (lo, hi, li, hi_incl)
            assert got.props()["hseq"] == 0
    assert used >= 3
    for op, v in (("==", ranges[0][0]), ("<", -49_000 if tt != T.bte else -99), (">=", 49_900 if tt != T.bte else 100)):
        got, want = gdk.BATthetaselect(b, None, v, op), ref.BATthetaselect(rb, None, v, op)
        assert same_bits(got.numpy(), want.numpy()), (op, v)
    # another hseqbase: against the definition
    if tt not in (T.hge,):
        b21 = gdk.BATfromhost(a, tt, hseq=21)
        gdk.BATorderidx(b21, True)
        lo, hi = ranges[0]
        got = gdk.BATselect(b21, None, lo, hi, True, True, False, False).numpy()
        assert gdk.lib.b200_algorithm() == b"select: orderidx"
        isnil = np.isnan(a) if tt in (T.flt, T.dbl) else a == np.iinfo(a.dtype).min
        assert np.array_equal(got, np.flatnonzero(~isnil & (a >= lo) & (a <= hi)).astype(np.uint64) + 21)
    # rewriting / freeing the column drops the index
    gdk.lib.b200_BATinvalidate(b._p)
    assert not gdk.BATcheckorderidx(b)
    with pytest.raises(G.GDKError, match="No order index on void"):
        gdk.BATorderidx(gdk.BATdense(0, 0, 10))
