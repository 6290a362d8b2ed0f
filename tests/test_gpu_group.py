"""BATgroup: CUDA path vs the real reference GDK — group ids (first-occurrence
numbering), extents and histogram bit-exact, for every fixed-width atom, with
and without candidates and prior grouping (gdk_group.c:657-1351)."""
import numpy as np
import pytest

from tests.util import ALL_TYPES, assert_same, both, random_col
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE


def _check(o, r):
    for a, b in zip(o, r):
        assert_same(a, b, exact_props=("nonil", "nil"))
    # properties of the ids we claim must be true; sortedness must match the reference exactly
    assert o[0].props()["sorted"] == r[0].props()["sorted"]
    assert o[0].props()["key"] == r[0].props()["key"]


@pytest.mark.parametrize("tt", ALL_TYPES)
@pytest.mark.parametrize("n,card", [(2, 2), (33, 5), (4097, 40), (100003, 300), (100003, 60000)])
def test_group_plain(gdk, ref, tt, n, card):
    rng = np.random.default_rng(n + card)
    hi = min(card, 120) if tt == T.bte else card
    a = random_col(rng, tt, n, nilfrac=0.05, lo=-(hi // 2), hi=hi // 2)
    if tt in (T.flt, T.dbl):
        a = np.round(a).astype(a.dtype)
        a[a == 0] = 1.0     # the reference hashes -0/+0 by bit pattern; steer around that quirk
    b, rb = both(gdk, ref, a, tt, hseq=7)
    _check(gdk.BATgroup(b), ref.BATgroup(rb))


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.dbl])
def test_group_subgroup(gdk, ref, tt):
    """refine a prior grouping (g, e, h given) — the Q1 returnflag/linestatus shape and wider"""
    rng = np.random.default_rng(5)
    n = 50021
    for card1, card2 in ((3, 2), (300, 7), (2, 3000)):
        a1 = random_col(rng, T.int, n, lo=0, hi=card1 - 1)
        a2 = random_col(rng, tt, n, nilfrac=0.02, lo=0, hi=min(card2 - 1, 100 if tt == T.bte else card2 - 1))
        b1, r1 = both(gdk, ref, a1, T.int)
        b2, r2 = both(gdk, ref, a2, tt)
        g1, rg1 = gdk.BATgroup(b1), ref.BATgroup(r1)
        _check(g1, rg1)
        _check(gdk.BATgroup(b2, None, g1[0], g1[1], g1[2]), ref.BATgroup(r2, None, rg1[0], rg1[1], rg1[2]))
        # without e/h the prior group count has to be found by scanning g
        _check(gdk.BATgroup(b2, None, g1[0], None, None), ref.BATgroup(r2, None, rg1[0], None, None))


@pytest.mark.parametrize("tt", [T.bte, T.int, T.lng])
def test_group_candidates(gdk, ref, tt):
    rng = np.random.default_rng(11)
    n = 30011
    a = random_col(rng, tt, n, nilfrac=0.03, lo=-50, hi=50)
    b, rb = both(gdk, ref, a, tt, hseq=100)
    o = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.uint64) + np.uint64(100)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    _check(gdk.BATgroup(b, s), ref.BATgroup(rb, rs))
    _check(gdk.BATgroup(b, gdk.BATdense(0, 1100, 5000)), ref.BATgroup(rb, ref.dense(0, 1100, 5000)))


@pytest.mark.parametrize("tt", [T.int, T.lng, T.date, T.oid])
def test_group_ranged_direct_table(gdk, ref, tt):
    """>= 2^18 rows of a 4- or 8-byte integer key that only spans a small range: one min/max pass, then a direct table
    indexed by value - min (nil has its own entry) instead of hashing — with nils, negative values, a candidate list,
    a prior grouping, a misaligned view, and a range too wide for the table (falls back to the hash path)"""
    rng = np.random.default_rng(int(tt) + 3)
    n = 300_007
    base = {T.int: -700, T.lng: -(1 << 40), T.date: 1_000_000, T.oid: 5}[tt]
    for span, nilfrac in ((900, 0.03), (3, 0.0), (70_000, 0.01), (5_000_000, 0.0)):
        a = (rng.integers(0, span, n) + base).astype(G.NP_OF[tt])
        if nilfrac and tt != T.oid:
            a[rng.random(n) < nilfrac] = np.iinfo(a.dtype).min
        b, rb = both(gdk, ref, a, tt, hseq=3)
        o, r = gdk.BATgroup(b), ref.BATgroup(rb)
        _check(o, r)
        if span == 900:
            assert "direct" in gdk.algorithm()
            # candidates (scalar gather path) and a prior grouping
            for first in (1003, 1004):      # a 16-byte aligned and a misaligned dense slice
                s, rs = gdk.BATdense(0, first, 280_001), ref.dense(0, first, 280_001)
                _check(gdk.BATgroup(b, s), ref.BATgroup(rb, rs))
            cand = np.sort(rng.choice(n, 270_000, replace=False)).astype(np.uint64) + np.uint64(3)
            s, rs = both(gdk, ref, cand, T.oid, sorted=True, key=True)
            _check(gdk.BATgroup(b, s), ref.BATgroup(rb, rs))
            a1 = rng.integers(0, 5, n).astype(np.int32)
            b1, r1 = both(gdk, ref, a1, T.int, hseq=3)
            g1, rg1 = gdk.BATgroup(b1), ref.BATgroup(r1)
            _check(gdk.BATgroup(b, None, g1[0], g1[1], g1[2]), ref.BATgroup(rb, None, rg1[0], rg1[1], rg1[2]))
        if span == 5_000_000:
            assert "hash" in gdk.algorithm()


def test_group_trivial_and_sorted(gdk, ref):
    a = np.arange(1000, dtype=np.int32)
    b, rb = both(gdk, ref, a, T.int, key=True, sorted=True)
    _check(gdk.BATgroup(b), ref.BATgroup(rb))           # 1 element per group
    a = np.zeros(5000, dtype=np.int64)
    b, rb = both(gdk, ref, a, T.lng)
    _check(gdk.BATgroup(b), ref.BATgroup(rb))           # single group
    a = np.repeat(np.arange(50, dtype=np.int32), 100)
    b, rb = both(gdk, ref, a, T.int, sorted=True)
    _check(gdk.BATgroup(b), ref.BATgroup(rb))           # compare-consecutive path of the reference
    e, re_ = both(gdk, ref, np.zeros(0, dtype=np.int32), T.int)
    _check(gdk.BATgroup(e), ref.BATgroup(re_))          # empty


def test_config1_pipeline(gdk, ref):
    """BASELINE.json config 1 (CPU GDK harness) at 1M rows: select -> project x2 -> group -> groupsum"""
    from monetdb_b200 import pipelines, workload
    n = 1_000_000
    k, v = workload.config1(0, n)
    bk, rk = both(gdk, ref, k, T.int)
    bv, rv = both(gdk, ref, v, T.lng)
    oo, ro = {}, {}
    so = pipelines.config1(gdk, {"k": bk, "v": bv}, oo)
    sr = pipelines.config1(ref, {"k": rk, "v": rv}, ro)
    for name in ("s", "g", "e", "h"):
        assert_same(oo[name], ro[name])
    assert_same(so, sr)


@pytest.mark.parametrize("tt", ALL_TYPES)
@pytest.mark.parametrize("skip_nils", [True, False])
def test_groupminmax(gdk, ref, tt, skip_nils):
    """BATgroupmin/max return the oid of the first row holding the extreme (gdk_aggr.c:3207-3240)"""
    rng = np.random.default_rng(77)
    n = 20011
    for ngrp in (1, 7, 900):
        a = random_col(rng, tt, n, nilfrac=0.1, lo=-40, hi=40)
        if tt in (T.flt, T.dbl):
            a = np.round(a).astype(a.dtype)
        gid = rng.integers(0, ngrp, n).astype(np.uint64)
        gid[:ngrp] = np.arange(ngrp, dtype=np.uint64)      # every group occurs
        b, rb = both(gdk, ref, a, tt, hseq=3)
        g, rg = both(gdk, ref, gid, T.oid, hseq=3)
        e, re_ = both(gdk, ref, np.arange(ngrp, dtype=np.uint64) + np.uint64(3), T.oid, sorted=True, key=True)
        assert_same(gdk.BATgroupmin(b, g, e, None, T.oid, skip_nils), ref.BATgroupmin(rb, rg, re_, None, T.oid, skip_nils))
        assert_same(gdk.BATgroupmax(b, g, e, None, T.oid, skip_nils), ref.BATgroupmax(rb, rg, re_, None, T.oid, skip_nils))
    # a group that never occurs and an all-nil group -> oid_nil
    a = random_col(rng, tt, 100, lo=1, hi=9)
    nilv = np.nan if tt in (T.flt, T.dbl) else None
    gid = np.array([0, 2] * 50, dtype=np.uint64)
    if tt == T.hge:
        a[1::2] = G.hge_from_ints([-(1 << 127)])[0]
    elif nilv is not None:
        a[1::2] = nilv
    else:
        a[1::2] = np.iinfo(a.dtype).min
    b, rb = both(gdk, ref, a, tt)
    g, rg = both(gdk, ref, gid, T.oid)
    e, re_ = both(gdk, ref, np.arange(3, dtype=np.uint64), T.oid, sorted=True, key=True)
    assert_same(gdk.BATgroupmin(b, g, e, None, T.oid, skip_nils), ref.BATgroupmin(rb, rg, re_, None, T.oid, skip_nils))
    assert_same(gdk.BATgroupmax(b, g, e, None, T.oid, skip_nils), ref.BATgroupmax(rb, rg, re_, None, T.oid, skip_nils))


def test_q1_pipeline_small(gdk, ref):
    """TPC-H Q1 shape (config 3) on 300k rows: every output bit-exact (avgs are exact-formula results)"""
    from monetdb_b200 import pipelines, workload
    n = 300_000
    cols = ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus")
    host = workload.lineitem(0, n, cols)
    types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
             "returnflag": T.bte, "linestatus": T.bte}
    dev, rf = {}, {}
    for k in cols:
        dev[k], rf[k] = both(gdk, ref, host[k], types[k])
    oo, ro = {}, {}
    o = pipelines.q1(gdk, dev, oo)
    r = pipelines.q1(ref, rf, ro)
    assert oo["nsel"] == ro["nsel"]
    assert_same(oo["groups"], ro["groups"])
    for k in o:
        assert_same(o[k], r[k])
    assert len(o["extents"]) == 4
