"""BATcalc arithmetic, BATconvert, BATsum, BATgroupsum/avg/count: CUDA path vs
the real reference GDK.  Integer results bit-exact, error texts identical;
flt/dbl sums within 1e-12 (dbl) / 1e-5 (flt) relative as BASELINE.json states."""
import numpy as np
import pytest

from tests.util import ALL_TYPES, INT_TYPES, assert_same, both, random_col, same_bits
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE
RANK = {T.bte: 0, T.sht: 1, T.int: 2, T.lng: 3, T.hge: 4, T.flt: 5, T.dbl: 6}


def _int_triples():
    for t1 in INT_TYPES:
        for t2 in INT_TYPES:
            for t3 in INT_TYPES:
                if RANK[t3] >= max(RANK[t1], RANK[t2]):
                    yield t1, t2, t3


def _cmp_or_err(gdk, ref, fo, fr):
    """run both sides; either both succeed with identical bits or both fail with the same text"""
    try:
        r = fr()
    except ref.RefError as e:
        with pytest.raises(G.GDKError) as ei:
            fo()
        assert str(ei.value).strip() == str(e).strip(), (str(ei.value), str(e))
        return None
    o = fo()
    assert_same(o, r)
    return o


@pytest.mark.parametrize("op", ["add", "sub", "mul", "div"])
def test_calc_int_matrix(gdk, ref, op):
    rng = np.random.default_rng(42)
    n = 4099
    for t1, t2, t3 in _int_triples():
        a1 = random_col(rng, t1, n, nilfrac=0.03, lo=-11, hi=11)
        a2 = random_col(rng, t2, n, nilfrac=0.03, lo=1 if op == "div" else -11, hi=11)
        b1, r1 = both(gdk, ref, a1, t1)
        b2, r2 = both(gdk, ref, a2, t2)
        _cmp_or_err(gdk, ref, lambda: gdk.BATcalc(op, b1, b2, None, None, t3), lambda: ref.BATcalc(op, r1, r2, None, None, t3))


@pytest.mark.parametrize("op", ["add", "sub", "mul"])
@pytest.mark.parametrize("tt", INT_TYPES)
def test_calc_overflow_message(gdk, ref, op, tt):
    """first overflowing row aborts with the reference's text"""
    rng = np.random.default_rng(1)
    n = 10007
    mx = (1 << (8 * G.NP_OF[tt].itemsize - 1)) - 1
    if tt == T.hge:
        a1 = [int(x) for x in rng.integers(-5, 5, n)]
        a2 = [int(x) for x in rng.integers(-5, 5, n)]
        a1[5003], a2[5003] = mx - 1, (-mx if op == "sub" else mx)
        a1[7001], a2[7001] = mx, mx
        a1, a2 = G.hge_from_ints(a1), G.hge_from_ints(a2)
    else:
        a1 = rng.integers(-5, 5, n).astype(G.NP_OF[tt])
        a2 = rng.integers(-5, 5, n).astype(G.NP_OF[tt])
        a1[5003], a2[5003] = mx - 1, (-mx if op == "sub" else mx)
        a1[7001], a2[7001] = mx, mx
    b1, r1 = both(gdk, ref, a1, tt)
    b2, r2 = both(gdk, ref, a2, tt)
    with pytest.raises(ref.RefError) as er:
        ref.BATcalc(op, r1, r2, None, None, tt)
    with pytest.raises(G.GDKError) as eo:
        gdk.BATcalc(op, b1, b2, None, None, tt)
    assert "22003!overflow in calculation" in str(eo.value)
    assert str(eo.value).strip() == str(er.value).strip()


def test_calc_div_by_zero(gdk, ref):
    a1 = np.array([4, 5, 6], dtype=np.int64)
    a2 = np.array([2, 0, 3], dtype=np.int64)
    b1, r1 = both(gdk, ref, a1, T.lng)
    b2, r2 = both(gdk, ref, a2, T.lng)
    with pytest.raises(G.GDKError, match="22012!division by zero"):
        gdk.BATcalc("div", b1, b2, None, None, T.lng)
    with pytest.raises(ref.RefError, match="22012!division by zero"):
        ref.BATcalc("div", r1, r2, None, None, T.lng)


@pytest.mark.parametrize("op", ["add", "sub", "mul", "div"])
def test_calc_float(gdk, ref, op):
    rng = np.random.default_rng(7)
    n = 5003
    combos = [(T.flt, T.flt, T.flt), (T.flt, T.flt, T.dbl), (T.dbl, T.dbl, T.dbl), (T.flt, T.dbl, T.dbl), (T.dbl, T.flt, T.dbl),
              (T.int, T.flt, T.flt), (T.lng, T.dbl, T.dbl), (T.dbl, T.int, T.dbl), (T.int, T.int, T.dbl), (T.lng, T.lng, T.flt),
              (T.hge, T.dbl, T.dbl), (T.bte, T.sht, T.flt)]
    for t1, t2, t3 in combos:
        a1 = random_col(rng, t1, n, nilfrac=0.03, lo=-100, hi=100)
        a2 = random_col(rng, t2, n, nilfrac=0.03, lo=1, hi=100)
        b1, r1 = both(gdk, ref, a1, t1)
        b2, r2 = both(gdk, ref, a2, t2)
        _cmp_or_err(gdk, ref, lambda: gdk.BATcalc(op, b1, b2, None, None, t3), lambda: ref.BATcalc(op, r1, r2, None, None, t3))


@pytest.mark.parametrize("op", ["add", "sub", "mul", "div"])
def test_calc_cst_and_candidates(gdk, ref, op):
    rng = np.random.default_rng(9)
    n = 30011
    a = random_col(rng, T.lng, n, nilfrac=0.02, lo=-1000, hi=1000)
    b, rb = both(gdk, ref, a, T.lng, hseq=4)
    o = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.uint64) + np.uint64(4)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    for cand, rcand in ((None, None), (s, rs), (gdk.BATdense(0, 104, 999), ref.dense(0, 104, 999))):
        for first in (False, True):
            for tp in (T.lng, T.hge):
                _cmp_or_err(gdk, ref, lambda: gdk.BATcalccst(op, b, 100, T.lng, cand, tp, cst_first=first),
                            lambda: ref.BATcalccst(op, rb, 100, T.lng, rcand, tp, cst_first=first))
    # two BATs through different candidate lists of equal size
    a2 = random_col(rng, T.int, n, lo=1, hi=50)
    b2, rb2 = both(gdk, ref, a2, T.int, hseq=4)
    _cmp_or_err(gdk, ref, lambda: gdk.BATcalc(op, b, b2, s, s, T.hge), lambda: ref.BATcalc(op, rb, rb2, rs, rs, T.hge))
    with pytest.raises(G.GDKError, match="inputs not the same size"):
        gdk.BATcalc(op, b, b2, s, None, T.hge)


@pytest.mark.parametrize("t1", ALL_TYPES)
@pytest.mark.parametrize("t2", ALL_TYPES)
def test_convert(gdk, ref, t1, t2):
    rng = np.random.default_rng(t1 * 16 + t2)
    a = random_col(rng, t1, 10007, nilfrac=0.03, lo=-120, hi=120)
    b, rb = both(gdk, ref, a, t1)
    _cmp_or_err(gdk, ref, lambda: gdk.BATconvert(b, None, t2), lambda: ref.BATconvert(rb, None, t2))
    a = random_col(rng, t1, 1009, nilfrac=0.0, lo=-30000, hi=30000) if RANK[t1] >= 1 else a
    b, rb = both(gdk, ref, a, t1)
    s, rs = gdk.BATdense(0, 10, 500), ref.dense(0, 10, 500)
    _cmp_or_err(gdk, ref, lambda: gdk.BATconvert(b, s, t2), lambda: ref.BATconvert(rb, rs, t2))


def _sum_types(t1):
    if t1 in INT_TYPES:
        return [t for t in INT_TYPES if RANK[t] >= RANK[t1]] + [T.dbl]
    return [T.flt, T.dbl] if t1 == T.flt else [T.dbl]


def _close(a, b, tt, scale=None):
    """|ours - reference| <= tol * max(|reference|, scale): 1e-12 (dbl) / 1e-5 (flt) relative, as BASELINE.json
    states.  `scale` (the mean |input| of the group) bounds the error of results that cancel to
    almost nothing — the reference's own running mean (AVERAGE_ITER_FLOAT) is only accurate to that."""
    tol = 1e-5 if tt == T.flt else 1e-12
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    ref = np.maximum(np.abs(b[m]), 1e-300)
    if scale is not None:
        ref = np.maximum(ref, np.broadcast_to(np.asarray(scale, dtype=np.float64), a.shape)[m])
    assert np.all(np.abs(a[m] - b[m]) <= tol * ref), (a[m][:5], b[m][:5])


@pytest.mark.parametrize("t1", ALL_TYPES)
def test_batsum(gdk, ref, t1):
    rng = np.random.default_rng(t1)
    for n in (0, 1, 1000, 200003):
        for nilfrac in (0.0, 0.1):
            a = random_col(rng, t1, n, nilfrac=nilfrac, lo=-100, hi=100)
            b, rb = both(gdk, ref, a, t1)
            for tp in _sum_types(t1):
                if n == 0 and t1 == T.flt and tp == T.dbl:
                    continue    # the reference stores a 4-byte flt nil into the 8-byte result here (gdk_aggr.c:1160-1163)
                for skip in (True, False):
                    try:
                        r = ref.BATsum(tp, rb, None, skip, True)
                    except ref.RefError as e:
                        with pytest.raises(G.GDKError) as ei:
                            gdk.BATsum(tp, b, None, skip, True)
                        assert str(ei.value).strip() == str(e).strip()
                        continue
                    o = gdk.BATsum(tp, b, None, skip, True)
                    if tp in (T.flt, T.dbl):
                        _close(o, r, tp)
                    else:
                        assert same_bits(o, r), (t1, tp, n, nilfrac, skip, o, r)


def test_batsum_overflow_and_transient(gdk, ref):
    """running-prefix overflow semantics of AGGR_SUM: a transient overflow errors even if the total fits"""
    mx = (1 << 63) - 1
    cases = [np.array([mx, 1, -5], dtype=np.int64),          # prefix overflows, total fits -> error
             np.array([mx - 10, -20, 15], dtype=np.int64),   # P > max but no prefix overflows -> ok
             np.array([-mx, -1, 5], dtype=np.int64),
             np.array([mx, -mx, mx, -mx, 3], dtype=np.int64)]
    for a in cases:
        b, rb = both(gdk, ref, a, T.lng)
        try:
            r = ref.BATsum(T.lng, rb)
        except ref.RefError as e:
            with pytest.raises(G.GDKError, match="22003!overflow in sum aggregate"):
                gdk.BATsum(T.lng, b)
            continue
        assert same_bits(gdk.BATsum(T.lng, b), r), a


@pytest.mark.parametrize("ngrp", [1, 3, 4, 8, 9, 500, 7000, 100000])
@pytest.mark.parametrize("t1", [T.bte, T.int, T.lng, T.hge, T.dbl])
def test_groupsum_avg_count(gdk, ref, ngrp, t1):
    rng = np.random.default_rng(ngrp + t1)
    n = 120011
    a = random_col(rng, t1, n, nilfrac=0.05, lo=-1000 if t1 != T.bte else -100, hi=1000 if t1 != T.bte else 100)
    gids = rng.integers(0, ngrp, n, dtype=np.uint64)
    b, rb = both(gdk, ref, a, t1, hseq=2)
    g, rg = both(gdk, ref, gids, T.oid, hseq=2)
    e, re_ = gdk.BATdense(0, 0, ngrp), ref.dense(0, 0, ngrp)   # extents: only count/hseqbase are used
    for skip in (True, False):
        for tp in _sum_types(t1):
            if tp == T.dbl and t1 in INT_TYPES:
                continue
            try:
                r = ref.BATgroupsum(rb, rg, re_, None, tp, skip)
            except ref.RefError as ex:
                with pytest.raises(G.GDKError) as ei:
                    gdk.BATgroupsum(b, g, e, None, tp, skip)
                assert str(ei.value).strip() == str(ex).strip()
                continue
            o = gdk.BATgroupsum(b, g, e, None, tp, skip)
            if tp in (T.flt, T.dbl):
                _close(o.numpy(), r.numpy(), tp)
            else:
                assert_same(o, r)
        r, rc = ref.BATgroupavg(rb, rg, re_, None, T.dbl, skip, 2)
        o, oc = gdk.BATgroupavg(b, g, e, None, T.dbl, skip, 2)
        if t1 in INT_TYPES:
            assert same_bits(o.numpy(), r.numpy())
        else:
            _close(o.numpy(), r.numpy(), T.dbl, scale=10.0)     # inputs are |x| <= 1000, scaled by 10^-2
        assert same_bits(oc.numpy(), rc.numpy())
        assert_same(gdk.BATgroupcount(b, g, e, None, T.lng, skip), ref.BATgroupcount(rb, rg, re_, None, T.lng, skip))
    # without extents: the group range comes from scanning g
    assert_same(gdk.BATgroupsum(b, g, None, None, T.hge if t1 in INT_TYPES else T.dbl, True),
                ref.BATgroupsum(rb, rg, None, None, T.hge if t1 in INT_TYPES else T.dbl, True)) if t1 in INT_TYPES else None


@pytest.mark.parametrize("ngrp", [9, 300, 1000])
@pytest.mark.parametrize("t1,tp,bits", [(T.int, T.lng, 31), (T.lng, T.hge, 62), (T.lng, T.lng, 40), (T.lng, T.lng, 62),
                                         (T.hge, T.hge, 100), (T.hge, T.hge, 126), (T.int, T.int, 20), (T.int, T.int, 30)])
def test_groupsum_wide_values_shared_table(gdk, ref, ngrp, t1, tp, bits):
    """no nils, 9..1000 groups: the shared-memory table of 32-bit halves (aggr_tab32_kernel).  Values span
    `bits` bits of both signs, so the low halves wrap many times and the high halves see negative digits; the
    narrow result types run the overflow-checked form and, for the widest values, raise the reference's error"""
    rng = np.random.default_rng(ngrp * 131 + bits)
    n = 200003
    if t1 == T.hge:
        a = G.hge_from_ints([int(x) * (1 << (bits - 62)) + int(y)
                             for x, y in zip(rng.integers(-(1 << 62), 1 << 62, n), rng.integers(0, 1 << 30, n))])
    else:
        a = rng.integers(-(1 << bits), 1 << bits, n, dtype=np.int64).astype(np.int32 if t1 == T.int else np.int64)
    gids = rng.integers(0, ngrp, n, dtype=np.uint64)
    gids[:ngrp] = np.arange(ngrp, dtype=np.uint64)
    b, rb = both(gdk, ref, a, t1, hseq=0)
    g, rg = both(gdk, ref, gids, T.oid, hseq=0)
    e, re_ = gdk.BATdense(0, 0, ngrp), ref.dense(0, 0, ngrp)
    _cmp_or_err(gdk, ref, lambda: gdk.BATgroupsum(b, g, e, None, tp, True), lambda: ref.BATgroupsum(rb, rg, re_, None, tp, True))
    if t1 != T.hge:
        # all-positive values: the sum of positives equals the sum, overflow is reached monotonically
        b2, rb2 = both(gdk, ref, np.abs(np.asarray(a)), t1, hseq=0)
        _cmp_or_err(gdk, ref, lambda: gdk.BATgroupsum(b2, g, e, None, tp, True), lambda: ref.BATgroupsum(rb2, rg, re_, None, tp, True))
    assert_same(gdk.BATgroupcount(b, g, e, None, T.lng, True), ref.BATgroupcount(rb, rg, re_, None, T.lng, True))


@pytest.mark.parametrize("ngrp", [1, 4, 9, 500, 7000])
@pytest.mark.parametrize("t1", [T.bte, T.sht, T.int, T.lng, T.hge])
def test_groupavg3_and_combine(gdk, ref, ngrp, t1):
    """mergetable's exact integer average: phase 1 triples per slice (BATgroupavg3, gdk_aggr.c:1996),
    phase 2 recombination over the concatenated partials (BATgroupavg3combine, :2634)"""
    rng = np.random.default_rng(31 * ngrp + t1)
    n = 60013
    # lng: the reference's combine_averages_lng multiplies avg * cnt in a lng before widening
    # (gdk_aggr.c:2531), so keep |avg| * cnt below 2^63 where both sides are exact
    big = {T.bte: 127, T.sht: 32767, T.int: 2**31 - 1, T.lng: 2**44, T.hge: 1 << 100}[t1]
    e, re_ = gdk.BATdense(0, 0, ngrp), ref.dense(0, 0, ngrp)
    parts = []
    for slice_no, (lo, hi, nilfrac) in enumerate([(-big, big, 0.05), (-1000 if t1 != T.bte else -100, 0, 0.0),
                                                  (0, big, 0.3)]):
        a = random_col(rng, t1, n, nilfrac=nilfrac, lo=lo, hi=hi)
        gids = rng.integers(0, ngrp, n, dtype=np.uint64)
        if ngrp > 4:
            gids[gids == 3] = 2          # one group stays empty
        b, rb = both(gdk, ref, a, t1, hseq=5)
        g, rg = both(gdk, ref, gids, T.oid, hseq=5)
        for skip in (True, False):
            o = gdk.BATgroupavg3(b, g, e, None, skip)
            r = ref.BATgroupavg3(rb, rg, re_, None, skip)
            for x, y in zip(o, r):
                assert_same(x, y)
        parts.append((gdk.BATgroupavg3(b, g, e, None, True), ref.BATgroupavg3(rb, rg, re_, None, True)))
    if t1 == T.hge:
        return                            # the recombination of hge partials is not on the device path
    # phase 2: concatenate the partial triples, group id of partial row i of every slice = i
    cat = [np.concatenate([p[1][k].numpy() for p in parts]) for k in range(3)]
    pg = np.concatenate([np.arange(ngrp, dtype=np.uint64) for _ in parts])
    if ngrp > 4:                          # empty group: avg nil, cnt 0 in every slice
        assert all(p[1][2].numpy()[3] == 0 for p in parts)
    av, rav = both(gdk, ref, cat[0], t1)
    rm, rrm = both(gdk, ref, cat[1], T.lng)
    ct, rct = both(gdk, ref, cat[2], T.lng)
    g2, rg2 = both(gdk, ref, pg, T.oid)
    for skip in (True, False):
        assert_same(gdk.BATgroupavg3combine(av, rm, ct, g2, e, skip), ref.BATgroupavg3combine(rav, rrm, rct, rg2, re_, skip))


@pytest.mark.parametrize("t1", ALL_TYPES)
def test_calcavg(gdk, ref, t1):
    rng = np.random.default_rng(77 + t1)
    for n, nilfrac in ((0, 0), (1, 0), (5, 1.0), (100003, 0.1), (1 << 20, 0.0)):
        a = random_col(rng, t1, n, nilfrac=nilfrac)
        b, rb = both(gdk, ref, a, t1, hseq=3)
        for scale in (0, 2):
            o, r = gdk.BATcalcavg(b, None, scale), ref.BATcalcavg(rb, None, scale)
            assert o[1] == r[1]
            if t1 in INT_TYPES:
                assert same_bits(np.float64(o[0]), np.float64(r[0])), (n, o, r)
            else:
                assert (np.isnan(o[0]) and np.isnan(r[0])) or abs(o[0] - r[0]) <= 1e-9 * 1000, (n, o, r)
    # the lng sum overflows a lng but not a hge: still (dbl) sum / n
    a = np.full(1000, 2**63 - 1, dtype=np.int64)
    b, rb = both(gdk, ref, a, T.lng)
    assert gdk.BATcalcavg(b) == ref.BATcalcavg(rb)
    # candidates
    a = random_col(rng, T.int, 50000, nilfrac=0.05)
    b, rb = both(gdk, ref, a, T.int, hseq=10)
    cand = np.sort(rng.choice(50000, 7000, replace=False)).astype(np.uint64) + 10
    s, rs = both(gdk, ref, cand, T.oid, sorted=True, key=True)
    assert gdk.BATcalcavg(b, s) == ref.BATcalcavg(rb, rs)
