"""BATgroupprod (gdk_aggr.c:1575) against the reference: every result type the reference's table allows, with
zeros, nils in front of / behind a group's first value (the AGGR_PROD `seen` quirk), both skip_nils settings,
candidate lists, missing extents, group ids outside the extents, and the overflow verdict
("22003!overflow in product aggregate.") — which depends on where a group's first zero sits.  Integer results are
bit-exact (order-independent exact form), flt/dbl results are bit-exact too (ordered replay per group)."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import INT_TYPES, assert_same, both, same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE

RESULTS = {
    T.bte: [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl],
    T.sht: [T.sht, T.int, T.lng, T.hge, T.flt, T.dbl],
    T.int: [T.int, T.lng, T.hge, T.flt, T.dbl],
    T.lng: [T.lng, T.hge, T.flt, T.dbl],
    T.hge: [T.hge, T.flt, T.dbl],
    T.flt: [T.flt, T.dbl],
    T.dbl: [T.dbl],
}


def column(rng, tt, n, small, zerofrac, nilfrac):
    """mostly +-1 with a few larger factors (so that big groups stay inside lng / hge), zeros and nils"""
    mags = np.where(rng.random(n) < small, rng.integers(2, 4, n), 1)
    v = mags * rng.choice([-1, 1], n)
    v[rng.random(n) < zerofrac] = 0
    nil = rng.random(n) < nilfrac
    if tt in (T.flt, T.dbl):
        a = (v * rng.choice([0.5, 1.0, 1.25, 3.0], n)).astype(G.NP_OF[tt])
        a[nil] = np.nan
        return a
    if tt == T.hge:
        return G.hge_from_ints([-(1 << 127) if nil[i] else int(x) for i, x in enumerate(v)])
    a = v.astype(G.NP_OF[tt])
    a[nil] = np.iinfo(G.NP_OF[tt]).min
    return a


def check(gdk, ref, b, rb, g, rg, e, re_, s, rs, tp, skip):
    try:
        r = ref.BATgroupprod(rb, rg, re_, rs, tp, skip)
    except ref.RefError as ex:
        with pytest.raises(G.GDKError) as ei:
            gdk.BATgroupprod(b, g, e, s, tp, skip)
        assert str(ei.value).strip() == str(ex).strip()
        return "error"
    o = gdk.BATgroupprod(b, g, e, s, tp, skip)
    assert same_bits(o.numpy(), r.numpy()), (tp, skip, o.numpy()[:8], r.numpy()[:8])
    assert_same(o, r)
    return "ok"


@pytest.mark.parametrize("t1", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
@pytest.mark.parametrize("shape", [(20011, 3000, 0.5, 0.02, 0.03), (100003, 5, 0.0004, 0.0, 0.0), (100003, 40, 0.004, 0.00005, 0.0001),
                                   (5003, 1, 0.004, 0.0, 0.001), (777, 700, 1.0, 0.1, 0.2)])
def test_groupprod(gdk, ref, t1, shape):
    n, ngrp, small, zf, nf = shape
    rng = np.random.default_rng(n + ngrp + t1)
    a = column(rng, t1, n, small, zf, nf)
    gids = rng.integers(0, ngrp, n, dtype=np.uint64)
    b, rb = both(gdk, ref, a, t1, hseq=4)
    g, rg = both(gdk, ref, gids, T.oid, hseq=4)
    e, re_ = gdk.BATdense(0, 0, ngrp), ref.dense(0, 0, ngrp)
    seen = set()
    for skip in (True, False):
        for tp in RESULTS[t1]:
            seen.add(check(gdk, ref, b, rb, g, rg, e, re_, None, None, tp, skip))
    if t1 in INT_TYPES:
        assert "ok" in seen
    # without extents (group range scanned from g) and with fewer extents than group ids (ids beyond are ignored)
    tp = RESULTS[t1][-1] if t1 in INT_TYPES else T.dbl
    check(gdk, ref, b, rb, g, rg, None, None, None, None, tp, True)
    if ngrp > 4:
        e2, re2 = gdk.BATdense(0, 0, ngrp // 2), ref.dense(0, 0, ngrp // 2)
        check(gdk, ref, b, rb, g, rg, e2, re2, None, None, tp, False)


@pytest.mark.parametrize("t1", [T.int, T.lng, T.dbl])
def test_groupprod_candidates(gdk, ref, t1):
    """the reference reads g at the candidate's ROW position (gids[oid - b->hseqbase], gdk_aggr.c:1192-1200) while it
    demands g to be aligned with the CANDIDATES, so only a candidate list that is a prefix of b is well-defined there
    (anything else reads g out of bounds); that form is compared, the others must give the same 'aligned' error"""
    rng = np.random.default_rng(t1)
    n, ngrp = 50021, 900
    a = column(rng, t1, n, 0.3, 0.01, 0.02)
    b, rb = both(gdk, ref, a, t1, hseq=10)
    m = n // 3
    gids = rng.integers(0, ngrp, m, dtype=np.uint64)
    g, rg = both(gdk, ref, gids, T.oid, hseq=10)
    e, re_ = gdk.BATdense(0, 0, ngrp), ref.dense(0, 0, ngrp)
    sd, rsd = gdk.BATdense(0, 10, m), ref.dense(0, 10, m)
    for skip in (True, False):
        for tp in RESULTS[t1][-2:]:
            assert check(gdk, ref, b, rb, g, rg, e, re_, sd, rsd, tp, skip) in ("ok", "error")
    # a candidate list that does not start where g starts: both sides refuse
    c = np.sort(rng.choice(np.arange(5, n), m, replace=False)).astype(np.uint64) + 10
    s, rs = gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True), ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)
    assert check(gdk, ref, b, rb, g, rg, e, re_, s, rs, RESULTS[t1][-1], True) == "error"


def test_nil_order_quirk_and_zero_cut(gdk, ref):
    """hand-made groups: nil before the first value (forgotten for results narrower than hge, final for hge),
    nil behind it (final), an overflowing tail behind a zero (no error), the same tail in front of it (error)"""
    nil = np.iinfo(np.int64).min
    big = 1 << 40
    rows = [  # (gid, value)
        (0, nil), (0, 3), (0, -5),              # leading nil
        (1, 7), (1, nil), (1, 2),               # nil behind the first value
        (2, 0), (2, big), (2, big),             # overflowing factors behind a zero
        (3, nil), (3, nil),                     # only nils
        (4, -1), (4, -1), (4, -1),
    ]
    gids = np.array([r[0] for r in rows], dtype=np.uint64)
    vals = np.array([r[1] for r in rows], dtype=np.int64)
    # interleave the groups round-robin; the order inside each group is kept
    rank = np.array([sum(1 for q in rows[:i] if q[0] == r[0]) for i, r in enumerate(rows)])
    order = np.lexsort((-gids.astype(np.int64), rank))
    gids, vals = gids[order], vals[order]
    b, rb = both(gdk, ref, vals, T.lng)
    g, rg = both(gdk, ref, gids, T.oid)
    e, re_ = gdk.BATdense(0, 0, 6), ref.dense(0, 0, 6)    # group 5 is empty
    for tp in (T.lng, T.hge, T.dbl):
        for skip in (True, False):
            assert check(gdk, ref, b, rb, g, rg, e, re_, None, None, tp, skip) == "ok"
    r = gdk.BATgroupprod(b, g, e, None, T.lng, False).numpy()
    assert r.tolist() == [-15, nil, 0, nil, -1, nil]
    # now the big factors come before the zero: the running product overflows lng before it is zeroed
    vals2 = vals.copy()
    z = np.flatnonzero((gids == 2))
    vals2[z] = [big, big, 0]
    b2, rb2 = both(gdk, ref, vals2, T.lng)
    assert check(gdk, ref, b2, rb2, g, rg, e, re_, None, None, T.lng, True) == "error"
    assert check(gdk, ref, b2, rb2, g, rg, e, re_, None, None, T.hge, True) == "ok"


def test_float_overflow_depends_on_order(gdk, ref):
    a = np.array([1e200, 1e-300, 1e200, 1e200, 1e200, 1e-300], dtype=np.float64)
    gids = np.array([0, 0, 0, 1, 1, 1], dtype=np.uint64)
    b, rb = both(gdk, ref, a, T.dbl)
    g, rg = both(gdk, ref, gids, T.oid)
    e, re_ = gdk.BATdense(0, 0, 2), ref.dense(0, 0, 2)
    assert check(gdk, ref, b, rb, g, rg, e, re_, None, None, T.dbl, True) == "error"     # group 1: 1e200 * 1e200 first
    # one extent: group 1 is out of range and ignored; group 0 shrinks in between and stays finite
    assert check(gdk, ref, b, rb, g, rg, gdk.BATdense(0, 0, 1), ref.dense(0, 0, 1), None, None, T.dbl, True) == "ok"


def test_trivial_forms(gdk, ref):
    rng = np.random.default_rng(3)
    a = rng.integers(-50, 50, 1000).astype(np.int32)
    b, rb = both(gdk, ref, a, T.int)
    # singleton groups -> BATconvert
    g, rg = gdk.BATdense(0, 0, 1000), ref.dense(0, 0, 1000)
    assert_same(gdk.BATgroupprod(b, g, None, None, T.lng, True), ref.BATgroupprod(rb, rg, None, None, T.lng, True))
    # no group column
    with pytest.raises(G.GDKError) as e1:
        gdk.BATgroupprod(b, None, None, None, T.lng, True)
    with pytest.raises(ref.RefError) as e2:
        ref.BATgroupprod(rb, None, None, None, T.lng, True)
    assert str(e1.value).strip() == str(e2.value).strip()
    # empty input, some extents
    eb, reb = both(gdk, ref, a[:0], T.int)
    eg, reg = both(gdk, ref, np.zeros(0, dtype=np.uint64), T.oid)
    e, re_ = gdk.BATdense(0, 0, 3), ref.dense(0, 0, 3)
    assert_same(gdk.BATgroupprod(eb, eg, e, None, T.lng, True), ref.BATgroupprod(reb, reg, re_, None, T.lng, True))
    # unsupported combination
    g5, rg5 = both(gdk, ref, rng.integers(0, 5, 1000, dtype=np.uint64), T.oid)
    with pytest.raises(G.GDKError) as e1:
        gdk.BATgroupprod(b, g5, None, None, T.bte, True)
    with pytest.raises(ref.RefError) as e2:
        ref.BATgroupprod(rb, rg5, None, None, T.bte, True)
    assert str(e1.value).strip() == str(e2.value).strip()


# ------------------------------------------------------------------ BATprod, the scalar form (gdk_aggr.c:1650)
@pytest.mark.parametrize("t1", [T.bte, T.int, T.lng, T.hge, T.flt, T.dbl])
def test_batprod(gdk, ref, t1):
    rng = np.random.default_rng(t1 + 55)
    for n, small, zf, nf in ((0, 0, 0, 0), (1, 1.0, 0, 0), (300_007, 0.00005, 0.0, 0.0), (300_007, 0.00005, 0.0, 0.00001),
                             (300_007, 0.0001, 0.00001, 0.0), (5000, 0.05, 0.0, 0.0), (64, 0.3, 0.02, 0.05)):
        a = column(rng, t1, n, small, zf, nf)
        b, rb = both(gdk, ref, a, t1, hseq=3)
        cands = [(None, None)]
        if n > 10:
            c = np.sort(rng.choice(n, n // 2, replace=False)).astype(np.uint64) + 3
            cands.append((gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True), ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)))
        for s, rs in cands:
            for tp in RESULTS[t1]:
                for skip in (True, False):
                    for nie in (True, False):
                        try:
                            want = ref.BATprod(tp, rb, rs, skip, nie)
                        except ref.RefError as ex:
                            with pytest.raises(G.GDKError) as ei:
                                gdk.BATprod(tp, b, s, skip, nie)
                            assert str(ei.value).strip() == str(ex).strip()
                            continue
                        got = gdk.BATprod(tp, b, s, skip, nie)
                        assert same_bits(got, want), (n, tp, skip, nie, got, want)
