"""The leftjoin family (gdk_join.c:4049 leftjoin; BATleftjoin/BATouterjoin/BATsemijoin/BATintersect/
BATdiff :4320-4405) and BATunique (gdk_unique.c) on the device vs the real reference GDK.

r1 is in l's candidate order on both sides; the order of several right matches of one left row is
not specified, so pair lists are compared after a canonical sort.  A semijoin may return any one
match per left row (the reference's own algorithms differ in which), so r1 is compared exactly and
r2 is checked to hold an equal value.  Candidate-list results (intersect, diff, unique) are compared
oid for oid."""
import numpy as np
import pytest

from tests.util import assert_props_true, both, random_col, same_bits
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE
NIL = np.uint64(1 << 63)


def _pairs(r1, r2):
    a, b = r1.numpy().astype(np.uint64), r2.numpy().astype(np.uint64)
    assert len(a) == len(b)
    o = np.lexsort((b, a))
    return a[o], b[o]


def _same_pairs(o1, o2, e1, e2):
    a, b = _pairs(o1, o2)
    c, d = _pairs(e1, e2)
    assert np.array_equal(a, c) and np.array_equal(b, d), (len(a), len(c))
    x = o1.numpy()
    assert np.all(x[:-1] <= x[1:])
    assert_props_true(o1)
    assert_props_true(o2)


def _cols(gdk, ref, tt, seed, nl=20011, nr=5003, hi=6000, nilfrac=0.03, lhseq=11, rhseq=5):
    rng = np.random.default_rng(seed)
    a = random_col(rng, tt, nl, nilfrac=nilfrac, lo=0, hi=100 if tt == T.bte else hi)
    b = random_col(rng, tt, nr, nilfrac=nilfrac, lo=0, hi=100 if tt == T.bte else hi)
    if tt in (T.flt, T.dbl):
        a, b = np.round(a).astype(a.dtype) + 1, np.round(b).astype(b.dtype) + 1
    l, rl = both(gdk, ref, a, tt, hseq=lhseq)
    r, rr = both(gdk, ref, b, tt, hseq=rhseq)
    return a, b, l, rl, r, rr


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
@pytest.mark.parametrize("nil_matches", [False, True])
def test_leftjoin_outerjoin(gdk, ref, tt, nil_matches):
    a, b, l, rl, r, rr = _cols(gdk, ref, tt, 100 + tt)
    _same_pairs(*gdk.BATleftjoin(l, r, None, None, nil_matches), *ref.BATleftjoin(rl, rr, None, None, nil_matches))
    o1, o2 = gdk.BATouterjoin(l, r, None, None, nil_matches)
    e1, e2 = ref.BATouterjoin(rl, rr, None, None, nil_matches)
    _same_pairs(o1, o2, e1, e2)
    # every row of l appears at least once; misses carry oid_nil
    assert len(np.unique(o1.numpy())) == len(a)
    miss = o2.numpy().astype(np.uint64) == NIL
    if tt != T.bte:
        assert miss.any()
    p = o2.props()
    assert p["nil"] == bool(miss.any()) and p["nonil"] == (not miss.any())


@pytest.mark.parametrize("tt", [T.bte, T.int, T.lng, T.hge, T.dbl])
@pytest.mark.parametrize("nil_matches", [False, True])
def test_semijoin_intersect_diff(gdk, ref, tt, nil_matches):
    a, b, l, rl, r, rr = _cols(gdk, ref, tt, 200 + tt)
    o1, o2 = gdk.BATsemijoin(l, r, None, None, nil_matches)
    e1, e2 = ref.BATsemijoin(rl, rr, None, None, nil_matches)
    x, y = o1.numpy().astype(np.uint64), o2.numpy().astype(np.uint64)
    assert np.array_equal(x, e1.numpy().astype(np.uint64))
    assert len(np.unique(x)) == len(x)
    # r2 points at a row of r that holds the same value
    va, vb = a[(x - np.uint64(11)).astype(np.int64)], b[(y - np.uint64(5)).astype(np.int64)]
    assert same_bits(va, vb)
    assert len(e2.numpy()) == len(y)
    i = gdk.BATintersect(l, r, None, None, nil_matches)
    assert np.array_equal(i.numpy().astype(np.uint64), ref.BATintersect(rl, rr, None, None, nil_matches).numpy().astype(np.uint64))
    assert np.array_equal(i.numpy().astype(np.uint64), x)
    assert_props_true(i)
    d = gdk.BATdiff(l, r, None, None, nil_matches)
    ed = ref.BATdiff(rl, rr, None, None, nil_matches)
    assert np.array_equal(d.numpy().astype(np.uint64), ed.numpy().astype(np.uint64))
    assert_props_true(d)
    # intersect and diff split l's rows
    assert len(i) + len(d) == len(a)


@pytest.mark.parametrize("rnil", [False, True])
def test_diff_not_in(gdk, ref, rnil):
    """NOT IN: a nil on the right empties the result, a nil on the left never qualifies"""
    a, b, l, rl, r, rr = _cols(gdk, ref, T.int, 7, nilfrac=0.0)
    a = a.copy()
    a[::97] = np.iinfo(np.int32).min
    if rnil:
        b = b.copy()
        b[13] = np.iinfo(np.int32).min
    l, rl = both(gdk, ref, a, T.int, hseq=11)
    r, rr = both(gdk, ref, b, T.int, hseq=5)
    for nm in (False, True):
        d = gdk.BATdiff(l, r, None, None, nm, True)
        ed = ref.BATdiff(rl, rr, None, None, nm, True)
        assert np.array_equal(d.numpy().astype(np.uint64), ed.numpy().astype(np.uint64))
        assert (len(d) == 0) == rnil


def test_leftjoin_family_candidates(gdk, ref):
    rng = np.random.default_rng(3)
    a, b, l, rl, r, rr = _cols(gdk, ref, T.lng, 9, nl=30011, nr=7001, hi=9000, lhseq=3, rhseq=9)
    ol = np.sort(rng.choice(30011, 9000, replace=False)).astype(np.uint64) + np.uint64(3)
    orr = np.sort(rng.choice(7001, 2000, replace=False)).astype(np.uint64) + np.uint64(9)
    sl, rsl = both(gdk, ref, ol, T.oid, sorted=True, key=True)
    sr, rsr = both(gdk, ref, orr, T.oid, sorted=True, key=True)
    for s, rs, t, rt in ((sl, rsl, sr, rsr), (gdk.BATdense(0, 1000, 5000), ref.dense(0, 1000, 5000), None, None),
                         (None, None, gdk.BATdense(0, 100, 800), ref.dense(0, 100, 800))):
        _same_pairs(*gdk.BATleftjoin(l, r, s, t), *ref.BATleftjoin(rl, rr, rs, rt))
        _same_pairs(*gdk.BATouterjoin(l, r, s, t), *ref.BATouterjoin(rl, rr, rs, rt))
        assert np.array_equal(gdk.BATintersect(l, r, s, t).numpy().astype(np.uint64),
                              ref.BATintersect(rl, rr, rs, rt).numpy().astype(np.uint64))
        assert np.array_equal(gdk.BATdiff(l, r, s, t).numpy().astype(np.uint64),
                              ref.BATdiff(rl, rr, rs, rt).numpy().astype(np.uint64))


def test_leftjoin_family_empty_sides(gdk, ref):
    """nomatch (gdk_join.c:302-360)"""
    a, b, l, rl, r, rr = _cols(gdk, ref, T.int, 5, nl=1000, nr=300)
    e, re_ = both(gdk, ref, np.zeros(0, dtype=np.int32), T.int)
    o1, o2 = gdk.BATouterjoin(l, e)
    q1, q2 = ref.BATouterjoin(rl, re_)
    assert np.array_equal(o1.numpy().astype(np.uint64), q1.numpy().astype(np.uint64))
    assert len(o2) == len(a) and np.all(o2.numpy().astype(np.uint64) == NIL)
    assert len(gdk.BATintersect(l, e)) == 0
    d = gdk.BATdiff(l, e)
    assert np.array_equal(d.numpy().astype(np.uint64), ref.BATdiff(rl, re_).numpy().astype(np.uint64))
    assert len(d) == len(a) and d.props()["type"] == T.void        # dense candidate list
    for fn in (gdk.BATleftjoin, gdk.BATouterjoin, gdk.BATsemijoin):
        x1, x2 = fn(e, r)
        assert len(x1) == 0 and len(x2) == 0
    assert len(gdk.BATdiff(e, r)) == 0


def test_foreign_key_shapes(gdk, ref):
    """unique build side: match_one outer join succeeds, semijoin of an FK keeps every row and comes back dense"""
    rng = np.random.default_rng(1)
    nr, nl = 50000, 200000
    okey = (rng.permutation(nr) + 1).astype(np.int32)
    lkey = rng.integers(1, nr + 1, nl).astype(np.int32)
    l, rl = both(gdk, ref, lkey, T.int)
    r, rr = both(gdk, ref, okey, T.int, key=True)
    _same_pairs(*gdk.BATouterjoin(l, r, None, None, False, True), *ref.BATouterjoin(rl, rr, None, None, False, True))
    i = gdk.BATintersect(l, r, None, None, False, True)
    assert len(i) == nl and i.props()["type"] == T.void
    assert len(gdk.BATdiff(l, r)) == 0
    # duplicates on the right: max_one / match_one must fail like the reference
    dup = np.concatenate([okey, okey[:10]])
    r2, rr2 = both(gdk, ref, dup, T.int)
    with pytest.raises(G.GDKError, match="more than one match"):
        gdk.BATouterjoin(l, r2, None, None, False, True)
    with pytest.raises(Exception, match="more than one match"):
        ref.BATouterjoin(rl, rr2, None, None, False, True)
    with pytest.raises(G.GDKError, match="more than one match"):
        gdk.BATsemijoin(l, r2, None, None, False, True)
    # a left value without partner: nil_on_miss wins over min_one (gdk_join.c:2853-2871), the row gets a nil
    lk2 = lkey.copy()
    lk2[77] = nr + 5
    l2, rl2 = both(gdk, ref, lk2, T.int)
    o1, o2 = gdk.BATouterjoin(l2, r, None, None, False, True)
    _same_pairs(o1, o2, *ref.BATouterjoin(rl2, rr, None, None, False, True))
    assert o2.numpy().astype(np.uint64)[77] == NIL


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
def test_unique(gdk, ref, tt):
    rng = np.random.default_rng(300 + tt)
    n = 40009
    a = random_col(rng, tt, n, nilfrac=0.02, lo=0, hi=100 if tt == T.bte else 3000)
    if tt in (T.flt, T.dbl):
        a = np.round(a).astype(a.dtype) + 1
    b, rb = both(gdk, ref, a, tt, hseq=7)
    u = gdk.BATunique(b)
    assert np.array_equal(u.numpy().astype(np.uint64), ref.BATunique(rb).numpy().astype(np.uint64))
    assert_props_true(u)
    ol = np.sort(rng.choice(n, 9000, replace=False)).astype(np.uint64) + np.uint64(7)
    s, rs = both(gdk, ref, ol, T.oid, sorted=True, key=True)
    assert np.array_equal(gdk.BATunique(b, s).numpy().astype(np.uint64), ref.BATunique(rb, rs).numpy().astype(np.uint64))
    # a key column is its own unique list
    k, rk = both(gdk, ref, np.arange(100, dtype=np.int32), T.int, key=True)
    uk = gdk.BATunique(k)
    assert len(uk) == 100 and np.array_equal(uk.numpy().astype(np.uint64), ref.BATunique(rk).numpy().astype(np.uint64))


# ------------------------------------------------------------------ dense right-hand side: mergejoin_void (gdk_join.c:563)
def _fk(gdk, ref, rng, nl, nr, rhseq, rtseq, lhseq, nilfrac, lkind):
    """l: an oid column of foreign keys into the dense r (some dangling, some nil), or a dense l"""
    if lkind == "dense":
        ltseq = max(0, rtseq - nl // 3)
        l, rl = gdk.BATdense(lhseq, ltseq, nl), ref.dense(lhseq, ltseq, nl)
        a = np.arange(ltseq, ltseq + nl, dtype=np.uint64)
    else:
        a = rng.integers(max(0, rtseq - nr // 4), rtseq + nr + nr // 4, nl).astype(np.uint64)
        if nilfrac:
            a[rng.random(nl) < nilfrac] = NIL
        if lkind == "sorted":
            a = np.sort(a)
        l, rl = both(gdk, ref, a, T.oid, hseq=lhseq, sorted=(lkind == "sorted" and not nilfrac))
    r, rr = gdk.BATdense(rhseq, rtseq, nr), ref.dense(rhseq, rtseq, nr)
    return a, l, rl, r, rr


def _cand_pairs(gdk, ref, rng, n, hseq):
    out = [(None, None)]
    lo, cnt = hseq + n // 5, n // 2
    out.append((gdk.BATdense(0, lo, cnt), ref.dense(0, lo, cnt)))
    c = np.sort(rng.choice(n, max(1, n // 3), replace=False)).astype(np.uint64) + hseq
    out.append((gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True), ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)))
    return out


@pytest.mark.parametrize("lkind", ["random", "sorted", "dense"])
@pytest.mark.parametrize("nilfrac", [0.0, 0.05])
def test_dense_right_fetch_joins(gdk, ref, lkind, nilfrac):
    if lkind == "dense" and nilfrac:
        pytest.skip("a dense column has no nils")
    rng = np.random.default_rng(len(lkind) * 7 + int(nilfrac * 100))
    nl, nr, rhseq, rtseq, lhseq = 60_011, 20_003, 9, 1000, 4
    a, l, rl, r, rr = _fk(gdk, ref, rng, nl, nr, rhseq, rtseq, lhseq, nilfrac, lkind)
    for sl, rsl in _cand_pairs(gdk, ref, rng, nl, lhseq):
        for sr, rsr in ((None, None), (gdk.BATdense(0, rhseq + 100, nr // 2), ref.dense(0, rhseq + 100, nr // 2))):
            _same_pairs(*gdk.BATjoin(l, r, sl, sr, False), *ref.BATjoin(rl, rr, rsl, rsr, False))
            o1, o2 = gdk.BATleftjoin(l, r, sl, sr, False)
            e1, e2 = ref.BATleftjoin(rl, rr, rsl, rsr, False)
            # one partner per row: the result is fully determined, compare it element for element
            assert np.array_equal(o1.numpy(), e1.numpy()) and np.array_equal(o2.numpy(), e2.numpy())
            assert_props_true(o1)
            assert_props_true(o2)
            o1, o2 = gdk.BATouterjoin(l, r, sl, sr, False)
            if lkind == "dense" and sl is not None and not sl.is_dense():
                # the reference crashes here (mergejoin_void's loop for a dense l with a non-dense candidate list,
                # gdk_join.c:795-860, runs past its buffers): check against the definition instead
                cl = sl.numpy().astype(np.int64)
                v = a[cl - lhseq].astype(np.int64)
                rlo = rtseq + (0 if sr is None else 100)
                rhi = rlo + (nr if sr is None else nr // 2)
                want = np.where((v >= rlo) & (v < rhi), (v - rtseq + rhseq).astype(np.uint64), NIL)
                assert np.array_equal(o1.numpy().astype(np.int64), cl) and np.array_equal(o2.numpy().astype(np.uint64), want)
            else:
                e1, e2 = ref.BATouterjoin(rl, rr, rsl, rsr, False)
                assert np.array_equal(o1.numpy(), e1.numpy()) and np.array_equal(o2.numpy().astype(np.uint64), e2.numpy().astype(np.uint64))
            miss = o2.numpy().astype(np.uint64) == NIL
            assert o2.props()["nil"] == bool(miss.any()) and o2.props()["nonil"] == (not miss.any())
            o1, o2 = gdk.BATsemijoin(l, r, sl, sr, False)
            e1, e2 = ref.BATsemijoin(rl, rr, rsl, rsr, False)
            assert np.array_equal(o1.numpy(), e1.numpy()) and np.array_equal(o2.numpy(), e2.numpy())
            assert np.array_equal(gdk.BATintersect(l, r, sl, sr, False).numpy(), ref.BATintersect(rl, rr, rsl, rsr, False).numpy())
            for not_in in (False, True):
                assert np.array_equal(gdk.BATdiff(l, r, sl, sr, False, not_in).numpy(), ref.BATdiff(rl, rr, rsl, rsr, False, not_in).numpy())
    # the partner really is the row whose (virtual) tail holds the key
    o1, o2 = gdk.BATleftjoin(l, r, None, None, False)
    x, y = o1.numpy().astype(np.int64) - lhseq, o2.numpy().astype(np.int64) - rhseq
    assert np.array_equal(a[x], (y + rtseq).astype(np.uint64))


def test_dense_sides_that_need_the_hash_join(gdk, ref):
    """dense r with NON-dense candidates, and a dense l against a real column: materialised and hash-joined"""
    rng = np.random.default_rng(77)
    nl, nr = 30_011, 9_001
    a, l, rl, r, rr = _fk(gdk, ref, rng, nl, nr, 3, 500, 2, 0.02, "random")
    c = np.sort(rng.choice(nr, nr // 2, replace=False)).astype(np.uint64) + 3
    sr, rsr = gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True), ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)
    _same_pairs(*gdk.BATjoin(l, r, None, sr, False), *ref.BATjoin(rl, rr, None, rsr, False))
    _same_pairs(*gdk.BATleftjoin(l, r, None, sr, False), *ref.BATleftjoin(rl, rr, None, rsr, False))
    assert np.array_equal(gdk.BATdiff(l, r, None, sr, False, False).numpy(), ref.BATdiff(rl, rr, None, rsr, False, False).numpy())
    # dense l, real r (values with duplicates)
    b = rng.integers(0, 12_000, nr).astype(np.uint64)
    r2, rr2 = both(gdk, ref, b, T.oid, hseq=8)
    dl, rdl = gdk.BATdense(1, 4000, 10_000), ref.dense(1, 4000, 10_000)
    _same_pairs(*gdk.BATjoin(dl, r2, None, None, False), *ref.BATjoin(rdl, rr2, None, None, False))
    _same_pairs(*gdk.BATleftjoin(dl, r2, None, None, False), *ref.BATleftjoin(rdl, rr2, None, None, False))
    assert np.array_equal(gdk.BATintersect(dl, r2, None, None, False).numpy(), ref.BATintersect(rdl, rr2, None, None, False).numpy())
