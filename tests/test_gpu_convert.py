"""BATconvert with decimal scale / precision (gdk_calc_convert.c:262-424): integer rescaling
with round-half-away division, integer -> floating point through 10^scale1, floating point -> integer through
`(long double) v * 10^scale2` + roundl — all bit-exact against the reference, overflow texts included."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import assert_same, both, random_col

pytestmark = pytest.mark.gpu
T = G.TYPE
INTS = [T.bte, T.sht, T.int, T.lng, T.hge]


def _try(fn):
    try:
        return fn(), None
    except Exception as ex:      # noqa: BLE001
        return None, str(ex).strip()


DIGITS = {T.bte: 2, T.sht: 4, T.int: 9, T.lng: 18, T.hge: 38}


def sane(t1, t2, s1, s2, prec):
    """the reference computes (TYPE) 10^k without checking that it fits (a wrapped 0 divides by zero): only
    scale / precision combinations whose powers of ten fit the types involved are meaningful"""
    if t2 in DIGITS and (max(0, s2 - s1) > DIGITS[t2] or prec > DIGITS[t2] or (t1 not in DIGITS and s2 > DIGITS[t2])):
        return False
    if t1 in DIGITS and (max(0, s1 - s2) > DIGITS[t1] or (t2 not in DIGITS and s1 > DIGITS[t1])):
        return False
    return True


def check(gdk, ref, b, rb, tp, s1, s2, prec, s=None, rs=None):
    if not sane(b.ttype, tp, s1, s2, prec):
        return
    got, e1 = _try(lambda: gdk.BATconvert(b, s, tp, s1, s2, prec))
    want, e2 = _try(lambda: ref.BATconvert(rb, rs, tp, s1, s2, prec))
    assert e1 == e2, (tp, s1, s2, prec, e1, e2)
    if want is not None:
        assert_same(got, want)


@pytest.mark.parametrize("t1", INTS)
@pytest.mark.parametrize("t2", INTS)
def test_int_rescale(gdk, ref, t1, t2):
    rng = np.random.default_rng(t1 * 11 + t2)
    hi = {T.bte: 120, T.sht: 30_000, T.int: 2_000_000_000, T.lng: 9 * 10**18, T.hge: 10**30}[t1]
    for n in (1, 257, 20_003):
        a = random_col(rng, t1, n, nilfrac=0.05, lo=-hi, hi=hi)
        b, rb = both(gdk, ref, a, t1)
        # small magnitudes too: rescaling up must not overflow for them
        sm = random_col(rng, t1, n, nilfrac=0.05, lo=-99, hi=99)
        c, rc = both(gdk, ref, sm, t1)
        for s1, s2, prec in ((2, 0, 0), (0, 2, 0), (3, 1, 0), (1, 4, 0), (2, 2, 0), (0, 0, 3), (2, 0, 2), (0, 1, 4), (5, 0, 9), (0, 9, 18)):
            check(gdk, ref, b, rb, t2, s1, s2, prec)
            check(gdk, ref, c, rc, t2, s1, s2, prec)
    # halves round away from zero
    a = np.array([5, 15, 25, -5, -15, -25, 149, 150, 151, -149, -150, -151], dtype=G.NP_OF[T.sht])
    if t1 == T.sht:
        b, rb = both(gdk, ref, a, t1)
        for s1 in (1, 2):
            check(gdk, ref, b, rb, t2, s1, 0, 0)
    # candidate list
    a = random_col(rng, t1, 5000, nilfrac=0.1, lo=-99, hi=99)
    b, rb = both(gdk, ref, a, t1, hseq=3)
    cand = np.sort(rng.choice(5000, 1200, replace=False)).astype(np.uint64) + 3
    s, rs = gdk.BATfromhost(cand, T.oid).setprops(sorted=True, key=True, nonil=True), ref.bat(cand, T.oid, sorted=True, key=True, nonil=True)
    check(gdk, ref, b, rb, t2, 1, 0, 0, s, rs)


@pytest.mark.parametrize("t1", INTS)
@pytest.mark.parametrize("t2", [T.flt, T.dbl])
def test_int_to_float_scale(gdk, ref, t1, t2):
    rng = np.random.default_rng(t1 * 3 + t2)
    hi = {T.bte: 120, T.sht: 30_000, T.int: 2_000_000_000, T.lng: 9 * 10**18, T.hge: 10**30}[t1]
    a = random_col(rng, t1, 10_007, nilfrac=0.05, lo=-hi, hi=hi)
    b, rb = both(gdk, ref, a, t1)
    for s1 in (0, 1, 2, 5) + ((9, 18) if t1 in (T.lng, T.hge) else ()):
        check(gdk, ref, b, rb, t2, s1, 0, 0)


@pytest.mark.parametrize("t1", [T.flt, T.dbl])
@pytest.mark.parametrize("t2", INTS)
def test_float_to_decimal(gdk, ref, t1, t2):
    rng = np.random.default_rng(t1 * 5 + t2)
    hi = {T.bte: 1.2, T.sht: 300.0, T.int: 2.0e7, T.lng: 9.0e15, T.hge: 1.0e15}[t2]   # x 10^18 stays inside the type
    n = 20_011
    a = rng.uniform(-hi, hi, n).astype(G.NP_OF[t1])
    # exact decimal halves and near-halves: the 64-bit-mantissa product decides them
    k = rng.integers(-10**6, 10**6, n // 2)
    a[: n // 2] = (k.astype(np.float64) / 200.0 + rng.choice([0.0, 0.005, -0.005, 0.0049999999, 0.0050000001], n // 2)).astype(G.NP_OF[t1])
    a[rng.random(n) < 0.05] = np.nan
    a = np.clip(a, -hi, hi)
    b, rb = both(gdk, ref, a, t1)
    for s2, prec in ((0, 0), (1, 0), (2, 0), (2, 9), (4, 0), (0, 2)) + (((9, 0), (18, 0), (18, 38)) if t2 in (T.lng, T.hge) else ()):
        check(gdk, ref, b, rb, t2, 0, s2, prec)


def test_float_float_scale(gdk, ref):
    rng = np.random.default_rng(1)
    a = rng.uniform(-1e6, 1e6, 5000)
    b, rb = both(gdk, ref, a, T.dbl)
    for s1, s2 in ((0, 0), (2, 0), (0, 2)):
        check(gdk, ref, b, rb, T.flt, s1, s2, 0)
        check(gdk, ref, b, rb, T.dbl, s1, s2, 0)
    f, rf = both(gdk, ref, a.astype(np.float32), T.flt)
    for s1 in (0, 3):
        check(gdk, ref, f, rf, T.dbl, s1, 0, 0)
        check(gdk, ref, f, rf, T.flt, s1, 0, 0)
