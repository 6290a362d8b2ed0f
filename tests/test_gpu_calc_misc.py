"""BATcalcmin/max(+_no_nil), BATcalcbetween (+cst forms) and BATcalcifthenelse (+cst forms) on the device
vs the real reference GDK (gdk_calc.c:976-2300, 3968-4290, 4376-4800): bit-exact tails and agreed
nil properties.  The constant forms are checked against the reference's BAT forms fed with constant
columns (the same BETWEEN / IFTHENELSELOOP macro runs underneath)."""
import itertools

import numpy as np
import pytest

from tests.util import ALL_TYPES, assert_same, both, random_col, same_bits
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE


def _col(rng, tt, n, nilfrac=0.05):
    a = random_col(rng, tt, n, nilfrac=nilfrac, lo=-50, hi=50)
    if tt in (T.flt, T.dbl):
        a = np.round(a).astype(a.dtype)
    return a


def _const(tt, v, n):
    if tt == T.hge:
        return G.hge_from_ints([v] * n)
    return np.full(n, v, dtype=G.NP_OF[tt])


@pytest.mark.parametrize("tt", ALL_TYPES)
@pytest.mark.parametrize("which", ["min", "min_no_nil", "max", "max_no_nil"])
def test_minmax(gdk, ref, tt, which):
    rng = np.random.default_rng(tt * 7 + len(which))
    n = 30011
    a, b = _col(rng, tt, n), _col(rng, tt, n)
    x, rx = both(gdk, ref, a, tt, hseq=5)
    y, ry = both(gdk, ref, b, tt, hseq=5)
    assert_same(gdk.BATcalcminmax(which, x, y), ref.BATcalcminmax(which, rx, ry))
    # through candidate lists on both sides
    o = np.sort(rng.choice(n, 7000, replace=False)).astype(np.uint64) + np.uint64(5)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    assert_same(gdk.BATcalcminmax(which, x, y, s, s), ref.BATcalcminmax(which, rx, ry, rs, rs))
    d, rd = gdk.BATdense(0, 105, 999), ref.dense(0, 105, 999)
    assert_same(gdk.BATcalcminmax(which, x, y, d, d), ref.BATcalcminmax(which, rx, ry, rd, rd))


def test_minmax_errors(gdk, ref):
    x, _ = both(gdk, ref, np.arange(5, dtype=np.int32), T.int)
    y, _ = both(gdk, ref, np.arange(5, dtype=np.int64), T.lng)
    z, _ = both(gdk, ref, np.arange(6, dtype=np.int32), T.int)
    with pytest.raises(G.GDKError, match="incompatible types"):
        gdk.BATcalcminmax("min", x, y)
    with pytest.raises(G.GDKError, match="not the same size"):
        gdk.BATcalcminmax("max", x, z)
    e, re_ = both(gdk, ref, np.zeros(0, dtype=np.int32), T.int)
    assert_same(gdk.BATcalcminmax("min", e, e), ref.BATcalcminmax("min", re_, re_))


FLAGS = list(itertools.product([False, True], repeat=5))


@pytest.mark.parametrize("tt", ALL_TYPES)
def test_between_all_flags(gdk, ref, tt):
    rng = np.random.default_rng(50 + tt)
    n = 20011
    v, lo, hi = _col(rng, tt, n), _col(rng, tt, n), _col(rng, tt, n)
    b, rb = both(gdk, ref, v, tt, hseq=3)
    l, rl = both(gdk, ref, lo, tt, hseq=3)
    h, rh = both(gdk, ref, hi, tt, hseq=3)
    for sym, linc, hinc, nf, anti in FLAGS:
        got = gdk.BATcalcbetween(b, l, h, None, None, None, sym, linc, hinc, nf, anti)
        want = ref.BATcalcbetween(rb, rl, rh, None, None, None, sym, linc, hinc, nf, anti)
        assert_same(got, want)


@pytest.mark.parametrize("tt", [T.int, T.lng, T.hge, T.dbl])
def test_between_constants_and_candidates(gdk, ref, tt):
    rng = np.random.default_rng(60 + tt)
    n = 20011
    v, lo, hi = _col(rng, tt, n), _col(rng, tt, n), _col(rng, tt, n)
    b, rb = both(gdk, ref, v, tt, hseq=3)
    l, rl = both(gdk, ref, lo, tt, hseq=3)
    h, rh = both(gdk, ref, hi, tt, hseq=3)
    _, rcl = both(gdk, ref, _const(tt, -10, n), tt, hseq=3)
    _, rch = both(gdk, ref, _const(tt, 20, n), tt, hseq=3)
    o = np.sort(rng.choice(n, 5000, replace=False)).astype(np.uint64) + np.uint64(3)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    for sym, linc, hinc, nf, anti in FLAGS[::3]:
        fl = (sym, linc, hinc, nf, anti)
        assert_same(gdk.BATcalcbetween(b, -10, 20, None, None, None, *fl), ref.BATcalcbetween(rb, rcl, rch, None, None, None, *fl))
        assert_same(gdk.BATcalcbetween(b, l, 20, None, None, None, *fl), ref.BATcalcbetween(rb, rl, rch, None, None, None, *fl))
        assert_same(gdk.BATcalcbetween(b, -10, h, None, None, None, *fl), ref.BATcalcbetween(rb, rcl, rh, None, None, None, *fl))
        assert_same(gdk.BATcalcbetween(b, l, h, s, s, s, *fl), ref.BATcalcbetween(rb, rl, rh, rs, rs, rs, *fl))
        assert_same(gdk.BATcalcbetween(b, -10, 20, s, None, None, *fl), ref.BATcalcbetween(rb, rcl, rch, rs, rs, rs, *fl))


def test_between_q6_shape(gdk, ref):
    """l_discount BETWEEN 5 AND 7 as batcalc.between emits it; agrees with the select"""
    rng = np.random.default_rng(1)
    d = rng.integers(0, 11, 100003).astype(np.int64)
    b, rb = both(gdk, ref, d, T.lng)
    m = gdk.BATcalcbetween(b, 5, 7).numpy()
    assert np.array_equal(m.astype(bool), (d >= 5) & (d <= 7))
    sel = gdk.BATselect(b, None, 5, 7, True, True, False, False).numpy()
    assert np.array_equal(np.nonzero(m)[0].astype(np.uint64), sel.astype(np.uint64))


@pytest.mark.parametrize("tt", ALL_TYPES)
def test_ifthenelse(gdk, ref, tt):
    rng = np.random.default_rng(70 + tt)
    n = 30011
    cond = rng.integers(0, 2, n).astype(np.int8)
    cond[rng.random(n) < 0.1] = -128                          # bit_nil selects the "else" side
    a, b = _col(rng, tt, n), _col(rng, tt, n)
    c, rc = both(gdk, ref, cond, T.bit, hseq=9)
    x, rx = both(gdk, ref, a, tt, hseq=9)
    y, ry = both(gdk, ref, b, tt, hseq=9)
    # the reference claims tnil = false even when nils were copied; compare bits and the props it derives
    got, want = gdk.BATcalcifthenelse(c, x, y), ref.BATcalcifthenelse(rc, rx, ry)
    assert same_bits(got.numpy(), want.numpy())
    pg, pw = got.props(), want.props()
    for k in ("count", "hseq", "nonil", "nil", "type"):
        assert pg[k] == pw[k], (k, pg, pw)
    # constant forms against constant columns
    _, rk1 = both(gdk, ref, _const(tt, 7, n), tt, hseq=9)
    _, rk2 = both(gdk, ref, _const(tt, -3, n), tt, hseq=9)
    assert same_bits(gdk.BATcalcifthenelse(c, x, -3).numpy(), ref.BATcalcifthenelse(rc, rx, rk2).numpy())
    assert same_bits(gdk.BATcalcifthenelse(c, 7, y).numpy(), ref.BATcalcifthenelse(rc, rk1, ry).numpy())
    k = gdk.BATcalcifthenelse(c, 7, -3, tt)
    assert same_bits(k.numpy(), ref.BATcalcifthenelse(rc, rk1, rk2).numpy())
    assert k.props()["nonil"]


def test_ifthenelse_errors(gdk, ref):
    c, _ = both(gdk, ref, np.ones(5, dtype=np.int8), T.bit)
    x, _ = both(gdk, ref, np.arange(5, dtype=np.int32), T.int)
    y, _ = both(gdk, ref, np.arange(5, dtype=np.int64), T.lng)
    z, _ = both(gdk, ref, np.arange(6, dtype=np.int32), T.int)
    with pytest.raises(G.GDKError, match="different types"):
        gdk.BATcalcifthenelse(c, x, y)
    with pytest.raises(G.GDKError, match="not the same size"):
        gdk.BATcalcifthenelse(c, x, z)
    with pytest.raises(G.GDKError, match="different types"):
        gdk.BATcalcifthenelse(x, x, x)                        # condition must be a bit column


@pytest.mark.parametrize("tt", ALL_TYPES)
@pytest.mark.parametrize("which", ["negate", "absolute", "iszero", "sign", "not"])
def test_unary(gdk, ref, tt, which):
    if which == "not" and tt in (T.flt, T.dbl):
        b, rb = both(gdk, ref, np.ones(3, dtype=G.NP_OF[tt]), tt)
        with pytest.raises(G.GDKError, match="not supported"):
            gdk.BATcalcunary("not", b)
        return
    rng = np.random.default_rng(80 + tt)
    n = 30011
    a = _col(rng, tt, n)
    b, rb = both(gdk, ref, a, tt, hseq=4)
    assert_same(gdk.BATcalcunary(which, b), ref.BATcalcunary(which, rb))
    o = np.sort(rng.choice(n, 6000, replace=False)).astype(np.uint64) + np.uint64(4)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    assert_same(gdk.BATcalcunary(which, b, s), ref.BATcalcunary(which, rb, rs))
    # sorted input: the order properties the reference derives
    srt = np.sort(a[:1000]) if tt != T.hge else a[:1]
    if tt != T.hge:
        sb, rsb = both(gdk, ref, srt, tt, sorted=True)
        got, want = gdk.BATcalcunary(which, sb), ref.BATcalcunary(which, rsb)
        assert_same(got, want)
        assert got.props()["sorted"] == want.props()["sorted"] and got.props()["revsorted"] == want.props()["revsorted"]


def test_not_bit_and_overflow(gdk, ref):
    c = np.array([0, 1, -128, 1, 0], dtype=np.int8)
    b, rb = both(gdk, ref, c, T.bit)
    assert_same(gdk.BATcalcunary("not", b), ref.BATcalcunary("not", rb))
    x, rx = both(gdk, ref, np.array([1, 2, 2147483647], dtype=np.int32), T.int)
    with pytest.raises(G.GDKError, match=r"overflow in calculation NOT\(2147483647\)"):
        gdk.BATcalcunary("not", x)
    with pytest.raises(Exception, match=r"overflow in calculation NOT\(2147483647\)"):
        ref.BATcalcunary("not", rx)
