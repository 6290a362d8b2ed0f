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


# ------------------------------------------------------------------ BATcalcmod (gdk_calc_mod.c)
_RANK = {T.bte: 0, T.sht: 1, T.int: 2, T.lng: 3, T.hge: 4, T.flt: 5, T.dbl: 6}


def _mod_cmp(gdk, ref, fo, fr):
    try:
        r = fr()
    except ref.RefError as e:
        with pytest.raises(G.GDKError) as ei:
            fo()
        assert str(ei.value).strip() == str(e).strip(), (str(ei.value), str(e))
        return
    assert_same(fo(), r)


def test_mod_int_matrix(gdk, ref):
    """every (t1, t2) -> t3 of MOD_3TYPE, including result types narrower than the left operand"""
    rng = np.random.default_rng(77)
    n = 4099
    ints = [T.bte, T.sht, T.int, T.lng, T.hge]
    for t1 in ints:
        for t2 in ints:
            a1 = random_col(rng, t1, n, nilfrac=0.03, lo=-120 if t1 == T.bte else -30000, hi=120 if t1 == T.bte else 30000)
            a2 = random_col(rng, t2, n, nilfrac=0.03, lo=1, hi=97)
            if t2 != T.hge:
                a2[::3] = -a2[::3]                      # negative divisors: C's sign-of-dividend rule
            b1, r1 = both(gdk, ref, a1, t1)
            b2, r2 = both(gdk, ref, a2, t2)
            for t3 in ints:
                _mod_cmp(gdk, ref, lambda: gdk.BATcalc("mod", b1, b2, None, None, t3),
                         lambda: ref.BATcalc("mod", r1, r2, None, None, t3))


def test_mod_float_matrix(gdk, ref):
    """FMOD_3TYPE: fmodf / fmod are exact, so the bits match the host libm's"""
    rng = np.random.default_rng(78)
    n = 4099
    for t1 in ALL_TYPES:
        for t2 in ALL_TYPES:
            if t1 not in (T.flt, T.dbl) and t2 not in (T.flt, T.dbl):
                continue
            a1 = random_col(rng, t1, n, nilfrac=0.03, lo=-120, hi=120)
            a2 = random_col(rng, t2, n, nilfrac=0.03, lo=1, hi=97)
            b1, r1 = both(gdk, ref, a1, t1)
            b2, r2 = both(gdk, ref, a2, t2)
            for t3 in (T.lng, T.flt, T.dbl):
                _mod_cmp(gdk, ref, lambda: gdk.BATcalc("mod", b1, b2, None, None, t3),
                         lambda: ref.BATcalc("mod", r1, r2, None, None, t3))


def test_mod_cst_candidates_and_zero(gdk, ref):
    rng = np.random.default_rng(79)
    n = 30011
    a = random_col(rng, T.lng, n, nilfrac=0.02, lo=-100000, hi=100000)
    b, rb = both(gdk, ref, a, T.lng, hseq=4)
    o = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.uint64) + np.uint64(4)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    for cand, rcand in ((None, None), (s, rs), (gdk.BATdense(0, 104, 999), ref.dense(0, 104, 999))):
        for first in (False, True):
            for cst, ct in ((100, T.lng), (-7, T.int), (0, T.int)):
                if first and cst == 0:
                    continue
                for tp in (T.int, T.lng, T.hge):
                    if first:
                        bb, rbb = both(gdk, ref, np.where(a == 0, 1, a), T.lng, hseq=4)
                    else:
                        bb, rbb = b, rb
                    _mod_cmp(gdk, ref, lambda: gdk.BATcalccst("mod", bb, cst, ct, cand, tp, cst_first=first),
                             lambda: ref.BATcalccst("mod", rbb, cst, ct, rcand, tp, cst_first=first))
    # a zero divisor anywhere aborts with the reference's text
    z = random_col(rng, T.int, n, lo=1, hi=50)
    z[20000] = 0
    b2, rb2 = both(gdk, ref, z, T.int, hseq=4)
    with pytest.raises(G.GDKError, match="22012!division by zero"):
        gdk.BATcalc("mod", b, b2, None, None, T.lng)
    with pytest.raises(Exception, match="22012!division by zero"):
        ref.BATcalc("mod", rb, rb2, None, None, T.lng)
    with pytest.raises(G.GDKError, match="inputs not the same size"):
        gdk.BATcalc("mod", b, b2, s, None, T.lng)


# ------------------------------------------------------------------ BATcalcand / or / xor (gdk_calc.c:2376-3032)
@pytest.mark.parametrize("op", ["and", "or", "xor"])
def test_logic_bit_three_valued(gdk, ref, op):
    vals = np.array([0, 1, -128], dtype=np.int8)
    a = np.repeat(vals, 3)
    b = np.tile(vals, 3)
    ba, ra = both(gdk, ref, a, T.bit)
    bb, rb = both(gdk, ref, b, T.bit)
    assert_same(gdk.BATcalclogic(op, ba, bb), ref.BATcalclogic(op, ra, rb))
    rng = np.random.default_rng(5)
    n = 100003
    a = rng.choice(vals, n)
    b = rng.choice(vals, n)
    ba, ra = both(gdk, ref, a, T.bit, hseq=7)
    bb, rb = both(gdk, ref, b, T.bit, hseq=7)
    assert_same(gdk.BATcalclogic(op, ba, bb), ref.BATcalclogic(op, ra, rb))
    o = np.sort(rng.choice(n, 20000, replace=False)).astype(np.uint64) + np.uint64(7)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    assert_same(gdk.BATcalclogic(op, ba, bb, s, s), ref.BATcalclogic(op, ra, rb, rs, rs))
    for cst in (0, 1, -128):
        for cand, rcand in ((None, None), (s, rs)):
            got, want = gdk.BATcalclogiccst(op, ba, cst, T.bit, cand), ref.BATcalclogiccst(op, ra, cst, T.bit, rcand)
            assert_same(got, want)
            assert got.props()["sorted"] == want.props()["sorted"] and got.props()["revsorted"] == want.props()["revsorted"]
            assert_same(gdk.BATcalclogiccst(op, ba, cst, T.bit, cand, cst_first=True), want)


@pytest.mark.parametrize("op", ["and", "or", "xor"])
@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge])
def test_logic_bitwise(gdk, ref, op, tt):
    rng = np.random.default_rng(90 + tt)
    n = 30011
    # non-negative operands cannot combine to the nil pattern
    a = random_col(rng, tt, n, nilfrac=0.04, lo=0, hi=127)
    b = random_col(rng, tt, n, nilfrac=0.04, lo=0, hi=127)
    ba, ra = both(gdk, ref, a, tt, hseq=2)
    bb, rb = both(gdk, ref, b, tt, hseq=2)
    assert_same(gdk.BATcalclogic(op, ba, bb), ref.BATcalclogic(op, ra, rb))
    o = np.sort(rng.choice(n, 7000, replace=False)).astype(np.uint64) + np.uint64(2)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    assert_same(gdk.BATcalclogic(op, ba, bb, s, s), ref.BATcalclogic(op, ra, rb, rs, rs))
    assert_same(gdk.BATcalclogiccst(op, ba, 85, tt, s), ref.BATcalclogiccst(op, ra, 85, tt, rs))
    with pytest.raises(G.GDKError, match="inputs not the same size"):
        gdk.BATcalclogic(op, ba, bb, s, None)
    other = T.int if tt != T.int else T.lng
    c, _ = both(gdk, ref, np.ones(n, dtype=G.NP_OF[other]), other, hseq=2)
    with pytest.raises(G.GDKError, match="incompatible input types"):
        gdk.BATcalclogic(op, ba, c)


@pytest.mark.parametrize("op", ["and", "xor"])
def test_logic_overflow_text(gdk, ref, op):
    """two non-nil operands that combine to the nil pattern abort with the reference's text"""
    for tt, bits in ((T.bte, 8), (T.int, 32), (T.lng, 64)):
        dt = G.NP_OF[tt]
        top = -(1 << (bits - 1))
        x = np.ones(5000, dtype=dt)
        y = np.ones(5000, dtype=dt)
        if op == "and":
            x[3333], y[3333] = top + 1, top + (1 << (bits - 2))       # 10..01 & 110..0 = 10..0
        else:
            x[3333], y[3333] = top + 1, 1                             # 10..01 ^ 0..01 = 10..0
        bx, rx = both(gdk, ref, x, tt)
        by, ry = both(gdk, ref, y, tt)
        with pytest.raises(ref.RefError) as want:
            ref.BATcalclogic(op, rx, ry)
        with pytest.raises(G.GDKError) as got:
            gdk.BATcalclogic(op, bx, by)
        assert str(got.value).strip() == str(want.value).strip()


# ------------------------------------------------------------------ BATcalc{min,max}cst[_no_nil] (gdk_calc.c:1403-2330)
@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.flt, T.dbl])
@pytest.mark.parametrize("which", ["min", "min_no_nil", "max", "max_no_nil"])
def test_minmax_cst(gdk, ref, tt, which):
    rng = np.random.default_rng(120 + tt)
    n = 30011
    a = _col(rng, tt, n)
    b, rb = both(gdk, ref, a, tt, hseq=3)
    o = np.sort(rng.choice(n, 6000, replace=False)).astype(np.uint64) + np.uint64(3)
    s, rs = both(gdk, ref, o, T.oid, sorted=True, key=True)
    nilv = float("nan") if tt in (T.flt, T.dbl) else int(np.iinfo(G.NP_OF[tt]).min)
    for cst in (7, -3, nilv):
        for cand, rcand in ((None, None), (s, rs)):
            assert_same(gdk.BATcalcminmaxcst(which, b, cst, tt, cand), ref.BATcalcminmaxcst(which, rb, cst, tt, rcand))
    other = T.int if tt != T.int else T.lng
    with pytest.raises(G.GDKError, match="inputs have incompatible types"):
        gdk.BATcalcminmaxcst(which, b, 1, other)


# ---------------------------------------------------------------------------- shifts (gdk_calc.c:3246-3770)
@pytest.mark.parametrize("op", ["lsh", "rsh"])
@pytest.mark.parametrize("t1", [T.bte, T.sht, T.int, T.lng, T.hge])
@pytest.mark.parametrize("t2", [T.bte, T.int, T.lng, T.hge])
def test_shifts(gdk, ref, op, t1, t2):
    """BATcalclsh / BATcalcrsh and the cst forms: result type = left operand, nil in nil out, the reference's
    three check variants (SHIFT_CHECK, LSH_CHECK, none for hge << bte) and its error text"""
    rng = np.random.default_rng(t1 * 11 + t2 + (op == "rsh"))
    bits = {T.bte: 8, T.sht: 16, T.int: 32, T.lng: 64, T.hge: 128}[t1]
    n = 20_003
    # small non-negative left values and small counts: everything passes for lsh as well
    a = random_col(rng, t1, n, nilfrac=0.05, lo=0, hi=3)
    cnt = random_col(rng, t2, n, nilfrac=0.05, lo=0, hi=bits - 4)
    b1, r1 = both(gdk, ref, a, t1)
    b2, r2 = both(gdk, ref, cnt, t2)
    assert_same(gdk.BATcalcshift(op, b1, b2), ref.BATcalcshift(op, r1, r2))
    assert_same(gdk.BATcalcshiftcst(op, b1, 2, t2), ref.BATcalcshiftcst(op, r1, 2, t2))
    assert_same(gdk.BATcalcshiftcst(op, b2, 1, t1, cst_first=True), ref.BATcalcshiftcst(op, r2, 1, t1, cst_first=True))
    # right shifts of negative values are fine; left shifts of negative / too large values are refused (except
    # where the reference checks only the count)
    neg = random_col(rng, t1, n, nilfrac=0.05, lo=-100, hi=100)
    bn, rn = both(gdk, ref, neg, t1)
    small = random_col(rng, t2, n, nilfrac=0.0, lo=0, hi=2)
    bs, rs = both(gdk, ref, small, t2)
    for x, rx, y, ry in ((bn, rn, bs, rs), (b1, r1, b2, r2)):
        try:
            want, err = ref.BATcalcshift(op, rx, ry), None
        except Exception as ex:     # noqa: BLE001
            want, err = None, str(ex).strip()
        if err is None:
            assert_same(gdk.BATcalcshift(op, x, y), want)
        else:
            with pytest.raises(G.GDKError) as e1:
                gdk.BATcalcshift(op, x, y)
            assert str(e1.value).strip() == err
    # a count beyond the width is refused with the reference's text (first failing row)
    if not (op == "lsh" and t1 == T.hge and t2 == T.bte):
        big = cnt.copy()
        if t2 == T.hge:
            big["lo"][777] = bits
            big["hi"][777] = 0
        else:
            big[777] = min(bits, 127) if t2 == T.bte else bits
        if not (t2 == T.bte and bits > 127):
            bb, rb = both(gdk, ref, big, t2)
            a2 = a.copy()
            if t1 == T.hge:
                a2["lo"][777], a2["hi"][777] = 1, 0
            else:
                a2[777] = 1
            ba, ra = both(gdk, ref, a2, t1)
            with pytest.raises(G.GDKError) as e1:
                gdk.BATcalcshift(op, ba, bb)
            with pytest.raises(Exception) as e2:
                ref.BATcalcshift(op, ra, rb)
            assert str(e1.value).strip() == str(e2.value).strip()
            assert "shift operand too large" in str(e1.value)
