"""GDKanalyticaldiff / windowbounds / sum / avg / count / min / max: CUDA vs the real reference GDK
(gdk_analytic_bounds.c, gdk_analytic_func.c, gdk_analytic_statistics.c).  Integer results, frame
bounds and framed floating-point sums (same fan-out-16 tree order) are compared bit for bit;
running floating-point sums/averages (frame types 3-5, sequential in the reference) within
1e-12 (dbl) / 1e-5 (flt) relative."""
import numpy as np
import pytest

from tests.util import both, has_nil, random_col, same_bits
from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE


def window(rng, n, nparts, tt=T.lng, nilfrac=0.0, step=3):
    """sorted partitions with an ascending order column (duplicates) -> pkey, ord"""
    cuts = np.sort(rng.choice(np.arange(1, n), size=min(nparts - 1, n - 1), replace=False)) if n > 1 and nparts > 1 else []
    pkey = np.zeros(n, dtype=np.int32)
    pkey[cuts] = 1
    pkey = np.cumsum(pkey).astype(np.int32)
    inc = rng.integers(0, step, n)
    o = np.cumsum(inc)
    start_of = np.zeros(n, dtype=np.int64)
    if n:
        # restart the order value in every partition
        start_of = np.maximum.accumulate(np.where(np.r_[True, pkey[1:] != pkey[:-1]], np.arange(n), 0))
        o = o - o[start_of]
    if tt in (T.flt, T.dbl):
        a = (o * 0.5).astype(G.NP_OF[tt])
        if nilfrac:
            # nils sort first within a partition: put them at the partition starts
            m = (rng.random(n) < nilfrac) & (np.arange(n) == start_of)
            a[m] = np.nan
        return pkey, a
    a = o.astype(G.NP_OF[tt]) if tt != T.hge else G.hge_from_i64(o)
    if nilfrac and tt != T.hge:
        m = (rng.random(n) < nilfrac * 20) & (np.arange(n) == start_of)
        a[m] = np.iinfo(G.NP_OF[tt]).min
    return pkey, a


def flags(gdk, ref, col, tt, prior=None, rprior=None):
    """partition / peer flags through GDKanalyticaldiff on both sides"""
    n = len(col)
    b, rb = both(gdk, ref, col, tt)
    r = gdk.GDKanalyticaldiff(gdk.colnew(T.bit, n), b, prior, None, tt)
    rr = ref.GDKanalyticaldiff(ref.colnew(T.bit, n), rb, rprior, None, tt)
    assert same_bits(r.numpy(), rr.numpy())
    return r, rr


def check_oid(o, r):
    a, b = o.numpy(), r.numpy()
    assert len(a) == len(b)
    assert same_bits(a, b), (a[:20], b[:20], np.flatnonzero(a != b)[:10])
    po, pr = o.props(), r.props()
    assert (po["nonil"], po["nil"]) == (pr["nonil"], pr["nil"])


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
@pytest.mark.parametrize("mode", ["none", "prior", "npbit"])
def test_diff(gdk, ref, tt, mode):
    rng = np.random.default_rng(tt * 7 + len(mode))
    for n in (0, 1, 2, 33, 4097, 50021):
        a = random_col(rng, tt, n, nilfrac=0.2, lo=0, hi=3)
        b, rb = both(gdk, ref, a, tt)
        prior = rprior = None
        npbit = None
        if mode == "prior":
            pf = (rng.random(n) < 0.1).astype(np.int8)
            prior, rprior = both(gdk, ref, pf, T.bit)
        elif mode == "npbit":
            npbit = 1
        r = gdk.GDKanalyticaldiff(gdk.colnew(T.bit, n), b, prior, npbit, tt)
        rr = ref.GDKanalyticaldiff(ref.colnew(T.bit, n), rb, rprior, npbit, tt)
        assert same_bits(r.numpy(), rr.numpy()), (tt, mode, n)
        assert r.props()["nonil"] and not r.props()["nil"]


@pytest.mark.parametrize("n,nparts", [(0, 1), (1, 1), (2, 2), (31, 3), (1025, 1), (4099, 40), (70001, 9), (70001, 30000)])
@pytest.mark.parametrize("preceding", [True, False])
def test_rows_bounds(gdk, ref, n, nparts, preceding):
    rng = np.random.default_rng(n + nparts)
    pkey, o = window(rng, n, nparts)
    p, rp = flags(gdk, ref, pkey, T.int) if n else both(gdk, ref, np.zeros(0, np.int8), T.bit)
    b, rb = both(gdk, ref, o, T.lng)
    for lim, tp2, half in ((0, T.lng, 1), (0, T.lng, 0), (3, T.int, 0), (100, T.lng, 1), (5, T.bte, 1), ((1 << 63) - 1, T.lng, 0),
                           (7, T.hge, 0)):
        r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, T.lng, tp2, 0, preceding, half)
        rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, T.lng, tp2, 0, preceding, half)
        check_oid(r, rr)
    # no partition column at all
    r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, None, None, 2, T.lng, T.lng, 0, preceding, 0)
    rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, None, None, 2, T.lng, T.lng, 0, preceding, 0)
    check_oid(r, rr)
    # dynamic bounds
    if n:
        for dt in (T.bte, T.int, T.lng, T.hge):
            lcol = random_col(rng, dt, n, lo=0, hi=9)
            l, rl = both(gdk, ref, lcol, dt)
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, l, None, T.lng, dt, 0, preceding, 1)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, rl, None, T.lng, dt, 0, preceding, 1)
            check_oid(r, rr)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
@pytest.mark.parametrize("preceding", [True, False])
def test_range_bounds(gdk, ref, tt, preceding):
    rng = np.random.default_rng(tt)
    for n, nparts, nilfrac in ((1, 1, 0), (40, 3, 0), (5003, 17, 0), (5003, 17, 0.5), (30011, 300, 0.5)):
        step = 2 if tt == T.bte else 3
        if tt == T.bte:
            n, nparts = min(n, 90), min(nparts, 3)
        pkey, o = window(rng, n, nparts, tt, nilfrac, step)
        p, rp = flags(gdk, ref, pkey, T.int)
        b, rb = both(gdk, ref, o, tt)
        if tt in (T.flt, T.dbl):
            cases = [(1.5, tt), (0.0, tt), (float(np.finfo(G.NP_OF[tt]).max), tt), (0.25, tt)]
        elif tt == T.hge:
            cases = [(5, T.hge), (0, T.hge), ((1 << 127) - 1, T.hge)]
        else:
            cases = [(5, T.lng), (0, T.lng), (1, T.int), (2, T.bte), (3, T.hge), ((1 << 63) - 1, T.lng), (127, T.bte)]
        for lim, tp2 in cases:
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, tt, tp2, 1, preceding, 0)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, tt, tp2, 1, preceding, 0)
            check_oid(r, rr)
        if tt not in (T.flt, T.dbl, T.hge):
            lcol = random_col(rng, T.int, n, lo=0, hi=6)
            l, rl = both(gdk, ref, lcol, T.int)
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, l, None, tt, T.int, 1, preceding, 0)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, rl, None, tt, T.int, 1, preceding, 0)
            check_oid(r, rr)


@pytest.mark.parametrize("preceding", [True, False])
def test_groups_bounds(gdk, ref, preceding):
    rng = np.random.default_rng(5)
    for n, nparts in ((1, 1), (50, 4), (6007, 23), (40009, 2000)):
        pkey, o = window(rng, n, nparts)
        p, rp = flags(gdk, ref, pkey, T.int)
        op, rop = flags(gdk, ref, o, T.lng, p, rp)       # peer starts (includes partition starts)
        for lim, tp2 in ((0, T.lng), (1, T.lng), (4, T.int), (1000, T.lng), ((1 << 63) - 1, T.lng)):
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), op, p, None, lim, T.bit, tp2, 2, preceding, 0)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rop, rp, None, lim, T.bit, tp2, 2, preceding, 0)
            check_oid(r, rr)
        lcol = random_col(rng, T.lng, n, lo=0, hi=5)
        l, rl = both(gdk, ref, lcol, T.lng)
        r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), op, p, l, None, T.bit, T.lng, 2, preceding, 0)
        rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rop, rp, rl, None, T.bit, T.lng, 2, preceding, 0)
        check_oid(r, rr)


def test_bound_errors(gdk, ref):
    rng = np.random.default_rng(1)
    n = 100
    pkey, o = window(rng, n, 3)
    p, rp = flags(gdk, ref, pkey, T.int)
    b, rb = both(gdk, ref, o, T.lng)
    for unit, word in ((0, "row"), (1, "range"), (2, "groups")):
        bb, rbb = (b, rb) if unit != 2 else (p, rp)
        with pytest.raises(G.GDKError) as e1:
            gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), bb, p, None, -1, T.lng, T.lng, unit, True, 0)
        with pytest.raises(Exception) as e2:
            ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rbb, rp, None, -1, T.lng, T.lng, unit, True, 0)
        assert str(e1.value).strip() == str(e2.value).strip() == f"42000!{word} frame bound must be non negative and non null."
    # dynamic bound with a negative entry
    lcol = np.full(n, 2, dtype=np.int64)
    lcol[37] = -4
    l, rl = both(gdk, ref, lcol, T.lng)
    with pytest.raises(G.GDKError) as e1:
        gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, l, None, T.lng, T.lng, 0, True, 0)
    with pytest.raises(Exception) as e2:
        ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, rl, None, T.lng, T.lng, 0, True, 0)
    assert str(e1.value).strip() == str(e2.value).strip()
    # RANGE overflow of the checked subtraction
    big = np.array([-(1 << 62) - 5, 0, (1 << 62) + 5], dtype=np.int64)
    bb, rbb = both(gdk, ref, big, T.lng)
    with pytest.raises(G.GDKError) as e1:
        gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, 3), bb, None, None, (1 << 63) - 2, T.lng, T.lng, 1, True, 0)
    with pytest.raises(Exception) as e2:
        ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, 3), rbb, None, None, (1 << 63) - 2, T.lng, T.lng, 1, True, 0)
    assert str(e1.value).strip() == str(e2.value).strip() == "22003!overflow in calculation."


# ------------------------------------------------------------------ framed aggregates
def frames(gdk, ref, n, nparts, rng, rows=(3, 2), nilfrac=0.0):
    """partition flags, peer flags and ROWS start/end columns on both sides"""
    pkey, o = window(rng, n, nparts)
    p, rp = flags(gdk, ref, pkey, T.int)
    op, rop = flags(gdk, ref, o, T.lng, p, rp)
    b, rb = both(gdk, ref, o, T.lng)
    s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, rows[0], T.lng, T.lng, 0, True, 0)
    rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, rows[0], T.lng, T.lng, 0, True, 0)
    e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, rows[1], T.lng, T.lng, 0, False, 1)
    re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, rows[1], T.lng, T.lng, 0, False, 1)
    return (p, op, s, e), (rp, rop, rs, re_)


def close(a, b, tt_out, exact, mag=1.0):
    """mag: magnitude of the input data — the reference's sequential flt running mean loses
    accuracy relative to the data it has seen, not to the (possibly small) current mean"""
    if exact:
        assert same_bits(a, b), (a[:10], b[:10], np.flatnonzero(a.view(np.uint8).reshape(len(a), -1) != b.view(np.uint8).reshape(len(b), -1))[:10])
        return
    tol = 1e-5 if tt_out == T.flt else 1e-12
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb)
    scale = np.maximum(np.abs(b[~nb]), mag)
    assert np.all(np.abs(a[~na] - b[~nb]) <= tol * scale), np.max(np.abs(a[~na] - b[~nb]) / scale)


def _check_avg_suffix(got, v, pf, of):
    """frame type 4: mean of the non-nil values from the start of the row's peer group to the end
    of its partition"""
    n = len(v)
    pstart = [i for i in range(n) if i == 0 or pf[i]] + [n]
    for a, z in zip(pstart[:-1], pstart[1:]):
        ostart = [i for i in range(a, z) if i == a or of[i]]
        for j, os_ in enumerate(ostart):
            oe = ostart[j + 1] if j + 1 < len(ostart) else z
            w = v[os_:z].astype(np.float64)
            w = w[~np.isnan(w)]
            want = w.mean() if len(w) else np.nan
            for k in range(os_, oe):
                assert (np.isnan(want) and np.isnan(got[k])) or abs(got[k] - want) <= 1e-9 * max(1.0, abs(want)), (k, got[k], want)


SHAPES = [(1, 1, (3, 2)), (37, 3, (3, 2)), (5000, 7, (40, 25)), (70001, 11, (300, 0)), (70001, 11, (5000, 5000)), (30000, 9000, (2, 2))]


@pytest.mark.parametrize("tp1,tp2", [(T.bte, T.bte), (T.bte, T.lng), (T.sht, T.int), (T.int, T.lng), (T.int, T.hge), (T.lng, T.lng),
                                     (T.lng, T.hge), (T.hge, T.hge), (T.flt, T.flt), (T.flt, T.dbl), (T.dbl, T.dbl)])
@pytest.mark.parametrize("frame_type", [0, 3, 4, 5, 6])
def test_sum(gdk, ref, tp1, tp2, frame_type):
    rng = np.random.default_rng(tp1 * 31 + tp2 + frame_type)
    for n, nparts, rows in SHAPES:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        lo, hi = (-1, 1) if tp2 == T.bte else (-3, 3) if tp2 == T.int else (-1000, 1000)
        if tp1 in (T.flt, T.dbl):
            v = np.round(random_col(rng, tp1, n, nilfrac=0.1, lo=lo, hi=hi) * 8) / 8
            v = v.astype(G.NP_OF[tp1])
        else:
            v = random_col(rng, tp1, n, nilfrac=0.1, lo=lo, hi=hi)
        if tp2 == T.bte and n > 100:
            continue
        b, rb = both(gdk, ref, v, tp1)
        r = gdk.GDKanalyticalsum(gdk.colnew(tp2, n), p, op, b, s, e, tp1, tp2, frame_type)
        rr = ref.GDKanalyticalsum(ref.colnew(tp2, n), rp, rop, rb, rs, re_, tp1, tp2, frame_type)
        exact = tp2 not in (T.flt, T.dbl) or frame_type in (0, 6)
        close(r.numpy(), rr.numpy(), tp2, exact)
        assert (r.props()["nonil"], r.props()["nil"]) == (rr.props()["nonil"], rr.props()["nil"])


def test_sum_overflow(gdk, ref):
    n = 40
    v = np.full(n, (1 << 62), dtype=np.int64)
    b, rb = both(gdk, ref, v, T.lng)
    o, ro = both(gdk, ref, np.zeros(n, dtype=np.int8), T.bit)
    for ft in (3, 5):
        with pytest.raises(G.GDKError) as e1:
            gdk.GDKanalyticalsum(gdk.colnew(T.lng, n), None, o, b, None, None, T.lng, T.lng, ft)
        with pytest.raises(Exception) as e2:
            ref.GDKanalyticalsum(ref.colnew(T.lng, n), None, ro, rb, None, None, T.lng, T.lng, ft)
        assert str(e1.value).strip() == str(e2.value).strip() == "22003!overflow in calculation."
    with pytest.raises(G.GDKError) as e1:
        gdk.GDKanalyticalsum(gdk.colnew(T.flt, n), None, None, b, None, None, T.lng, T.flt, 5)
    assert "not supported" in str(e1.value)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.flt, T.dbl])
@pytest.mark.parametrize("frame_type", [0, 3, 4, 5, 6])
def test_avg(gdk, ref, tt, frame_type):
    rng = np.random.default_rng(tt * 31 + frame_type)
    for n, nparts, rows in SHAPES:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        if tt in (T.flt, T.dbl):
            v = (np.round(random_col(rng, tt, n, nilfrac=0.1, lo=-1000, hi=1000) * 8) / 8).astype(G.NP_OF[tt])
        else:
            v = random_col(rng, tt, n, nilfrac=0.1)
        b, rb = both(gdk, ref, v, tt)
        r = gdk.GDKanalyticalfn("avg", gdk.colnew(T.dbl, n), p, op, b, s, e, tt, frame_type)
        if tt in (T.flt, T.dbl) and frame_type == 4:
            # reference quirk: ANALYTICAL_AVG_IMP_FP_CURRENT_ROW_TILL_UNBOUNDED never assigns curval
            # (gdk_analytic_statistics.c:228-250), so the reference returns nil for every row.  We
            # return the mean the frame defines; check it against a direct evaluation instead.
            if n <= 5000:
                _check_avg_suffix(r.numpy(), v, p.numpy(), op.numpy())
            continue
        rr = ref.GDKanalyticalfn("avg", ref.colnew(T.dbl, n), rp, rop, rb, rs, re_, tt, frame_type)
        # integers: exact formula on both sides; floats: framed = same tree order (bit-exact),
        # running means are sequential in the reference -> tolerance
        exact = tt not in (T.flt, T.dbl) or frame_type in (0, 6)
        if tt == T.flt and not exact:
            close(r.numpy(), rr.numpy(), T.flt, False, mag=1000.0)
        else:
            close(r.numpy(), rr.numpy(), T.dbl, exact)
        assert (r.props()["nonil"], r.props()["nil"]) == (rr.props()["nonil"], rr.props()["nil"])


@pytest.mark.parametrize("tt", [T.bte, T.int, T.lng, T.hge, T.dbl])
@pytest.mark.parametrize("frame_type", [0, 3, 4, 5, 6])
@pytest.mark.parametrize("ignore_nils", [True, False])
def test_count(gdk, ref, tt, frame_type, ignore_nils):
    rng = np.random.default_rng(tt * 31 + frame_type)
    for n, nparts, rows in SHAPES[:5]:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        v = random_col(rng, tt, n, nilfrac=0.3, lo=0, hi=9)
        b, rb = both(gdk, ref, v, tt)
        r = gdk.GDKanalyticalcount(gdk.colnew(T.lng, n), p, op, b, s, e, ignore_nils, tt, frame_type)
        rr = ref.GDKanalyticalcount(ref.colnew(T.lng, n), rp, rop, rb, rs, re_, ignore_nils, tt, frame_type)
        close(r.numpy(), rr.numpy(), T.lng, True)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
@pytest.mark.parametrize("frame_type", [0, 3, 4, 5, 6])
@pytest.mark.parametrize("fn", ["min", "max"])
def test_minmax(gdk, ref, tt, frame_type, fn):
    rng = np.random.default_rng(tt * 31 + frame_type)
    for n, nparts, rows in SHAPES[:5]:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        v = random_col(rng, tt, n, nilfrac=0.3)
        if tt in (T.flt, T.dbl):
            v[v == 0] = 1      # keep away from -0/+0 ties
        b, rb = both(gdk, ref, v, tt)
        r = gdk.GDKanalyticalfn(fn, gdk.colnew(tt, n), p, op, b, s, e, tt, frame_type)
        rr = ref.GDKanalyticalfn(fn, ref.colnew(tt, n), rp, rop, rb, rs, re_, tt, frame_type)
        a, c = r.numpy(), rr.numpy()
        if tt in (T.flt, T.dbl):
            assert np.array_equal(np.isnan(a), np.isnan(c)) and np.array_equal(a[~np.isnan(a)], c[~np.isnan(c)])
        else:
            close(a, c, tt, True)
        assert (r.props()["nonil"], r.props()["nil"]) == (rr.props()["nonil"], rr.props()["nil"])


def test_config5_shape(gdk, ref):
    """config 5 at a size the oracle finishes quickly: ROWS 100 PRECEDING..CURRENT ROW and
    RANGE 50 PRECEDING..50 FOLLOWING, moving sum (lng -> hge) and avg"""
    from monetdb_b200 import workload
    n = 300_000
    cols = workload.window(0, n, lens=(100, 1000, 20000))
    pk, rpk = both(gdk, ref, cols["pkey"], T.lng)
    p = gdk.GDKanalyticaldiff(gdk.colnew(T.bit, n), pk, None, None, T.lng)
    rp = ref.GDKanalyticaldiff(ref.colnew(T.bit, n), rpk, None, None, T.lng)
    o, ro = both(gdk, ref, cols["ord"], T.lng)
    v, rv = both(gdk, ref, cols["val"], T.lng)
    for unit, lo, hi, half in ((0, 100, 0, 1), (1, 50, 50, 0)):
        s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, lo, T.lng, T.lng, unit, True, 0)
        rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, lo, T.lng, T.lng, unit, True, 0)
        e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, hi, T.lng, T.lng, unit, False, half)
        re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, hi, T.lng, T.lng, unit, False, half)
        check_oid(s, rs)
        check_oid(e, re_)
        r = gdk.GDKanalyticalsum(gdk.colnew(T.hge, n), p, None, v, s, e, T.lng, T.hge, unit)
        rr = ref.GDKanalyticalsum(ref.colnew(T.hge, n), rp, None, rv, rs, re_, T.lng, T.hge, unit)
        close(r.numpy(), rr.numpy(), T.hge, True)
        r = gdk.GDKanalyticalfn("avg", gdk.colnew(T.dbl, n), p, None, v, s, e, T.lng, unit)
        rr = ref.GDKanalyticalfn("avg", ref.colnew(T.dbl, n), rp, None, rv, rs, re_, T.lng, unit)
        close(r.numpy(), rr.numpy(), T.dbl, True)


@pytest.mark.parametrize("name,ncases", [("analytics00", 50), ("analytics02", 36)])
def test_analytics_golden(gdk, name, ncases):
    """the reference's own expected columns for its window queries — sql/test/analytics/Tests/analytics00.test (default
    frames) and analytics02.test (ROWS BETWEEN n PRECEDING AND m FOLLOWING), fixtures tests/golden/analytics0N.json; the
    oracle is pinned to the same fixtures in tests/test_cpu_oracle.py — through BATsort -> BATproject ->
    GDKanalyticaldiff -> GDKanalyticalwindowbounds -> GDKanalytical<fn> on the device"""
    import json
    from pathlib import Path
    from tests.window_sql import run_case
    d = json.loads((Path(__file__).parent / "golden" / (name + ".json")).read_text())
    assert len(d["cases"]) == ncases

    def mk(arr, tt):
        hn = bool(has_nil(tt, arr))
        return gdk.BATfromhost(arr, tt, nonil=not hn, nil=hn)

    for c in d["cases"]:
        assert run_case(gdk, mk, T, d, c) == c["expected"], c["sql"]


def test_tree_twin_reuse_and_invalidation(gdk, ref):
    """the inner tree levels of an aggregate are kept as a twin of the value column, keyed by the partition list:
    reused for another frame over the same partitions, rebuilt for other partitions, another aggregate, and after the
    column has been rewritten (here: a window sum written into the same result BAT twice)"""
    rng = np.random.default_rng(99)
    n = 120_000
    pk1 = np.repeat(np.arange(6), n // 6).astype(np.int64)
    pk2 = np.repeat(np.arange(40), n // 40).astype(np.int64)
    vals = rng.integers(-10**9, 10**9, n, dtype=np.int64)
    v, rv = both(gdk, ref, vals, T.lng)
    o, ro = both(gdk, ref, np.arange(n, dtype=np.int64), T.lng)

    def part(pk):
        a, ra = both(gdk, ref, pk, T.lng)
        return (gdk.GDKanalyticaldiff(gdk.colnew(T.bit, n), a, None, None, T.lng),
                ref.GDKanalyticaldiff(ref.colnew(T.bit, n), ra, None, None, T.lng))

    def bounds(p, rp, lo, hi):
        s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, lo, T.lng, T.lng, 0, True, 0)
        e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, hi, T.lng, T.lng, 0, False, 1)
        rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, lo, T.lng, T.lng, 0, True, 0)
        re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, hi, T.lng, T.lng, 0, False, 1)
        return s, e, rs, re_

    def check(fn, b, rb, p, rp, fr, out_t=T.dbl):
        s, e, rs, re_ = fr
        r = gdk.GDKanalyticalfn(fn, gdk.colnew(out_t, n), p, None, b, s, e, T.lng, 0)
        rr = ref.GDKanalyticalfn(fn, ref.colnew(out_t, n), rp, None, rb, rs, re_, T.lng, 0)
        assert same_bits(r.numpy(), rr.numpy()), fn

    p1, rp1 = part(pk1)
    p2, rp2 = part(pk2)
    f1a, f1b, f2 = bounds(p1, rp1, 700, 0), bounds(p1, rp1, 40, 5000), bounds(p2, rp2, 700, 300)
    check("avg", v, rv, p1, rp1, f1a)
    check("avg", v, rv, p1, rp1, f1b)           # same partitions: the twin is reused
    check("avg", v, rv, p2, rp2, f2)            # other partitions: rebuilt
    check("avg", v, rv, p1, rp1, f1a)           # and back
    check("min", v, rv, p1, rp1, f1b, T.lng)    # another aggregate over the same column
    check("avg", v, rv, p1, rp1, f1b)
    # a column this library rewrites in place: the twin must not survive the rewrite
    small, rsmall = both(gdk, ref, (vals % 1000).astype(np.int64), T.lng)
    w, rw = gdk.colnew(T.lng, n), ref.colnew(T.lng, n)
    for src, rsrc in ((small, rsmall), (o, ro)):
        s, e, rs, re_ = f1a
        w = gdk.GDKanalyticalsum(w, p1, None, src, s, e, T.lng, T.lng, 0)
        rw = ref.GDKanalyticalsum(rw, rp1, None, rsrc, rs, re_, T.lng, T.lng, 0)
        assert same_bits(w.numpy(), rw.numpy())
        check("avg", w, rw, p1, rp1, f1b)


@pytest.mark.parametrize("tt", [T.sht, T.int, T.lng])
@pytest.mark.parametrize("preceding", [True, False])
def test_range_bounds_staged_search(gdk, ref, tt, preceding):
    """static limit, no nils, ascending partitions: the shared-memory staged search (an_range_staged_kernel), with
    frames inside the halo, longer than the halo (the search continues in global memory), whole partitions, runs of
    duplicates longer than a tile, and partitions from one row to many tiles"""
    rng = np.random.default_rng(7 * tt + preceding)
    for n, nparts, step in ((1, 1, 2), (1024, 1, 2), (1025, 2, 3), (20011, 1, 2), (20011, 7, 2), (20011, 900, 4), (60013, 5, 1)):
        pkey, o = window(rng, n, nparts, tt, 0, step)
        p, rp = flags(gdk, ref, pkey, T.int)
        b, rb = both(gdk, ref, o, tt)
        for lim in (1, 40, 400, 1500, 30000):
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, tt, T.lng, 1, preceding, 0)
            assert "staged" in gdk.algorithm()
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, tt, T.lng, 1, preceding, 0)
            check_oid(r, rr)
    # without a partition column
    o = np.cumsum(rng.integers(0, 3, 5000)).astype(G.NP_OF[tt])
    b, rb = both(gdk, ref, o, tt)
    r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, 5000), b, None, None, 9, tt, T.lng, 1, preceding, 0)
    rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, 5000), rb, None, None, 9, tt, T.lng, 1, preceding, 0)
    check_oid(r, rr)


def test_range_bounds_staged_overflow(gdk, ref):
    """the first row outside a frame is subtracted with the overflow check, also when it sits tiles away"""
    n = 5000
    for preceding in (True, False):
        o = np.arange(n, dtype=np.int64)
        if preceding:
            o[0] = -(1 << 63) + 2          # rows whose frame start reaches row 0 subtract it: overflow
        else:
            o[-1] = (1 << 63) - 1
            o[:-1] -= 10000                # negative values: v - max overflows
        b, rb = both(gdk, ref, o, T.lng)
        for lim in (3, 3000):
            with pytest.raises(Exception) as e2:
                ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, None, None, lim, T.lng, T.lng, 1, preceding, 0)
            with pytest.raises(G.GDKError) as e1:
                gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, None, None, lim, T.lng, T.lng, 1, preceding, 0)
            assert str(e1.value).strip() == str(e2.value).strip() == "22003!overflow in calculation."


@pytest.mark.parametrize("tt", [T.int, T.lng, T.dbl])
@pytest.mark.parametrize("preceding", [True, False])
@pytest.mark.parametrize("order", ["random", "desc"])
def test_range_bounds_unsorted(gdk, ref, tt, preceding, order):
    """order columns that are NOT ascending inside the partitions take the per-row walk
    (the reference itself never requires sortedness)"""
    rng = np.random.default_rng(tt + len(order))
    for n, nparts, nilfrac in ((300, 4, 0.0), (20011, 40, 0.0), (20011, 40, 0.1)):
        pkey, o = window(rng, n, nparts, tt)
        if order == "random":
            o = random_col(rng, tt, n, nilfrac=nilfrac, lo=0, hi=40)
        else:
            o = o[::-1].copy()
        p, rp = flags(gdk, ref, pkey, T.int)
        b, rb = both(gdk, ref, o, tt)
        for lim in ((1.5, 7.0) if tt == T.dbl else (1, 7)):
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, tt, tt, 1, preceding, 0)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, tt, tt, 1, preceding, 0)
            check_oid(r, rr)
        assert "walk" in gdk.algorithm()


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng])
@pytest.mark.parametrize("frame_type", [0, 3, 4, 5, 6])
def test_avginteger(gdk, ref, tt, frame_type):
    """GDKanalyticalavginteger: integer-typed average with the reference's rounding, bit-exact"""
    rng = np.random.default_rng(tt * 17 + frame_type)
    for n, nparts, rows in SHAPES:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        v = random_col(rng, tt, n, nilfrac=0.1)
        b, rb = both(gdk, ref, v, tt)
        r = gdk.GDKanalyticalfn("avginteger", gdk.colnew(tt, n), p, op, b, s, e, tt, frame_type)
        rr = ref.GDKanalyticalfn("avginteger", ref.colnew(tt, n), rp, rop, rb, rs, re_, tt, frame_type)
        close(r.numpy(), rr.numpy(), tt, True)
        assert (r.props()["nonil"], r.props()["nil"]) == (rr.props()["nonil"], rr.props()["nil"])


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng])
@pytest.mark.parametrize("fn", ["avg", "avginteger"])
def test_integer_averages_nil_free_framed(gdk, ref, tt, fn):
    """framed integer averages over a column without nils: level 0 and the inner tree levels both come from running
    totals (per-level totals of the node values), bit-exact with the reference's node walk"""
    rng = np.random.default_rng(tt * 5 + len(fn))
    for n, nparts, rows in SHAPES:
        (p, op, s, e), (rp, rop, rs, re_) = frames(gdk, ref, n, nparts, rng, rows)
        v = random_col(rng, tt, n, nilfrac=0.0)
        b, rb = both(gdk, ref, v, tt)
        out_t = T.dbl if fn == "avg" else tt
        for _ in range(2):          # the second run takes the tree and its totals from the column's twin
            r = gdk.GDKanalyticalfn(fn, gdk.colnew(out_t, n), p, op, b, s, e, tt, 0)
            rr = ref.GDKanalyticalfn(fn, ref.colnew(out_t, n), rp, rop, rb, rs, re_, tt, 0)
            assert same_bits(r.numpy(), rr.numpy())
            assert (r.props()["nonil"], r.props()["nil"]) == (rr.props()["nonil"], rr.props()["nil"])


# ---------------------------------------------------------------------------- temporal RANGE frames (a20)
def _dates(rng, n, span_days=4000):
    """packed GDK dates (gdk_time.c:17-31) of random days after 1990-01-01, incl. month ends and leap days"""
    import datetime as dt
    base = dt.date(1990, 1, 1)
    days = np.sort(rng.integers(0, span_days, n))
    out = np.empty(n, dtype=np.int32)
    for i, d in enumerate(days):
        x = base + dt.timedelta(days=int(d))
        out[i] = (((x.year + 4712) * 12 + x.month - 1) << 5) | x.day
    return out


def _timestamps(rng, n):
    d = _dates(rng, n).astype(np.int64)
    usec = rng.integers(0, 86_400_000_000, n).astype(np.int64)
    ts = (d << 37) | usec
    return np.sort(ts)


@pytest.mark.parametrize("tp1", [T.date, T.timestamp, T.daytime])
@pytest.mark.parametrize("preceding", [True, False])
@pytest.mark.parametrize("order", ["sorted", "random"])
def test_temporal_range_bounds(gdk, ref, tp1, preceding, order):
    """RANGE frames on date / timestamp / daytime order columns with month (int) and millisecond (lng)
    interval bounds (gdk_analytic_bounds.c:458-611), static and per-row bounds, with and without nils"""
    rng = np.random.default_rng(tp1 * 7 + preceding)
    for n, nparts, nilfrac in ((1, 1, 0.0), (300, 4, 0.0), (6007, 40, 0.0), (6007, 40, 0.15)):
        pkey = np.sort(rng.integers(0, nparts, n)).astype(np.int32)
        if tp1 == T.date:
            o = _dates(rng, n)
        elif tp1 == T.timestamp:
            o = _timestamps(rng, n)
        else:
            o = np.sort(rng.integers(0, 86_400_000_000, n)).astype(np.int64)
        # ascending inside every partition (what ORDER BY produces), or shuffled
        if order == "sorted":
            for k in range(nparts):
                m = pkey == k
                o[m] = np.sort(o[m])
        else:
            o = rng.permutation(o)
        if nilfrac:
            nil = np.iinfo(o.dtype).min
            mask = rng.random(n) < nilfrac
            o = o.copy()
            o[mask] = nil
            if order == "sorted":                       # nil sorts first inside a partition
                for k in range(nparts):
                    m = pkey == k
                    o[m] = np.sort(o[m])
        p, rp = flags(gdk, ref, pkey, T.int)
        b, rb = both(gdk, ref, o, tp1)
        cases = []
        if tp1 != T.daytime:
            cases += [(1, T.int), (3, T.int), (14, T.int), (0, T.int), (2**31 - 1, T.int), (400000, T.int)]
        cases += [(86_400_000, T.lng), (3 * 86_400_000 + 5, T.lng), (3_600_000, T.lng), (1, T.lng), (0, T.lng), ((1 << 63) - 1, T.lng),
                  (10**15, T.lng)]
        for lim, tp2 in cases:
            r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, tp1, tp2, 1, preceding, 0)
            rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, tp1, tp2, 1, preceding, 0)
            check_oid(r, rr)
        if n > 1:
            for tp2, hi in (((T.int, 20),) if tp1 != T.daytime else ()) + ((T.lng, 5 * 86_400_000),):
                lcol = random_col(rng, tp2, n, lo=0, hi=hi)
                l, rl = both(gdk, ref, lcol, tp2)
                r = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, l, None, tp1, tp2, 1, preceding, 0)
                rr = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, rl, None, tp1, tp2, 1, preceding, 0)
                check_oid(r, rr)
            if order == "sorted" and not nilfrac:
                assert "galloping" in gdk.algorithm()


def test_temporal_range_bound_errors(gdk, ref):
    rng = np.random.default_rng(3)
    n = 200
    o = _dates(rng, n)
    pkey = np.zeros(n, dtype=np.int32)
    p, rp = flags(gdk, ref, pkey, T.int)
    b, rb = both(gdk, ref, o, T.date)
    for lim, tp2 in ((-1, T.int), (-5, T.lng), (3, T.sht), (1.5, T.dbl)):
        with pytest.raises(G.GDKError) as e1:
            gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), b, p, None, lim, T.date, tp2, 1, True, 0)
        with pytest.raises(Exception) as e2:
            ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rb, rp, None, lim, T.date, tp2, 1, True, 0)
        assert str(e1.value).strip() == str(e2.value).strip(), (lim, tp2)
    # daytime takes no month interval
    d, rd = both(gdk, ref, np.sort(rng.integers(0, 86_400_000_000, n)).astype(np.int64), T.daytime)
    with pytest.raises(G.GDKError) as e1:
        gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), d, p, None, 2, T.daytime, T.int, 1, True, 0)
    with pytest.raises(Exception) as e2:
        ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), rd, rp, None, 2, T.daytime, T.int, 1, True, 0)
    assert str(e1.value).strip() == str(e2.value).strip()


# ---------------------------------------------------------------------------- frame types 1 (RANGE) and 2 (GROUPS)
def frames_unit(gdk, ref, n, nparts, rng, unit):
    """partition / peer flags and start / end columns made by RANGE (unit 1) or GROUPS (unit 2) bounds"""
    pkey, o = window(rng, n, nparts)
    p, rp = flags(gdk, ref, pkey, T.int)
    op, rop = flags(gdk, ref, o, T.lng, p, rp)
    b, rb = both(gdk, ref, o, T.lng)
    if unit == 1:
        args, rargs, tp1 = (b, p, None), (rb, rp, None), T.lng
    else:
        args, rargs, tp1 = (op, p, None), (rop, rp, None), T.bit
    s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), *args, 2, tp1, T.lng, unit, True, 0)
    rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), *rargs, 2, tp1, T.lng, unit, True, 0)
    e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), *args, 1, tp1, T.lng, unit, False, 0)
    re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), *rargs, 1, tp1, T.lng, unit, False, 0)
    check_oid(s, rs)
    check_oid(e, re_)
    return (p, op, s, e), (rp, rop, rs, re_)


@pytest.mark.parametrize("frame_type", [1, 2])
@pytest.mark.parametrize("tt", [T.int, T.lng, T.dbl])
def test_aggregates_on_range_and_groups_frames(gdk, ref, frame_type, tt):
    """sum / avg / avginteger / count / min / max with frame_type 1 (RANGE) and 2 (GROUPS): the start / end
    columns come from the corresponding bounds (the reference evaluates them like ROWS frames, through its
    segment tree: gdk_analytic_func.c:1959, gdk_analytic_statistics.c:364)"""
    rng = np.random.default_rng(tt * 7 + frame_type)
    for n, nparts in ((1, 1), (300, 4), (40001, 23)):
        (p, op, s, e), (rp, rop, rs, re_) = frames_unit(gdk, ref, n, nparts, rng, frame_type)
        if tt == T.dbl:
            v = (np.round(random_col(rng, tt, n, nilfrac=0.1, lo=-1000, hi=1000) * 8) / 8).astype(np.float64)
        else:
            v = random_col(rng, tt, n, nilfrac=0.1)
        b, rb = both(gdk, ref, v, tt)
        tp2 = T.dbl if tt == T.dbl else T.hge
        r = gdk.GDKanalyticalsum(gdk.colnew(tp2, n), p, op, b, s, e, tt, tp2, frame_type)
        rr = ref.GDKanalyticalsum(ref.colnew(tp2, n), rp, rop, rb, rs, re_, tt, tp2, frame_type)
        close(r.numpy(), rr.numpy(), tp2, True)
        r = gdk.GDKanalyticalfn("avg", gdk.colnew(T.dbl, n), p, op, b, s, e, tt, frame_type)
        rr = ref.GDKanalyticalfn("avg", ref.colnew(T.dbl, n), rp, rop, rb, rs, re_, tt, frame_type)
        close(r.numpy(), rr.numpy(), T.dbl, True)
        if tt != T.dbl:
            r = gdk.GDKanalyticalfn("avginteger", gdk.colnew(tt, n), p, op, b, s, e, tt, frame_type)
            rr = ref.GDKanalyticalfn("avginteger", ref.colnew(tt, n), rp, rop, rb, rs, re_, tt, frame_type)
            close(r.numpy(), rr.numpy(), tt, True)
        for ign in (True, False):
            r = gdk.GDKanalyticalcount(gdk.colnew(T.lng, n), p, op, b, s, e, ign, tt, frame_type)
            rr = ref.GDKanalyticalcount(ref.colnew(T.lng, n), rp, rop, rb, rs, re_, ign, tt, frame_type)
            close(r.numpy(), rr.numpy(), T.lng, True)
        for fn in ("min", "max"):
            r = gdk.GDKanalyticalfn(fn, gdk.colnew(tt, n), p, op, b, s, e, tt, frame_type)
            rr = ref.GDKanalyticalfn(fn, ref.colnew(tt, n), rp, rop, rb, rs, re_, tt, frame_type)
            a, c = r.numpy(), rr.numpy()
            if tt == T.dbl:
                assert np.array_equal(np.isnan(a), np.isnan(c)) and np.array_equal(a[~np.isnan(a)], c[~np.isnan(c)])
            else:
                close(a, c, tt, True)


def test_window_sum_node_overflow_matches_reference(gdk, ref):
    """With lng sums of 2^60-sized values a level-1 node of the fan-out-16 segment tree (16 consecutive rows)
    overflows although no FRAME does (ROWS 1 PRECEDING .. CURRENT ROW adds two values).  The reference fails the
    whole call (gdk_analytic_func.c:1762-1796 builds every node with ADD_WITH_CHECK); the device path builds the
    same nodes with the same check and must fail with the same text — an earlier note in BASELINE.md §5 claimed a
    deviation here; this test is what settled that there is none."""
    n = 64
    v = np.full(n, 1 << 60, dtype=np.int64)
    pkey = np.zeros(n, dtype=np.int32)
    p, rp = flags(gdk, ref, pkey, T.int)
    b, rb = both(gdk, ref, v, T.lng)
    o, ro = both(gdk, ref, np.arange(n, dtype=np.int64), T.lng)
    s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, 1, T.lng, T.lng, 0, True, 0)
    e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, n), o, p, None, 0, T.lng, T.lng, 0, False, 1)
    rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, 1, T.lng, T.lng, 0, True, 0)
    re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, n), ro, rp, None, 0, T.lng, T.lng, 0, False, 1)
    with pytest.raises(Exception) as e2:
        ref.GDKanalyticalsum(ref.colnew(T.lng, n), rp, None, rb, rs, re_, T.lng, T.lng, 0)
    with pytest.raises(G.GDKError) as e1:
        gdk.GDKanalyticalsum(gdk.colnew(T.lng, n), p, None, b, s, e, T.lng, T.lng, 0)
    assert str(e1.value).strip() == str(e2.value).strip() == "22003!overflow in calculation."
    # partial nodes count too: the reference builds every level up to a single root (GDKrebuild_segment_tree,
    # gdk_analytic_func.c:34-45), so 15 rows make ONE level-1 node of 15 values that overflows; 7 rows do not
    for m, fails in ((15, True), (7, False), (17, True)):
        b, rb = both(gdk, ref, v[:m], T.lng)
        p, rp = flags(gdk, ref, pkey[:m], T.int)
        o, ro = both(gdk, ref, np.arange(m, dtype=np.int64), T.lng)
        s = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, m), o, p, None, 1, T.lng, T.lng, 0, True, 0)
        e = gdk.GDKanalyticalwindowbounds(gdk.colnew(T.oid, m), o, p, None, 0, T.lng, T.lng, 0, False, 1)
        rs = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, m), ro, rp, None, 1, T.lng, T.lng, 0, True, 0)
        re_ = ref.GDKanalyticalwindowbounds(ref.colnew(T.oid, m), ro, rp, None, 0, T.lng, T.lng, 0, False, 1)
        if fails:
            with pytest.raises(Exception) as e2:
                ref.GDKanalyticalsum(ref.colnew(T.lng, m), rp, None, rb, rs, re_, T.lng, T.lng, 0)
            with pytest.raises(G.GDKError) as e1:
                gdk.GDKanalyticalsum(gdk.colnew(T.lng, m), p, None, b, s, e, T.lng, T.lng, 0)
            assert str(e1.value).strip() == str(e2.value).strip() == "22003!overflow in calculation."
        else:
            want = ref.GDKanalyticalsum(ref.colnew(T.lng, m), rp, None, rb, rs, re_, T.lng, T.lng, 0).numpy()
            r = gdk.GDKanalyticalsum(gdk.colnew(T.lng, m), p, None, b, s, e, T.lng, T.lng, 0)
            assert np.array_equal(r.numpy(), want)
            assert want.tolist() == [1 << 60] + [1 << 61] * (m - 1)
