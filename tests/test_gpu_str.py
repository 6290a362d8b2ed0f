"""Dictionary-encoded string columns (first slice) against the reference: the str BAT is
built by the real GDK (BUNappend: its heap, duplicate elimination and offset width), shipped to the device as
it is stored (offsets + heap), and BATgroup (gdk_group.c:897-919), the (anti-)equality selects
(gdk_select.c:608-656, 2102-2127) and BATproject (gdk_project.c:681-695) must give the reference's results:
oids and group ids bit-exact, projected strings equal value by value."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE
NIL = b"\x80"


def make(gdk, ref, values, hseq=0):
    rb = ref.strbat(values, hseq=hseq)
    off, heap, elim = ref.str_parts(rb)
    b = gdk.BATstrfromhost(off, heap, hseq=hseq, elim_doubles=elim)
    return b, rb, (off, heap, elim)


def words(rng, n, distinct, nilfrac=0.0, wlen=(1, 12)):
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz0123456789 _-", dtype=np.uint8)
    pool = []
    seen = set()
    while len(pool) < distinct:
        w = bytes(rng.choice(alphabet, int(rng.integers(wlen[0], wlen[1] + 1))))
        if w not in seen:
            seen.add(w)
            pool.append(w)
    idx = rng.integers(0, distinct, n)
    vals = [pool[i] for i in idx]
    if nilfrac:
        for i in np.nonzero(rng.random(n) < nilfrac)[0]:
            vals[i] = NIL
    return vals, pool


CASES = [
    # n, distinct, nil fraction, word lengths -> offset width / duplicate-free heap as GDK decides
    (1, 1, 0.0, (1, 5)),
    (40, 6, 0.2, (1, 5)),          # width 1
    (5000, 12, 0.05, (1, 8)),       # width 1
    (70_001, 900, 0.02, (3, 12)),   # width 2, duplicate-free
    (70_001, 3000, 0.0, (3, 7)),    # width 2, heap close to the 64 KiB duplicate-elimination limit
    (60_000, 30_000, 0.01, (6, 16)),  # heap beyond 64 KiB: duplicates possible, width 4
]


@pytest.mark.parametrize("case", CASES)
def test_roundtrip_and_layout(gdk, ref, case):
    n, d, nf, wl = case
    rng = np.random.default_rng(n + d)
    vals, _ = words(rng, n, d, nf, wl)
    b, rb, (off, heap, elim) = make(gdk, ref, vals, hseq=3)
    o2, h2 = gdk.str_parts(b)
    assert same_bits(o2, off) and h2 == heap
    assert gdk.str_decode(o2, h2) == vals == ref.str_values(rb)
    assert b.props()["hseq"] == 3 and len(b) == n


@pytest.mark.parametrize("case", CASES[:5])
def test_group(gdk, ref, case):
    n, d, nf, wl = case
    rng = np.random.default_rng(7 * n + d)
    vals, _ = words(rng, n, d, nf, wl)
    b, rb, (off, heap, elim) = make(gdk, ref, vals, hseq=11)
    assert elim
    g, e, h = gdk.BATgroup(b)
    rg, re_, rh = ref.BATgroup(rb)
    assert same_bits(g.numpy(), rg.numpy())
    assert same_bits(e.numpy(), re_.numpy())
    assert same_bits(h.numpy(), rh.numpy())
    # refinement by a second str column and a candidate list
    vals2, _ = words(rng, n, 3, 0.1, (1, 3))
    b2, rb2, _ = make(gdk, ref, vals2, hseq=11)
    g2, e2, h2 = gdk.BATgroup(b2, None, g, e, h)
    rg2, re2, rh2 = ref.BATgroup(rb2, None, rg, re_, rh)
    assert same_bits(g2.numpy(), rg2.numpy())
    assert same_bits(e2.numpy(), re2.numpy())
    assert same_bits(h2.numpy(), rh2.numpy())
    if n > 10:
        c = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.uint64) + 11
        s = gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True)
        rs = ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)
        g3, e3, h3 = gdk.BATgroup(b, s)
        rg3, re3, rh3 = ref.BATgroup(rb, rs)
        assert same_bits(g3.numpy(), rg3.numpy())
        assert same_bits(e3.numpy(), re3.numpy())
        assert same_bits(h3.numpy(), rh3.numpy())


def test_group_needs_unique_heap(gdk, ref):
    rng = np.random.default_rng(5)
    vals, _ = words(rng, 60_000, 30_000, 0.0, (6, 16))
    b, rb, (off, heap, elim) = make(gdk, ref, vals)
    assert not elim and off.dtype.itemsize == 4
    with pytest.raises(G.GDKError, match="not on the device path"):
        gdk.BATgroup(b)


def cands(gdk, ref, rng, n, hseq):
    """(ours, theirs) candidate lists: none, dense slice, sparse oid list"""
    out = [(None, None)]
    if n > 4:
        lo, cnt = hseq + n // 4, n // 2
        out.append((gdk.BATdense(0, lo, cnt), ref.dense(0, lo, cnt)))
        c = np.sort(rng.choice(n, max(1, n // 5), replace=False)).astype(np.uint64) + hseq
        out.append((gdk.BATfromhost(c, T.oid, sorted=True, key=True, nonil=True),
                    ref.bat(c, ref.T.oid, sorted=True, key=True, nonil=True)))
    return out


@pytest.mark.parametrize("case", CASES)
def test_thetaselect(gdk, ref, case):
    n, d, nf, wl = case
    rng = np.random.default_rng(3 * n + d)
    vals, pool = words(rng, n, d, nf, wl)
    b, rb, _ = make(gdk, ref, vals, hseq=5)
    probes = [pool[0], pool[-1], pool[0][:-1] + b"~", pool[0] + b"x", b"", NIL]
    for s, rs in cands(gdk, ref, rng, n, 5):
        for v in probes:
            for op in ("==", "=", "!=", "<>", "eq", "ne"):
                r = gdk.BATthetaselect(b, s, v, op)
                rr = ref.BATthetaselect(rb, rs, v, op)
                assert same_bits(r.numpy(), rr.numpy()), (op, v, len(r), len(rr))
                assert r.props()["hseq"] == rr.props()["hseq"] == 0


@pytest.mark.parametrize("case", CASES[1:4])
def test_select_forms(gdk, ref, case):
    n, d, nf, wl = case
    rng = np.random.default_rng(n ^ d)
    vals, pool = words(rng, n, d, max(nf, 0.1), wl)
    b, rb, _ = make(gdk, ref, vals, hseq=0)
    for s, rs in cands(gdk, ref, rng, n, 0):
        for tl in (pool[1], NIL):
            for th in (None, tl):
                for anti in (False, True):
                    for nm in (False, True):
                        r = gdk.BATselect(b, s, tl, th, True, True, anti, nm)
                        rr = ref.BATselect(rb, rs, tl, th, True, True, anti, nm)
                        assert same_bits(r.numpy(), rr.numpy()), (tl, th, anti, nm, len(r), len(rr))


def test_select_ranges_fail_loudly(gdk, ref):
    b, rb, _ = make(gdk, ref, [b"a", b"b", b"c"])
    with pytest.raises(G.GDKError, match="only equality / inequality selects on str columns"):
        gdk.BATselect(b, None, b"a", b"b", True, True, False, False)
    with pytest.raises(G.GDKError, match="not on the device path"):
        gdk.BATthetaselect(b, None, b"b", "<")


@pytest.mark.parametrize("case", CASES[1:])
def test_project(gdk, ref, case):
    n, d, nf, wl = case
    rng = np.random.default_rng(n + 17 * d)
    vals, pool = words(rng, n, d, nf, wl)
    b, rb, _ = make(gdk, ref, vals, hseq=9)
    # through a select result, a random oid list and a dense slice
    r = gdk.BATthetaselect(b, None, pool[0], "!=")
    rr = ref.BATthetaselect(rb, None, pool[0], "!=")
    p, rp = gdk.BATproject(r, b), ref.BATproject(rr, rb)
    assert gdk.str_decode(*gdk.str_parts(p)) == ref.str_values(rp)
    l = (rng.integers(0, n, 2 * n + 1).astype(np.uint64) + 9)
    lb, rlb = gdk.BATfromhost(l, T.oid, nonil=True), ref.bat(l, ref.T.oid, nonil=True)
    p, rp = gdk.BATproject(lb, b), ref.BATproject(rlb, rb)
    assert gdk.str_decode(*gdk.str_parts(p)) == ref.str_values(rp)
    assert (p.props()["hseq"], len(p)) == (rp.props()["hseq"], len(rp))
    ld, rld = gdk.BATdense(2, 9 + n // 3, n // 2), ref.dense(2, 9 + n // 3, n // 2)
    p, rp = gdk.BATproject(ld, b), ref.BATproject(rld, rb)
    assert gdk.str_decode(*gdk.str_parts(p)) == ref.str_values(rp)
    assert p.props()["hseq"] == rp.props()["hseq"] == 2
    # the projected column is a str column again: group it / select on it
    if ref.str_parts(rp)[2]:
        g, e, h = gdk.BATgroup(p)
        rg, re_, rh = ref.BATgroup(rp)
        assert same_bits(g.numpy(), rg.numpy()) and same_bits(h.numpy(), rh.numpy())
    q, rq = gdk.BATthetaselect(p, None, pool[-1], "=="), ref.BATthetaselect(rp, None, pool[-1], "==")
    assert same_bits(q.numpy(), rq.numpy())
    # owner freed before the view: the shared offsets heap stays alive
    b.free()
    assert gdk.str_decode(*gdk.str_parts(p)) == ref.str_values(rp)


def test_q1_style_string_keys(gdk, ref):
    """GROUP BY two string flags + SUM: the Q1 shape with the keys as the str columns they are in TPC-H"""
    rng = np.random.default_rng(42)
    n = 200_000
    rf = [bytes([c]) for c in rng.choice(np.frombuffer(b"ANR", dtype=np.uint8), n)]
    ls = [bytes([c]) for c in rng.choice(np.frombuffer(b"FO", dtype=np.uint8), n)]
    qty = rng.integers(1, 51, n).astype(np.int32)
    b1, r1, _ = make(gdk, ref, rf)
    b2, r2, _ = make(gdk, ref, ls)
    q, rq = gdk.BATfromhost(qty, T.int, nonil=True), ref.bat(qty, ref.T.int, nonil=True)
    g, e, h = gdk.BATgroup(b1)
    g, e, h = gdk.BATgroup(b2, None, g, e, h)
    rg, re_, rh = ref.BATgroup(r1)
    rg, re_, rh = ref.BATgroup(r2, None, rg, re_, rh)
    assert same_bits(g.numpy(), rg.numpy()) and same_bits(e.numpy(), re_.numpy())
    s, rs = gdk.BATgroupsum(q, g, e, None, T.lng), ref.BATgroupsum(rq, rg, re_, None, ref.T.lng)
    assert same_bits(s.numpy(), rs.numpy())
    k1, rk1 = gdk.BATproject(e, b1), ref.BATproject(re_, r1)
    k2, rk2 = gdk.BATproject(e, b2), ref.BATproject(re_, r2)
    assert gdk.str_decode(*gdk.str_parts(k1)) == ref.str_values(rk1)
    assert gdk.str_decode(*gdk.str_parts(k2)) == ref.str_values(rk2)
