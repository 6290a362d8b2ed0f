"""The operator sequences of BASELINE.json's configs, written once against the GDK
entry-point names.  ``X`` is any module exposing the GDK API
mirror (``monetdb_b200.gdk`` for the CUDA path; the tests and the CPU-baseline
leg of bench.py pass the reference driver), so both sides run the identical
MAL-equivalent plan."""
from __future__ import annotations

from .workload import todate

# GDK type ids (gdk.h:429-451)
T_bte, T_int, T_oid, T_dbl, T_lng, T_hge = 3, 5, 6, 9, 10, 11

D19940101 = todate(1994, 1, 1)
D19950101 = todate(1995, 1, 1)
D19980902 = todate(1998, 9, 2)


def q6_product(X, col, counts=None):
    """Q6 up to the per-row product: 3 selects -> 2 projects -> mul (hge)"""
    c1 = X.BATselect(col["shipdate"], None, D19940101, D19950101, True, False, False, False)
    c2 = X.BATselect(col["discount"], c1, 5, 7, True, True, False, False)
    c3 = X.BATthetaselect(col["quantity"], c2, 2400, "<")
    pe = X.BATproject(c3, col["extprice"])
    pd = X.BATproject(c3, col["discount"])
    x = X.BATcalc("mul", pe, pd, None, None, T_hge)
    if counts is not None:
        counts.update(n1=len(c1), n2=len(c2), n3=len(c3))
    return x


def q6(X, col, counts=None):
    """TPC-H Q6 shape (config 2): 3 selects -> 2 projects -> mul -> sum(hge).
    col: dict of BATs shipdate(date) discount(lng) quantity(lng) extprice(lng)."""
    return X.BATsum(T_hge, q6_product(X, col, counts), None, True, True)


def q6_bytes(n, n1, n2, n3):
    """algorithmic bytes of one Q6 pass, per operator (SURVEY.md §8d formulas)"""
    return {
        "select1": 4 * n + 8 * n1,
        "select2": (8 + 8) * n1 + 8 * n2,
        "select3": (8 + 8) * n2 + 8 * n3,
        "project": 2 * (8 + 8 + 8) * n3,
        "mul": (8 + 8 + 16) * n3,
        "sum": 16 * n3,
    }


def config1(X, col, out=None):
    """config 1: select(50%) -> 2 projects -> group -> groupsum(hge)"""
    s = X.BATselect(col["k"], None, 100, 599, True, True, False, False)
    pk = X.BATproject(s, col["k"])
    pv = X.BATproject(s, col["v"])
    g, e, h = X.BATgroup(pk, None, None, None, None)
    sums = X.BATgroupsum(pv, g, e, None, T_hge, True)
    if out is not None:
        out.update(s=s, g=g, e=e, h=h, sums=sums)
    return sums


def q1_groups_and_calcs(X, col, o):
    """Q1 up to the aggregates: select -> 6 projects -> 2-level group -> 4 calcs; the BATs go into o"""
    c = X.BATthetaselect(col["shipdate"], None, D19980902, "<=")
    qty = X.BATproject(c, col["quantity"])
    price = X.BATproject(c, col["extprice"])
    disc = X.BATproject(c, col["discount"])
    tax = X.BATproject(c, col["tax"])
    rf = X.BATproject(c, col["returnflag"])
    ls = X.BATproject(c, col["linestatus"])
    g1, e1, h1 = X.BATgroup(rf, None, None, None, None)
    g, e, h = X.BATgroup(ls, None, g1, e1, h1)
    t1 = X.BATcalccst("sub", disc, 100, T_lng, None, T_lng, cst_first=True)
    dp = X.BATcalc("mul", price, t1, None, None, T_hge)
    t2 = X.BATcalccst("add", tax, 100, T_lng, None, T_lng, cst_first=True)
    ch = X.BATcalc("mul", dp, t2, None, None, T_hge)
    o.update(cand=c, qty=qty, price=price, disc=disc, tax=tax, rf=rf, ls=ls, g=g, e=e, h=h, dp=dp, ch=ch)


def q1(X, col, out=None):
    """TPC-H Q1 shape (config 3): select -> 6 projects -> 2-level group -> 4 calcs -> 8 aggregates"""
    o = {}
    q1_groups_and_calcs(X, col, o)
    g, e, h, qty, price, disc, dp, ch = (o[k] for k in ("g", "e", "h", "qty", "price", "disc", "dp", "ch"))
    res = {
        "sum_qty": X.BATgroupsum(qty, g, e, None, T_hge, True),
        "sum_base_price": X.BATgroupsum(price, g, e, None, T_hge, True),
        "sum_disc_price": X.BATgroupsum(dp, g, e, None, T_hge, True),
        "sum_charge": X.BATgroupsum(ch, g, e, None, T_hge, True),
        "avg_qty": X.BATgroupavg(qty, g, e, None, T_dbl, True, 2)[0],
        "avg_price": X.BATgroupavg(price, g, e, None, T_dbl, True, 2)[0],
        "avg_disc": X.BATgroupavg(disc, g, e, None, T_dbl, True, 2)[0],
        "count_order": X.BATgroupcount(qty, g, e, None, T_lng, True),
        "extents": e, "histo": h,
    }
    if out is not None:
        out.update(nsel=len(o["cand"]), groups=g, cand=o["cand"], rf=o["rf"], ls=o["ls"], disc=disc)
    return res


def q1_bytes(n, s):
    """algorithmic bytes of one Q1 pass with s selected rows (4 + 390 sigma per row)"""
    return 4 * n + 8 * s + 116 * s + 26 * s + 104 * s + 136 * s


def q1_bytes_sharded(n, s):
    """the sharded plan's phase 1 (shard.q1_phase1): the three BATgroupavg passes (3 x (8+8)) are replaced by
    one more BATgroupsum (l_discount: 8+8); the counts of the non-nil values are BATgroup's histogram when the
    columns hold no nils (the headline tables), three BATgroupcount (8 each) otherwise; the projections of the
    extents and phase 2 touch a handful of rows"""
    return q1_bytes(n, s) - 3 * 16 * s + 16 * s - 8 * s
