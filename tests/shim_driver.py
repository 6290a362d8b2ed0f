"""Runs BASELINE config 1, the Q6 plan, a join, error cases and a window chain through the REFERENCE driver
(oracle/refgdk.py: BATs created by the real GDK, calls made through the reference's own entry-point names) and
dumps tails, properties and GDKerrbuf texts as JSON.  tests/test_gpu_shim.py runs it twice — plain
libbat_ref, and with REFGDK_SHIM=integration/_build/libbat_b200shim.so interposed — and diffs the dumps."""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from oracle import refgdk as R          # noqa: E402
from monetdb_b200 import pipelines, workload   # noqa: E402  (no CUDA library is loaded by these two)


def dump(b, order_free=False):
    a = b.numpy()
    if order_free:
        a = np.sort(a)
    p = b.props()
    return {"sha": hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest(), "count": p["count"], "hseq": p["hseq"],
            "type": p["type"] if p["type"] != 0 else 6, "nonil": p["nonil"], "nil": p["nil"],
            # sortedness / key flags may be conservatively false on either side; a TRUE claim must hold (checked below)
            "head": a[:4].tolist() if a.dtype.names is None else [int(x) for x in a["lo"][:4]]}


def main():
    n = int(sys.argv[1])
    R.init(1)
    out = {}
    T = R.T
    # ---- config 1
    k, v = workload.config1(0, n)
    c1 = {"k": R.bat(k, T.int, nonil=True), "v": R.bat(v, T.lng, nonil=True)}
    o = {}
    sums = pipelines.config1(R, c1, o)
    out["config1"] = {x: dump(o[x]) for x in ("s", "g", "e", "h", "sums")}
    for x in ("s", "g", "e", "h", "sums"):
        o[x].assert_props()
    # ---- Q6
    h = workload.lineitem_fast(0, n)
    types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng}
    cols = {c: R.bat(h[c], types[c], nonil=True) for c in types}
    r = pipelines.q6(R, cols)
    out["q6"] = str((int(r["hi"][0]) << 64) | int(r["lo"][0]))
    # ---- Q1 (groupavg, cst calcs, byte keys)
    h1 = workload.lineitem_fast(0, n, ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus"))
    t1 = dict(types, tax=T.lng, returnflag=T.bte, linestatus=T.bte)
    q1 = pipelines.q1(R, {c: R.bat(h1[c], t1[c], nonil=True) for c in t1})
    out["q1"] = {kk: dump(vv) for kk, vv in q1.items()}
    # ---- join (pair multiset)
    rng = np.random.default_rng(7)
    rk = rng.permutation(np.arange(1, n // 4 + 1)).astype(np.int32)
    lk = rng.integers(1, n // 4 + 1, n).astype(np.int32)
    r1, r2 = R.BATjoin(R.bat(lk, T.int, nonil=True), R.bat(rk, T.int, nonil=True), None, None, False)
    a1, a2 = r1.numpy(), r2.numpy()
    idx = np.lexsort((a2, a1))
    out["join"] = {"pairs": len(a1), "sha": hashlib.sha256(np.ascontiguousarray(np.stack([a1[idx], a2[idx]])).tobytes()).hexdigest()}
    # ---- errors: texts must be the reference's
    errs = {}
    big = R.bat(np.full(1000, 2**62, dtype=np.int64), T.lng, nonil=True)
    for name, fn in (("mul_overflow", lambda: R.BATcalc("mul", big, big, None, None, T.lng)),
                     ("div_zero", lambda: R.BATcalccst("div", big, 0, T.lng, None, T.lng)),
                     ("project_miss", lambda: R.BATproject(R.bat(np.array([5, 2000], dtype=np.uint64), T.oid), big)),
                     ("sum_overflow", lambda: R.BATsum(T.lng, big, None, True, True)),
                     ("size_mismatch", lambda: R.BATcalc("add", big, c1["v"], None, None, T.lng))):
        try:
            fn()
            errs[name] = "no error"
        except R.RefError as ex:
            errs[name] = str(ex).strip()
    out["errors"] = errs
    # ---- window chain: diff -> ROWS bounds -> sum / avg
    w = workload.window(0, 2 * workload.window_layout())
    m = len(w["pkey"])
    pk, od, vl = R.bat(w["pkey"], T.lng, nonil=True), R.bat(w["ord"], T.lng, nonil=True), R.bat(w["val"], T.lng, nonil=True)
    p = R.GDKanalyticaldiff(R.colnew(T.bit, m), pk, None, None, T.lng)
    s = R.GDKanalyticalwindowbounds(R.colnew(T.oid, m), od, p, None, 100, T.lng, T.lng, 0, True, 0)
    e = R.GDKanalyticalwindowbounds(R.colnew(T.oid, m), od, p, None, 0, T.lng, T.lng, 0, False, 1)
    ws = R.GDKanalyticalsum(R.colnew(T.hge, m), p, None, vl, s, e, T.lng, T.hge, 0)
    wa = R.GDKanalyticalfn("avg", R.colnew(T.dbl, m), p, None, vl, s, e, T.lng, 0)
    out["window"] = {"p": dump(p), "s": dump(s), "e": dump(e), "sum": dump(ws), "avg": dump(wa)}
    # ---- ORDER BY / LIMIT, FK fetch join, DISTINCT, CAST: sort, firstn cascade, leftjoin family, unique, convert
    rng = np.random.default_rng(13)
    m = min(n, 300_000)
    c1 = R.bat(rng.integers(0, 50, m).astype(np.int32), T.int, nonil=True)
    c2 = R.bat(rng.integers(0, 1000, m).astype(np.int64), T.lng, nonil=True)
    so, oo, go = R.BATsort(c1, None, None, False, False, True)
    s2, o2, g2 = R.BATsort(c2, oo, go, True, True, False)
    s1f, g1f = R.BATfirstn(c1, None, 100, True, False, None, False, True)
    s2f, g2f = R.BATfirstn(c2, s1f, 100, True, False, g1f, False, True)
    fk = R.bat(rng.integers(0, m + m // 10, m).astype(np.uint64), T.oid, nonil=True)
    dim = R.dense(0, 0, m)
    j1, j2 = R.BATleftjoin(fk, dim, None, None, False)
    oj1, oj2 = R.BATouterjoin(fk, dim, None, None, False)
    sj1, sj2 = R.BATsemijoin(c1, R.bat(np.arange(10, 30, dtype=np.int32), T.int, nonil=True), None, None, False)
    out["order"] = {"sorted": dump(so), "order": dump(oo), "groups": dump(go), "sorted2": dump(s2), "groups2": dump(g2),
                    "topn1": dump(s1f), "gids1": dump(g1f), "topn2": dump(s2f), "gids2": dump(g2f),
                    "fk1": dump(j1), "fk2": dump(j2), "outer1": dump(oj1), "outer2": dump(oj2), "semi1": dump(sj1),
                    "dangling": dump(R.BATdiff(fk, dim, None, None, False, False)),
                    "unique": dump(R.BATunique(c1, None)),
                    "convert": dump(R.BATconvert(c2, None, T.int, 0, 0, 0)), "decimal": dump(R.BATconvert(c1, None, T.lng, 0, 2, 0))}
    # ---- CASE / BETWEEN / logic / unary / min-max / mod / shifts
    x = R.bat(rng.integers(-1000, 1000, m).astype(np.int32), T.int, nonil=True)
    y = R.bat(rng.integers(1, 50, m).astype(np.int32), T.int, nonil=True)
    lo = R.bat(rng.integers(-500, 0, m).astype(np.int32), T.int, nonil=True)
    hi = R.bat(rng.integers(0, 500, m).astype(np.int32), T.int, nonil=True)
    cond = R.BATcalcbetween(x, lo, hi)
    out["calc"] = {
        "between": dump(cond), "case": dump(R.BATcalcifthenelse(cond, x, y)),
        "mod": dump(R.BATcalc("mod", x, y, None, None, T.int)), "modcst": dump(R.BATcalccst("mod", x, 7, T.int, None, T.int)),
        "and": dump(R.BATcalclogic("and", x, y)), "xorcst": dump(R.BATcalclogiccst("xor", x, 0x55, T.int)),
        "not": dump(R.BATcalcunary("not", cond)), "neg": dump(R.BATcalcunary("negate", x)), "abs": dump(R.BATcalcunary("absolute", x)),
        "sign": dump(R.BATcalcunary("sign", x)), "iszero": dump(R.BATcalcunary("iszero", x)),
        "isnil": dump(R.BATcalcisnil(x)), "min": dump(R.BATcalcminmax("min", x, y)), "maxcst": dump(R.BATcalcminmaxcst("max", x, 3, T.int)),
        "lsh": dump(R.BATcalcshiftcst("lsh", y, 3, T.int)), "rsh": dump(R.BATcalcshift("rsh", y, R.bat(np.full(m, 2, dtype=np.int32), T.int, nonil=True))),
    }
    # ---- dictionary strings: GROUP BY two str flags + equality selects on a str column created by the real GDK
    rng = np.random.default_rng(11)
    ns = min(n, 200_000)
    flags = [b"A", b"N", b"R", b"\x80"]
    rf = R.strbat([flags[i] for i in rng.integers(0, 4, ns)])
    ls = R.strbat([(b"F", b"O")[i] for i in rng.integers(0, 2, ns)])
    g, e, hh = R.BATgroup(rf)
    g, e, hh = R.BATgroup(ls, None, g, e, hh)
    sel = {op + ":" + v.decode("latin1"): dump(R.BATthetaselect(rf, None, v, op))
           for op in ("==", "!=", "eq", "ne") for v in (b"N", b"\x80", b"zz")}
    out["str"] = {"g": dump(g), "e": dump(e), "h": dump(hh), "sel": sel}
    try:
        out["shim_stats"] = list(R.shim_stats())
    except AttributeError:
        out["shim_stats"] = None
    print(json.dumps(out))


if __name__ == "__main__":
    main()
