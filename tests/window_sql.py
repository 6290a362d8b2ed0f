"""The GDK call chain MonetDB's SQL layer emits for  fn(col) OVER ([PARTITION BY x ORDER BY x] ROWS BETWEEN lo
PRECEDING AND hi FOLLOWING)  (sql/backends/monet5/rel_bin.c window projections -> sql_rank.c: BATsort (stable) ->
BATproject -> GDKanalyticaldiff -> GDKanalyticalwindowbounds x2 -> GDKanalytical<fn>), written once against the operator
interface so that the same code runs on the reference driver (`oracle.refgdk`) and on the device mirror
(`monetdb_b200.gdk`).  Used with the golden vectors of tests/golden/analytics02.json."""
import numpy as np

NP = {"int": np.int32, "lng": np.int64}


def column(table, name):
    """(numpy array with the in-band nil, GDK type name) of one column of the fixture's table"""
    idx = table["columns"].index(name)
    tname = table["types"][idx]
    dt = NP[tname]
    nil = np.iinfo(dt).min
    return np.array([nil if r[idx] is None else r[idx] for r in table["rows"]], dtype=dt), tname


def run_case(X, mk, T, table, case):
    """-> the result column as the strings a sqllogictest file holds (NULL, integers, reals with 3 decimals), sorted
    when the case says rowsort.  X: operator module; mk(array, type) -> BAT of that backend; T: its TYPE table.

    case["frame"] == "rows":    ROWS BETWEEN preceding PRECEDING AND following FOLLOWING (frame type 0, bounds columns)
    case["frame"] == "default": no ORDER BY -> the whole partition (frame type 5); ORDER BY -> unbounded preceding to
                                the current row's last peer (frame type 3, peers = rows with the same order value)"""
    n = len(table["rows"])
    vname = case["col"] if case["col"] != "*" else "aa"
    varr, vt = column(table, vname)
    tp = getattr(T, vt)
    vb = mk(varr, tp)
    p, o, ob, otp = None, None, vb, tp
    key = case["partition"] or case.get("order")
    if key:
        karr, kt = column(table, key)
        ktp = getattr(T, kt)
        desc = bool(case.get("desc"))
        # ascending: nils first; descending: nils last (they are the smallest value) — both stable
        srt, order, _ = X.BATsort(mk(karr, ktp), None, None, desc, desc, True)
        vb = X.BATproject(order, vb)
        if case["partition"]:
            p = X.GDKanalyticaldiff(X.colnew(T.bit, n), srt, None, None, ktp)
        if case.get("order") and case["frame"] == "default":
            o = X.GDKanalyticaldiff(X.colnew(T.bit, n), srt, p, None, ktp)      # peers: a new order value or partition
        ob, otp = srt, ktp
    if case["frame"] == "rows":
        s = X.GDKanalyticalwindowbounds(X.colnew(T.oid, n), ob, p, None, case["preceding"], otp, T.lng, 0, True, 0)
        e = X.GDKanalyticalwindowbounds(X.colnew(T.oid, n), ob, p, None, case["following"], otp, T.lng, 0, False, 1)
        ft = 0
    else:
        s = e = None
        ft = 3 if case.get("order") else 5
    fn = case["fn"]
    if fn == "sum":
        r = X.GDKanalyticalsum(X.colnew(T.lng, n), p, o, vb, s, e, tp, T.lng, ft)
        nil = np.iinfo(np.int64).min
    elif fn == "avg":
        r = X.GDKanalyticalfn("avg", X.colnew(T.dbl, n), p, o, vb, s, e, tp, ft)
        nil = None
    elif fn in ("min", "max"):
        r = X.GDKanalyticalfn(fn, X.colnew(tp, n), p, o, vb, s, e, tp, ft)
        nil = np.iinfo(NP[vt]).min
    else:
        r = X.GDKanalyticalcount(X.colnew(T.lng, n), p, o, vb, s, e, case["col"] != "*", tp, ft)
        nil = np.iinfo(np.int64).min
    out = []
    for v in r.numpy()[:n]:
        if fn == "avg":
            out.append("NULL" if np.isnan(v) else "%.3f" % float(v))
        else:
            out.append("NULL" if int(v) == nil else str(int(v)))
    return sorted(out) if case["sort"] == "rowsort" else out
