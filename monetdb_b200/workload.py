"""Synthetic TPC-H-shaped columns: splitmix64, one stream per
column, seed = 0x5EED0000 + column index, value = f(z of row i) so that any row
range can be generated independently on any GPU or host."""
from __future__ import annotations

import numpy as np

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed: int, lo: int, hi: int) -> np.ndarray:
    """z values of rows [lo, hi) of the stream with the given seed"""
    with np.errstate(over="ignore"):
        i = np.arange(lo + 1, hi + 1, dtype=np.uint64)
        z = np.uint64(seed) + i * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def col_seed(idx: int) -> int:
    return 0x5EED0000 + idx


def todate(year: int, month: int, day: int) -> int:
    """MonetDB's packed date (gdk_time.c:17-28): ((y+4712)*12 + m-1) << 5 | d"""
    return (((year + 4712) * 12 + (month - 1)) << 5) | day


_MDAYS = [31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31]


def _date_table(ndays: int = 2526) -> np.ndarray:
    """packed dates of the ndays days starting 1992-01-02"""
    out = np.empty(ndays, dtype=np.int32)
    y, m, d = 1992, 1, 2
    for i in range(ndays):
        out[i] = todate(y, m, d)
        d += 1
        leap = y % 4 == 0 and (y % 100 != 0 or y % 400 == 0)
        md = _MDAYS[m - 1] + (1 if (m == 2 and leap) else 0)
        if d > md:
            d = 1
            m += 1
            if m > 12:
                m = 1
                y += 1
    return out


DATES = _date_table()
# column indices (seed offsets)
C_SHIPDATE, C_DISCOUNT, C_QUANTITY, C_EXTPRICE, C_TAX, C_RETFLAG, C_K, C_V, C_LKEY, C_OKEY = range(10)


LINEITEM_DTYPES = {"shipdate": np.int32, "discount": np.int64, "quantity": np.int64, "extprice": np.int64,
                   "tax": np.int64, "returnflag": np.int8, "linestatus": np.int8}
CUT_LS = todate(1995, 6, 17)
CUT_BAND = todate(1995, 6, 10)


def _nil(z, vals, nils: bool, nil):
    """the 1 %-nil variant of SURVEY.md §8(d): a value is nil where its own z % 100 == 0"""
    if nils:
        vals = vals.copy()
        vals[z % np.uint64(100) == 0] = nil
    return vals


def lineitem(lo: int, hi: int, cols=("shipdate", "discount", "quantity", "extprice"), nils: bool = False) -> dict:
    """rows [lo,hi) of the synthetic lineitem table (configs 2 and 3); numpy form = the definition"""
    out = {}
    I64NIL, I32NIL = np.iinfo(np.int64).min, np.iinfo(np.int32).min
    if "shipdate" in cols or "returnflag" in cols or "linestatus" in cols:
        z = splitmix64(col_seed(C_SHIPDATE), lo, hi)
        day = (z % np.uint64(2526)).astype(np.int64)
        ship = DATES[day]
        if "shipdate" in cols:
            out["shipdate"] = _nil(z, ship, nils, I32NIL)
        if "returnflag" in cols or "linestatus" in cols:
            ls = (ship > CUT_LS).astype(np.int8)
            z2 = splitmix64(col_seed(C_RETFLAG), lo, hi)
            # A/F (0,0), R/F (2,0) before the cut; N/O (1,1) after; a thin N/F band (1,0) just before it
            band = (ship > CUT_BAND) & (ls == 0)
            rf = np.where(ls == 1, 1, np.where(band, 1, np.where(z2 % np.uint64(2) == 1, 0, 2))).astype(np.int8)
            out["returnflag"] = rf
            out["linestatus"] = ls
    if "discount" in cols:
        z = splitmix64(col_seed(C_DISCOUNT), lo, hi)
        out["discount"] = _nil(z, (z % np.uint64(11)).astype(np.int64), nils, I64NIL)
    if "quantity" in cols:
        z = splitmix64(col_seed(C_QUANTITY), lo, hi)
        out["quantity"] = _nil(z, ((z % np.uint64(50) + np.uint64(1)) * np.uint64(100)).astype(np.int64), nils, I64NIL)
    if "extprice" in cols:
        z = splitmix64(col_seed(C_EXTPRICE), lo, hi)
        out["extprice"] = _nil(z, (z % np.uint64(10400001) + np.uint64(90000)).astype(np.int64), nils, I64NIL)
    if "tax" in cols:
        z = splitmix64(col_seed(C_TAX), lo, hi)
        out["tax"] = _nil(z, (z % np.uint64(9)).astype(np.int64), nils, I64NIL)
    return out


_WL = None


def _wl():
    """libworkload.so (hostsrc/workload_host.c): the same generator in threaded C"""
    global _WL
    if _WL is None:
        import ctypes as C
        from pathlib import Path
        so = Path(__file__).resolve().parent / "libworkload.so"
        if not so.exists():
            from .build import build_workload
            build_workload()
        lib = C.CDLL(str(so))
        lib.wl_window.restype = C.c_int
        lib.wl_window.argtypes = [C.c_uint64, C.c_uint64, C.c_int] + [C.c_uint64] * 3 + [C.c_void_p] * 3
        lib.wl_lineitem.restype = C.c_int
        lib.wl_lineitem.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int32, C.c_int32] + \
            [C.c_void_p] * 7
        _WL = lib
    return _WL


def lineitem_into(lo: int, hi: int, out: dict, nils: bool = False, threads: int | None = None) -> dict:
    """fill caller-provided arrays (e.g. page-locked staging memory) with rows [lo,hi); `out` maps
    column name -> numpy array of hi-lo items of LINEITEM_DTYPES[name]"""
    import os
    order = ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus")
    ptrs = []
    for k in order:
        a = out.get(k)
        if a is None:
            ptrs.append(None)
            continue
        assert a.dtype == LINEITEM_DTYPES[k] and len(a) == hi - lo and a.flags.c_contiguous, k
        ptrs.append(a.ctypes.data)
    dates = np.ascontiguousarray(DATES, dtype=np.int32)
    _wl().wl_lineitem(lo, hi, threads or min(64, os.cpu_count() or 1), 1 if nils else 0, dates.ctypes.data, len(dates),
                      CUT_LS, CUT_BAND, *ptrs)
    return out


def lineitem_fast(lo: int, hi: int, cols=("shipdate", "discount", "quantity", "extprice"), nils: bool = False) -> dict:
    """lineitem() through the threaded C generator (bit-identical; tests/test_cpu_shard.py checks it)"""
    out = {k: np.empty(hi - lo, dtype=LINEITEM_DTYPES[k]) for k in cols}
    return lineitem_into(lo, hi, out, nils)


def config1(lo: int, hi: int):
    """config 1: k:int = z%1000, v:lng = (z>>32)%100000"""
    zk = splitmix64(col_seed(C_K), lo, hi)
    zv = splitmix64(col_seed(C_V), lo, hi)
    return (zk % np.uint64(1000)).astype(np.int32), ((zv >> np.uint64(32)) % np.uint64(100000)).astype(np.int64)


C_WPERM, C_WORD, C_WVAL = 10, 11, 12
_PERMS = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]


def window_layout(lens=(1000, 10000, 100000)):
    """config 5 partitioning: rows come in blocks of sum(lens) rows; block k holds one partition of
    each length, in the order given by permutation z(k) % 6 — so the partition of any row can be
    computed without generating the rows before it"""
    return sum(lens)


def window(lo: int, hi: int, lens=(1000, 10000, 100000)) -> dict:
    """rows [lo,hi) of the window-analytics table (config 5): pkey (lng, ascending), ord (lng,
    ascending within a partition with duplicates: prev + z%3), val (lng = z%10000 - 5000).
    lo must be a multiple of sum(lens) (shards start on block = partition boundaries)."""
    blk = sum(lens)
    assert lo % blk == 0, "window shards must start on a block boundary"
    n = hi - lo
    b0, b1 = lo // blk, (hi + blk - 1) // blk
    zperm = splitmix64(col_seed(C_WPERM), b0, b1) % np.uint64(6)
    starts = np.empty((b1 - b0) * 3, dtype=np.int64)     # partition start rows (absolute)
    for i, zp in enumerate(zperm):
        off = (b0 + i) * blk
        for j, which in enumerate(_PERMS[int(zp)]):
            starts[3 * i + j] = off
            off += lens[which]
    rows = np.arange(lo, hi, dtype=np.int64)
    pidx = np.searchsorted(starts, rows, side="right") - 1
    pkey = (3 * b0 + pidx).astype(np.int64)
    inc = (splitmix64(col_seed(C_WORD), lo, hi) % np.uint64(3)).astype(np.int64)
    c = np.cumsum(inc)
    pstart = starts[pidx] - lo
    ordv = c - c[pstart]
    val = (splitmix64(col_seed(C_WVAL), lo, hi) % np.uint64(10000)).astype(np.int64) - 5000
    return {"pkey": pkey, "ord": ordv, "val": val, "n": n}


def window_fast(lo: int, hi: int, lens=(1000, 10000, 100000), threads: int | None = None) -> dict:
    """window() through the threaded C generator (bit-identical; tests/test_cpu_shard.py checks it)"""
    import os
    n = hi - lo
    out = {k: np.empty(n, dtype=np.int64) for k in ("pkey", "ord", "val")}
    rc = _wl().wl_window(lo, hi, threads or min(64, os.cpu_count() or 1), *lens, out["pkey"].ctypes.data, out["ord"].ctypes.data,
                         out["val"].ctypes.data)
    assert rc == 0, "window shards must start on a block boundary"
    out["n"] = n
    return out
