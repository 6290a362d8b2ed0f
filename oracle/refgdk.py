"""TEST INFRASTRUCTURE ONLY — ctypes driver of the REAL reference GDK.

``oracle/_ref/libbat_ref.so`` is the reference's own gdk/*.c compiled by
``oracle/Makefile`` (target ``ref``) from the sources under /root/reference;
``libref_harness.so`` adds the few helpers of ``ref_harness.c``.  This module lets
the tests, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs call the reference's entry points (BATselect,
BATproject, BATgroup, BATgroupsum, BATjoin, GDKanalytical*, ...) on in-memory
BATs built from numpy arrays.  Nothing in ``monetdb_b200/`` may import it.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent / "_ref"
LIBBAT = _DIR / "libbat_ref.so"
LIBH = _DIR / "libref_harness.so"


def available() -> bool:
    return LIBBAT.exists() and LIBH.exists()


class T:
    void, msk, bit, bte, sht, int, oid, ptr, flt, dbl, lng, hge, date, daytime, timestamp = range(15)


hge_dtype = np.dtype([("lo", "<u8"), ("hi", "<i8")])
NP_OF = {
    T.bit: np.dtype(np.int8), T.bte: np.dtype(np.int8), T.sht: np.dtype(np.int16),
    T.int: np.dtype(np.int32), T.oid: np.dtype(np.uint64), T.flt: np.dtype(np.float32),
    T.dbl: np.dtype(np.float64), T.lng: np.dtype(np.int64), T.hge: hge_dtype,
    T.date: np.dtype(np.int32), T.daytime: np.dtype(np.int64), T.timestamp: np.dtype(np.int64),
    T.void: np.dtype(np.uint64), T.msk: np.dtype(np.uint32),
}
oid_nil = 1 << 63
BUN_NONE = (1 << 63) - 1


class _Info(C.Structure):
    _fields_ = [("type", C.c_int32), ("ccand", C.c_int32), ("hseq", C.c_uint64), ("count", C.c_uint64),
                ("tseq", C.c_uint64), ("sorted", C.c_uint8), ("revsorted", C.c_uint8), ("key", C.c_uint8),
                ("nonil", C.c_uint8), ("nil", C.c_uint8), ("pad", C.c_uint8 * 3)]


_bat = None
_h = None


class RefError(RuntimeError):
    pass


class _Chain:
    """attribute lookup through several CDLLs in order"""

    def __init__(self, libs):
        self._libs = libs

    def __getattr__(self, name):
        for lib in self._libs:
            try:
                f = getattr(lib, name)
            except AttributeError:
                continue
            setattr(self, name, f)
            return f
        raise AttributeError(name)


def shim_stats():
    """(device calls, forwarded calls, inputs served by a device twin, inputs staged) of the loaded shim"""
    out = (C.c_ulong * 4)()
    _bat.b200shim_stats(out)
    return tuple(out)


def _load(nthreads: int = 0):
    global _bat, _h
    if _h is not None:
        return
    if not available():
        raise RefError(f"reference GDK not built: run `make -C oracle ref` (needs /root/reference); looked in {_DIR}")
    shim = os.environ.get("REFGDK_SHIM")
    if shim:
        # boundary test (tests/test_gpu_shim.py): integration/_build/libbat_b200shim.so is loaded FIRST and
        # globally, so the GDK entry points it exports interpose the reference's for the harness (PLT lookups)
        # and for this driver (names are resolved in the shim first); everything else comes from libbat_ref
        _shimlib = C.CDLL(shim, mode=C.RTLD_GLOBAL)
        _bat = _Chain([_shimlib, C.CDLL(str(LIBBAT), mode=C.RTLD_GLOBAL)])
    else:
        _bat = C.CDLL(str(LIBBAT), mode=C.RTLD_GLOBAL)
    _h = C.CDLL(str(LIBH), mode=C.RTLD_GLOBAL)
    V, P = C.c_void_p, C.c_void_p
    _h.refh_init.restype = C.c_int
    _h.refh_init.argtypes = [C.c_int]
    _h.refh_errbuf.restype = C.c_char_p
    _h.refh_algorithm.restype = C.c_char_p
    _h.refh_bat_new.restype = P
    _h.refh_bat_new.argtypes = [C.c_int, C.c_uint64, C.c_uint64, V] + [C.c_int] * 5
    _h.refh_msk_new.restype = P
    _h.refh_msk_new.argtypes = [C.c_uint64, C.c_uint64, V]
    _h.refh_dense.restype = P
    _h.refh_dense.argtypes = [C.c_uint64] * 3
    _h.refh_tloc.restype = V
    _h.refh_tloc.argtypes = [P]
    _h.refh_info.argtypes = [P, C.POINTER(_Info)]
    _h.refh_tail.argtypes = [P, V]
    _h.refh_free.argtypes = [P]
    _h.refh_assert_props.argtypes = [P]
    _h.refh_calc_cst.restype = P
    _h.refh_calc_cst.argtypes = [C.c_int, C.c_int, P, V, C.c_int, P, C.c_int]
    _h.refh_logic_cst.restype = P
    _h.refh_logic_cst.argtypes = [C.c_int, P, V, C.c_int, P]
    _h.refh_minmax_cst.restype = P
    _h.refh_minmax_cst.argtypes = [C.c_int, P, V, C.c_int, P]
    _h.refh_cst_cmp.restype = P
    _h.refh_cst_cmp.argtypes = [C.c_int, V, C.c_int, P, P, C.c_int]
    _h.refh_cmp_cst.restype = P
    _h.refh_cmp_cst.argtypes = [C.c_int, P, V, C.c_int, P, C.c_int]
    _h.refh_shift_cst.restype = P
    _h.refh_shift_cst.argtypes = [C.c_int, C.c_int, P, V, C.c_int, P]
    _h.refh_usec.restype = C.c_int64
    rc = _h.refh_init(nthreads)
    if rc != 0:
        raise RefError(f"GDKinit failed ({rc})")
    B = C.c_bool
    sigs = {
        "BATselect": (P, [P, P, V, V, B, B, B, B]),
        "BATthetaselect": (P, [P, P, V, C.c_char_p]),
        "BATproject": (P, [P, P]),
        "BATproject2": (P, [P, P, P]),
        "BATmergecand": (P, [P, P]),
        "BATintersectcand": (P, [P, P]),
        "BATdiffcand": (P, [P, P]),
        "BATnegcands": (P, [C.c_uint64, C.c_uint64, P]),
        "BATmaskedcands": (P, [C.c_uint64, C.c_uint64, P, B]),
        "BATgroup": (C.c_int, [C.POINTER(P), C.POINTER(P), C.POINTER(P), P, P, P, P, P]),
        "BATgroupsum": (P, [P, P, P, P, C.c_int, B]),
        "BATgroupprod": (P, [P, P, P, P, C.c_int, B]),
        "BATgroupcount": (P, [P, P, P, P, C.c_int, B]),
        "BATgroupmin": (P, [P, P, P, P, C.c_int, B]),
        "BATgroupmax": (P, [P, P, P, P, C.c_int, B]),
        "BATgroupavg": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, P, C.c_int, B, C.c_int]),
        "BATsum": (C.c_int, [V, C.c_int, P, P, B, B]),
        "BATprod": (C.c_int, [V, C.c_int, P, P, B, B]),
        "BATgroupavg3": (C.c_int, [C.POINTER(P), C.POINTER(P), C.POINTER(P), P, P, P, P, B]),
        "BATgroupavg3combine": (P, [P, P, P, P, P, B]),
        "BATcalcavg": (C.c_int, [P, P, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int]),
        "BATjoin": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, P, B, C.c_uint64]),
        "BAThash": (C.c_int, [P]),
        "BATcalcnot": (P, [P, P]),
        "BATcalcnegate": (P, [P, P]),
        "BATcalcabsolute": (P, [P, P]),
        "BATcalciszero": (P, [P, P]),
        "BATcalcsign": (P, [P, P]),
        "BATcalcmin": (P, [P, P, P, P]),
        "BATcalcmin_no_nil": (P, [P, P, P, P]),
        "BATcalcmax": (P, [P, P, P, P]),
        "BATcalcmax_no_nil": (P, [P, P, P, P]),
        "BATcalcbetween": (P, [P, P, P, P, P, P, B, B, B, B, B]),
        "BATcalcifthenelse": (P, [P, P, P]),
        "BATleftjoin": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, P, B, C.c_uint64]),
        "BATouterjoin": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, P, B, B, C.c_uint64]),
        "BATsemijoin": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, P, B, B, C.c_uint64]),
        "BATintersect": (P, [P, P, P, P, B, B, C.c_uint64]),
        "BATdiff": (P, [P, P, P, P, B, B, C.c_uint64]),
        "BATunique": (P, [P, P]),
        "BATconvert": (P, [P, P, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8]),
        "BATcalcisnil": (P, [P, P]),
        "BATcalcisnotnil": (P, [P, P]),
        "GDKanalyticaldiff": (C.c_int, [P, P, P, V, C.c_int]),
        "GDKanalyticalwindowbounds": (C.c_int, [P, P, P, P, V, C.c_int, C.c_int, C.c_int, B, C.c_uint64]),
        "GDKanalyticalsum": (C.c_int, [P, P, P, P, P, P, C.c_int, C.c_int, C.c_int]),
        "GDKanalyticalavg": (C.c_int, [P, P, P, P, P, P, C.c_int, C.c_int]),
        "GDKanalyticalavginteger": (C.c_int, [P, P, P, P, P, P, C.c_int, C.c_int]),
        "GDKanalyticalcount": (C.c_int, [P, P, P, P, P, P, C.c_int8, C.c_int, C.c_int]),
        "GDKanalyticalmin": (C.c_int, [P, P, P, P, P, P, C.c_int, C.c_int]),
        "GDKanalyticalmax": (C.c_int, [P, P, P, P, P, P, C.c_int, C.c_int]),
        "COLnew": (P, [C.c_uint64, C.c_int, C.c_uint64, C.c_int]),
        "BATsort": (C.c_int, [C.POINTER(P), C.POINTER(P), C.POINTER(P), P, P, P, B, B, B]),
        "BATfirstn": (C.c_int, [C.POINTER(P), C.POINTER(P), P, P, P, C.c_uint64, B, B, B]),
    }
    for op in ("add", "sub", "mul", "div", "mod"):
        sigs["BATcalc" + op] = (P, [P, P, P, P, C.c_int])
    for op in ("lt", "le", "gt", "ge", "and", "or", "xor", "lsh", "rsh"):
        sigs["BATcalc" + op] = (P, [P, P, P, P])
    for op in ("eq", "ne"):
        sigs["BATcalc" + op] = (P, [P, P, P, P, B])
    for name, (res, args) in sigs.items():
        f = getattr(_bat, name)
        f.restype = res
        f.argtypes = args


def init(nthreads: int = 0):
    _load(nthreads)


def errbuf() -> str:
    """the GDKerror text of the last failure, without the tracer's "!ERROR: function: " prefix"""
    raw = (_h.refh_errbuf() or b"").decode()
    lines = [ln for ln in raw.splitlines() if ln.strip()]
    if not lines:
        return ""
    msg = lines[-1]
    if msg.startswith("!ERROR: "):
        msg = msg[len("!ERROR: "):]
        head, sep, tail = msg.partition(": ")
        if sep and " " not in head:
            msg = tail
    return msg


def clrerr():
    _h.refh_clrerr()


def algorithm() -> str:
    return (_h.refh_algorithm() or b"").decode()


def usec() -> int:
    return int(_h.refh_usec())


class RBAT:
    """a reference BAT* (owned)"""

    def __init__(self, p, what="BAT"):
        if not p:
            msg = errbuf().strip()
            clrerr()
            raise RefError(msg or f"{what}: NULL")
        self.p = p

    def free(self):
        if self.p:
            _h.refh_free(self.p)
            self.p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def info(self) -> _Info:
        i = _Info()
        _h.refh_info(self.p, C.byref(i))
        return i

    def props(self) -> dict:
        i = self.info()
        return dict(type=i.type, hseq=i.hseq, count=i.count, tseq=i.tseq, sorted=bool(i.sorted),
                    revsorted=bool(i.revsorted), key=bool(i.key), nonil=bool(i.nonil), nil=bool(i.nil),
                    ccand=i.ccand)

    def __len__(self):
        return int(self.info().count)

    def numpy(self) -> np.ndarray:
        i = self.info()
        if i.type == T.void:
            out = np.empty(i.count, dtype=np.uint64)
        elif i.type == T.msk:
            out = np.empty((i.count + 31) // 32, dtype=np.uint32)
        else:
            out = np.empty(i.count, dtype=NP_OF[i.type])
        if i.count:
            _h.refh_tail(self.p, out.ctypes.data_as(C.c_void_p))
        return out

    def assert_props(self):
        _h.refh_assert_props(self.p)


def _p(b):
    return b.p if b is not None else None


def _vp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _val(tt, v):
    if v is None:
        return None
    st = {T.bit: T.bte, T.date: T.int, T.daytime: T.lng, T.timestamp: T.lng}.get(tt, tt)
    if st == T.hge:
        a = np.empty(1, dtype=hge_dtype)
        a["lo"][0] = int(v) & 0xFFFFFFFFFFFFFFFF
        a["hi"][0] = int(v) >> 64
        return a
    if st in (T.flt, T.dbl):
        return np.array([v], dtype=NP_OF[st])
    if st in (T.oid, T.void):
        return np.array([int(v)], dtype=np.uint64)
    return np.array([int(v)], dtype=np.int64).astype(NP_OF[st])


def bat(arr: np.ndarray, tt: int, hseq: int = 0, sorted=False, revsorted=False, key=False,
        nonil=False, nil=False) -> RBAT:
    _load()
    arr = np.ascontiguousarray(arr)
    assert arr.dtype.itemsize == NP_OF[tt].itemsize, (arr.dtype, tt)
    return RBAT(_h.refh_bat_new(tt, len(arr), hseq, _vp(arr), sorted, revsorted, key, nonil, nil))


def bat_uninit(tt: int, n: int, hseq: int = 0, nonil=False, nil=False):
    """(BAT of n uninitialised items, numpy view of its tail heap): lets a generator write straight into the
    GDK heap instead of going through a second host copy (bench.py's full-table reference arm)"""
    _load()
    b = RBAT(_h.refh_bat_new(tt, n, hseq, None, False, False, False, nonil, nil))
    dt = NP_OF[tt]
    _h.refh_tloc.restype = C.c_void_p
    addr = _h.refh_tloc(b.p)
    view = np.frombuffer((C.c_char * (n * dt.itemsize)).from_address(addr), dtype=dt, count=n) if n else np.empty(0, dt)
    return b, view


def dense(hseq, tseq, cnt) -> RBAT:
    _load()
    return RBAT(_h.refh_dense(hseq, tseq, cnt))


def colnew(tt: int, cap: int, hseq: int = 0) -> RBAT:
    """caller-allocated result BAT for the GDKanalytical* functions (sql_rank.c:161)"""
    _load()
    return RBAT(_bat.COLnew(hseq, tt, cap, 1))


def negcands(seq, cnt, exceptions) -> RBAT:
    _load()
    e = np.ascontiguousarray(exceptions, dtype=np.uint64)
    eb = bat(e, T.oid, sorted=True, key=True, nonil=True)
    return RBAT(_bat.BATnegcands(seq, cnt, eb.p))


def maskedcands(hseq, nrows, words, selected=True) -> RBAT:
    _load()
    w = np.ascontiguousarray(words, dtype=np.uint32)
    m = RBAT(_h.refh_msk_new(hseq, nrows, _vp(w)))
    return RBAT(_bat.BATmaskedcands(hseq, nrows, m.p, selected))


def BATunmask(b) -> RBAT:
    _load()
    if not hasattr(_bat.BATunmask, "_set"):
        _bat.BATunmask.restype = C.c_void_p
        _bat.BATunmask.argtypes = [C.c_void_p]
        _bat.BATunmask._set = True
    return RBAT(_bat.BATunmask(b.p), "BATunmask")


def bincopy_decode(kind: int, raw: bytes, n: int) -> np.ndarray:
    """kind 0 date / 1 daytime / 2 timestamp records decoded with libbat's date_create / daytime_create /
    timestamp_create (the loops of sql_bincopyconvert.c:168-280)"""
    _load()
    src = np.frombuffer(raw, dtype=np.uint8)
    out = np.empty(n, dtype=np.int32 if kind == 0 else np.int64)
    fn = (_h.refh_bincopy_date, _h.refh_bincopy_time, _h.refh_bincopy_timestamp)[kind]
    fn(out.ctypes.data_as(C.c_void_p), src.ctypes.data_as(C.c_void_p), C.c_uint64(n))
    return out


def bincopy_encode(kind: int, vals: np.ndarray) -> bytes:
    _load()
    vals = np.ascontiguousarray(vals)
    out = np.empty(len(vals) * (4, 8, 12)[kind], dtype=np.uint8)
    _h.refh_bincopy_encode(C.c_int(kind), out.ctypes.data_as(C.c_void_p), vals.ctypes.data_as(C.c_void_p), C.c_uint64(len(vals)))
    return out.tobytes()


STR = 16
str_nil = b"\x80"
_STR_OFF = {1: np.dtype(np.uint8), 2: np.dtype(np.uint16), 4: np.dtype(np.uint32), 8: np.dtype(np.uint64)}


def strbat(values, hseq: int = 0) -> RBAT:
    """a TYPE_str BAT built with BUNappend from bytes values (b"\\x80" = nil)"""
    _load()
    packed = b"".join(v + b"\0" for v in values)
    _h.refh_str_new.restype = C.c_void_p
    _h.refh_str_new.argtypes = [C.c_uint64, C.c_uint64, C.c_char_p]
    return RBAT(_h.refh_str_new(hseq, len(values), packed), "strbat")


def str_parts(b: RBAT):
    """(offsets as stored, heap bytes, heap is duplicate-free) of a str BAT"""
    w, free, elim = C.c_int(), C.c_uint64(), C.c_int()
    _h.refh_str_parts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    _h.refh_str_parts(b.p, C.byref(w), C.byref(free), C.byref(elim))
    n = len(b)
    off = np.empty(n, dtype=_STR_OFF[w.value])
    if n:
        _h.refh_tail(b.p, off.ctypes.data_as(C.c_void_p))
    heap = C.create_string_buffer(free.value + 1)
    _h.refh_str_heap.argtypes = [C.c_void_p, C.c_void_p]
    _h.refh_str_heap(b.p, heap)
    return off, heap.raw[:free.value], bool(elim.value)


def str_values(b: RBAT) -> list:
    off, heap, _ = str_parts(b)
    add = 8192 if off.dtype.itemsize <= 2 else 0
    return [heap[o + add:heap.index(b"\0", o + add)] for o in off.tolist()]


def BATselect(b, s, tl, th, li, hi, anti, nil_matches=False) -> RBAT:
    tt = b.info().type
    if tt == STR:
        return RBAT(_bat.BATselect(b.p, _p(s), tl, th, li, hi, anti, nil_matches), "BATselect")
    vl, vh = _val(tt, tl), _val(tt, th)
    return RBAT(_bat.BATselect(b.p, _p(s), _vp(vl), _vp(vh), li, hi, anti, nil_matches), "BATselect")


def BATthetaselect(b, s, val, op) -> RBAT:
    if b.info().type == STR:
        return RBAT(_bat.BATthetaselect(b.p, _p(s), val, op.encode()), "BATthetaselect")
    v = _val(b.info().type, val)
    return RBAT(_bat.BATthetaselect(b.p, _p(s), _vp(v), op.encode()), "BATthetaselect")


def BATproject(l, r) -> RBAT:
    return RBAT(_bat.BATproject(l.p, r.p), "BATproject")


def BATproject2(l, r1, r2) -> RBAT:
    return RBAT(_bat.BATproject2(l.p, r1.p, _p(r2)), "BATproject2")


def BATmergecand(a, b) -> RBAT:
    return RBAT(_bat.BATmergecand(a.p, b.p))


def BATintersectcand(a, b) -> RBAT:
    return RBAT(_bat.BATintersectcand(a.p, b.p))


def BATdiffcand(a, b) -> RBAT:
    return RBAT(_bat.BATdiffcand(a.p, b.p))


_OPS = {"add": 0, "sub": 1, "mul": 2, "div": 3, "mod": 4}
_CMPS = {"lt": 0, "le": 1, "gt": 2, "ge": 3, "eq": 4, "ne": 5}


def BATcalc(op, b1, b2, s1, s2, tp) -> RBAT:
    return RBAT(getattr(_bat, "BATcalc" + op)(b1.p, b2.p, _p(s1), _p(s2), tp), "BATcalc" + op)


def BATcalccst(op, b, v, vtype, s, tp, cst_first=False) -> RBAT:
    a = _val(vtype, v)
    return RBAT(_h.refh_calc_cst(_OPS[op], 1 if cst_first else 0, b.p, _vp(a), vtype, _p(s), tp), "BATcalc" + op + "cst")


def BATcalccmp(op, b1, b2, s1=None, s2=None, nil_matches=False) -> RBAT:
    f = getattr(_bat, "BATcalc" + op)
    if op in ("eq", "ne"):
        return RBAT(f(b1.p, b2.p, _p(s1), _p(s2), nil_matches), "BATcalc" + op)
    return RBAT(f(b1.p, b2.p, _p(s1), _p(s2)), "BATcalc" + op)


def BATcalccmpcst(op, b, v, vtype, s=None, nil_matches=False, cst_first=False) -> RBAT:
    a = _val(vtype, v)
    if cst_first:
        return RBAT(_h.refh_cst_cmp(_CMPS[op], _vp(a), vtype, b.p, _p(s), nil_matches), "BATcalccst" + op)
    return RBAT(_h.refh_cmp_cst(_CMPS[op], b.p, _vp(a), vtype, _p(s), nil_matches), "BATcalc" + op + "cst")


def BATcalclogic(op, b1, b2, s1=None, s2=None) -> RBAT:
    return RBAT(getattr(_bat, "BATcalc" + op)(b1.p, b2.p, _p(s1), _p(s2)), "BATcalc" + op)


def BATcalclogiccst(op, b, v, vtype, s=None) -> RBAT:
    a = _val(vtype, v)
    return RBAT(_h.refh_logic_cst({"and": 0, "or": 1, "xor": 2}[op], b.p, _vp(a), vtype, _p(s)), "BATcalc" + op + "cst")


def BATcalcshift(op, b1, b2, s1=None, s2=None) -> RBAT:
    return RBAT(getattr(_bat, "BATcalc" + op)(b1.p, b2.p, _p(s1), _p(s2)), "BATcalc" + op)


def BATcalcshiftcst(op, b, v, vtype, s=None, cst_first=False) -> RBAT:
    a = _val(vtype, v)
    return RBAT(_h.refh_shift_cst(0 if op == "lsh" else 1, 1 if cst_first else 0, b.p, _vp(a), vtype, _p(s)), "BATcalc" + op + "cst")


def BATcalcunary(which, b, s=None) -> RBAT:
    return RBAT(getattr(_bat, "BATcalc" + which)(b.p, _p(s)), "BATcalc" + which)


def BATcalcminmaxcst(which, b, v, vtype, s=None) -> RBAT:
    a = _val(vtype, v)
    w = {"min": 0, "min_no_nil": 1, "max": 2, "max_no_nil": 3}[which]
    return RBAT(_h.refh_minmax_cst(w, b.p, _vp(a), vtype, _p(s)), "BATcalc" + which + "cst")


def BATcalcminmax(which, b1, b2, s1=None, s2=None) -> RBAT:
    return RBAT(getattr(_bat, "BATcalc" + which)(b1.p, b2.p, _p(s1), _p(s2)), "BATcalc" + which)


def BATcalcbetween(b, lo, hi, s=None, slo=None, shi=None, symmetric=False, linc=True, hinc=True, nils_false=False,
                   anti=False) -> RBAT:
    return RBAT(_bat.BATcalcbetween(b.p, lo.p, hi.p, _p(s), _p(slo), _p(shi), symmetric, linc, hinc, nils_false, anti),
                "BATcalcbetween")


def BATcalcifthenelse(b, b1, b2) -> RBAT:
    return RBAT(_bat.BATcalcifthenelse(b.p, b1.p, b2.p), "BATcalcifthenelse")


def BATcalcisnil(b, s=None) -> RBAT:
    return RBAT(_bat.BATcalcisnil(b.p, _p(s)))


def BATcalcisnotnil(b, s=None) -> RBAT:
    return RBAT(_bat.BATcalcisnotnil(b.p, _p(s)))


def BATconvert(b, s, tp, scale1=0, scale2=0, precision=0) -> RBAT:
    return RBAT(_bat.BATconvert(b.p, _p(s), tp, scale1, scale2, precision), "BATconvert")


def BATgroup(b, s=None, g=None, e=None, h=None):
    pg, pe, ph = C.c_void_p(), C.c_void_p(), C.c_void_p()
    rc = _bat.BATgroup(C.byref(pg), C.byref(pe), C.byref(ph), b.p, _p(s), _p(g), _p(e), _p(h))
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATgroup failed")
    return RBAT(pg.value), RBAT(pe.value), RBAT(ph.value)


def _grp(name, b, g, e, s, tp, skip_nils) -> RBAT:
    return RBAT(getattr(_bat, name)(b.p, _p(g), _p(e), _p(s), tp, skip_nils), name)


def BATgroupsum(b, g, e, s, tp, skip_nils=True) -> RBAT:
    return _grp("BATgroupsum", b, g, e, s, tp, skip_nils)


def BATgroupprod(b, g, e, s, tp, skip_nils=True) -> RBAT:
    return _grp("BATgroupprod", b, g, e, s, tp, skip_nils)


def BATgroupcount(b, g, e, s, tp=T.lng, skip_nils=True) -> RBAT:
    return _grp("BATgroupcount", b, g, e, s, tp, skip_nils)


def BATgroupmin(b, g, e, s, tp=T.oid, skip_nils=True) -> RBAT:
    return _grp("BATgroupmin", b, g, e, s, tp, skip_nils)


def BATgroupmax(b, g, e, s, tp=T.oid, skip_nils=True) -> RBAT:
    return _grp("BATgroupmax", b, g, e, s, tp, skip_nils)


def BATgroupavg(b, g, e, s, tp=T.dbl, skip_nils=True, scale=0, want_counts=True):
    pa, pc = C.c_void_p(), C.c_void_p()
    rc = _bat.BATgroupavg(C.byref(pa), C.byref(pc) if want_counts else None, b.p, _p(g), _p(e), _p(s), tp, skip_nils, scale)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATgroupavg failed")
    return RBAT(pa.value), (RBAT(pc.value) if want_counts else None)


def BATgroupavg3(b, g, e, s, skip_nils=True):
    pa, pr, pc = C.c_void_p(), C.c_void_p(), C.c_void_p()
    rc = _bat.BATgroupavg3(C.byref(pa), C.byref(pr), C.byref(pc), b.p, _p(g), _p(e), _p(s), skip_nils)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATgroupavg3 failed")
    return RBAT(pa.value), RBAT(pr.value), RBAT(pc.value)


def BATgroupavg3combine(avg, rem, cnt, g, e, skip_nils=True):
    r = _bat.BATgroupavg3combine(avg.p, rem.p, cnt.p, _p(g), _p(e), skip_nils)
    if not r:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATgroupavg3combine failed")
    return RBAT(r)


def BATcalcavg(b, s=None, scale=0):
    avg, n = C.c_double(), C.c_uint64()
    if _bat.BATcalcavg(b.p, _p(s), C.byref(avg), C.byref(n), scale) != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATcalcavg failed")
    return avg.value, n.value


def BATcalcavgsum(sums, cnts, scale=0) -> RBAT:
    """phase 2 of the two-phase average: exact Python integers, then BATgroupavg's closing formula
    floor(S/n) + (dbl) rem / n, / 10^scale (gdk_aggr.c:1741-1748,1952-1957).  The reference has no single
    entry point for this (mergetable recombines in MAL); the CUDA path's b200_BATcalcavgsum is checked
    against this and against one BATgroupavg over all rows."""
    sa, ca = sums.numpy(), cnts.numpy()
    out = np.empty(len(ca), dtype=np.float64)
    for i in range(len(ca)):
        S = ((int(sa["hi"][i]) << 64) | int(sa["lo"][i])) if sa.dtype.names else int(sa[i])
        n = int(ca[i])
        nil_s = S == -(1 << 127) if sa.dtype.names else S == -(1 << 63)
        if n <= 0 or nil_s:
            out[i] = np.nan
            continue
        a, rem = divmod(S, n)
        r = float(a) + float(rem) / float(n)
        out[i] = r / float(10.0 ** scale) if scale else r
    return bat(out, T.dbl)


def BATsum(tp, b, s=None, skip_nils=True, nil_if_empty=True):
    out = np.zeros(1, dtype=NP_OF[tp])
    rc = _bat.BATsum(_vp(out), tp, b.p, _p(s), skip_nils, nil_if_empty)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATsum failed")
    return out


def BATprod(tp, b, s=None, skip_nils=True, nil_if_empty=True):
    out = np.zeros(1, dtype=NP_OF[tp])
    rc = _bat.BATprod(_vp(out), tp, b.p, _p(s), skip_nils, nil_if_empty)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATprod failed")
    return out


def BATsort(b, o=None, g=None, reverse=False, nilslast=False, stable=False, want=(True, True, True)):
    ps, po, pg = C.c_void_p(), C.c_void_p(), C.c_void_p()
    rc = _bat.BATsort(C.byref(ps) if want[0] else None, C.byref(po) if want[1] else None, C.byref(pg) if want[2] else None,
                      b.p, _p(o), _p(g), reverse, nilslast, stable)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATsort failed")
    return (RBAT(ps.value) if want[0] else None), (RBAT(po.value) if want[1] else None), (RBAT(pg.value) if want[2] else None)


def BATfirstn(b, s, n, asc=True, nilslast=False, g=None, distinct=False, want_gids=False):
    """BATfirstn (gdk_firstn.c:1280); returns topn, or (topn, gids) with want_gids"""
    p, q = C.c_void_p(), C.c_void_p()
    if _bat.BATfirstn(C.byref(p), C.byref(q) if want_gids else None, b.p, _p(s), _p(g), n, asc, nilslast, distinct) != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATfirstn failed")
    return (RBAT(p.value), RBAT(q.value)) if want_gids else RBAT(p.value)


def BATjoin(l, r, sl=None, sr=None, nil_matches=False, estimate=BUN_NONE):
    p1, p2 = C.c_void_p(), C.c_void_p()
    rc = _bat.BATjoin(C.byref(p1), C.byref(p2), l.p, r.p, _p(sl), _p(sr), nil_matches, estimate)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or "BATjoin failed")
    return RBAT(p1.value), RBAT(p2.value)


def _join2(fn, what, l, r, sl, sr, *flags):
    p1, p2 = C.c_void_p(), C.c_void_p()
    rc = fn(C.byref(p1), C.byref(p2), l.p, r.p, _p(sl), _p(sr), *flags, BUN_NONE)
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or what + " failed")
    return RBAT(p1.value), RBAT(p2.value)


def BATleftjoin(l, r, sl=None, sr=None, nil_matches=False):
    return _join2(_bat.BATleftjoin, "BATleftjoin", l, r, sl, sr, nil_matches)


def BATouterjoin(l, r, sl=None, sr=None, nil_matches=False, match_one=False):
    return _join2(_bat.BATouterjoin, "BATouterjoin", l, r, sl, sr, nil_matches, match_one)


def BATsemijoin(l, r, sl=None, sr=None, nil_matches=False, max_one=False):
    return _join2(_bat.BATsemijoin, "BATsemijoin", l, r, sl, sr, nil_matches, max_one)


def _cand1(p, what):
    if not p:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or what + " failed")
    return RBAT(p)


def BATintersect(l, r, sl=None, sr=None, nil_matches=False, max_one=False):
    return _cand1(_bat.BATintersect(l.p, r.p, _p(sl), _p(sr), nil_matches, max_one, BUN_NONE), "BATintersect")


def BATdiff(l, r, sl=None, sr=None, nil_matches=False, not_in=False):
    return _cand1(_bat.BATdiff(l.p, r.p, _p(sl), _p(sr), nil_matches, not_in, BUN_NONE), "BATdiff")


def BATunique(b, s=None):
    return _cand1(_bat.BATunique(b.p, _p(s)), "BATunique")


def _ok(rc, what):
    if rc != 1:
        msg = errbuf().strip()
        clrerr()
        raise RefError(msg or what + " failed")


def GDKanalyticaldiff(r, b, p, npbit, tpe):
    nb = np.array([npbit], dtype=np.int8) if npbit is not None else None
    _ok(_bat.GDKanalyticaldiff(r.p, b.p, _p(p), _vp(nb), tpe), "GDKanalyticaldiff")
    return r


def GDKanalyticalwindowbounds(r, b, p, l, bound, tp1, tp2, unit, preceding, first_half):
    bv = _val(tp2, bound) if bound is not None else None
    _ok(_bat.GDKanalyticalwindowbounds(r.p, _p(b), _p(p), _p(l), _vp(bv), tp1, tp2, unit, preceding, first_half),
        "GDKanalyticalwindowbounds")
    return r


def GDKanalyticalsum(r, p, o, b, s, e, tp1, tp2, frame_type):
    _ok(_bat.GDKanalyticalsum(r.p, _p(p), _p(o), b.p, _p(s), _p(e), tp1, tp2, frame_type), "GDKanalyticalsum")
    return r


def GDKanalyticalfn(name, r, p, o, b, s, e, tpe, frame_type):
    _ok(getattr(_bat, "GDKanalytical" + name)(r.p, _p(p), _p(o), b.p, _p(s), _p(e), tpe, frame_type), name)
    return r


def GDKanalyticalcount(r, p, o, b, s, e, ignore_nils, tpe, frame_type):
    _ok(_bat.GDKanalyticalcount(r.p, _p(p), _p(o), b.p, _p(s), _p(e), 1 if ignore_nils else 0, tpe, frame_type), "count")
    return r
