"""ctypes mirror of include/gdk_b200.h — the host side of the drop-in boundary.

Function names, argument order and error behaviour follow the reference's GDK
entry points (gdk.h:2245-2274, gdk_calc.h:36-147, gdk_analytic.h:21-42): a
BAT-returning function that fails raises :class:`GDKError` carrying the
reference's message text (GDKerror/GDKerrbuf, gdk.h:1710-1717).

There is no CPU fallback here: if ``libgdk_b200.so`` is missing the import
fails, and if no CUDA device is usable every call raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
_LIB = _PKG / "libgdk_b200.so"
if not _LIB.exists():
    raise ImportError(
        f"{_LIB} not found: build it with `python -m monetdb_b200.build` "
        "(nvcc, sm_100a). There is no CPU fallback for this package.")
lib = C.CDLL(str(_LIB))


class GDKError(RuntimeError):
    """GDK_FAIL / NULL result; the message is the GDKerrbuf text."""


# GDK type ids (gdk.h:429-451, HAVE_HGE build)
class TYPE:
    void, msk, bit, bte, sht, int, oid, ptr, flt, dbl, lng, hge, date, daytime, timestamp = range(15)
    str = 16


oid_nil = 1 << 63
BUN_NONE = (1 << 63) - 1
hge_dtype = np.dtype([("lo", "<u8"), ("hi", "<i8")])

NP_OF = {
    TYPE.bit: np.dtype(np.int8), TYPE.bte: np.dtype(np.int8), TYPE.sht: np.dtype(np.int16),
    TYPE.int: np.dtype(np.int32), TYPE.oid: np.dtype(np.uint64), TYPE.flt: np.dtype(np.float32),
    TYPE.dbl: np.dtype(np.float64), TYPE.lng: np.dtype(np.int64), TYPE.hge: hge_dtype,
    TYPE.date: np.dtype(np.int32), TYPE.daytime: np.dtype(np.int64), TYPE.timestamp: np.dtype(np.int64),
    TYPE.void: np.dtype(np.uint64), TYPE.msk: np.dtype(np.uint32),
}
TYPE_NAME = {v: k for k, v in vars(TYPE).items() if not k.startswith("_")}


def nil_of(tt: int):
    """in-band nil of a fixed-width atom (gdk_atoms.h:149-218)"""
    if tt in (TYPE.flt, TYPE.dbl):
        return float("nan")
    if tt in (TYPE.oid, TYPE.void):
        return oid_nil
    if tt == TYPE.hge:
        return -(1 << 127)
    return int(np.iinfo(NP_OF[tt]).min)


def hge_from_ints(vals) -> np.ndarray:
    a = np.empty(len(vals), dtype=hge_dtype)
    for i, v in enumerate(vals):
        v = int(v)
        a["lo"][i] = v & 0xFFFFFFFFFFFFFFFF
        a["hi"][i] = v >> 64
    return a


def hge_to_ints(a: np.ndarray) -> list:
    return [(int(h) << 64) | int(l) for l, h in zip(a["lo"], a["hi"])]


def hge_from_i64(a: np.ndarray) -> np.ndarray:
    a = np.asarray(a, dtype=np.int64)
    r = np.empty(a.shape, dtype=hge_dtype)
    r["lo"] = a.view(np.uint64)
    r["hi"] = a >> 63
    return r


class _bat(C.Structure):
    _fields_ = [
        ("ttype", C.c_int32), ("device", C.c_int32),
        ("hseqbase", C.c_uint64), ("count", C.c_uint64), ("tseqbase", C.c_uint64),
        ("base", C.c_void_p), ("capacity", C.c_uint64),
        ("ccand", C.c_int32), ("firstbit", C.c_uint32),
        ("vheap", C.c_void_p), ("nvheap", C.c_uint64),
        ("tsorted", C.c_uint8), ("trevsorted", C.c_uint8), ("tkey", C.c_uint8),
        ("tnonil", C.c_uint8), ("tnil", C.c_uint8), ("owns", C.c_uint8),
        ("oidrange_known", C.c_uint8), ("pad", C.c_uint8),
        ("first_oid", C.c_uint64), ("last_oid", C.c_uint64),
        ("parent", C.c_void_p), ("producer", C.c_uint32), ("pad2", C.c_uint32),
    ]


_P = C.POINTER(_bat)
_PP = C.POINTER(_P)


def _sig(name, res, *args):
    f = getattr(lib, name)
    f.restype = res
    f.argtypes = list(args)
    return f


_sig("b200_init", C.c_int, C.c_int)
_sig("b200_device_count", C.c_int)
_sig("b200_errbuf", C.c_char_p)
_sig("b200_clrerr", None)
_sig("b200_algorithm", C.c_char_p)
_sig("b200_kernel_launches", C.c_uint64)
_sig("b200_sync", None)
_sig("b200_stream", C.c_void_p)
_sig("b200_heap_cache_trim", None)
_sig("b200_heap_cache_bytes", C.c_uint64)
_sig("b200_pinned_alloc", C.c_void_p, C.c_size_t)
_sig("b200_pinned_free", None, C.c_void_p)
_sig("b200_COLnew", _P, C.c_uint64, C.c_int, C.c_uint64)
_sig("b200_BATdense", _P, C.c_uint64, C.c_uint64, C.c_uint64)
_sig("b200_BATfromhost", _P, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64)
_sig("b200_BATwrap", _P, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64)
_sig("b200_BATnegcands_fromhost", _P, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64)
_sig("b200_BATmaskedcands_fromhost", _P, C.c_uint64, C.c_uint64, C.c_void_p, C.c_bool)
_sig("b200_BATunmask", _P, _P)
_sig("b200_BATextend", C.c_int, _P, C.c_uint64)
_sig("b200_BATsetprops", None, _P, C.c_bool, C.c_bool, C.c_bool, C.c_bool, C.c_bool)
_sig("b200_BATtohost", C.c_int, _P, C.c_void_p)
_sig("b200_BATfree", None, _P)
_sig("b200_BATinvalidate", None, _P)
_sig("b200_BATstrfromhost", _P, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_bool)
_sig("b200_BATstrheap_tohost", C.c_int, _P, C.c_void_p)
_sig("b200_BATstrwidth", C.c_int, _P)
_sig("b200_bincopy_decode", _P, C.c_int, C.c_void_p, C.c_uint64, C.c_bool, C.c_int, C.c_char_p, C.c_uint64)
_sig("b200_bincopy_encode", C.c_int, _P, C.c_void_p, C.c_bool)
_sig("b200_HEAPload", _P, C.c_char_p, C.c_int, C.c_uint64, C.c_uint64)
_sig("b200_HEAPsave", C.c_int, _P, C.c_char_p, C.c_bool)
_sig("b200_select_lng_noalloc", C.c_int, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
     C.c_void_p)
NOALLOC_MAX_ROWS = 1 << 24
_sig("b200_BATselect", _P, _P, _P, C.c_void_p, C.c_void_p, C.c_bool, C.c_bool, C.c_bool, C.c_bool)
_sig("b200_BATthetaselect", _P, _P, _P, C.c_void_p, C.c_char_p)
for _n in ("b200_BATmergecand", "b200_BATintersectcand", "b200_BATdiffcand", "b200_BATproject"):
    _sig(_n, _P, _P, _P)
_sig("b200_BATproject2", _P, _P, _P, _P)


def errbuf() -> str:
    return (lib.b200_errbuf() or b"").decode()


def _fail(what: str):
    msg = errbuf().strip()
    raise GDKError(msg or f"{what}: GDK_FAIL")


def init(device: int = 0) -> None:
    if lib.b200_init(device) != 0:
        _fail("b200_init")


def sync() -> None:
    lib.b200_sync()


def kernel_launches() -> int:
    return int(lib.b200_kernel_launches())


def algorithm() -> str:
    return (lib.b200_algorithm() or b"").decode()


class BAT:
    """Owner of one ``b200_bat`` descriptor (a device-resident column or candidate list)."""

    __slots__ = ("_p", "_keep")

    def __init__(self, p, keep=None):
        if not p:
            _fail("BAT")
        self._p = p
        self._keep = keep

    # descriptor fields
    def __getattr__(self, name):
        if name in ("_p", "_keep"):
            raise AttributeError(name)
        return getattr(self._p.contents, name)

    def __len__(self):
        return int(self._p.contents.count)

    @property
    def ptr(self):
        return self._p

    def free(self):
        if self._p:
            lib.b200_BATfree(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def setprops(self, sorted=False, revsorted=False, key=False, nonil=False, nil=False):
        lib.b200_BATsetprops(self._p, sorted, revsorted, key, nonil, nil)
        return self

    def is_dense(self) -> bool:
        b = self._p.contents
        return b.ccand == 0 and b.ttype == TYPE.void and b.tseqbase != oid_nil

    def numpy(self) -> np.ndarray:
        """logical tail on the host (void / candidate encodings materialised as oids)"""
        b = self._p.contents
        tt = b.ttype
        if tt == TYPE.void or b.ccand != 0:
            out = np.empty(b.count, dtype=np.uint64)
        elif tt == TYPE.msk:
            out = np.empty((b.count + 31) // 32, dtype=np.uint32)
        else:
            out = np.empty(b.count, dtype=NP_OF[tt])
        if b.count and lib.b200_BATtohost(self._p, out.ctypes.data_as(C.c_void_p)) != 0:
            _fail("BATtohost")
        return out

    def props(self) -> dict:
        b = self._p.contents
        return dict(type=b.ttype, hseq=b.hseqbase, count=b.count, tseq=b.tseqbase,
                    sorted=bool(b.tsorted), revsorted=bool(b.trevsorted), key=bool(b.tkey),
                    nonil=bool(b.tnonil), nil=bool(b.tnil), ccand=b.ccand)


def _ptr(b):
    return b._p if b is not None else None


def _val(tt: int, v):
    """a C buffer holding one atom value of storage type tt (None -> NULL)"""
    if v is None:
        return None
    if tt == TYPE.hge:
        a = hge_from_ints([v])
    elif tt in (TYPE.flt, TYPE.dbl):
        a = np.array([v], dtype=NP_OF[tt])
    else:
        a = np.array([int(v)], dtype=np.int64).astype(NP_OF[tt]) if tt not in (TYPE.oid, TYPE.void) \
            else np.array([int(v)], dtype=np.uint64)
    return a


def _vp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


_ATOMS: dict = {}
_NOATOM = (None, None)


def _valp(tt: int, v):
    """(buffer, pointer) of one atom value, memoised: the per-call numpy/ctypes work of _val + _vp
    is a measurable part of a short operator call"""
    if v is None:
        return _NOATOM
    key = (tt, v)
    try:
        return _ATOMS[key]
    except (KeyError, TypeError):
        a = _val(tt, v)
        r = (a, _vp(a))
        try:
            if len(_ATOMS) < 4096:
                _ATOMS[key] = r
        except TypeError:
            pass
        return r


# ------------------------------------------------------------------ construction
def BATfromhost(arr: np.ndarray, tt: int | None = None, hseq: int = 0, **props) -> BAT:
    """upload a tail heap (the device heap tier's load path)"""
    arr = np.ascontiguousarray(arr)
    if tt is None:
        tt = {np.dtype(np.int8): TYPE.bte, np.dtype(np.int16): TYPE.sht, np.dtype(np.int32): TYPE.int,
              np.dtype(np.int64): TYPE.lng, np.dtype(np.uint64): TYPE.oid, np.dtype(np.float32): TYPE.flt,
              np.dtype(np.float64): TYPE.dbl, hge_dtype: TYPE.hge}[arr.dtype]
    assert arr.dtype.itemsize == NP_OF[tt].itemsize, (arr.dtype, tt)
    b = BAT(lib.b200_BATfromhost(_vp(arr), tt, len(arr), hseq))
    if props:
        b.setprops(**props)
    return b


def BATdense(hseq: int, tseq: int, cnt: int) -> BAT:
    return BAT(lib.b200_BATdense(hseq, tseq, cnt))


def COLnew(hseq: int, tt: int, cap: int) -> BAT:
    return BAT(lib.b200_COLnew(hseq, tt, cap))


def BATwrap(devptr: int, tt: int, cnt: int, hseq: int = 0, keep=None, **props) -> BAT:
    b = BAT(lib.b200_BATwrap(C.c_void_p(devptr), tt, cnt, hseq), keep=keep)
    if props:
        b.setprops(**props)
    return b


def BATnegcands(seq: int, cnt: int, exceptions) -> BAT:
    e = np.ascontiguousarray(exceptions, dtype=np.uint64)
    return BAT(lib.b200_BATnegcands_fromhost(seq, cnt, _vp(e), len(e)))


def BATmaskedcands(hseq: int, nrows: int, words, selected: bool = True) -> BAT:
    w = np.ascontiguousarray(words, dtype=np.uint32)
    return BAT(lib.b200_BATmaskedcands_fromhost(hseq, nrows, _vp(w), selected))


def BATunmask(b: BAT) -> BAT:
    return BAT(lib.b200_BATunmask(b._p))


def BATextend(b: BAT, newcap: int) -> BAT:
    if lib.b200_BATextend(b._p, newcap) != 1:
        _fail("BATextend")
    return b


# ------------------------------------------------------------------ select / project / cands
def _storage(tt: int) -> int:
    return {TYPE.bit: TYPE.bte, TYPE.date: TYPE.int, TYPE.daytime: TYPE.lng, TYPE.timestamp: TYPE.lng}.get(tt, tt)


_OPS = {o: o.encode() for o in ("=", "==", "!=", "<>", "<", "<=", ">", ">=", "eq", "ne")}


str_nil = b"\x80"
GDK_VAROFFSET = 8192
_STR_OFF = {1: np.dtype(np.uint8), 2: np.dtype(np.uint16), 4: np.dtype(np.uint32), 8: np.dtype(np.uint64)}


def _cstr(v):
    """a str atom as the NUL-terminated bytes GDK passes around (None stays NULL, nil is "\\200")"""
    if v is None:
        return None
    return C.c_char_p(v if isinstance(v, bytes) else v.encode())


def BATstrfromhost(offsets: np.ndarray, heap: bytes, hseq: int = 0, elim_doubles: bool = True) -> BAT:
    """a str column as GDK stores it: the tail of offsets (uint8/16/32/64; the two narrow forms are stored
    minus GDK_VAROFFSET) and the string heap (gdk_atoms.h:370-377)"""
    offsets = np.ascontiguousarray(offsets)
    if offsets.dtype not in _STR_OFF.values():
        raise TypeError("offsets must be uint8 / uint16 / uint32 / uint64")
    hb = (C.c_char * len(heap)).from_buffer_copy(heap)
    return BAT(lib.b200_BATstrfromhost(offsets.ctypes.data_as(C.c_void_p), offsets.dtype.itemsize, len(offsets), hseq,
                                       C.cast(hb, C.c_void_p), len(heap), elim_doubles))


def str_parts(b: BAT):
    """(offsets as stored, heap bytes) of a device str column"""
    w = lib.b200_BATstrwidth(b._p)
    if w not in _STR_OFF:
        raise GDKError("not a str column")
    off = np.empty(len(b), dtype=_STR_OFF[w])
    if len(b) and lib.b200_BATtohost(b._p, off.ctypes.data_as(C.c_void_p)) != 0:
        _fail("BATtohost")
    heap = C.create_string_buffer(int(b.nvheap) + 1)
    if lib.b200_BATstrheap_tohost(b._p, C.cast(heap, C.c_void_p)) != 0:
        _fail("BATstrheap_tohost")
    return off, heap.raw[:int(b.nvheap)]


def str_decode(offsets: np.ndarray, heap: bytes) -> list:
    """the strings a tail of offsets names (bytes; nil = b"\\x80")"""
    add = GDK_VAROFFSET if offsets.dtype.itemsize <= 2 else 0
    out = []
    for o in offsets.tolist():
        o += add
        out.append(heap[o:heap.index(b"\0", o)])
    return out


BIN_RECORD = {TYPE.date: 4, TYPE.daytime: 8, TYPE.timestamp: 12}


def bincopy_decode(tt: int, raw, count: int, byteswap: bool = False, width: int = 0, filename: str = "<buffer>", hseq: int = 0) -> BAT:
    """binary COPY INTO records (bytes / numpy buffer) -> device column (sql_bincopyconvert.c)"""
    buf = np.frombuffer(raw, dtype=np.uint8) if not isinstance(raw, np.ndarray) else np.ascontiguousarray(raw).view(np.uint8)
    return BAT(lib.b200_bincopy_decode(tt, buf.ctypes.data_as(C.c_void_p), count, byteswap, width, filename.encode(), hseq))


def bincopy_encode(b: BAT, byteswap: bool = False) -> bytes:
    rs = BIN_RECORD.get(b.ttype) or NP_OF[b.ttype].itemsize
    out = np.empty(len(b) * rs, dtype=np.uint8)
    if lib.b200_bincopy_encode(b._p, out.ctypes.data_as(C.c_void_p), byteswap) != 0:
        _fail("bincopy_encode")
    return out.tobytes()


def HEAPload(path: str, tt: int, count: int, hseq: int = 0) -> BAT:
    return BAT(lib.b200_HEAPload(str(path).encode(), tt, count, hseq))


def HEAPsave(b: BAT, path: str, dosync: bool = False) -> None:
    if lib.b200_HEAPsave(b._p, str(path).encode(), dosync) != 0:
        _fail("HEAPsave")


def BATselect(b: BAT, s: BAT | None, tl, th, li: bool, hi: bool, anti: bool, nil_matches: bool = False) -> BAT:
    if b.ttype == TYPE.str:
        return BAT(lib.b200_BATselect(b._p, _ptr(s), _cstr(tl), _cstr(th), li, hi, anti, nil_matches))
    st = _storage(b.ttype)
    vl, vh = _valp(st, tl), _valp(st, th)
    return BAT(lib.b200_BATselect(b._p, _ptr(s), vl[1], vh[1], li, hi, anti, nil_matches))


def BATthetaselect(b: BAT, s: BAT | None, val, op: str) -> BAT:
    if b.ttype == TYPE.str:
        return BAT(lib.b200_BATthetaselect(b._p, _ptr(s), _cstr(val), _OPS.get(op) or op.encode()))
    v = _valp(_storage(b.ttype), val)
    return BAT(lib.b200_BATthetaselect(b._p, _ptr(s), v[1], _OPS.get(op) or op.encode()))


def BATproject(l: BAT, r: BAT) -> BAT:
    return BAT(lib.b200_BATproject(l._p, r._p))


def BATproject2(l: BAT, r1: BAT, r2: BAT | None) -> BAT:
    return BAT(lib.b200_BATproject2(l._p, r1._p, _ptr(r2)))


def BATmergecand(a: BAT, b: BAT) -> BAT:
    return BAT(lib.b200_BATmergecand(a._p, b._p))


def BATintersectcand(a: BAT, b: BAT) -> BAT:
    return BAT(lib.b200_BATintersectcand(a._p, b._p))


def BATdiffcand(a: BAT, b: BAT) -> BAT:
    return BAT(lib.b200_BATdiffcand(a._p, b._p))


# ------------------------------------------------------------------ calc
for _op in ("add", "sub", "mul", "div", "mod"):
    _sig(f"b200_BATcalc{_op}", _P, _P, _P, _P, _P, C.c_int)
    _sig(f"b200_BATcalc{_op}cst", _P, _P, C.c_void_p, C.c_int, _P, C.c_int)
    _sig(f"b200_BATcalccst{_op}", _P, C.c_void_p, C.c_int, _P, _P, C.c_int)
_sig("b200_BATconvert", _P, _P, _P, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8)


def BATcalc(op: str, b1: BAT, b2: BAT, s1: BAT | None, s2: BAT | None, tp: int) -> BAT:
    """BATcalcadd/sub/mul/div/mod (gdk_calc.h:36-50)"""
    return BAT(getattr(lib, f"b200_BATcalc{op}")(b1._p, b2._p, _ptr(s1), _ptr(s2), tp))


def BATcalccst(op: str, b: BAT, v, vtype: int, s: BAT | None, tp: int, cst_first: bool = False) -> BAT:
    """BATcalc<op>cst (b OP v) or BATcalccst<op> (v OP b); the ValRecord is (value, type id)"""
    a = _valp(_storage(vtype), v)
    if cst_first:
        return BAT(getattr(lib, f"b200_BATcalccst{op}")(a[1], vtype, b._p, _ptr(s), tp))
    return BAT(getattr(lib, f"b200_BATcalc{op}cst")(b._p, a[1], vtype, _ptr(s), tp))


def BATcalcadd(b1, b2, s1, s2, tp): return BATcalc("add", b1, b2, s1, s2, tp)
def BATcalcsub(b1, b2, s1, s2, tp): return BATcalc("sub", b1, b2, s1, s2, tp)
def BATcalcmul(b1, b2, s1, s2, tp): return BATcalc("mul", b1, b2, s1, s2, tp)
def BATcalcdiv(b1, b2, s1, s2, tp): return BATcalc("div", b1, b2, s1, s2, tp)
def BATcalcmod(b1, b2, s1, s2, tp): return BATcalc("mod", b1, b2, s1, s2, tp)


def BATconvert(b: BAT, s: BAT | None, tp: int, scale1: int = 0, scale2: int = 0, precision: int = 0) -> BAT:
    return BAT(lib.b200_BATconvert(b._p, _ptr(s), tp, scale1, scale2, precision))


# ------------------------------------------------------------------ compares (gdk_calc.h:66-83)
for _op in ("lt", "le", "gt", "ge"):
    _sig(f"b200_BATcalc{_op}", _P, _P, _P, _P, _P)
    _sig(f"b200_BATcalc{_op}cst", _P, _P, C.c_void_p, C.c_int, _P)
    _sig(f"b200_BATcalccst{_op}", _P, C.c_void_p, C.c_int, _P, _P)
for _op in ("eq", "ne"):
    _sig(f"b200_BATcalc{_op}", _P, _P, _P, _P, _P, C.c_bool)
    _sig(f"b200_BATcalc{_op}cst", _P, _P, C.c_void_p, C.c_int, _P, C.c_bool)
    _sig(f"b200_BATcalccst{_op}", _P, C.c_void_p, C.c_int, _P, _P, C.c_bool)
_sig("b200_BATcalcisnil", _P, _P, _P)
_sig("b200_BATcalcisnotnil", _P, _P, _P)


def BATcalccmp(op: str, b1: BAT, b2: BAT, s1: BAT | None = None, s2: BAT | None = None, nil_matches: bool = False) -> BAT:
    """BATcalclt/le/gt/ge/eq/ne -> bit column"""
    f = getattr(lib, f"b200_BATcalc{op}")
    if op in ("eq", "ne"):
        return BAT(f(b1._p, b2._p, _ptr(s1), _ptr(s2), nil_matches))
    return BAT(f(b1._p, b2._p, _ptr(s1), _ptr(s2)))


def BATcalccmpcst(op: str, b: BAT, v, vtype: int, s: BAT | None = None, nil_matches: bool = False,
                  cst_first: bool = False) -> BAT:
    """BATcalc<op>cst (b OP v) or, with cst_first, BATcalccst<op> (v OP b; gdk_calc_compare.h:928)"""
    a = _val(_storage(vtype), v)
    if cst_first:
        f = getattr(lib, f"b200_BATcalccst{op}")
        if op in ("eq", "ne"):
            return BAT(f(_vp(a), vtype, b._p, _ptr(s), nil_matches))
        return BAT(f(_vp(a), vtype, b._p, _ptr(s)))
    f = getattr(lib, f"b200_BATcalc{op}cst")
    if op in ("eq", "ne"):
        return BAT(f(b._p, _vp(a), vtype, _ptr(s), nil_matches))
    return BAT(f(b._p, _vp(a), vtype, _ptr(s)))


def BATcalcisnil(b: BAT, s: BAT | None = None) -> BAT:
    return BAT(lib.b200_BATcalcisnil(b._p, _ptr(s)))


def BATcalcisnotnil(b: BAT, s: BAT | None = None) -> BAT:
    return BAT(lib.b200_BATcalcisnotnil(b._p, _ptr(s)))


# ------------------------------------------------------------------ min/max, between, ifthenelse (gdk_calc.h:24-33,87-95)
for _n in ("min", "min_no_nil", "max", "max_no_nil"):
    _sig(f"b200_BATcalc{_n}", _P, _P, _P, _P, _P)
_B5 = (C.c_bool,) * 5
_sig("b200_BATcalcbetween", _P, _P, _P, _P, _P, _P, _P, *_B5)
_sig("b200_BATcalcbetweencstcst", _P, _P, C.c_void_p, C.c_int, C.c_void_p, C.c_int, _P, *_B5)
_sig("b200_BATcalcbetweenbatcst", _P, _P, _P, C.c_void_p, C.c_int, _P, _P, *_B5)
_sig("b200_BATcalcbetweencstbat", _P, _P, C.c_void_p, C.c_int, _P, _P, _P, *_B5)
_sig("b200_BATcalcifthenelse", _P, _P, _P, _P)
_sig("b200_BATcalcifthenelsecst", _P, _P, _P, C.c_void_p, C.c_int)
_sig("b200_BATcalcifthencstelse", _P, _P, C.c_void_p, C.c_int, _P)
_sig("b200_BATcalcifthencstelsecst", _P, _P, C.c_void_p, C.c_int, C.c_void_p, C.c_int)


for _n in ("not", "negate", "absolute", "iszero", "sign"):
    _sig(f"b200_BATcalc{_n}", _P, _P, _P)


for _n in ("and", "or", "xor"):
    _sig(f"b200_BATcalc{_n}", _P, _P, _P, _P, _P)
    _sig(f"b200_BATcalc{_n}cst", _P, _P, C.c_void_p, C.c_int, _P)
    _sig(f"b200_BATcalccst{_n}", _P, C.c_void_p, C.c_int, _P, _P)


for _n in ("lsh", "rsh"):
    _sig(f"b200_BATcalc{_n}", _P, _P, _P, _P, _P)
    _sig(f"b200_BATcalc{_n}cst", _P, _P, C.c_void_p, C.c_int, _P)
    _sig(f"b200_BATcalccst{_n}", _P, C.c_void_p, C.c_int, _P, _P)


def BATcalcshift(op: str, b1: BAT, b2: BAT, s1: BAT | None = None, s2: BAT | None = None) -> BAT:
    """BATcalclsh / BATcalcrsh (gdk_calc.c:3246, 3596); op = "lsh" | "rsh" """
    return BAT(getattr(lib, f"b200_BATcalc{op}")(b1._p, b2._p, _ptr(s1), _ptr(s2)))


def BATcalcshiftcst(op: str, b: BAT, v, vtype: int, s: BAT | None = None, cst_first: bool = False) -> BAT:
    """b OP v, or with cst_first v OP b (b holds the shift counts)"""
    a = _valp(_storage(vtype), v)
    if cst_first:
        return BAT(getattr(lib, f"b200_BATcalccst{op}")(a[1], vtype, b._p, _ptr(s)))
    return BAT(getattr(lib, f"b200_BATcalc{op}cst")(b._p, a[1], vtype, _ptr(s)))


def BATcalclogic(op: str, b1: BAT, b2: BAT, s1: BAT | None = None, s2: BAT | None = None) -> BAT:
    """BATcalcand / or / xor (gdk_calc.h:51-58): three-valued on bit, bitwise on the integer atoms"""
    return BAT(getattr(lib, f"b200_BATcalc{op}")(b1._p, b2._p, _ptr(s1), _ptr(s2)))


def BATcalclogiccst(op: str, b: BAT, v, vtype: int, s: BAT | None = None, cst_first: bool = False) -> BAT:
    """BATcalc<op>cst / BATcalccst<op> for and / or / xor"""
    a = _valp(_storage(vtype), v)
    if cst_first:
        return BAT(getattr(lib, f"b200_BATcalccst{op}")(a[1], vtype, b._p, _ptr(s)))
    return BAT(getattr(lib, f"b200_BATcalc{op}cst")(b._p, a[1], vtype, _ptr(s)))


def BATcalcunary(which: str, b: BAT, s: BAT | None = None) -> BAT:
    """BATcalcnot / negate / absolute / iszero / sign"""
    return BAT(getattr(lib, f"b200_BATcalc{which}")(b._p, _ptr(s)))


for _n in ("mincst", "mincst_no_nil", "maxcst", "maxcst_no_nil"):
    _sig(f"b200_BATcalc{_n}", _P, _P, C.c_void_p, C.c_int, _P)


def BATcalcminmaxcst(which: str, b: BAT, v, vtype: int, s: BAT | None = None) -> BAT:
    """BATcalcmincst / mincst_no_nil / maxcst / maxcst_no_nil (gdk_calc.c:1403,1563,2086,2219); which = "min", "max_no_nil", ..."""
    a = _valp(_storage(vtype), v)
    name = which.replace("min", "mincst", 1) if which.startswith("min") else which.replace("max", "maxcst", 1)
    return BAT(getattr(lib, f"b200_BATcalc{name}")(b._p, a[1], vtype, _ptr(s)))


def BATcalcminmax(which: str, b1: BAT, b2: BAT, s1: BAT | None = None, s2: BAT | None = None) -> BAT:
    """BATcalcmin / min_no_nil / max / max_no_nil"""
    return BAT(getattr(lib, f"b200_BATcalc{which}")(b1._p, b2._p, _ptr(s1), _ptr(s2)))


def BATcalcbetween(b: BAT, lo, hi, s: BAT | None = None, slo: BAT | None = None, shi: BAT | None = None,
                   symmetric: bool = False, linc: bool = True, hinc: bool = True, nils_false: bool = False,
                   anti: bool = False) -> BAT:
    """BATcalcbetween and its batcst / cstbat / cstcst forms: lo and hi are BATs or plain values of b's type"""
    fl = (symmetric, linc, hinc, nils_false, anti)
    st = _storage(b.ttype)
    lb, hb = isinstance(lo, BAT), isinstance(hi, BAT)
    if lb and hb:
        return BAT(lib.b200_BATcalcbetween(b._p, lo._p, hi._p, _ptr(s), _ptr(slo), _ptr(shi), *fl))
    if lb:
        return BAT(lib.b200_BATcalcbetweenbatcst(b._p, lo._p, _valp(st, hi)[1], b.ttype, _ptr(s), _ptr(slo), *fl))
    if hb:
        return BAT(lib.b200_BATcalcbetweencstbat(b._p, _valp(st, lo)[1], b.ttype, hi._p, _ptr(s), _ptr(shi), *fl))
    return BAT(lib.b200_BATcalcbetweencstcst(b._p, _valp(st, lo)[1], b.ttype, _valp(st, hi)[1], b.ttype, _ptr(s), *fl))


def BATcalcifthenelse(b: BAT, v1, v2, vtype: int | None = None) -> BAT:
    """BATcalcifthenelse and its cst forms: v1 / v2 are BATs or plain values (of type vtype when both are values)"""
    b1, b2 = isinstance(v1, BAT), isinstance(v2, BAT)
    tt = v1.ttype if b1 else v2.ttype if b2 else vtype
    st = _storage(tt)
    if b1 and b2:
        return BAT(lib.b200_BATcalcifthenelse(b._p, v1._p, v2._p))
    if b1:
        return BAT(lib.b200_BATcalcifthenelsecst(b._p, v1._p, _valp(st, v2)[1], tt))
    if b2:
        return BAT(lib.b200_BATcalcifthencstelse(b._p, _valp(st, v1)[1], tt, v2._p))
    return BAT(lib.b200_BATcalcifthencstelsecst(b._p, _valp(st, v1)[1], tt, _valp(st, v2)[1], tt))


# ------------------------------------------------------------------ grouping
_sig("b200_BATgroup", C.c_int, _PP, _PP, _PP, _P, _P, _P, _P, _P)


def BATgroup(b: BAT, s: BAT | None = None, g: BAT | None = None, e: BAT | None = None, h: BAT | None = None,
             want_extents: bool = True, want_histo: bool = True):
    """BATgroup (gdk.h:1447) -> (groups, extents, histo)"""
    pg, pe, ph = _P(), _P(), _P()
    rc = lib.b200_BATgroup(C.byref(pg), C.byref(pe) if want_extents else None, C.byref(ph) if want_histo else None,
                           b._p, _ptr(s), _ptr(g), _ptr(e), _ptr(h))
    if rc != 1:
        _fail("BATgroup")
    return BAT(pg), (BAT(pe) if want_extents else None), (BAT(ph) if want_histo else None)


# ------------------------------------------------------------------ sort / top-N (gdk.h BATsort, BATfirstn)
_sig("b200_BATsort", C.c_int, _PP, _PP, _PP, _P, _P, _P, C.c_bool, C.c_bool, C.c_bool)
_sig("b200_BATorderidx", C.c_int, _P, C.c_bool)
_sig("b200_BATcheckorderidx", C.c_int, _P)


def BATorderidx(b: BAT, stable: bool = True) -> None:
    """BATorderidx (gdk_orderidx.c:184): build and keep the column's order index"""
    if lib.b200_BATorderidx(b._p, stable) != 1:
        _fail("BATorderidx")


def BATcheckorderidx(b: BAT) -> bool:
    return bool(lib.b200_BATcheckorderidx(b._p))


_sig("b200_BATfirstn", C.c_int, _PP, _PP, _P, _P, _P, C.c_uint64, C.c_bool, C.c_bool, C.c_bool)


def BATsort(b: BAT, o: BAT | None = None, g: BAT | None = None, reverse: bool = False, nilslast: bool = False,
            stable: bool = False, want=(True, True, True)):
    """BATsort (gdk_batop.c:2342) -> (sorted, order, groups); entries of `want` that are False come back None"""
    ps, po, pg = _P(), _P(), _P()
    rc = lib.b200_BATsort(C.byref(ps) if want[0] else None, C.byref(po) if want[1] else None, C.byref(pg) if want[2] else None,
                          b._p, _ptr(o), _ptr(g), reverse, nilslast, stable)
    if rc != 1:
        _fail("BATsort")
    return (BAT(ps) if want[0] else None), (BAT(po) if want[1] else None), (BAT(pg) if want[2] else None)


def BATfirstn(b: BAT, s: BAT | None, n: int, asc: bool = True, nilslast: bool = False, g: BAT | None = None,
              distinct: bool = False, want_gids: bool = False):
    """BATfirstn (gdk_firstn.c:1280) -> candidate list of the n smallest / largest values, or (topn, gids) with
    want_gids; g / distinct / gids are the refinement forms of a multi-column ORDER BY ... LIMIT"""
    p, q = _P(), _P()
    if lib.b200_BATfirstn(C.byref(p), C.byref(q) if want_gids else None, b._p, _ptr(s), _ptr(g), n, asc, nilslast, distinct) != 1:
        _fail("BATfirstn")
    return (BAT(p), BAT(q)) if want_gids else BAT(p)


# ------------------------------------------------------------------ hash + join (gdk_hash.h:36, gdk.h:2266)
_sig("b200_BAThash", C.c_int, _P)
_sig("b200_BATjoin", C.c_int, _PP, _PP, _P, _P, _P, _P, C.c_bool, C.c_uint64)


def BAThash(b: BAT) -> None:
    if lib.b200_BAThash(b._p) != 1:
        _fail("BAThash")


def BATjoin(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
            estimate: int = BUN_NONE, want_r2: bool = True):
    """BATjoin -> (r1, r2): aligned oid lists of matching rows of l and r"""
    p1, p2 = _P(), _P()
    rc = lib.b200_BATjoin(C.byref(p1), C.byref(p2) if want_r2 else None, l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, estimate)
    if rc != 1:
        _fail("BATjoin")
    return BAT(p1), (BAT(p2) if want_r2 else None)


_sig("b200_BATleftjoin", C.c_int, _PP, _PP, _P, _P, _P, _P, C.c_bool, C.c_uint64)
_sig("b200_BATouterjoin", C.c_int, _PP, _PP, _P, _P, _P, _P, C.c_bool, C.c_bool, C.c_uint64)
_sig("b200_BATsemijoin", C.c_int, _PP, _PP, _P, _P, _P, _P, C.c_bool, C.c_bool, C.c_uint64)
_sig("b200_BATintersect", _P, _P, _P, _P, _P, C.c_bool, C.c_bool, C.c_uint64)
_sig("b200_BATdiff", _P, _P, _P, _P, _P, C.c_bool, C.c_bool, C.c_uint64)
_sig("b200_BATunique", _P, _P, _P)


def BATleftjoin(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
                estimate: int = BUN_NONE, want_r2: bool = True):
    """BATleftjoin (gdk.h:2254) -> (r1, r2), r1 in l's order"""
    p1, p2 = _P(), _P()
    rc = lib.b200_BATleftjoin(C.byref(p1), C.byref(p2) if want_r2 else None, l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, estimate)
    if rc != 1:
        _fail("BATleftjoin")
    return BAT(p1), (BAT(p2) if want_r2 else None)


def BATouterjoin(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
                 match_one: bool = False, estimate: int = BUN_NONE):
    """BATouterjoin (gdk.h:2258) -> (r1, r2); r2 holds oid_nil where a row of l has no match"""
    p1, p2 = _P(), _P()
    rc = lib.b200_BATouterjoin(C.byref(p1), C.byref(p2), l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, match_one, estimate)
    if rc != 1:
        _fail("BATouterjoin")
    return BAT(p1), BAT(p2)


def BATsemijoin(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
                max_one: bool = False, estimate: int = BUN_NONE, want_r2: bool = True):
    """BATsemijoin (gdk.h:2262) -> (r1, r2): at most one pair per row of l"""
    p1, p2 = _P(), _P()
    rc = lib.b200_BATsemijoin(C.byref(p1), C.byref(p2) if want_r2 else None, l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, max_one,
                              estimate)
    if rc != 1:
        _fail("BATsemijoin")
    return BAT(p1), (BAT(p2) if want_r2 else None)


def BATintersect(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
                 max_one: bool = False, estimate: int = BUN_NONE) -> BAT:
    """BATintersect (gdk.h:2264) -> candidate list of the rows of l whose value occurs in r"""
    p = lib.b200_BATintersect(l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, max_one, estimate)
    if not p:
        _fail("BATintersect")
    return BAT(p)


def BATdiff(l: BAT, r: BAT, sl: BAT | None = None, sr: BAT | None = None, nil_matches: bool = False,
            not_in: bool = False, estimate: int = BUN_NONE) -> BAT:
    """BATdiff (gdk.h:2265) -> candidate list of the rows of l whose value does not occur in r"""
    p = lib.b200_BATdiff(l._p, r._p, _ptr(sl), _ptr(sr), nil_matches, not_in, estimate)
    if not p:
        _fail("BATdiff")
    return BAT(p)


def BATunique(b: BAT, s: BAT | None = None) -> BAT:
    """BATunique (gdk.h:2279) -> candidate list of the first row of each distinct value"""
    p = lib.b200_BATunique(b._p, _ptr(s))
    if not p:
        _fail("BATunique")
    return BAT(p)


# ------------------------------------------------------------------ aggregates
for _n in ("sum", "prod", "count", "min", "max"):
    if hasattr(lib, f"b200_BATgroup{_n}"):
        _sig(f"b200_BATgroup{_n}", _P, _P, _P, _P, _P, C.c_int, C.c_bool)
_sig("b200_BATgroupavg", C.c_int, _PP, _PP, _P, _P, _P, _P, C.c_int, C.c_bool, C.c_int)
_sig("b200_BATsum", C.c_int, C.c_void_p, C.c_int, _P, _P, C.c_bool, C.c_bool)
_sig("b200_BATgroupavg3", C.c_int, _PP, _PP, _PP, _P, _P, _P, _P, C.c_bool)
_sig("b200_BATgroupavg3combine", _P, _P, _P, _P, _P, _P, C.c_bool)
_sig("b200_BATcalcavg", C.c_int, _P, _P, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int)


def BATgroupsum(b, g, e, s, tp, skip_nils=True) -> BAT:
    return BAT(lib.b200_BATgroupsum(b._p, _ptr(g), _ptr(e), _ptr(s), tp, skip_nils))


def BATgroupprod(b, g, e, s, tp, skip_nils=True) -> BAT:
    """BATgroupprod (gdk_aggr.c:1575)"""
    return BAT(lib.b200_BATgroupprod(b._p, _ptr(g), _ptr(e), _ptr(s), tp, skip_nils))


def BATgroupcount(b, g, e, s, tp=TYPE.lng, skip_nils=True) -> BAT:
    return BAT(lib.b200_BATgroupcount(b._p, _ptr(g), _ptr(e), _ptr(s), tp, skip_nils))


def BATgroupmin(b, g, e, s, tp=TYPE.oid, skip_nils=True) -> BAT:
    return BAT(lib.b200_BATgroupmin(b._p, _ptr(g), _ptr(e), _ptr(s), tp, skip_nils))


def BATgroupmax(b, g, e, s, tp=TYPE.oid, skip_nils=True) -> BAT:
    return BAT(lib.b200_BATgroupmax(b._p, _ptr(g), _ptr(e), _ptr(s), tp, skip_nils))


def BATgroupavg(b, g, e, s, tp=TYPE.dbl, skip_nils=True, scale=0, want_counts=True):
    pa, pc = _P(), _P()
    rc = lib.b200_BATgroupavg(C.byref(pa), C.byref(pc) if want_counts else None, b._p, _ptr(g), _ptr(e), _ptr(s),
                              tp, skip_nils, scale)
    if rc != 1:
        _fail("BATgroupavg")
    return BAT(pa), (BAT(pc) if want_counts else None)


def BATgroupavg3(b, g, e, s, skip_nils=True):
    """(avg, rem, cnt) triples, gdk_aggr.c:1996"""
    pa, pr, pc = _P(), _P(), _P()
    if lib.b200_BATgroupavg3(C.byref(pa), C.byref(pr), C.byref(pc), b._p, _ptr(g), _ptr(e), _ptr(s), skip_nils) != 1:
        _fail("BATgroupavg3")
    return BAT(pa), BAT(pr), BAT(pc)


def BATgroupavg3combine(avg, rem, cnt, g, e, skip_nils=True) -> BAT:
    return BAT(lib.b200_BATgroupavg3combine(avg._p, rem._p, cnt._p, _ptr(g), _ptr(e), skip_nils))


def BATcalcavg(b, s=None, scale=0):
    """(avg as float, number of non-nil values), gdk_aggr.c:2987"""
    avg, n = C.c_double(), C.c_uint64()
    if lib.b200_BATcalcavg(b._p, _ptr(s), C.byref(avg), C.byref(n), scale) != 1:
        _fail("BATcalcavg")
    return avg.value, n.value


_sig("b200_BATcalcavgsum", _P, _P, _P, C.c_int)


def BATcalcavgsum(sums: BAT, cnts: BAT, scale: int = 0) -> BAT:
    """dbl averages from merged (sum, count) with BATgroupavg's closing formula (phase 2 of a two-phase avg)"""
    return BAT(lib.b200_BATcalcavgsum(sums._p, cnts._p, scale))


def BATsum(tp: int, b: BAT, s: BAT | None = None, skip_nils: bool = True, nil_if_empty: bool = True) -> np.ndarray:
    """returns a 1-element array of the result type (hge as (lo,hi))"""
    out = np.zeros(1, dtype=NP_OF[tp])
    if lib.b200_BATsum(_vp(out), tp, b._p, _ptr(s), skip_nils, nil_if_empty) != 1:
        _fail("BATsum")
    return out


_sig("b200_BATprod", C.c_int, C.c_void_p, C.c_int, _P, _P, C.c_bool, C.c_bool)


def BATprod(tp: int, b: BAT, s: BAT | None = None, skip_nils: bool = True, nil_if_empty: bool = True) -> np.ndarray:
    """BATprod (gdk_aggr.c:1650): 1-element array of the result type"""
    out = np.zeros(1, dtype=NP_OF[tp])
    if lib.b200_BATprod(_vp(out), tp, b._p, _ptr(s), skip_nils, nil_if_empty) != 1:
        _fail("BATprod")
    return out


# ------------------------------------------------------------------ window functions (gdk_analytic.h:21-42)
_sig("b200_GDKanalyticaldiff", C.c_int, _P, _P, _P, C.c_void_p, C.c_int)
_sig("b200_GDKanalyticalwindowbounds", C.c_int, _P, _P, _P, _P, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_bool, C.c_uint64)
_sig("b200_GDKanalyticalsum", C.c_int, _P, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int)
for _n in ("avg", "avginteger", "min", "max"):
    _sig(f"b200_GDKanalytical{_n}", C.c_int, _P, _P, _P, _P, _P, _P, C.c_int, C.c_int)
_sig("b200_GDKanalyticalcount", C.c_int, _P, _P, _P, _P, _P, _P, C.c_int8, C.c_int, C.c_int)


def colnew(tt: int, cap: int, hseq: int = 0) -> BAT:
    """caller-allocated result BAT for the GDKanalytical* functions (sql_rank.c:161)"""
    return COLnew(hseq, tt, cap)


def _ok(rc, what):
    if rc != 1:
        _fail(what)


def GDKanalyticaldiff(r, b, p, npbit, tpe):
    nb = np.array([npbit], dtype=np.int8) if npbit is not None else None
    _ok(lib.b200_GDKanalyticaldiff(r._p, b._p, _ptr(p), _vp(nb), tpe), "GDKanalyticaldiff")
    return r


def GDKanalyticalwindowbounds(r, b, p, l, bound, tp1, tp2, unit, preceding, first_half):
    bv = _val(_storage(tp2), bound) if bound is not None else None
    _ok(lib.b200_GDKanalyticalwindowbounds(r._p, _ptr(b), _ptr(p), _ptr(l), _vp(bv), tp1, tp2, unit, preceding, first_half),
        "GDKanalyticalwindowbounds")
    return r


def GDKanalyticalsum(r, p, o, b, s, e, tp1, tp2, frame_type):
    _ok(lib.b200_GDKanalyticalsum(r._p, _ptr(p), _ptr(o), b._p, _ptr(s), _ptr(e), tp1, tp2, frame_type), "GDKanalyticalsum")
    return r


def GDKanalyticalfn(name, r, p, o, b, s, e, tpe, frame_type):
    """GDKanalyticalavg / min / max"""
    _ok(getattr(lib, "b200_GDKanalytical" + name)(r._p, _ptr(p), _ptr(o), b._p, _ptr(s), _ptr(e), tpe, frame_type), name)
    return r


def GDKanalyticalcount(r, p, o, b, s, e, ignore_nils, tpe, frame_type):
    _ok(lib.b200_GDKanalyticalcount(r._p, _ptr(p), _ptr(o), b._p, _ptr(s), _ptr(e), 1 if ignore_nils else 0, tpe, frame_type),
        "GDKanalyticalcount")
    return r


# ------------------------------------------------------------------ multi-GPU exchange step
_sig("b200_hash_partition", C.c_int, _P, _P, C.c_int, C.c_bool, C.c_void_p, C.c_void_p, C.c_void_p)


def hash_partition(b: BAT, s: BAT | None, nparts: int, keys_ptr: int | None = None, oids_ptr: int | None = None,
                   keep_nil: bool = False) -> np.ndarray:
    """counts per destination of hash(key) mod nparts; with device pointers also scatters the (key, oid)
    tuples into nparts contiguous runs (destination d at offset sum(counts[:d]))"""
    counts = np.zeros(nparts, dtype=np.uint64)
    rc = lib.b200_hash_partition(b._p, _ptr(s), nparts, keep_nil, C.c_void_p(keys_ptr) if keys_ptr else None,
                                 C.c_void_p(oids_ptr) if oids_ptr else None, _vp(counts))
    if rc != 1:
        _fail("hash_partition")
    return counts


_sig("b200_BATpartition", C.c_int, _PP, _PP, _P, _P, C.c_int, C.c_bool, C.c_void_p)


def BATpartition(b: BAT, s: BAT | None, nparts: int, keep_nil: bool = False):
    """(keys, oids, counts): the (key, oid) tuples of b (restricted to s) as `nparts` contiguous send runs"""
    counts = np.zeros(nparts, dtype=np.uint64)
    pk, po = _P(), _P()
    if lib.b200_BATpartition(C.byref(pk), C.byref(po), b._p, _ptr(s), nparts, keep_nil, _vp(counts)) != 1:
        _fail("BATpartition")
    return BAT(pk), BAT(po), counts


# ------------------------------------------------------------------ multi-GPU: NCCL inside the library (comm.cu)
_sig("b200_comm_unique_id", C.c_int, C.c_void_p)
_sig("b200_comm_init", C.c_int, C.c_int, C.c_int, C.c_void_p)
_sig("b200_comm_rank", C.c_int)
_sig("b200_comm_world", C.c_int)
_sig("b200_comm_nccl_version", C.c_int)
_sig("b200_comm_destroy", None)
_sig("b200_comm_pack", C.c_int, C.c_int, C.POINTER(_P), C.POINTER(_P), C.c_uint64)
_sig("b200_comm_exchange", C.c_int, C.c_int, C.POINTER(_P), C.c_void_p, C.POINTER(_P), C.c_void_p)


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    if lib.b200_comm_unique_id(buf) != 0:
        _fail("comm_unique_id")
    return buf.raw


def comm_init(rank: int, world: int, uid: bytes | None) -> None:
    if lib.b200_comm_init(rank, world, uid) != 0:
        _fail("comm_init")


def comm_init_torch(dist) -> None:
    """bootstrap through an initialised torch.distributed group: rank 0's NCCL id is broadcast once"""
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [comm_unique_id() if rank == 0 and world > 1 else None]
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    comm_init(rank, world, box[0])


def comm_world() -> int:
    return int(lib.b200_comm_world())


def comm_rank() -> int:
    return int(lib.b200_comm_rank())


def comm_destroy() -> None:
    lib.b200_comm_destroy()


def comm_pack(bats, max_rows: int = 256):
    """mat.pack across the ranks: list of small BATs -> list of their rank-order concatenations"""
    k = len(bats)
    inp = (_P * k)(*[b._p for b in bats])
    out = (_P * k)()
    if lib.b200_comm_pack(k, inp, out, max_rows) != 1:
        _fail("comm_pack")
    return [BAT(out[i]) for i in range(k)]


def comm_exchange(bats, send_counts):
    """all-to-all of parallel columns laid out as world runs -> (received columns, recv_counts)"""
    k = len(bats)
    inp = (_P * k)(*[b._p for b in bats])
    out = (_P * k)()
    sc = np.ascontiguousarray(send_counts, dtype=np.uint64)
    rc = np.zeros(len(sc), dtype=np.uint64)
    if lib.b200_comm_exchange(k, inp, _vp(sc), out, _vp(rc)) != 1:
        _fail("comm_exchange")
    return [BAT(out[i]) for i in range(k)], rc


# ------------------------------------------------------------------ bench support
_sig("b200_timing_enable", None, C.c_int)
_sig("b200_timing_read", C.c_int, C.c_char_p, C.c_size_t)
_sig("b200_BAThostwrap", _P, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64)


_sig("b200_stage_cache", None, C.c_int)
_sig("b200_stage_cache_clear", None)


def stage_cache(on: bool = True) -> None:
    """keep staged copies of HOST heaps until stage_cache_clear(): a host column used by several
    operators of one query crosses PCIe once"""
    lib.b200_stage_cache(1 if on else 0)


def stage_cache_clear() -> None:
    lib.b200_stage_cache_clear()


def timing_enable(on: bool = True) -> None:
    lib.b200_timing_enable(1 if on else 0)


def timing_read() -> dict:
    """{kernel name: (launches, total_ms, input_bytes)} since the last read"""
    buf = C.create_string_buffer(1 << 16)
    lib.b200_timing_read(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms, by = line.split()
        out[name] = (int(n), float(ms), int(by))
    return out


def pinned_empty(n: int, dtype) -> np.ndarray:
    """numpy array over page-locked host memory (freed with the process)"""
    dt = np.dtype(dtype)
    p = lib.b200_pinned_alloc(max(1, n * dt.itemsize))
    if not p:
        _fail("pinned_alloc")
    buf = (C.c_char * (n * dt.itemsize)).from_address(p)
    return np.frombuffer(buf, dtype=dt, count=n)


def BAThostwrap(arr: np.ndarray, tt: int, hseq: int = 0, **props) -> BAT:
    """a BAT whose heap lives in HOST memory: operators stage it to the device on every use"""
    arr = np.ascontiguousarray(arr)
    b = BAT(lib.b200_BAThostwrap(_vp(arr), tt, len(arr), hseq), keep=arr)
    if props:
        b.setprops(**props)
    return b
