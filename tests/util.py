"""helpers shared by the parity tests: build the same column on both sides,
compare logical tails bit-exactly, and check that the properties we claim are true."""
import numpy as np

from monetdb_b200 import gdk as G

INT_TYPES = [G.TYPE.bte, G.TYPE.sht, G.TYPE.int, G.TYPE.lng, G.TYPE.hge]
FLT_TYPES = [G.TYPE.flt, G.TYPE.dbl]
ALL_TYPES = INT_TYPES + FLT_TYPES


def hge_arr(ints):
    return G.hge_from_ints(ints)


def random_col(rng, tt, n, nilfrac=0.0, lo=None, hi=None):
    """random values of type tt with a fraction of nils; returns the numpy array in storage form"""
    if tt in (G.TYPE.flt, G.TYPE.dbl):
        a = rng.uniform(lo if lo is not None else -1000, hi if hi is not None else 1000, n).astype(G.NP_OF[tt])
        if nilfrac:
            a[rng.random(n) < nilfrac] = np.nan
        return a
    if tt == G.TYPE.hge:
        l = lo if lo is not None else -(1 << 100)
        h = hi if hi is not None else (1 << 100)
        vals = [int(rng.integers(0, 1 << 62)) * (h - l) // (1 << 62) + l for _ in range(n)]
        if nilfrac:
            m = rng.random(n) < nilfrac
            vals = [-(1 << 127) if m[i] else v for i, v in enumerate(vals)]
        return G.hge_from_ints(vals)
    dt = G.NP_OF[tt]
    info = np.iinfo(dt)
    l = lo if lo is not None else max(info.min + 1, -1000)
    h = hi if hi is not None else min(info.max, 1000)
    a = rng.integers(l, h + 1, n, dtype=np.int64).astype(dt)
    if nilfrac:
        a[rng.random(n) < nilfrac] = info.min
    return a


def has_nil(tt, a):
    if tt in (G.TYPE.flt, G.TYPE.dbl):
        return bool(np.isnan(a).any())
    if tt == G.TYPE.hge:
        return bool(((a["lo"] == 0) & (a["hi"] == np.iinfo(np.int64).min)).any())
    if tt == G.TYPE.oid:
        return bool((a == G.oid_nil).any())
    return bool((a == np.iinfo(G.NP_OF[tt]).min).any())


def both(gdk, ref, arr, tt, hseq=0, **props):
    """the same column as a device BAT and a reference BAT; nonil/nil are set truthfully"""
    hn = has_nil(tt, arr) if len(arr) else False
    props.setdefault("nonil", not hn)
    props.setdefault("nil", hn)
    return gdk.BATfromhost(arr, tt, hseq=hseq, **props), ref.bat(arr, tt, hseq=hseq, **props)


def same_bits(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.dtype.itemsize == b.dtype.itemsize and a.shape == b.shape and a.tobytes() == b.tobytes()


def keyview(tt, a):
    """array usable with numpy comparisons (hge -> python ints as object array)"""
    if tt == G.TYPE.hge:
        return np.array(G.hge_to_ints(a), dtype=object)
    return a


def assert_props_true(bat, arr=None):
    """every property we claim must hold on the data (BATassertProps, gdk_bat.c:2699)"""
    p = bat.props()
    if arr is None:
        arr = bat.numpy()
    tt = p["type"] if p["type"] != G.TYPE.void else G.TYPE.oid
    n = len(arr)
    assert n == p["count"]
    if n == 0:
        return
    hn = has_nil(tt, arr)
    if p["nonil"]:
        assert not hn, "tnonil claimed but nil present"
    if p["nil"]:
        assert hn, "tnil claimed but no nil"
    v = keyview(tt, arr)
    if tt in (G.TYPE.flt, G.TYPE.dbl) or hn:
        return  # ordering with nils/NaN: nil sorts first in GDK; skip the strict check
    if p["sorted"]:
        assert all(v[i] <= v[i + 1] for i in range(min(n - 1, 5000))), "tsorted claimed"
    if p["revsorted"]:
        assert all(v[i] >= v[i + 1] for i in range(min(n - 1, 5000))), "trevsorted claimed"
    if p["key"] and n <= 100000:
        assert len(set(v.tolist())) == n, "tkey claimed"


def assert_same(ours, theirs, props=("count", "hseq"), exact_props=("nonil", "nil")):
    """bit-exact logical tail + agreed properties"""
    a, b = ours.numpy(), theirs.numpy()
    po, pt = ours.props(), theirs.props()
    for k in props:
        assert po[k] == pt[k], (k, po, pt)
    assert same_bits(a, b), (a[:10], b[:10], po, pt)
    for k in exact_props:
        assert po[k] == pt[k], (k, po, pt)
    assert_props_true(ours, a)
