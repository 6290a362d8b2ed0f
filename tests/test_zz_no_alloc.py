"""b200_select_lng_noalloc — the allocation-free range select (no counterpart in the reference)
— through the C-ABI against the existing entry point b200_BATselect: candidate
oids bit-exact, on two calls in a row that share the static scratch (the second call smaller than the first, so
stale mask words and chunk counts from the first call would show), plus the size limit.
Collected last on purpose (zz)."""
import ctypes as C

import numpy as np
import pytest

from monetdb_b200 import gdk as G

pytestmark = pytest.mark.gpu
T = G.TYPE


def run_noalloc(gdk, b, n, hseq, lo, hi, out, cnt):
    """inputs and outputs are device BATs made with the ordinary entry points; the call itself allocates nothing"""
    rc = gdk.lib.b200_select_lng_noalloc(C.c_void_p(b.base), n, hseq, lo, hi, C.c_void_p(out.base), C.c_void_p(cnt.base),
                                         C.c_void_p(gdk.lib.b200_stream()))
    assert rc == 0, gdk.lib.b200_errbuf()
    gdk.lib.b200_sync()
    k = np.empty(1, dtype=np.uint64)
    assert gdk.lib.b200_BATtohost(cnt._p, k.ctypes.data_as(C.c_void_p)) == 0
    oids = np.empty(n, dtype=np.uint64)
    if n:
        assert gdk.lib.b200_BATtohost(out._p, oids.ctypes.data_as(C.c_void_p)) == 0
    return oids[:int(k[0])], int(k[0])


def outputs(gdk, n):
    out = G.BAT(gdk.lib.b200_COLnew(0, T.oid, max(n, 1)))
    cnt = G.BAT(gdk.lib.b200_COLnew(0, T.oid, 1))
    # b200_BATtohost copies `count` items
    gdk.lib.b200_BATextend(out._p, max(n, 1))
    out._p.contents.count = n
    cnt._p.contents.count = 1
    return out, cnt


def test_two_calls_in_a_row_match_batselect(gdk):
    rng = np.random.default_rng(2024)
    nil = np.iinfo(np.int64).min
    cases = [(3_000_017, 7, -500, 700), (70_001, 123, 0, 10)]     # the second call is smaller and far more selective
    for n, hseq, lo, hi in cases:
        a = rng.integers(-1000, 1001, n).astype(np.int64)
        a[rng.random(n) < 0.01] = nil
        b = gdk.BATfromhost(a, T.lng, hseq=hseq)
        out, cnt = outputs(gdk, n)
        got, k = run_noalloc(gdk, b, n, hseq, lo, hi, out, cnt)
        want = gdk.BATselect(b, None, lo, hi, True, True, False, False).numpy()
        assert k == len(want)
        assert np.array_equal(got, want)
        # and against plain numpy, so the check does not rest on one implementation alone
        assert np.array_equal(got, np.flatnonzero((a != nil) & (a >= lo) & (a <= hi)).astype(np.uint64) + hseq)
    # same input again right away: the shared scratch keeps the last call's contents, the result must not change
    got2, k2 = run_noalloc(gdk, b, n, hseq, lo, hi, out, cnt)
    assert k2 == k and np.array_equal(got2, got)
    # nothing selected / everything selected / empty input
    got, k = run_noalloc(gdk, b, n, hseq, 5000, 6000, out, cnt)
    assert k == 0
    got, k = run_noalloc(gdk, b, n, hseq, -(1 << 62), 1 << 62, out, cnt)
    assert np.array_equal(got, np.flatnonzero(a != nil).astype(np.uint64) + hseq)
    got, k = run_noalloc(gdk, b, 0, hseq, 0, 1, out, cnt)
    assert k == 0


def test_limit_and_bad_arguments_are_refused(gdk):
    out, cnt = outputs(gdk, 8)
    b = gdk.BATfromhost(np.arange(8, dtype=np.int64), T.lng)
    st = C.c_void_p(gdk.lib.b200_stream())
    # one row beyond the static scratch: refused before anything is touched
    rc = gdk.lib.b200_select_lng_noalloc(C.c_void_p(b.base), G.NOALLOC_MAX_ROWS + 1, 0, 0, 1, C.c_void_p(out.base), C.c_void_p(cnt.base), st)
    assert rc == -2
    assert b"B200_NOALLOC_MAX_ROWS" in gdk.lib.b200_errbuf()
    gdk.lib.b200_clrerr()
    rc = gdk.lib.b200_select_lng_noalloc(C.c_void_p(b.base), 8, 0, np.iinfo(np.int64).min, 1, C.c_void_p(out.base), C.c_void_p(cnt.base), st)
    assert rc == -1 and b"nil bounds" in gdk.lib.b200_errbuf()
    gdk.lib.b200_clrerr()
    rc = gdk.lib.b200_select_lng_noalloc(None, 8, 0, 0, 1, C.c_void_p(out.base), C.c_void_p(cnt.base), st)
    assert rc == -1
    gdk.lib.b200_clrerr()


def test_exactly_the_maximum_size(gdk):
    n = G.NOALLOC_MAX_ROWS
    a = (np.arange(n, dtype=np.int64) * 2654435761) % 1000
    b = gdk.BATfromhost(a, T.lng)
    out, cnt = outputs(gdk, n)
    got, k = run_noalloc(gdk, b, n, 0, 10, 19, out, cnt)
    want = gdk.BATselect(b, None, 10, 19, True, True, False, False).numpy()
    assert k == len(want) and np.array_equal(got, want)
