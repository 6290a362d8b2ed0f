"""The data-format edges of the device tier: binary COPY INTO records
(sql_bincopyconvert.c) decoded / encoded on the device against the reference's own constructors
(date_create / daytime_create / timestamp_create of libbat, driven by the harness with the decoders' loops),
byte-swapped atoms against numpy's byteswap, the validations' SQLSTATE texts, and tail heap files
(HEAPload / HEAPsave, gdk_heap.c:729,860) written by one side and read by the other.  All bit-exact."""
import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import random_col, same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE


def date_records(rng, n):
    """valid dates, impossible ones (Feb 30, month 13, day 0, year out of range) and the nil record"""
    r = np.zeros((n, 4), dtype=np.uint8)
    r[:, 0] = rng.integers(0, 33, n)
    r[:, 1] = rng.integers(0, 14, n)
    year = rng.choice([-4713, -4712, -1, 0, 1, 1600, 1900, 1999, 2000, 2024, 2100, 9999, 32767], n).astype(np.int16)
    r[:, 2:4] = year.view(np.uint8).reshape(n, 2)
    r[rng.random(n) < 0.05] = 0xFF
    return r


def time_records(rng, n):
    r = np.zeros((n, 8), dtype=np.uint8)
    ms = rng.choice([0, 1, 999_999, 1_000_000, 2**31, 2**32 - 1], n).astype(np.uint32)
    good = rng.random(n) < 0.7
    ms[good] = rng.integers(0, 1_000_000, int(good.sum())).astype(np.uint32)
    r[:, 0:4] = ms.view(np.uint8).reshape(n, 4)
    r[:, 4] = rng.integers(0, 63, n)      # seconds: 60 is allowed, 61+ is not
    r[:, 5] = rng.integers(0, 62, n)
    r[:, 6] = rng.integers(0, 26, n)
    r[:, 7] = rng.integers(0, 256, n)     # padding is ignored
    r[rng.random(n) < 0.05] = 0xFF
    return r


@pytest.mark.parametrize("n", [0, 1, 1000, 300_007])
def test_temporal_decode_and_roundtrip(gdk, ref, n):
    rng = np.random.default_rng(n + 5)
    d, t = date_records(rng, n), time_records(rng, n)
    ts = np.concatenate([t, d], axis=1)
    for kind, tt, rec in ((0, T.date, d), (1, T.daytime, t), (2, T.timestamp, ts)):
        raw = rec.tobytes()
        want = ref.bincopy_decode(kind, raw, n)
        b = gdk.bincopy_decode(tt, raw, n, hseq=3)
        got = b.numpy()
        assert same_bits(got, want), (kind, got[:8], want[:8])
        p = b.props()
        nil = np.iinfo(want.dtype).min
        assert p["hseq"] == 3 and p["nil"] == bool((want == nil).any()) and p["nonil"] == (not (want == nil).any())
        # the way back: what the reference's encoders write for the decoded values
        assert gdk.bincopy_encode(b) == ref.bincopy_encode(kind, want)
        # foreign endianness: swap year / microseconds in the records, decode with byteswap
        sw = rec.copy()
        if kind != 1:
            o = 2 if kind == 0 else 10
            sw[:, [o, o + 1]] = sw[:, [o + 1, o]]
        if kind != 0:
            sw[:, [0, 1, 2, 3]] = sw[:, [3, 2, 1, 0]]
        assert same_bits(gdk.bincopy_decode(tt, sw.tobytes(), n, byteswap=True).numpy(), want)
        enc = np.frombuffer(gdk.bincopy_encode(b, byteswap=True), dtype=np.uint8).reshape(n, rec.shape[1]).copy()
        if kind != 1:
            o = 2 if kind == 0 else 10
            enc[:, [o, o + 1]] = enc[:, [o + 1, o]]
        if kind != 0:
            enc[:, [0, 1, 2, 3]] = enc[:, [3, 2, 1, 0]]
        assert enc.tobytes() == ref.bincopy_encode(kind, want)


@pytest.mark.parametrize("tt", [T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl])
def test_atoms_and_byteswap(gdk, tt):
    rng = np.random.default_rng(tt)
    n = 100_003
    a = random_col(rng, tt, n, nilfrac=0.02)
    b = gdk.bincopy_decode(tt, a.tobytes(), n)
    assert same_bits(b.numpy(), a)
    nils = G_has_nil(tt, a)
    assert b.props()["nil"] == nils and b.props()["nonil"] == (not nils)
    assert gdk.bincopy_encode(b) == a.tobytes()
    w = a.dtype.itemsize
    swapped = np.frombuffer(a.tobytes(), dtype=np.uint8).reshape(n, w)[:, ::-1].tobytes()
    assert same_bits(gdk.bincopy_decode(tt, swapped, n, byteswap=True).numpy(), a)
    assert gdk.bincopy_encode(b, byteswap=True) == swapped


def G_has_nil(tt, a):
    from tests.util import has_nil
    return has_nil(tt, a)


def test_validations(gdk):
    bits = np.array([0, 1, 0x80, 1, 0], dtype=np.uint8)
    b = gdk.bincopy_decode(T.bit, bits.tobytes(), 5)
    assert b.numpy().view(np.uint8).tolist() == bits.tolist() and b.props()["nil"]
    bits[3] = 7
    with pytest.raises(G.GDKError) as e:
        gdk.bincopy_decode(T.bit, bits.tobytes(), 5, filename="flags.bin")
    assert str(e.value).strip() == "22003!invalid boolean byte value 7 in flags.bin"
    dec = np.array([999, -999, np.iinfo(np.int32).min, 0], dtype=np.int32)        # DECIMAL(3): |v| < 1000, nil passes
    assert same_bits(gdk.bincopy_decode(T.int, dec.tobytes(), 4, width=3).numpy(), dec)
    for bad in (1000, -1000):
        dec[1] = bad
        with pytest.raises(G.GDKError) as e:
            gdk.bincopy_decode(T.int, dec.tobytes(), 4, width=3, filename="price.bin")
        assert str(e.value).strip() == "22003!decimal out of range in price.bin"
    with pytest.raises(G.GDKError, match="no fixed-width binary record"):
        gdk.bincopy_decode(G.TYPE.str, b"abc\0", 1)


@pytest.mark.parametrize("tt", [T.bte, T.int, T.lng, T.hge, T.dbl, T.date, T.timestamp])
def test_heap_files(gdk, ref, tt, tmp_path):
    rng = np.random.default_rng(tt + 100)
    n = 9_000_011 if tt == T.lng else 70_001          # lng: 72 MB, three staging chunks
    a = random_col(rng, tt, n, nilfrac=0.01)
    # a heap file as GDK has it: the tail bytes of a BAT made by the real GDK
    rb = ref.bat(a, tt, hseq=0)
    f = tmp_path / "col.tail"
    rb.numpy().tofile(f)
    b = gdk.HEAPload(f, tt, n, hseq=5)
    assert same_bits(b.numpy(), a) and b.props()["hseq"] == 5 and b.props()["type"] == tt
    # the loaded column is an ordinary device column
    if tt == T.lng:
        s = gdk.BATselect(b, None, 0, 500, True, True, False, False)
        rs = ref.BATselect(rb, None, 0, 500, True, True, False, False)
        assert same_bits(s.numpy() - 5, rs.numpy())
    # and back: what HEAPsave writes is what the reference's tail holds
    g = tmp_path / "out.tail"
    gdk.HEAPsave(b, g, dosync=True)
    assert g.read_bytes() == f.read_bytes()
    # a prefix of the file is a valid (shorter) heap; a file that is too short is refused
    assert same_bits(gdk.HEAPload(f, tt, n // 2).numpy(), a[: n // 2])
    with pytest.raises(G.GDKError, match="needed"):
        gdk.HEAPload(f, tt, n + 1)
    with pytest.raises(G.GDKError, match="cannot open"):
        gdk.HEAPload(tmp_path / "missing.tail", tt, 1)
