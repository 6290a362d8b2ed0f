import numpy as np, sys
sys.path.insert(0, '.')
from monetdb_b200 import gdk as G
from oracle import refgdk as R
G.init(0); R.init(1)
rng = np.random.default_rng(3)
n = 20011
w = rng.integers(0, 1 << 32, (n + 31) // 32, dtype=np.uint64).astype(np.uint32)
a, ra = G.BATmaskedcands(0, n, w, True), R.maskedcands(0, n, w, True)
bits = np.unpackbits(w.view(np.uint8), bitorder='little')[:n]
exp_a = np.nonzero(bits)[0].astype(np.uint64)
print("a: ours", len(a.numpy()), "ref", len(ra.numpy()), "numpy", len(exp_a), np.array_equal(a.numpy(), exp_a), np.array_equal(ra.numpy(), exp_a))
lo, cnt = n // 7, n // 2
b, rb = G.BATdense(0, lo, cnt), R.dense(0, lo, cnt)
exp = np.union1d(exp_a, np.arange(lo, lo + cnt, dtype=np.uint64))
o, r = G.BATmergecand(a, b).numpy(), R.BATmergecand(ra, rb).numpy()
print("merge: ours", len(o), "ref", len(r), "numpy", len(exp), np.array_equal(o, exp), np.array_equal(r, exp))
if not np.array_equal(r, exp):
    print("ref missing", np.setdiff1d(exp, r), "ref extra", np.setdiff1d(r, exp))
if not np.array_equal(o, exp):
    print("ours missing", np.setdiff1d(exp, o), "ours extra", np.setdiff1d(o, exp))
