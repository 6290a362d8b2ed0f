import numpy as np, sys
sys.path.insert(0, '.')
from monetdb_b200 import gdk as G
from oracle import refgdk as R
from tests.util import random_col, both
G.init(0); R.init(1)
T = G.TYPE
for t1 in (T.bte, T.sht):
    rng = np.random.default_rng(t1)
    for n in (0, 1, 1000, 200003):
        for nilfrac in (0.0, 0.1):
            a = random_col(rng, t1, n, nilfrac=nilfrac, lo=-100, hi=100)
            b, rb = both(G, R, a, t1)
            for tp in (T.bte, T.sht, T.int, T.lng, T.hge, T.dbl):
                if tp < t1 and tp != T.dbl: continue
                for skip in (True, False):
                    try:
                        r = R.BATsum(tp, rb, None, skip, True); re = None
                    except R.RefError as e:
                        r = None; re = str(e).strip()
                    try:
                        o = G.BATsum(tp, b, None, skip, True); oe = None
                    except G.GDKError as e:
                        o = None; oe = str(e).strip()
                    if re != oe or (r is not None and r.tobytes() != o.tobytes()):
                        print("MISMATCH", t1, n, nilfrac, tp, skip, "ref", r, re, "ours", o, oe, "nils in a", int((a == np.iinfo(a.dtype).min).sum()), "first nil at", int(np.argmax(a == np.iinfo(a.dtype).min)) if nilfrac else None, a[:12])
print("done")
