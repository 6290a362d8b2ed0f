"""per-operator wall time (with a sync after every call) of the Q1 plan at a given row count: shows where the
step's time goes outside the timed kernels.  usage: exp_optime.py ROWS [twophase]"""
import sys, time, collections
sys.path.insert(0, ".")
import numpy as np
from monetdb_b200 import gdk as G, pipelines as P, workload as W, shard
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 600_037_902
two = len(sys.argv) > 2
G.init(0)
G.comm_init(0, 1, None)
T = G.TYPE
types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng, "returnflag": T.bte, "linestatus": T.bte}
pinned = {k: G.pinned_empty(rows, W.LINEITEM_DTYPES[k]) for k in types}
W.lineitem_into(0, rows, pinned)
dev = {k: G.BATfromhost(pinned[k], types[k], nonil=True) for k in types}
log = collections.OrderedDict()
class Wrap:
    def __getattr__(self, name):
        f = getattr(G, name)
        if not callable(f) or not (name.startswith("BAT") or name.startswith("comm_")):
            return f
        def g(*a, **k):
            G.sync(); t = time.perf_counter()
            r = f(*a, **k)
            G.sync(); dt = (time.perf_counter() - t) * 1e3
            key = name + ":" + (str(a[0]) if isinstance(a[0], str) else "")
            log.setdefault(key, []).append(dt)
            return r
        return g
X = Wrap()
run = (lambda x: shard.q1_sharded(x, dev, shard.device_pack(x))) if two else (lambda x: P.q1(x, dev))
for it in range(4):
    log.clear()
    G.sync(); t0 = time.perf_counter()
    r = run(X)
    G.sync(); tot = (time.perf_counter() - t0) * 1e3
    del r
print("total ms (synced after every call)", tot)
for k, v in log.items():
    print(f"{k:28s} n={len(v):2d} total={sum(v):8.3f} ms  each={[round(x,3) for x in v]}")
for it in range(5):
    G.sync(); t0 = time.perf_counter(); r = run(G); G.sync(); print("plain step ms", (time.perf_counter() - t0) * 1e3); del r
