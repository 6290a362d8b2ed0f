"""per-operator wall time of config 1 with a device sync after each operator (diagnostic, not a bench):
the difference between an operator's wall time and its kernel time is launch + sync + allocation overhead"""
import time
import numpy as np
from monetdb_b200 import gdk as G, workload

T = G.TYPE
n = 59986052
k, v = workload.config1(0, n)
col = {"k": G.BATfromhost(k, T.int, nonil=True), "v": G.BATfromhost(v, T.lng, nonil=True)}
sync = G.lib.b200_sync if hasattr(G.lib, "b200_sync") else None
acc = {}
for rep in range(13):
    t = [time.perf_counter()]
    s = G.BATselect(col["k"], None, 100, 599, True, True, False, False); t.append(time.perf_counter())
    pk = G.BATproject(s, col["k"]); pv = G.BATproject(s, col["v"])
    len(G.BATselect(pk, None, 0, 0, True, True, False, False)); t.append(time.perf_counter())   # tiny select = a sync
    g, e, h = G.BATgroup(pk, None, None, None, None); t.append(time.perf_counter())
    sums = G.BATgroupsum(pv, g, e, None, T.hge, True); t.append(time.perf_counter())
    r = sums.numpy(); t.append(time.perf_counter())
    del s, pk, pv, g, e, h, sums; t.append(time.perf_counter())
    if rep >= 3:
        for name, a, b in zip(("select", "projects(+1 select of pk)", "group", "groupsum", "numpy", "free"), t, t[1:]):
            acc.setdefault(name, []).append((b - a) * 1e3)
for name, xs in acc.items():
    print(f"{name:28s} {np.median(xs):.4f} ms")
