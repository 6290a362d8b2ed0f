import cProfile, pstats, time, sys
sys.path.insert(0, ".")
import numpy as np
from monetdb_b200 import gdk as G, pipelines as P, workload as W
G.init(0)
n = 200000
cols = W.lineitem(0, n)
col = {k: G.BATfromhost(v, G.TYPE.date if k == "shipdate" else None, nonil=True) for k, v in cols.items()}
for _ in range(50):
    P.q6(G, col)
G.sync()
t = time.perf_counter()
for _ in range(1000):
    P.q6(G, col)
G.sync()
print("per q6 us", (time.perf_counter() - t) * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(1000):
    P.q6(G, col)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
