"""Row-range sharding of the hot path over the GPUs of one box.

Rank g of G holds rows [lo_g, hi_g) of every column of a table with ``hseqbase = lo_g``, so the
candidate lists, extents and min/max oids each rank produces are GLOBAL oids and the
concatenation of per-rank candidate lists in rank order is the global list.  select / project /
calc / compare need no exchange.  What does follows mergetable's two-phase plans
(opt_mergetable.c:1204-1215): phase 1 runs the operator chain on the rank's own rows and ends in a
small partial table; ``pack`` (mat.pack across the ranks: ``b200_comm_pack``, one ncclAllGather
inside the library) concatenates the ranks' tables in rank order; phase 2 is ordinary GDK calls on
the packed BATs, identical on every rank:

* scalar aggregates (Q6): per-rank ``BATgroupsum(x, g=NULL)`` (one row), pack, ``BATsum``;
* low-cardinality GROUP BY (Q1): per-rank keys / first-occurrence oid / partial sums / counts of
  each local group, pack, ``BATgroup`` on the packed keys — which reproduces BATgroup's
  first-occurrence numbering (gdk_group.c:74-100) because rank r's rows all precede rank r+1's and
  every rank's table is in local first-occurrence order — then ``BATgroupsum`` of the partials and
  the exact average from the merged (sum, count) with BATgroupavg's own closing formula
  (``BATcalcavgsum``; gdk_aggr.c:1741-1748,1952-1957);
* the non-broadcast hash join (config 4): ``BATpartition`` by hash(key) mod G, ``comm_exchange``
  (grouped ncclSend/ncclRecv inside the library), local ``BATjoin``, positions mapped back to the
  global oids the tuples carry.

``X`` is any module with the GDK API mirror: monetdb_b200.gdk on the GPUs; the CPU tests and the
reference arm of bench.py pass the reference driver, with a host ``pack`` (gloo all-gather, or the
in-process concatenation of mitosis slices).  Nothing here computes on column data itself.
"""
from __future__ import annotations

from . import pipelines
from .pipelines import T_hge, T_lng


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """contiguous row range of `rank`: [rank*n/world, (rank+1)*n/world)"""
    return n * rank // world, n * (rank + 1) // world


def hge_int(a, i: int = 0) -> int:
    return (int(a["hi"][i]) << 64) | int(a["lo"][i])


def device_pack(X, max_rows: int = 256):
    """pack = b200_comm_pack on the calling rank's library communicator"""
    return lambda bats: X.comm_pack(bats, max_rows)


def host_pack(R, gather=None):
    """pack for the reference driver: `gather(list of numpy arrays) -> list over ranks of such lists`
    (gloo all_gather_object in the CPU tests); the arrays are concatenated in rank order"""
    import numpy as np

    def pack(bats):
        mine = [(b.numpy(), b.props()["type"]) for b in bats]
        tables = gather(mine) if gather is not None else [mine]
        out = []
        for i in range(len(bats)):
            tt = mine[i][1]
            arr = np.concatenate([t[i][0] for t in tables])
            out.append(R.bat(arr, tt if tt != 0 else R.T.oid))
        return out
    return pack


def pack_tables(R, tables):
    """mat.pack of per-slice partial tables held in ONE process (the reference arm's mitosis slices)"""
    import numpy as np
    out = []
    for i in range(len(tables[0])):
        tt = tables[0][i].props()["type"]
        out.append(R.bat(np.concatenate([t[i].numpy() for t in tables]), tt if tt != 0 else R.T.oid))
    return out


# ---------------------------------------------------------------------------- Q6
def _scalar_bat(X, arr, tt):
    """mat.pack's view of a scalar aggregate: a one-row BAT.  On the CUDA path the row stays in host memory
    (BAThostwrap); comm_pack stages it with the record it builds anyway."""
    return X.BAThostwrap(arr, tt) if hasattr(X, "BAThostwrap") else X.bat(arr, tt)


def q6_phase1(X, cols, counts=None):
    """config 2 on this rank's rows, ending in the one-row partial sum(hge) (nil when nothing qualified)"""
    return [_scalar_bat(X, pipelines.q6(X, cols, counts), T_hge)]


def q6_phase2(X, packed):
    """BATsum over the ranks' partials -> 1-element hge array"""
    return X.BATsum(T_hge, packed[0], None, True, True)


def q6_sharded(X, cols, pack, counts=None) -> int:
    return hge_int(q6_phase2(X, pack(q6_phase1(X, cols, counts))))


# ---------------------------------------------------------------------------- Q1
Q1_SUMS = ("sum_qty", "sum_base_price", "sum_disc_price", "sum_charge")
Q1_PARTIAL = ("rf", "ls", "first", "histo", "sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "sum_disc",
              "cnt_qty", "cnt_price", "cnt_disc")


def q1_phase1(X, cols, out=None):
    """phase 1: this rank's partial group table as len(Q1_PARTIAL) small BATs, one row per local group in
    local first-occurrence order.  Averages are split into (sum, count of non-nil values) as mergetable
    does; the three per-rank BATgroupavg passes of the unsharded plan are not needed."""
    o = {}
    pipelines.q1_groups_and_calcs(X, cols, o)
    g, e, h = o["g"], o["e"], o["h"]
    gsum = lambda b: X.BATgroupsum(b, g, e, None, T_hge, True)
    # count(col) on a column without nils is count(*): the histogram BATgroup already produced
    gcnt = lambda b: h if b.props()["nonil"] else X.BATgroupcount(b, g, e, None, T_lng, True)
    table = [
        X.BATproject(e, o["rf"]), X.BATproject(e, o["ls"]), X.BATproject(e, o["cand"]), h,
        gsum(o["qty"]), gsum(o["price"]), gsum(o["dp"]), gsum(o["ch"]), gsum(o["disc"]),
        gcnt(o["qty"]), gcnt(o["price"]), gcnt(o["disc"]),
    ]
    if out is not None:
        out.update(nsel=len(o["cand"]))
    return table


def q1_phase2(X, packed) -> dict:
    """phase 2, identical on every rank: the Q1 result (BATs, BATgroup's group order) from the packed tables"""
    P = dict(zip(Q1_PARTIAL, packed))
    g1, e1, h1 = X.BATgroup(P["rf"], None, None, None, None)
    g, e, h = X.BATgroup(P["ls"], None, g1, e1, h1)
    tot = lambda name, tp: X.BATgroupsum(P[name], g, e, None, tp, True)
    res = {k: tot(k, T_hge) for k in Q1_SUMS}
    n_qty, n_price, n_disc = tot("cnt_qty", T_lng), tot("cnt_price", T_lng), tot("cnt_disc", T_lng)
    res.update(
        avg_qty=X.BATcalcavgsum(res["sum_qty"], n_qty, 2),
        avg_price=X.BATcalcavgsum(res["sum_base_price"], n_price, 2),
        avg_disc=X.BATcalcavgsum(tot("sum_disc", T_hge), n_disc, 2),
        count_order=n_qty,
        extents=X.BATproject(e, P["first"]),
        histo=tot("histo", T_lng),
        rf=X.BATproject(e, P["rf"]), ls=X.BATproject(e, P["ls"]),
    )
    return res


def q1_sharded(X, cols, pack, out=None) -> dict:
    """config 3 on this rank's rows, pack, phase 2"""
    return q1_phase2(X, pack(q1_phase1(X, cols, out)))


def q1_host(res: dict) -> dict:
    """a Q1 result (sharded or not) as plain Python values, for comparisons"""
    out = {}
    for k, v in res.items():
        a = v.numpy()
        if a.dtype.names:                         # hge as (lo, hi)
            out[k] = [hge_int(a, i) for i in range(len(a))]
        else:
            out[k] = a.tolist()
    return out


# ---------------------------------------------------------------------------- config 4: repartitioned hash join
def join_repartitioned(X, l, r, timings=None):
    """lineitem JOIN orders across the GPUs of one box: every rank hash-partitions its
    shard of both key columns by hash(key) mod world (BATpartition), one grouped NCCL send/recv per side
    (comm_exchange) moves the (key, global oid) tuples to the rank that owns the hash range, and each rank
    builds and probes locally (BATjoin).  Returns this rank's slice of the result as two aligned device
    BATs of GLOBAL oids (r1: rows of l, r2: rows of r); the concatenation over ranks is the join, in no
    particular order (BATjoin promises none, gdk_join.c:56-61)."""
    import time
    world = X.comm_world()

    def mark(name, t0):
        if timings is not None:
            X.sync()
            timings[name] = (time.perf_counter() - t0) * 1e3      # the latest call

    t0 = time.perf_counter()
    lk, lo, lc = X.BATpartition(l, None, world)
    rk, ro, rc = X.BATpartition(r, None, world)
    mark("partition_ms", t0)
    t0 = time.perf_counter()
    (lk2, lo2), _ = X.comm_exchange([lk, lo], lc)
    (rk2, ro2), _ = X.comm_exchange([rk, ro], rc)
    mark("exchange_ms", t0)
    if timings is not None:
        w = X.NP_OF[X._storage(lk.ttype)].itemsize + 8
        me = X.comm_rank()
        timings["exchange_bytes_sent"] = int((lc.sum() - lc[me]) * w + (rc.sum() - rc[me]) * w)
    t0 = time.perf_counter()
    r1, r2 = X.BATjoin(lk2, rk2, None, None, False)
    g1 = X.BATproject(r1, lo2)
    g2 = X.BATproject(r2, ro2)
    mark("join_ms", t0)
    return g1, g2


# ---------------------------------------------------------------------------- high-cardinality GROUP BY
def groupby_sum_repartitioned(X, key, val, tp=T_hge):
    """GROUP BY key, SUM(val) when the number of groups is too large for a packed partial table:
    every rank hash-partitions its (key, global oid) tuples by hash(key) mod world (BATpartition), fetches the
    values of those rows (BATproject through the oids), one grouped NCCL send/recv moves keys, oids and values to
    the rank owning the hash range (comm_exchange), and that rank groups and aggregates locally.  Every key
    lives on exactly one rank, so the ranks' results are disjoint and complete; they stay distributed:
    returns (keys, first, sums, counts) of this rank's groups, where `first` is the smallest GLOBAL oid of the
    group — BATgroup's first occurrence, so the concatenation over the ranks ordered by `first` is the
    unsharded BATgroup numbering.  nil keys are kept (they form a group, gdk_group.c:56-60)."""
    world = X.comm_world()
    k, oids, counts = X.BATpartition(key, None, world, keep_nil=True)
    v = X.BATproject(oids, val)
    (k2, o2, v2), _ = X.comm_exchange([k, oids, v], counts)
    g, e, h = X.BATgroup(k2, None, None, None, None)
    return (X.BATproject(e, k2), X.BATproject(X.BATgroupmin(o2, g, e, None), o2), X.BATgroupsum(v2, g, e, None, tp, True),
            X.BATgroupcount(v2, g, e, None, T_lng, True))


def _local_join(X, lk_ptr, lo_ptr, nl, rk_ptr, ro_ptr, nr, tt, keep=None):
    """BATjoin of received (key, oid) runs; positions are mapped back to the global oids they carry"""
    T = X.TYPE
    lkb = X.BATwrap(lk_ptr, tt, nl, 0, keep=keep)
    rkb = X.BATwrap(rk_ptr, tt, nr, 0, keep=keep)
    r1, r2 = X.BATjoin(lkb, rkb, None, None, False)
    g1 = X.BATproject(r1, X.BATwrap(lo_ptr, T.oid, nl, 0, keep=keep))
    g2 = X.BATproject(r2, X.BATwrap(ro_ptr, T.oid, nr, 0, keep=keep))
    return g1, g2


def join_partitioned_single_gpu(X, l, r, nparts):
    """the same plan on ONE GPU with the exchange replaced by slicing the send runs: partition both
    sides into `nparts` hash ranges, join range d of l with range d of r, concatenate.  Used by the
    parity tests (no NCCL needed) to pin the partitioning + local-join logic against BATjoin."""
    import numpy as np
    T = X.TYPE
    out1, out2 = [], []
    parts = []
    for col in (l, r):
        counts = X.hash_partition(col, None, nparts)
        total = int(counts.sum())
        tt = X._storage(col.ttype)
        keys = X.COLnew(0, tt, max(total, 1))
        oids = X.COLnew(0, T.oid, max(total, 1))
        if total:
            X.hash_partition(col, None, nparts, keys.base, oids.base)
        parts.append((counts, keys, oids, tt))
    (lc, lkeys, loids, tt), (rc, rkeys, roids, _) = parts
    w = X.NP_OF[tt].itemsize
    lo_off = ro_off = 0
    for d in range(nparts):
        nl, nr = int(lc[d]), int(rc[d])
        if nl and nr:
            g1, g2 = _local_join(X, lkeys.base + lo_off * w, loids.base + lo_off * 8, nl,
                                 rkeys.base + ro_off * w, roids.base + ro_off * 8, nr, tt, keep=(lkeys, loids, rkeys, roids))
            out1.append(g1.numpy())
            out2.append(g2.numpy())
        lo_off += nl
        ro_off += nr
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0, dtype=np.uint64)
    return cat(out1), cat(out2)
