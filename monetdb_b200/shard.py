"""Row-range sharding of the hot path over the GPUs of one box.

Rank g of G holds rows [lo_g, hi_g) of every column of a table with ``hseqbase = lo_g``, so the
candidate lists, extents and min/max oids each rank produces are GLOBAL oids and the
concatenation of per-rank candidate lists in rank order is the global list.  select / project /
calc / compare need no exchange.  What does:

* scalar aggregates (Q6): one all-reduce of the partial hge, carried as four 32-bit digits in
  int64 lanes so the sum cannot overflow for any world size <= 2^31;
* low-cardinality GROUP BY (Q1): mergetable's two-phase plan (opt_mergetable.c:1204-1215) — every
  rank groups and aggregates its own rows, the <= ngrp-entry partial tables are all-gathered and
  every rank performs the same deterministic merge.  The global group id of a key is the rank of
  its smallest first-occurrence oid, which is exactly BATgroup's first-occurrence numbering
  (gdk_group.c:74-100); averages are recomputed from the merged exact (sum, count) with the
  reference's formula floor(S/n) + rem/n, then / 10^scale (gdk_aggr.c:1717-1753,1952-1957).

``X`` is any module with the GDK API mirror (monetdb_b200.gdk on the GPUs; the CPU tests pass the
reference driver to exercise this logic under gloo).  ``dist`` is torch.distributed or None.
"""
from __future__ import annotations

from . import pipelines


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """contiguous row range of `rank`: [rank*n/world, (rank+1)*n/world)"""
    return n * rank // world, n * (rank + 1) // world


def hge_int(a, i: int = 0) -> int:
    return (int(a["hi"][i]) << 64) | int(a["lo"][i])


def hge_digits(v: int) -> list[int]:
    """signed 128-bit value -> four int64 lanes holding 32-bit digits (top digit signed)"""
    u = v & ((1 << 128) - 1)
    d = [(u >> (32 * i)) & 0xFFFFFFFF for i in range(4)]
    if v < 0:
        d[3] -= 1 << 32
    return d


def hge_from_digits(d) -> int:
    return sum(int(x) << (32 * i) for i, x in enumerate(d))


def allreduce_hge(partial: int, dist=None, device=None) -> int:
    """global sum of one hge per rank (the Q6 merge)"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return partial
    import torch
    t = torch.tensor(hge_digits(partial), dtype=torch.int64, device=device)
    dist.all_reduce(t)
    return hge_from_digits(t.tolist())


def q6_sharded(X, cols, dist=None, device=None, counts=None) -> int:
    """config 2 on this rank's rows + the hge all-reduce"""
    return allreduce_hge(hge_int(pipelines.q6(X, cols, counts)), dist, device)


# ---------------------------------------------------------------------------- Q1
Q1_SUMS = ("sum_qty", "sum_base_price", "sum_disc_price", "sum_charge")


def q1_partial(X, cols) -> list[dict]:
    """phase 1: this rank's partial group table, one dict per local group"""
    out = {}
    res = pipelines.q1(X, cols, out)
    e = res["extents"]
    if len(e) == 0:
        return []
    # key values and GLOBAL first-occurrence oid of each local group
    rf = X.BATproject(e, out["rf"]).numpy()
    ls = X.BATproject(e, out["ls"]).numpy()
    first = X.BATproject(e, out["cand"]).numpy()
    sum_disc = X.BATgroupsum(out["disc"], out["groups"], e, None, pipelines.T_hge, True).numpy()
    cnt = res["count_order"].numpy()
    sums = {k: res[k].numpy() for k in Q1_SUMS}
    table = []
    for i in range(len(e)):
        row = {"rf": int(rf[i]), "ls": int(ls[i]), "first": int(first[i]), "n": int(cnt[i]),
               "sum_disc": hge_int(sum_disc, i)}
        for k in Q1_SUMS:
            row[k] = hge_int(sums[k], i)
        table.append(row)
    return table


def exact_avg(total: int, n: int, scale: int) -> float:
    """BATgroupavg on integers: floor(S/n) + rem/n as dbl, then / 10^scale"""
    if n == 0:
        return float("nan")
    a, rem = divmod(total, n)          # Python floors, rem in [0, n)
    r = float(a) + float(rem) / float(n)
    if scale:
        r /= float(10 ** scale)
    return r


def q1_merge(tables: list[list[dict]]) -> dict:
    """phase 2: identical on every rank.  Returns the Q1 result in BATgroup's group order."""
    merged: dict[tuple[int, int], dict] = {}
    for t in tables:
        for row in t:
            k = (row["rf"], row["ls"])
            m = merged.get(k)
            if m is None:
                merged[k] = dict(row)
                continue
            m["first"] = min(m["first"], row["first"])
            m["n"] += row["n"]
            m["sum_disc"] += row["sum_disc"]
            for s in Q1_SUMS:
                m[s] += row[s]
    groups = sorted(merged.values(), key=lambda r: r["first"])
    lim = (1 << 127) - 1
    for g in groups:
        for s in Q1_SUMS:
            if abs(g[s]) > lim:
                raise OverflowError("22003!overflow in sum aggregate.")
    return {
        "keys": [(g["rf"], g["ls"]) for g in groups],
        "extents": [g["first"] for g in groups],
        "count_order": [g["n"] for g in groups],
        **{s: [g[s] for g in groups] for s in Q1_SUMS},
        "avg_qty": [exact_avg(g["sum_qty"], g["n"], 2) for g in groups],
        "avg_price": [exact_avg(g["sum_base_price"], g["n"], 2) for g in groups],
        "avg_disc": [exact_avg(g["sum_disc"], g["n"], 2) for g in groups],
    }


def q1_sharded(X, cols, dist=None) -> dict:
    """config 3 on this rank's rows, all-gather of the partial tables, merge"""
    mine = q1_partial(X, cols)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return q1_merge([mine])
    tables = [None] * dist.get_world_size()
    dist.all_gather_object(tables, mine)
    return q1_merge(tables)


# ---------------------------------------------------------------------------- config 4: repartitioned hash join
_TORCH_OF = {3: "int8", 4: "int16", 5: "int32", 10: "int64", 6: "int64", 8: "float32", 9: "float64"}


def _partition_to_torch(X, col, world, torch, device):
    """hash-partition one side into `world` contiguous send runs held in torch tensors"""
    counts = X.hash_partition(col, None, world)
    total = int(counts.sum())
    tt = X._storage(col.ttype)
    keys = torch.empty(max(total, 1), dtype=getattr(torch, _TORCH_OF[tt]), device=device)
    oids = torch.empty(max(total, 1), dtype=torch.int64, device=device)
    torch.cuda.current_stream().synchronize()
    if total:
        c2 = X.hash_partition(col, None, world, keys.data_ptr(), oids.data_ptr())
        assert (c2 == counts).all()
    X.sync()
    return counts.astype("int64"), keys, oids, tt


def join_repartitioned(X, l, r, dist, torch, device, timings=None):
    """lineitem JOIN orders across the GPUs of one box: every rank hash-partitions its
    shard of both key columns by hash(key) mod world (b200_hash_partition), one NCCL all-to-all per
    side moves (key, global oid) tuples to the rank that owns the hash range, and each rank then
    builds and probes locally (BAThash + BATjoin).  Returns this rank's slice of the result as two
    aligned device BATs of GLOBAL oids (r1: rows of l, r2: rows of r); the concatenation over ranks is
    the join, in no particular order (BATjoin promises none, gdk_join.c:56-61)."""
    import time
    world = dist.get_world_size()

    def mark(name, t0):
        if timings is not None:
            torch.cuda.synchronize()
            timings[name] = (time.perf_counter() - t0) * 1e3      # the latest call

    t0 = time.perf_counter()
    sides = [_partition_to_torch(X, c, world, torch, device) for c in (l, r)]
    mark("partition_ms", t0)
    t0 = time.perf_counter()
    recv = []
    for counts, keys, oids, tt in sides:
        send_counts = torch.tensor(counts, dtype=torch.int64, device=device)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts)
        rc = recv_counts.tolist()
        sc = [int(x) for x in counts]
        nrecv = sum(rc)
        rk = torch.empty(max(nrecv, 1), dtype=keys.dtype, device=device)
        ro = torch.empty(max(nrecv, 1), dtype=torch.int64, device=device)
        dist.all_to_all_single(rk[:nrecv], keys[:sum(sc)], rc, sc)
        dist.all_to_all_single(ro[:nrecv], oids[:sum(sc)], rc, sc)
        recv.append((rk, ro, nrecv, tt))
    mark("exchange_ms", t0)
    t0 = time.perf_counter()
    (lk, lo, nl, tt), (rk, ro, nr, _) = recv
    res = _local_join(X, lk.data_ptr(), lo.data_ptr(), nl, rk.data_ptr(), ro.data_ptr(), nr, tt, keep=(lk, lo, rk, ro))
    X.sync()
    mark("join_ms", t0)
    return res


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
