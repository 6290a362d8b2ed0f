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
