"""Host-side multi-GPU logic on CPU: the row-range sharding + merge of monetdb_b200/shard.py run
under torch.distributed/gloo with world_size 2, each rank computing its partials with the oracle
(the reference GDK), compared with the unsharded reference run."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


Q1_COLS = ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus")


def _cols(R, lo, hi):
    from monetdb_b200 import workload
    T = R.T
    types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
             "returnflag": T.bte, "linestatus": T.bte}
    h = workload.lineitem(lo, hi, cols=Q1_COLS)
    return {c: R.bat(h[c], types[c], hseq=lo, nonil=True) for c in types}


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import refgdk as R
    from monetdb_b200 import shard
    R.init(1)
    lo, hi = shard.shard_range(n, rank, world)
    cols = _cols(R, lo, hi)
    q1 = shard.q1_sharded(R, cols, dist)
    q6 = shard.q6_sharded(R, cols, dist)
    if rank == 0:
        q.put((q1, q6))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_q1_q6_sharded_world2(ref):
    from monetdb_b200 import pipelines, shard
    n, world = 120_000, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    q1, q6 = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # the unsharded reference run
    cols = _cols(ref, 0, n)
    out = {}
    want = pipelines.q1(ref, cols, out)
    assert q6 == shard.hge_int(pipelines.q6(ref, cols))
    hi = lambda a: [shard.hge_int(a, i) for i in range(len(a))]
    assert q1["count_order"] == want["count_order"].numpy().tolist()
    for s in shard.Q1_SUMS:
        assert q1[s] == hi(want[s].numpy()), s
    # extents are oids in the projected column; map them through the candidate list to global oids
    cand = out["cand"].numpy()
    assert q1["extents"] == cand[want["extents"].numpy().astype(np.int64)].tolist()
    for a in ("avg_qty", "avg_price", "avg_disc"):
        assert np.array(q1[a]).tobytes() == want[a].numpy().tobytes(), a        # bit-exact


def test_hge_digits_roundtrip():
    from monetdb_b200 import shard
    for v in (0, 1, -1, (1 << 127) - 1, -(1 << 127) + 1, 34525887377650, -987654321987654321987):
        assert shard.hge_from_digits(shard.hge_digits(v)) == v
        # digit-wise sums of many partials stay exact
        assert shard.hge_from_digits([8 * d for d in shard.hge_digits(v)]) == 8 * v
    assert shard.shard_range(10, 0, 3) == (0, 3) and shard.shard_range(10, 2, 3) == (6, 10)
    assert shard.exact_avg(-7, 2, 0) == -3.5 and shard.exact_avg(7, 2, 2) == 0.035
