"""Host-side multi-GPU logic on CPU: the two-phase plans of monetdb_b200/shard.py (phase 1 per rank, pack,
phase 2 as ordinary GDK calls on the packed BATs) run under torch.distributed/gloo with world_size 2, each rank
computing with the oracle (the reference GDK), compared with the unsharded reference run on all the rows."""
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


def _cols(R, lo, hi, nils=False):
    from monetdb_b200 import workload
    T = R.T
    types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
             "returnflag": T.bte, "linestatus": T.bte}
    h = workload.lineitem(lo, hi, cols=Q1_COLS, nils=nils)
    nonil = lambda c: not nils or c in ("returnflag", "linestatus")
    return {c: R.bat(h[c], types[c], hseq=lo, nonil=nonil(c), nil=not nonil(c)) for c in types}


def _worker(rank, world, port, n, nils, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import refgdk as R
    from monetdb_b200 import shard
    R.init(1)

    def gather(mine):
        tables = [None] * world
        dist.all_gather_object(tables, mine)
        return tables

    pack = shard.host_pack(R, gather)
    lo, hi = shard.shard_range(n, rank, world)
    cols = _cols(R, lo, hi, nils)
    q1 = shard.q1_host(shard.q1_sharded(R, cols, pack))
    q6 = shard.q6_sharded(R, cols, pack)
    if rank == 0:
        q.put((q1, q6))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("nils", [False, True])
def test_q1_q6_sharded_world2(ref, nils):
    from monetdb_b200 import pipelines, shard
    n, world = 120_000, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, nils, q)) for r in range(world)]
    for p in procs:
        p.start()
    q1, q6 = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    _check_against_unsharded(ref, q1, q6, n, nils)


def _check_against_unsharded(ref, q1, q6, n, nils):
    from monetdb_b200 import pipelines, shard
    cols = _cols(ref, 0, n, nils)
    out = {}
    want = pipelines.q1(ref, cols, out)
    assert q6 == shard.hge_int(pipelines.q6(ref, cols))
    hi = lambda a: [shard.hge_int(a, i) for i in range(len(a))]
    assert q1["count_order"] == want["count_order"].numpy().tolist()
    assert q1["histo"] == want["histo"].numpy().tolist()
    for s in shard.Q1_SUMS:
        assert q1[s] == hi(want[s].numpy()), s
    # extents are oids in the projected column; map them through the candidate list to global oids
    cand = out["cand"].numpy()
    assert q1["extents"] == cand[want["extents"].numpy().astype(np.int64)].tolist()
    for a in ("avg_qty", "avg_price", "avg_disc"):
        assert np.array(q1[a]).tobytes() == want[a].numpy().tobytes(), a        # bit-exact


def test_mitosis_slices_in_one_process(ref):
    """the reference arm of bench.py: per-slice phase 1, mat.pack in process, phase 2"""
    from monetdb_b200 import shard
    n, nsl = 150_000, 5
    tables = []
    sums = []
    for i in range(nsl):
        lo, hi = shard.shard_range(n, i, nsl)
        cols = _cols(ref, lo, hi)
        tables.append(shard.q1_phase1(ref, cols))
        sums.append(shard.q6_phase1(ref, cols))
    q1 = shard.q1_host(shard.q1_phase2(ref, shard.pack_tables(ref, tables)))
    q6 = shard.hge_int(shard.q6_phase2(ref, shard.pack_tables(ref, sums)))
    _check_against_unsharded(ref, q1, q6, n, False)


def test_shard_range_and_workload_generator():
    from monetdb_b200 import shard, workload
    assert shard.shard_range(10, 0, 3) == (0, 3) and shard.shard_range(10, 2, 3) == (6, 10)
    # the threaded C generator (libworkload.so) is bit-identical to the numpy definition, with and without nils
    for nils in (False, True):
        a = workload.lineitem(777, 777 + 200_000, Q1_COLS, nils)
        b = workload.lineitem_fast(777, 777 + 200_000, Q1_COLS, nils)
        for k in Q1_COLS:
            assert a[k].dtype == b[k].dtype and (a[k] == b[k]).all(), k
    blk = 6100
    for lo, hi in ((0, 3 * blk), (2 * blk, 5 * blk + 777)):
        a, b = workload.window(lo, hi, lens=(100, 1000, 5000)), workload.window_fast(lo, hi, lens=(100, 1000, 5000))
        for k in ("pkey", "ord", "val"):
            assert (a[k] == b[k]).all(), k
    if True:
        z = workload.lineitem(0, 100_000, ("discount",), nils=True)["discount"]
        assert 0.005 < (z == np.iinfo(np.int64).min).mean() < 0.02
