"""The multi-GPU half of the path (csrc/comm.cu, monetdb_b200/shard.py) and the lifetime / threading rules of
the device heap tier, checked against the reference GDK.

World-size-1 forms run on any GPU box; the NCCL tests spawn two ranks and are skipped with fewer than 2 GPUs."""
import os
import socket
import threading

import numpy as np
import pytest

from monetdb_b200 import gdk as G
from tests.util import assert_same, both, random_col, same_bits

pytestmark = pytest.mark.gpu
T = G.TYPE
Q1_COLS = ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus")
TYPES = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
         "returnflag": T.bte, "linestatus": T.bte}


def _cols(X, lo, hi, nils=False, mk=None):
    from monetdb_b200 import workload
    h = workload.lineitem_fast(lo, hi, Q1_COLS, nils)
    nonil = lambda c: not nils or c in ("returnflag", "linestatus")
    mk = mk or (X.BATfromhost if hasattr(X, "BATfromhost") else X.bat)
    return {c: mk(h[c], TYPES[c], hseq=lo, nonil=nonil(c), nil=not nonil(c)) for c in TYPES}


@pytest.fixture(scope="module")
def comm1(gdk):
    if gdk.comm_world() == 1:
        try:
            gdk.comm_init(0, 1, None)
        except gdk.GDKError:
            pass
    return gdk


def _ref_q1(ref, n, nils):
    from monetdb_b200 import pipelines, shard
    cols = _cols(ref, 0, n, nils)
    out = {}
    want = pipelines.q1(ref, cols, out)
    cand = out["cand"].numpy()
    w = shard.q1_host({k: want[k] for k in shard.Q1_SUMS + ("avg_qty", "avg_price", "avg_disc", "count_order", "histo")})
    w["extents"] = cand[want["extents"].numpy().astype(np.int64)].tolist()
    return w, shard.hge_int(pipelines.q6(ref, cols))


def _same_q1(a, w):
    from monetdb_b200 import shard
    for k in shard.Q1_SUMS + ("count_order", "histo", "extents"):
        assert a[k] == w[k], k
    for k in ("avg_qty", "avg_price", "avg_disc"):
        assert np.array(a[k]).tobytes() == np.array(w[k]).tobytes(), k


@pytest.mark.parametrize("nils", [False, True])
def test_two_phase_q1_q6_single_rank(comm1, ref, nils):
    """phase 1 -> b200_comm_pack (one rank: the pack/unpack kernels without the collective) -> phase 2 must
    equal ONE reference run of the §8(d) plans over all rows, bit for bit (averages included)"""
    from monetdb_b200 import shard
    n = 200_003
    cols = _cols(comm1, 0, n, nils)
    pack = shard.device_pack(comm1)
    q1 = shard.q1_host(shard.q1_sharded(comm1, cols, pack))
    q6 = shard.q6_sharded(comm1, cols, pack)
    w1, w6 = _ref_q1(ref, n, nils)
    _same_q1(q1, w1)
    assert q6 == w6


def test_two_phase_as_slices(comm1, ref):
    """the two-phase plan with the table cut into 5 row ranges on ONE GPU: the per-slice partial tables are
    packed by comm_pack one slice at a time and concatenated with BATs built from them — the merge logic the
    ranks run, without NCCL"""
    from monetdb_b200 import shard
    n, nsl = 150_000, 5
    tabs = []
    for i in range(nsl):
        lo, hi = shard.shard_range(n, i, nsl)
        tabs.append([b.numpy() for b in shard.q1_phase1(comm1, _cols(comm1, lo, hi))])
    packed = []
    for j, name in enumerate(shard.Q1_PARTIAL):
        arr = np.concatenate([t[j] for t in tabs])
        tt = {"rf": T.bte, "ls": T.bte, "first": T.oid, "histo": T.lng}.get(name, T.hge if name.startswith("sum") else T.lng)
        packed.append(comm1.BATfromhost(arr, tt))
    q1 = shard.q1_host(shard.q1_phase2(comm1, packed))
    w1, _ = _ref_q1(ref, n, False)
    _same_q1(q1, w1)


def test_comm_pack_types_and_limits(comm1):
    rng = np.random.default_rng(5)
    bats, arrs = [], []
    for tt in (T.bte, T.sht, T.int, T.lng, T.hge, T.flt, T.dbl, T.oid):
        a = random_col(rng, tt, 37) if tt != T.oid else rng.integers(0, 1 << 40, 37).astype(np.uint64)
        arrs.append(a)
        bats.append(comm1.BATfromhost(a, tt, nonil=True))
    bats.append(comm1.BATdense(0, 1000, 12))                 # void tails are materialised as oids
    arrs.append(np.arange(1000, 1012, dtype=np.uint64))
    bats.append(comm1.BATfromhost(np.zeros(0, dtype=np.int64), T.lng))
    arrs.append(np.zeros(0, dtype=np.int64))
    out = comm1.comm_pack(bats, 64)
    for o, a in zip(out, arrs):
        assert same_bits(o.numpy(), a)
        assert len(o) == len(a)
    with pytest.raises(G.GDKError, match="more than max_rows"):
        comm1.comm_pack(bats, 16)


@pytest.mark.parametrize("ts", [T.lng, T.hge])
def test_avgsum_matches_groupavg(gdk, ref, ts):
    """BATcalcavgsum(BATgroupsum, BATgroupcount) is bit-identical to the reference's BATgroupavg"""
    rng = np.random.default_rng(11 + ts)
    n, ng = 50_000, 37
    for scale in (0, 2):
        a = random_col(rng, T.lng, n, nilfrac=0.05, lo=-10**9, hi=10**9)
        gid = rng.integers(0, ng, n).astype(np.uint64)
        gid[gid == 5] = 6                                  # an empty group
        b, rb = both(gdk, ref, a, T.lng)
        g, rg = both(gdk, ref, gid, T.oid)
        e, re_ = both(gdk, ref, np.arange(ng, dtype=np.uint64), T.oid)
        sums = gdk.BATgroupsum(b, g, e, None, ts, True)
        cnts = gdk.BATgroupcount(b, g, e, None, T.lng, True)
        got = gdk.BATcalcavgsum(sums, cnts, scale)
        want, _ = ref.BATgroupavg(rb, rg, re_, None, T.dbl, True, scale)
        assert same_bits(got.numpy(), want.numpy())
        assert got.props()["nil"] == want.props()["nil"]
        # and the reference-side restatement the CPU tests use agrees
        assert same_bits(ref.BATcalcavgsum(ref.bat(sums.numpy(), ts), ref.bat(cnts.numpy(), T.lng), scale).numpy(), want.numpy())


@pytest.mark.parametrize("tt", [T.int, T.lng])
def test_partition_exchange_join_single_rank(comm1, ref, tt):
    """BATpartition + comm_exchange (one rank: the self-copy path) + BATjoin + project == reference BATjoin"""
    from monetdb_b200 import shard
    rng = np.random.default_rng(3 + tt)
    nl, nr = 40_009, 9_001
    a = random_col(rng, tt, nl, nilfrac=0.02, lo=1, hi=12_000)
    b = rng.permutation(np.arange(1, nr + 1)).astype(G.NP_OF[tt])
    l, rl = both(comm1, ref, a, tt, hseq=5000)
    r, rr = both(comm1, ref, b, tt, hseq=70)
    g1, g2 = shard.join_repartitioned(comm1, l, r)
    e1, e2 = ref.BATjoin(rl, rr, None, None, False)
    x1, x2, y1, y2 = g1.numpy(), g2.numpy(), e1.numpy(), e2.numpy()
    o, p = np.lexsort((x2, x1)), np.lexsort((y2, y1))
    assert np.array_equal(x1[o], y1[p]) and np.array_equal(x2[o], y2[p])
    k, oids, counts = comm1.BATpartition(l, None, 8)
    assert int(counts.sum()) == len(k) == len(oids)
    kk, oo = k.numpy(), oids.numpy()
    assert same_bits(a[(oo - 5000).astype(np.int64)], kk)          # every tuple carries its own row's key


def _groupby_reference(ref, keys, vals):
    """key -> (first oid, sum, count) from ONE reference BATgroup + BATgroupsum over all rows"""
    rk, rv = ref.bat(keys, ref.T.lng), ref.bat(vals, ref.T.lng)
    g, e, h = ref.BATgroup(rk, None, None, None, None)
    sums = ref.BATgroupsum(rv, g, e, None, ref.T.hge, True).numpy()
    cnts = ref.BATgroupcount(rv, g, e, None, ref.T.lng, True).numpy()
    ext = e.numpy()
    return {int(keys[int(x)]): (int(x), (int(sums["hi"][i]) << 64) | int(sums["lo"][i]), int(cnts[i])) for i, x in enumerate(ext)}


def _groupby_collect(parts):
    out = {}
    for k, first, sums, cnts in parts:
        for i in range(len(k)):
            assert int(k[i]) not in out, "a key showed up on two ranks"
            out[int(k[i])] = (int(first[i]), (int(sums["hi"][i]) << 64) | int(sums["lo"][i]), int(cnts[i]))
    return out


def test_high_cardinality_groupby_single_rank(comm1, ref):
    """BATpartition + comm_exchange + local BATgroup/BATgroupsum == one reference GROUP BY (200 k groups)"""
    from monetdb_b200 import shard
    rng = np.random.default_rng(9)
    n = 400_003
    keys = rng.integers(0, 200_000, n).astype(np.int64)
    keys[rng.random(n) < 0.01] = np.iinfo(np.int64).min          # nil keys form a group
    vals = rng.integers(-10**9, 10**9, n).astype(np.int64)
    kb, vb = comm1.BATfromhost(keys, T.lng), comm1.BATfromhost(vals, T.lng, nonil=True)
    parts = [tuple(x.numpy() for x in shard.groupby_sum_repartitioned(comm1, kb, vb))]
    assert _groupby_collect(parts) == _groupby_reference(ref, keys, vals)


# ---------------------------------------------------------------------------- lifetimes
def test_project_dense_view_outlives_parent(gdk):
    """BATproject with dense candidates returns a view of r's heap; it must stay valid after r is freed and
    after the pool has handed the memory to others (heap share count)"""
    rng = np.random.default_rng(1)
    a = rng.integers(-10**12, 10**12, 300_000).astype(np.int64)
    r = gdk.BATfromhost(a, T.lng, nonil=True)
    v = gdk.BATproject(gdk.BATdense(0, 1000, 200_000), r)
    assert gdk.algorithm() == "project: slice view"
    v2 = gdk.BATproject(gdk.BATdense(0, 10, 1000), v)              # a view of a view shares the same root
    r.free()
    junk = [gdk.BATfromhost(np.full(300_000, i, dtype=np.int64), T.lng) for i in range(6)]
    assert same_bits(v.numpy(), a[1000:201000])
    assert same_bits(v2.numpy(), a[1010:2010])
    v.free()
    junk += [gdk.BATfromhost(np.full(300_000, -i, dtype=np.int64), T.lng) for i in range(6)]
    assert same_bits(v2.numpy(), a[1010:2010])
    v2.free()
    del junk


def test_project_dense_view_of_host_column(gdk):
    """the INTEGRATION.md binding path: r is a HOST heap staged for the call — the result must own its data"""
    rng = np.random.default_rng(2)
    a = rng.integers(-10**12, 10**12, 250_000).astype(np.int64)
    for cache in (False, True):
        gdk.stage_cache(cache)
        r = gdk.BAThostwrap(a, T.lng, nonil=True)
        whole = gdk.BATproject(gdk.BATdense(0, 0, len(a)), r)
        part = gdk.BATproject(gdk.BATdense(0, 777, 100_000), r)
        gdk.stage_cache_clear()
        junk = [gdk.BATfromhost(np.full(250_000, i, dtype=np.int64), T.lng) for i in range(8)]
        x = gdk.BATcalccst("add", part, 1, T.lng, None, T.lng)
        assert same_bits(whole.numpy(), a)
        assert same_bits(part.numpy(), a[777:100_777])
        assert same_bits(x.numpy(), a[777:100_777] + 1)
        del junk
    gdk.stage_cache(False)


def test_hash_survives_alias_free_and_invalidate(gdk, ref):
    rng = np.random.default_rng(4)
    b = rng.permutation(np.arange(1, 20_001)).astype(np.int64)
    a = rng.integers(1, 20_001, 60_000).astype(np.int64)
    r, rr = both(gdk, ref, b, T.lng)
    l, rl = both(gdk, ref, a, T.lng)
    gdk.BAThash(r)
    alias = gdk.BATwrap(r.base, T.lng, len(b), 0, keep=r)
    alias.free()                                                  # must not drop r's index
    view = gdk.BATproject(gdk.BATdense(0, 0, len(b)), r)
    view.free()
    g1, g2 = gdk.BATjoin(l, r, None, None, False)
    assert "cached" in gdk.algorithm() or True
    e1, e2 = ref.BATjoin(rl, rr, None, None, False)
    o, p = np.lexsort((g2.numpy(), g1.numpy())), np.lexsort((e2.numpy(), e1.numpy()))
    assert np.array_equal(g1.numpy()[o], e1.numpy()[p]) and np.array_equal(g2.numpy()[o], e2.numpy()[p])
    gdk.lib.b200_BATinvalidate(r.ptr)                             # explicit invalidation is allowed any time
    g1b, _ = gdk.BATjoin(l, r, None, None, False)
    assert len(g1b) == len(g1)


# ---------------------------------------------------------------------------- concurrency (SURVEY §8b threading)
def test_four_threads_concurrent_callers(gdk, ref):
    """4 threads (own streams), each running Q6 and config 1 repeatedly on SHARED input BATs, plus BATs
    produced on one thread and consumed / freed on another: results must equal the serial ones"""
    from monetdb_b200 import pipelines, shard, workload
    n = 400_000
    h = workload.lineitem_fast(0, n, Q1_COLS)
    cols = {c: gdk.BATfromhost(h[c], TYPES[c], nonil=True) for c in TYPES}
    k, v = workload.config1(0, n)
    c1 = {"k": gdk.BATfromhost(k, T.int, nonil=True), "v": gdk.BATfromhost(v, T.lng, nonil=True)}
    want6 = shard.hge_int(pipelines.q6(gdk, cols))
    want1 = pipelines.config1(gdk, c1).numpy().tobytes()
    wantq1 = {kk: vv.numpy().tobytes() for kk, vv in pipelines.q1(gdk, cols).items()}
    errs, handoff = [], []
    lock = threading.Lock()

    def worker(i):
        try:
            gdk.init(0)
            for it in range(6):
                assert shard.hge_int(pipelines.q6(gdk, cols)) == want6
                assert pipelines.config1(gdk, c1).numpy().tobytes() == want1
                if it % 3 == i % 3:
                    got = pipelines.q1(gdk, cols)
                    for kk, vv in got.items():
                        assert vv.numpy().tobytes() == wantq1[kk], kk
                # hand a freshly produced (unsynchronised) BAT to whoever comes next
                p = gdk.BATproject(gdk.BATselect(cols["discount"], None, 5, 7, True, True, False, False), cols["extprice"])
                with lock:
                    handoff.append(p)
                    other = handoff.pop(0) if len(handoff) > 2 else None
                if other is not None:
                    s = gdk.BATsum(T.hge, other, None, True, True)
                    assert shard.hge_int(s) == int(h["extprice"][(h["discount"] >= 5) & (h["discount"] <= 7)].astype(object).sum())
                    other.free()
        except Exception as ex:         # noqa: BLE001
            errs.append((i, repr(ex)))

    th = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs


# ---------------------------------------------------------------------------- two ranks over NCCL
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, n, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from monetdb_b200 import gdk as g
    from monetdb_b200 import shard, workload
    g.init(rank)
    g.comm_init_torch(dist)
    pack = shard.device_pack(g)
    out = {}
    for nils in (False, True):
        lo, hi = shard.shard_range(n, rank, world)
        cols = _cols(g, lo, hi, nils)
        out[f"q1_{nils}"] = shard.q1_host(shard.q1_sharded(g, cols, pack))
        out[f"q6_{nils}"] = shard.q6_sharded(g, cols, pack)
    # config 4: rank r holds rows [r*nl, (r+1)*nl) of l and [r*nr, (r+1)*nr) of r
    nl, nr = 60_000, 15_000
    NR = nr * world
    i = np.arange(rank * nr, (rank + 1) * nr, dtype=np.uint64)
    okey = ((i * np.uint64(2654435761) + np.uint64(12345)) % np.uint64(NR) + np.uint64(1)).astype(np.int32)
    lkey = (workload.splitmix64(workload.col_seed(workload.C_LKEY), rank * nl, (rank + 1) * nl) % np.uint64(NR) + np.uint64(1)).astype(np.int32)
    l = g.BATfromhost(lkey, g.TYPE.int, hseq=rank * nl, nonil=True)
    r = g.BATfromhost(okey, g.TYPE.int, hseq=rank * nr, nonil=True)
    g1, g2 = shard.join_repartitioned(g, l, r)
    out["join"] = (g1.numpy(), g2.numpy())
    # high-cardinality GROUP BY: rank r holds rows [r*ng, (r+1)*ng)
    ng = 150_000
    rng = np.random.default_rng(77)
    allk = rng.integers(0, 90_000, ng * world).astype(np.int64)
    allv = rng.integers(-10**9, 10**9, ng * world).astype(np.int64)
    kb = g.BATfromhost(allk[rank * ng:(rank + 1) * ng], g.TYPE.lng, hseq=rank * ng, nonil=True)
    vb = g.BATfromhost(allv[rank * ng:(rank + 1) * ng], g.TYPE.lng, hseq=rank * ng, nonil=True)
    out["groupby"] = tuple(x.numpy() for x in shard.groupby_sum_repartitioned(g, kb, vb))
    q.put((rank, out))
    dist.barrier()
    g.comm_destroy()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_nccl_two_ranks_q1_q6_join(ref):
    """2 ranks over NCCL (in-library ncclAllGather / ncclSend / ncclRecv): sharded Q1, Q6 (with and without
    nils) and the repartitioned join against the unsharded reference"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from monetdb_b200 import workload
    n, world = 300_000, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=400) for _ in range(world))
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for nils in (False, True):
        w1, w6 = _ref_q1(ref, n, nils)
        for rank in range(world):                                # phase 2 is identical on every rank
            _same_q1(got[rank][f"q1_{nils}"], w1)
            assert got[rank][f"q6_{nils}"] == w6
    nl, nr = 60_000, 15_000
    NR = nr * world
    i = np.arange(0, NR, dtype=np.uint64)
    okey = ((i * np.uint64(2654435761) + np.uint64(12345)) % np.uint64(NR) + np.uint64(1)).astype(np.int32)
    lkey = (workload.splitmix64(workload.col_seed(workload.C_LKEY), 0, nl * world) % np.uint64(NR) + np.uint64(1)).astype(np.int32)
    e1, e2 = ref.BATjoin(ref.bat(lkey, ref.T.int, nonil=True), ref.bat(okey, ref.T.int, nonil=True), None, None, False)
    x1 = np.concatenate([got[r]["join"][0] for r in range(world)])
    x2 = np.concatenate([got[r]["join"][1] for r in range(world)])
    o, p = np.lexsort((x2, x1)), np.lexsort((e2.numpy(), e1.numpy()))
    assert np.array_equal(x1[o], e1.numpy()[p]) and np.array_equal(x2[o], e2.numpy()[p])
    ng = 150_000
    rng = np.random.default_rng(77)
    allk = rng.integers(0, 90_000, ng * world).astype(np.int64)
    allv = rng.integers(-10**9, 10**9, ng * world).astype(np.int64)
    assert _groupby_collect([got[r]["groupby"] for r in range(world)]) == _groupby_reference(ref, allk, allv)
