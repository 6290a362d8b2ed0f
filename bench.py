#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on the B200 GDK path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one pass of the TPC-H Q6 operator pipeline (config 2 of
BASELINE.json: 3 selects -> 2 projects -> mul -> sum over a synthetic SF10
lineitem, 59 986 052 rows) per GPU.  N > 1 (torchrun, one rank per GPU) shards
lineitem by contiguous row range, SF10 rows per rank (weak scaling), and merges the
per-rank hge partial sums with one NCCL all-reduce.  `value` = rows processed by
all ranks / device time (columns resident in HBM); `e2e` = the same pipeline
through the C-ABI with HOST (pinned) column heaps, so every operator's
host->device staging and the result read-back are inside the timed region.
`--impl reference` times the real reference GDK (oracle/_ref, built from
/root/reference by oracle/Makefile) on the host cores with mitosis-style
slicing on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_SF10 = 59_986_052
CPU_SAMPLE_ROWS = 6_000_000      # ~SF1: bounded sample for the CPU legs
METRIC = "GDK Q6 pipeline rows/s (TPC-H Q6 shape, SF10 lineitem per GPU)"


def measured_hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs"""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons)}


def hge_int(a) -> int:
    return (int(a["hi"][0]) << 64) | int(a["lo"][0])


# ---------------------------------------------------------------------------- reference (CPU) arm
def reference_q6(rows: int, steps: int, warmup: int, threads: int):
    """the real reference GDK on host cores, mitosis-style: the sample is cut into
    min(4*T, rows/100000) contiguous slices (opt_mitosis.c:18-19,204-205) that run the
    per-slice operator chain in T threads; partial sums are added (mergetable phase 2)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import refgdk as R
    from monetdb_b200 import pipelines, workload
    R.init(threads)
    nsl = max(1, min(4 * threads, rows // 100000))
    bounds = [rows * i // nsl for i in range(nsl + 1)]
    slices = []
    for i in range(nsl):
        lo, hi = bounds[i], bounds[i + 1]
        c = workload.lineitem(lo, hi)
        slices.append({
            "shipdate": R.bat(c["shipdate"], R.T.date, hseq=lo, nonil=True),
            "discount": R.bat(c["discount"], R.T.lng, hseq=lo, nonil=True),
            "quantity": R.bat(c["quantity"], R.T.lng, hseq=lo, nonil=True),
            "extprice": R.bat(c["extprice"], R.T.lng, hseq=lo, nonil=True),
        })
    registered = threading.local()

    def run_slice(col):
        if not getattr(registered, "ok", False):
            R._h.refh_thread_attach(b"bench")
            registered.ok = True
        return hge_int(pipelines.q6(R, col))

    times, total = [], 0
    with ThreadPoolExecutor(max_workers=threads) as ex:
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            total = sum(ex.map(run_slice, slices))
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    return sum(times), total, nsl


# ---------------------------------------------------------------------------- B200 arm
def b200_arm(args):
    import torch
    import torch.distributed as dist
    from monetdb_b200 import gdk as G
    from monetdb_b200 import pipelines, workload

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or G.lib.b200_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    G.init(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.rows
    lo, hi = rank * n, (rank + 1) * n                      # contiguous row range of this rank (weak scaling)
    host = workload.lineitem(lo, hi)
    types = {"shipdate": G.TYPE.date, "discount": G.TYPE.lng, "quantity": G.TYPE.lng, "extprice": G.TYPE.lng}
    # pinned host copies (e2e leg) and device-resident BATs (device heap tier)
    pinned = {}
    for k, a in host.items():
        p = G.pinned_empty(len(a), a.dtype)
        p[:] = a
        pinned[k] = p
    dev = {k: G.BATfromhost(pinned[k], types[k], hseq=lo, nonil=True) for k in types}
    hostbats = {k: G.BAThostwrap(pinned[k], types[k], hseq=lo, nonil=True) for k in types}
    stream = torch.cuda.ExternalStream(G.lib.b200_stream())
    red = torch.zeros(4, dtype=torch.int64, device="cuda")
    red_in = torch.zeros(4, dtype=torch.int64).pin_memory()
    red_out = torch.zeros(4, dtype=torch.int64).pin_memory()
    red_in_np, red_out_np = red_in.numpy(), red_out.numpy()

    def merge(partial: int) -> int:
        """global hge sum: all-reduce of the four 32-bit digits of each rank's partial (pinned staging
        buffers, one synchronisation per step)"""
        if world == 1:
            return partial
        u = partial & ((1 << 128) - 1)
        for i in range(4):
            red_in_np[i] = (u >> (32 * i)) & 0xFFFFFFFF
        if partial < 0:
            red_in_np[3] -= 1 << 32
        red.copy_(red_in, non_blocking=True)
        dist.all_reduce(red)
        red_out.copy_(red, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return sum(int(v) << (32 * i) for i, v in enumerate(red_out_np))

    def barrier():
        G.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    counts = {}

    def step(cols):
        G.stage_cache_clear()       # e2e leg: every step uploads each host column again (once per step)
        return merge(hge_int(pipelines.q6(G, cols, counts)))      # merge == shard.allreduce_hge with a reused buffer

    def timed(cols, steps, warmup, sample_clocks=False):
        for _ in range(warmup):
            res = step(cols)
        barrier()
        G.timing_read()          # drop the warm-up launches from the per-kernel event log
        cs = ClockSampler(local) if sample_clocks else None
        if cs:
            cs.start()
        l0 = G.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for _ in range(steps):
            res = step(cols)
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        clocks = cs.stop() if cs else None
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall, res, G.kernel_launches() - l0, clocks

    # ---- device-resident leg (value); per-kernel CUDA events on the launching stream are
    # recorded over the same timed region (the warm-up records are discarded inside timed())
    ms, wall, res, launches, clocks = timed(dev, args.steps, args.warmup, sample_clocks=not args.device_only)
    if args.device_only:
        # profiling aid (ncu launch lists): only the device-resident steps run, so every launch the
        # profiler sees belongs to the timed region's plan; the line printed is not a bench value
        if rank == 0:
            print(json.dumps({"device_only": True, "ms_per_step": ms / args.steps, "gpu_launches": launches}))
        return
    # a second pass over the same region with the per-kernel CUDA events switched on (the events
    # cost host time, so they stay out of the pass `value` comes from)
    G.timing_enable(True)
    timed(dev, args.steps, 1)
    ktimes = G.timing_read()
    G.timing_enable(False)
    # ---- e2e leg: host heaps through the C-ABI
    e2e_steps = max(1, min(args.steps, 5))
    G.stage_cache(True)             # a host column used by two operators of the plan (l_discount) is staged once per step
    ms_e, wall_e, res_e, _, _ = timed(hostbats, e2e_steps, 1)
    assert res_e == res, "e2e and device-resident results differ"

    if rank != 0:
        return
    nsel = dict(counts)
    by = pipelines.q6_bytes(n, nsel["n1"], nsel["n2"], nsel["n3"])
    peak, peak_src = measured_hbm_peak()
    # dominant operator: BATselect on l_shipdate (dense candidates) = mask scan + offset scan + expand;
    # its first kernel (select_dense) is the dominant kernel of the step
    nl, kms, _ = ktimes.get("select_dense", (0, 0.0, 0))
    per_launch_ms = kms / nl if nl else None
    mask_bytes = 4 * n + n // 8                         # column read once + one mask bit per row written
    achieved = mask_bytes / (per_launch_ms * 1e-3) / 1e9 if per_launch_ms else None
    total_rows = n * world
    value = total_rows * args.steps / (ms * 1e-3)
    # e2e host->device bytes per step: l_shipdate and l_discount are staged whole (dense scan / 14 % candidate
    # density); l_quantity and l_extprice are only touched through short candidate lists (3.9 % / 1.8 % of the
    # rows) and are read in place from the pinned host heap, 8 useful bytes per touched row
    h2d = 4 * n + 8 * n + 8 * nsel["n2"] + 8 * nsel["n3"]
    line = {
        "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32/int64 compare, int64 x int64 -> int128 multiply-add", "data": "synthetic",
        "config": {"workload": "TPC-H Q6 shape, SF10 synthetic lineitem (59986052 rows per GPU): 3-predicate select + "
                               "2 project + calc mul + sum(hge)", "rows_per_gpu": n,
                   "selected": nsel, "l2": "inputs (1.7 GB) larger than L2; no flush needed",
                   "pipeline_algorithmic_bytes": sum(by.values()),
                   "pipeline_gbs": sum(by.values()) * world * args.steps / (ms * 1e-3) / 1e9,
                   "kernel_ms": {k: round(v[1] / max(1, v[0]), 4) for k, v in ktimes.items()},
                   "kernel_launch_counts": {k: v[0] for k, v in ktimes.items()}},
        "gpu_launches": int(launches),
        "result_hge": str(res),
        "clocks": clocks,
        "e2e": {"value": total_rows * e2e_steps / (ms_e * 1e-3), "unit": "rows/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 16 + 7 * 32, "ms_per_step": ms_e / e2e_steps, "steps": e2e_steps},
        "roofline": {"bound": "hbm", "kernel": "select_dense_kernel<int> (BATselect on l_shipdate)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
                     "traffic": 245877248, "traffic_source": "ncu --set full, profiles/r1_q6_kernels_v2.md (dram read+write per launch)",
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": mask_bytes,
                     "launch_ms": per_launch_ms},
    }
    if world == 1 and not args.no_cpu:
        thr = os.cpu_count() or 1
        rows = min(CPU_SAMPLE_ROWS, n)
        t, total, nsl = reference_q6(rows, 3, 1, thr)
        line["cpu_baseline"] = {"value": rows * 3 / t, "unit": "rows/s", "cores": thr, "kind": "reference",
                                "sample": f"first {rows} rows of the same lineitem, {nsl} mitosis slices, 3 passes"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------- other BASELINE configs
def other_workload(args):
    """--workload q1 | window | config1 | join: the other BASELINE configs on the B200 path (same
    timing rules as the Q6 line: warm-up, CUDA events on the library's stream, max over ranks).
    They are parity-test cases first; these lines record their throughput for profiles/."""
    import torch
    import torch.distributed as dist
    from monetdb_b200 import gdk as G
    from monetdb_b200 import pipelines, shard, workload

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or G.lib.b200_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    G.init(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    T = G.TYPE
    n = args.rows
    w = args.workload
    info = {}

    def upload(gen, types, lo, hi, chunk=50_000_000):
        """generate rows [lo,hi) chunk by chunk on the host, one upload per column"""
        bufs = {k: np.empty(hi - lo, dtype=G.NP_OF[t]) for k, t in types.items()}
        for a in range(lo, hi, chunk):
            b = min(hi, a + chunk)
            part = gen(a, b)
            for k in types:
                bufs[k][a - lo:b - lo] = part[k]
        return {k: G.BATfromhost(bufs[k], types[k], hseq=lo, nonil=True) for k in types}

    if w == "q1":
        types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
                 "returnflag": T.bte, "linestatus": T.bte}
        lo, hi = rank * n, (rank + 1) * n
        dev = upload(lambda a, b: workload.lineitem(a, b, cols=tuple(types)), types, lo, hi)

        def step():
            r = shard.q1_sharded(G, dev, dist if world > 1 else None)
            info["nsel"] = sum(r["count_order"])
            return r["sum_charge"]
        metric = "GDK Q1 pipeline rows/s (TPC-H Q1 shape, synthetic lineitem)"
        bytes_of = lambda: pipelines.q1_bytes(n * world, info["nsel"])
        label = f"TPC-H Q1 shape: select + 6 project + 2-level group + 4 calc + 8 aggregates, {n} lineitem rows per GPU"
    elif w == "config1":
        k, v = workload.config1(0, n)
        dev = {"k": G.BATfromhost(k, T.int, nonil=True), "v": G.BATfromhost(v, T.lng, nonil=True)}

        def step():
            out = {}
            r = pipelines.config1(G, dev, out)
            info["sel"] = len(out["s"])
            return shard.hge_int(r.numpy())
        metric = "GDK config-1 pipeline rows/s (select + 2 project + group + groupsum)"
        bytes_of = lambda: 4 * n + 76 * info["sel"]
        label = f"config 1: BATselect range + BATgroup + BATgroupsum on a {n}-row int/lng column pair"
    elif w == "join":
        # config 4 shape: l_orderkey = 1 + z % NR (about 4 matches per key), o_orderkey = a fixed bijection of
        # 1..NR (unique, unsorted: forces hashjoin, gdk_join.c:4568).  Rank g holds rows [g*nl,(g+1)*nl) of
        # lineitem and [g*nr,(g+1)*nr) of orders with global hseqbase; NR = world * nr.
        nl, nr = n, max(1, n // 4)
        NR = nr * world
        A = 2654435761                                  # odd and coprime to NR for the sizes used here
        while np.gcd(A, NR) != 1:
            A += 2
        i = np.arange(rank * nr, (rank + 1) * nr, dtype=np.uint64)
        okey = ((i * np.uint64(A) + np.uint64(12345)) % np.uint64(NR) + np.uint64(1)).astype(np.int32)
        lkey = (workload.splitmix64(workload.col_seed(workload.C_LKEY), rank * nl, (rank + 1) * nl) % np.uint64(NR)
                + np.uint64(1)).astype(np.int32)
        dev = {"l": G.BATfromhost(lkey, T.int, hseq=rank * nl, nonil=True),
               "r": G.BATfromhost(okey, T.int, hseq=rank * nr, nonil=True, key=True)}
        phases = {}

        def step():
            if world == 1:
                r1, r2 = G.BATjoin(dev["l"], dev["r"], None, None, False)
            else:
                r1, r2 = shard.join_repartitioned(G, dev["l"], dev["r"], dist, torch, torch.device("cuda", local), phases)
            info["pairs_rank0"] = len(r1)
            return len(r1)
        metric = "GDK hash join rows/s (lineitem JOIN orders shape, probe rows)"
        # per rank: partition pass 2*(4+8) per tuple of both sides + exchange + build 4+8+8 per build row + probe
        bytes_of = lambda: world * ((4 + 8 + 8) * nr + (4 + 2 * 8 + 16) * nl + (0 if world == 1 else 2 * (4 + 8) * (nl + nr)))
        label = (f"config 4 shape: {nl} int32 probe keys x {nr} unique build keys per GPU (BAThash + BATjoin)"
                 + ("" if world == 1 else "; hash repartition of both sides + NCCL all-to-all of (key, oid) tuples"))
        info["phases_ms_last_step"] = phases
    elif w == "window":
        blk = workload.window_layout()
        n = max(blk, n // blk * blk)
        lo, hi = rank * n, (rank + 1) * n
        types = {"pkey": T.lng, "ord": T.lng, "val": T.lng}
        dev = upload(lambda a, b: workload.window(a, b), types, lo, hi, chunk=blk * 400)

        def step():
            p = G.GDKanalyticaldiff(G.colnew(T.bit, n), dev["pkey"], None, None, T.lng)
            tot = 0
            for unit, plo, phi, half in ((0, 100, 0, 1), (1, 50, 50, 0)):
                s = G.GDKanalyticalwindowbounds(G.colnew(T.oid, n), dev["ord"], p, None, plo, T.lng, T.lng, unit, True, 0)
                e = G.GDKanalyticalwindowbounds(G.colnew(T.oid, n), dev["ord"], p, None, phi, T.lng, T.lng, unit, False, half)
                r = G.GDKanalyticalsum(G.colnew(T.hge, n), p, None, dev["val"], s, e, T.lng, T.hge, unit)
                a = G.GDKanalyticalfn("avg", G.colnew(T.dbl, n), p, None, dev["val"], s, e, T.lng, unit)
                tot += len(r) + len(a)
            return tot
        metric = "GDK window analytics rows/s (ROWS + RANGE frames, moving sum + avg)"
        # diff 9; ROWS bounds 2 x 9; RANGE bounds 2 x 17; per frame kind sum 8+16+16 and avg 8+16+8
        bytes_of = lambda: n * world * (9 + 18 + 34 + 2 * (40 + 32))
        label = f"config 5: {n} rows per GPU, partitions of 1e3/1e4/1e5 rows, ROWS 100 PRECEDING..CURRENT and RANGE 50..50"
    else:
        raise SystemExit(f"unknown workload {w}")

    stream = torch.cuda.ExternalStream(G.lib.b200_stream())

    def barrier():
        G.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(steps, warmup):
        for _ in range(warmup):
            res = step()
        barrier()
        G.timing_read()
        l0 = G.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        walls = []
        for _ in range(steps):
            tw = time.perf_counter()
            res = step()
            walls.append(round((time.perf_counter() - tw) * 1e3, 3))
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        info["host_ms_per_step"] = walls
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, res, G.kernel_launches() - l0

    ms, res, launches = timed(args.steps, max(3, args.warmup))
    G.timing_enable(True)
    timed(min(args.steps, 3), 1)
    kt = G.timing_read()
    G.timing_enable(False)
    if rank == 0:
        peak, peak_src = measured_hbm_peak()
        rows = n * world
        by = bytes_of()
        gbs = by * args.steps / (ms * 1e-3) / 1e9
        print(json.dumps({
            "metric": metric, "value": rows * args.steps / (ms * 1e-3), "unit": "rows/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64/int128 integer", "data": "synthetic",
            "config": {"workload": label, "info": info, "pipeline_algorithmic_bytes": by, "pipeline_gbs": gbs,
                       "pipeline_frac_of_hbm_peak": gbs / (peak * world), "peak_source": peak_src,
                       "kernel_ms": {k: round(v[1] / max(1, v[0]), 4) for k, v in kt.items()},
                       "kernel_launch_counts": {k: v[0] for k, v in kt.items()},
                       "kernel_gbs": {k: round(v[2] / (v[1] * 1e-3) / 1e9, 1) for k, v in kt.items() if v[1] > 0}},
            "gpu_launches": int(launches), "result": str(res)[:200]}))
    if world > 1:
        dist.destroy_process_group()


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    thr = os.cpu_count() or 1
    rows = min(CPU_SAMPLE_ROWS, args.rows)
    t, total, nsl = reference_q6(rows, args.steps, args.warmup, thr)
    v = rows * args.steps / t
    sample = f"first {rows} rows of the SF10 lineitem, {nsl} mitosis slices on {thr} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "rows/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32/int64 compare, int64 x int64 -> int128 multiply-add",
        "data": "synthetic", "config": {"workload": "TPC-H Q6 shape, SF10 synthetic lineitem (bounded sample)", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "rows/s", "cores": thr, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result_hge": str(total),
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=N_SF10, help="lineitem rows per GPU (default SF10)")
    ap.add_argument("--device-only", action="store_true", help="run only the device-resident steps (for ncu launch lists)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--workload", default="q6", choices=["q6", "q1", "config1", "join", "window"],
                    help="q6 (default) is the contract line; the others record the remaining BASELINE configs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        reference_arm(args)
    elif args.workload != "q6":
        other_workload(args)
    else:
        if args.warmup < 3:
            args.warmup = 3
        b200_arm(args)


if __name__ == "__main__":
    main()
