#!/usr/bin/env python
"""bench.py — BASELINE.json's metric ("GDK Q1/Q6 pipeline rows/s and % HBM peak at 1/2/4/8 B200 vs CPU GDK").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--rows TOTAL]

The contract line is BASELINE config 3: the TPC-H Q1 operator pipeline over a synthetic SF100 lineitem
(600 037 902 rows), STRONG-scaled: N ranks (torchrun, one per GPU) each hold a contiguous 1/N row range of
every column.  A "step" is one pass of the Q1 plan over the whole table:
  N = 1   the §8(d) plan (select, 6 projects, 2-level group, 4 calcs, 4 sums, 3 avgs, count);
  N > 1   mergetable's two-phase form of it (monetdb_b200/shard.py): phase 1 on the rank's rows ends in a
          <= 4-row partial table, b200_comm_pack (one ncclAllGather inside the library) concatenates the
          ranks' tables, phase 2 is ordinary GDK calls on the packed BATs.
`value` = table rows / device time (columns resident in HBM, CUDA events on the library's stream, max over
ranks).  `config.q6` carries the same measurement for the Q6 plan (config 2's shape) on the same table.
`e2e` = the Q1 step through the C-ABI with HOST (page-locked) column heaps: every operator stages its inputs
inside the timed region and the result tables are read back.
`--impl reference` times the real reference GDK (oracle/_ref, compiled from /root/reference by
oracle/Makefile) on the host cores with mitosis slicing + mergetable phase 2, on the same table (or, if K
steps of that would take longer than the time budget, on the longest prefix of it that fits; the line says
which).  The B200 arm's cpu_baseline leg runs the same reference once and asserts that the GPU results are
identical (parity_checked).
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

N_SF100 = 600_037_902
N_SF10 = 59_986_052
METRIC = "GDK Q1 pipeline rows/s (TPC-H Q1 shape, SF100 synthetic lineitem, strong scaling; Q6 in config.q6)"
DTYPE = "int32/int64 compare, int64 x int64 -> int128 multiply-add, exact 192-bit sums"
Q1_COLS = ("shipdate", "discount", "quantity", "extprice", "tax", "returnflag", "linestatus")
Q6_COLS = ("shipdate", "discount", "quantity", "extprice")
REF_BUDGET_S = 240.0           # wall budget of the reference arm's K + W steps


def measured_hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def workload_label(total: int) -> str:
    sf = "SF100" if total == N_SF100 else "SF10" if total == N_SF10 else f"{total} rows"
    return (f"TPC-H Q1 shape, {sf} synthetic lineitem ({total} rows): select + 6 project + 2-level group + 4 calc + "
            f"8 aggregates; Q6 shape (3-predicate select + 2 project + calc mul + sum) on the same table in config.q6")


class ClockSampler:
    """SM clock and throttle reasons sampled through NVML from a thread WHILE the timed region runs.  (A
    `nvidia-smi -lms 100` child process was used first; its polling slowed a 25-sync Q1 step from 39 to 66 ms
    on the B200 box, so the sampler now makes two NVML calls every 50 ms in-process instead.)"""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period: float = 0.05):
        self.index, self.period = index, period
        self.sm, self.mx, self.reasons = [], None, set()
        self._stop = threading.Event()
        self._thr = None
        self.err = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates all GPUs of the box, CUDA only the visible ones: map through the UUID order
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis:
                ent = vis.split(",")[self.index].strip()
                if ent.isdigit():
                    idx = int(ent)
                else:
                    h = pynvml.nvmlDeviceGetHandleByUUID(ent.encode() if hasattr(ent, "encode") else ent)
                    idx = pynvml.nvmlDeviceGetIndex(h)
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as ex:            # noqa: BLE001
            self.err = repr(ex)
            return
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def _sample(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:                  # noqa: BLE001
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for bit, name in self.REASONS.items():
            if r & bit:
                self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                self._sample()
            except Exception as ex:        # noqa: BLE001
                self.err = repr(ex)
                return
            self._stop.wait(self.period)

    def stop(self) -> dict:
        if self._thr is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [f"NVML unavailable: {self.err}"]}
        try:
            self._sample()                 # at least one sample inside the region's tail
        except Exception:                  # noqa: BLE001
            pass
        self._stop.set()
        self._thr.join(1.0)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "how": "NVML, in-process thread, 50 ms"}


def hge_int(a, i=0) -> int:
    return (int(a["hi"][i]) << 64) | int(a["lo"][i])


def q1_unsharded_host(X, res, cand) -> dict:
    """the §8(d) Q1 plan's result as plain values, extents mapped through the candidate list to table oids
    (the form shard.q1_host gives for the two-phase plan)"""
    from monetdb_b200 import shard
    out = {}
    for k in shard.Q1_SUMS:
        a = res[k].numpy()
        out[k] = [hge_int(a, i) for i in range(len(a))]
    for k in ("avg_qty", "avg_price", "avg_disc", "count_order", "histo"):
        out[k] = res[k].numpy().tolist()
    out["extents"] = X.BATproject(res["extents"], cand).numpy().tolist()
    return out


def same_q1(a: dict, b: dict) -> bool:
    from monetdb_b200 import shard
    keys = shard.Q1_SUMS + ("count_order", "histo", "extents")
    if any(a[k] != b[k] for k in keys):
        return False
    for k in ("avg_qty", "avg_price", "avg_disc"):
        if np.array(a[k], dtype=np.float64).tobytes() != np.array(b[k], dtype=np.float64).tobytes():
            return False
    return True


# ---------------------------------------------------------------------------- reference (CPU) arm
class ReferenceTable:
    """rows [0, rows) of lineitem as mitosis slices of reference BATs: P = min(4T, rows/100000) contiguous row
    ranges (opt_mitosis.c:18-19,204-205); the generator writes straight into the GDK heaps"""

    def __init__(self, rows: int, threads: int, nils: bool = False):
        from oracle import refgdk as R
        from monetdb_b200 import shard, workload
        R.init(threads)
        # the slices' operator chains are driven from Python threads; the calls into the reference release the
        # GIL, but with the default 5 ms switch interval the hand-over between 16+ workers dominates short
        # operators (measured: Q6 157 -> 434 M rows/s on 8 cores)
        sys.setswitchinterval(1e-4)
        self.R, self.rows, self.threads = R, rows, threads
        self.nsl = max(1, min(4 * threads, rows // 100000))
        T = R.T
        types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
                 "returnflag": T.bte, "linestatus": T.bte}
        self.slices = []
        for i in range(self.nsl):
            lo, hi = shard.shard_range(rows, i, self.nsl)
            bats, views = {}, {}
            for k, tt in types.items():
                nonil = not nils or k in ("returnflag", "linestatus")
                bats[k], views[k] = R.bat_uninit(tt, hi - lo, lo, nonil=nonil, nil=not nonil)
            workload.lineitem_into(lo, hi, views, nils, threads)
            self.slices.append(bats)
        self._tl = threading.local()

    def _attach(self):
        if not getattr(self._tl, "ok", False):
            self.R._h.refh_thread_attach(b"bench")
            self._tl.ok = True

    def run(self, ex, first: int | None = None):
        """one pass of Q1 then Q6 over the first `first` slices (default all): per-slice operator chains in
        the thread pool, mat.pack of the partial tables, phase 2 (mergetable).  -> (q1 dict, q6 int, t_q1, t_q6)"""
        from monetdb_b200 import shard
        R = self.R
        sl = self.slices if first is None else self.slices[:first]

        def p1(cols):
            self._attach()
            return shard.q1_phase1(R, cols)

        def p6(cols):
            self._attach()
            return shard.q6_phase1(R, cols)

        t0 = time.perf_counter()
        tables = list(ex.map(p1, sl))
        q1 = shard.q1_host(shard.q1_phase2(R, shard.pack_tables(R, tables)))
        t1 = time.perf_counter()
        del tables
        sums = list(ex.map(p6, sl))
        q6 = hge_int(shard.q6_phase2(R, shard.pack_tables(R, sums)))
        t2 = time.perf_counter()
        return q1, q6, t1 - t0, t2 - t1

    def rows_of(self, first: int | None) -> int:
        from monetdb_b200 import shard
        return self.rows if first is None else shard.shard_range(self.rows, first - 1, self.nsl)[1]


def host_mem_ok(table_bytes: int) -> bool:
    try:
        import psutil
        return psutil.virtual_memory().available > 2.2 * table_bytes + (8 << 30)
    except Exception:
        return True


def table_bytes(rows: int) -> int:
    return rows * (4 + 4 * 8 + 2)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from concurrent.futures import ThreadPoolExecutor
    thr = host_threads()
    total = args.rows
    rows = total
    while not host_mem_ok(table_bytes(rows)) and rows > 10_000_000:
        rows //= 2
    tab = ReferenceTable(rows, thr)
    steps, warmup = args.steps, max(args.warmup, 1)
    first = None
    t_q1, t_q6 = [], []
    with ThreadPoolExecutor(max_workers=thr) as ex:
        q1, q6, a, b = tab.run(ex)                      # warm-up step 1, on everything that was generated
        est = (a + b) * (steps + warmup - 1)
        if est > REF_BUDGET_S:
            first = max(1, int(tab.nsl * REF_BUDGET_S / est))
        for it in range(1, warmup + steps):
            q1, q6, a, b = tab.run(ex, first)
            if it >= warmup:
                t_q1.append(a)
                t_q6.append(b)
    used = tab.rows_of(first)
    v1 = used * steps / sum(t_q1)
    v6 = used * steps / sum(t_q6)
    sample = (f"rows [0, {used}) of the {total}-row table ({'the whole table' if used == total else 'bounded prefix'}), "
              f"{tab.nsl if first is None else first} mitosis slices of {rows // tab.nsl} rows on {thr} threads, "
              f"per-slice operator chains + mergetable phase 2")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v1, "unit": "rows/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": steps, "warmup": warmup, "ms_per_step": sum(t_q1) / steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": workload_label(total), "rows": total,
                   "q6": {"value": v6, "unit": "rows/s", "ms_per_step": sum(t_q6) / steps * 1e3}},
        "same_rows": used == total, "rows_per_step": used,
        "cpu_baseline": {"value": v1, "unit": "rows/s", "cores": thr, "kind": "reference", "sample": sample},
        "e2e": {"value": v1, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result": {"q1_sum_charge": [str(x) for x in q1["sum_charge"]], "q1_count_order": q1["count_order"], "q6_hge": str(q6)},
    }))


# ---------------------------------------------------------------------------- B200 arm
def load_dram_table(workload: str, rows_per_gpu: int):
    """per-kernel dram__bytes (read + write) per launch from the committed ncu capture of the same plan at the
    same per-GPU row count (profiles/r2_dram_bytes.json, written by scripts/ncu_dram_table.py); None otherwise"""
    p = ROOT / "profiles" / "r2_dram_bytes.json"
    if not p.exists():
        return None, None
    try:
        d = json.loads(p.read_text())
        for ent in d.get(workload, []):
            if ent.get("rows_per_gpu") == rows_per_gpu:
                return ent["kernels"], ent.get("source")
    except Exception:
        pass
    return None, None


def b200_arm(args):
    import torch
    import torch.distributed as dist
    from monetdb_b200 import gdk as G
    from monetdb_b200 import pipelines, shard, workload

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or G.lib.b200_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    G.init(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        G.comm_init_torch(dist)
    else:
        G.comm_init(0, 1, None)
    total = args.rows
    lo, hi = shard.shard_range(total, rank, world)           # contiguous row range of this rank (strong scaling)
    n = hi - lo
    T = G.TYPE
    types = {"shipdate": T.date, "discount": T.lng, "quantity": T.lng, "extprice": T.lng, "tax": T.lng,
             "returnflag": T.bte, "linestatus": T.bte}
    thr = max(1, host_threads() // world)
    t_setup = time.perf_counter()
    pinned = {k: G.pinned_empty(n, workload.LINEITEM_DTYPES[k]) for k in types}
    workload.lineitem_into(lo, hi, pinned, False, thr)

    def upload(nils=False):
        nonil = lambda k: not nils or k in ("returnflag", "linestatus")
        dev = {k: G.BATfromhost(pinned[k], types[k], hseq=lo, nonil=nonil(k), nil=not nonil(k)) for k in types}
        hostb = {k: G.BAThostwrap(pinned[k], types[k], hseq=lo, nonil=nonil(k), nil=not nonil(k)) for k in types}
        return dev, hostb

    dev, hostbats = upload()
    setup_s = time.perf_counter() - t_setup
    stream = torch.cuda.ExternalStream(G.lib.b200_stream())
    pack = shard.device_pack(G)
    info = {}

    def q1_step(cols):
        if world == 1:
            out = {}
            res = pipelines.q1(G, cols, out)
            info["nsel"] = out["nsel"]
            info["cand"] = out["cand"]
            return res
        out = {}
        res = shard.q1_sharded(G, cols, pack, out)
        info["nsel"] = out["nsel"]
        return res

    counts = {}

    def q6_step(cols):
        if world == 1:
            return hge_int(pipelines.q6(G, cols, counts))
        return shard.q6_sharded(G, cols, pack, counts)

    def barrier():
        G.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, cols, steps, warmup, sample_clocks=False, e2e=False):
        res = None
        for _ in range(warmup):
            if e2e:
                G.stage_cache_clear()
            res = step(cols)
        barrier()
        G.timing_read()          # drop the warm-up launches from the per-kernel event log
        cs = ClockSampler(local) if sample_clocks else None
        if cs:
            cs.start()
        l0 = G.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            if e2e:
                G.stage_cache_clear()       # e2e leg: every step uploads each host column again (once per step)
            res = step(cols)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = cs.stop() if cs else None
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, res, G.kernel_launches() - l0, clocks

    def ktable(step, cols, steps):
        """a second pass over the same region with the per-kernel CUDA events switched on (the events cost
        host time, so they stay out of the pass `value` comes from)"""
        G.timing_enable(True)
        timed(step, cols, steps, 1)
        kt = G.timing_read()
        G.timing_enable(False)
        return {k: {"launches_per_step": v[0] / steps, "ms_per_launch": v[1] / max(1, v[0]),
                    "ms_per_step": v[1] / steps, "algorithmic_bytes_per_launch": v[2] // max(1, v[0])} for k, v in kt.items()}

    # ---- Q1: device-resident leg (value)
    if args.device_only:
        # profiling aid (ncu launch lists): only the device-resident steps of one plan run, so every launch the
        # profiler sees belongs to that plan; the line printed is not a bench value
        msd, _, ld, _ = timed(q6_step if args.device_only == "q6" else q1_step, dev, args.steps, args.warmup)
        if rank == 0:
            print(json.dumps({"device_only": args.device_only, "ms_per_step": msd / args.steps, "gpu_launches": ld}))
        return
    ms1, res1, launches1, clocks = timed(q1_step, dev, args.steps, args.warmup, sample_clocks=True)
    if world == 1:
        q1_res = q1_unsharded_host(G, res1, info["cand"])
    else:
        q1_res = shard.q1_host(res1)
    nsel_local = info["nsel"]
    k1 = ktable(q1_step, dev, min(args.steps, 3))
    # ---- Q6 on the same table
    ms6, res6, launches6, _ = timed(q6_step, dev, args.steps, args.warmup)
    k6 = ktable(q6_step, dev, min(args.steps, 5))
    nsel6 = dict(counts)
    info.clear()
    # ---- e2e leg: host heaps through the C-ABI (Q1; then Q6)
    for b in dev.values():
        b.free()
    e2e_steps = max(1, min(args.steps, 3))
    G.stage_cache(True)             # a host column used by several operators of the plan is staged once per step
    ms1e, res1e, _, _ = timed(q1_step, hostbats, e2e_steps, 1, e2e=True)
    q1_res_e = q1_unsharded_host(G, res1e, info["cand"]) if world == 1 else shard.q1_host(res1e)
    assert same_q1(q1_res_e, q1_res), "e2e and device-resident Q1 results differ"
    info.clear()
    ms6e, res6e, _, _ = timed(q6_step, hostbats, e2e_steps, 1, e2e=True)
    assert res6e == res6, "e2e and device-resident Q6 results differ"
    G.stage_cache_clear()
    G.stage_cache(False)

    # ---- totals over ranks
    if world > 1:
        t = torch.tensor([nsel_local, nsel6["n1"], nsel6["n2"], nsel6["n3"]], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        nsel_tot, n1, n2, n3 = (int(x) for x in t.tolist())
    else:
        nsel_tot, n1, n2, n3 = nsel_local, nsel6["n1"], nsel6["n2"], nsel6["n3"]

    # ---- the 1 %-nil second run: same plans, every value column nil where z % 100 == 0
    nil_run = None
    if not args.no_nil_run:
        workload.lineitem_into(lo, hi, pinned, True, thr)
        devn, _ = upload(nils=True)
        nsteps = max(1, min(args.steps, 5))
        msn1, resn1, _, _ = timed(q1_step, devn, nsteps, 1)
        qn1 = q1_unsharded_host(G, resn1, info["cand"]) if world == 1 else shard.q1_host(resn1)
        msn6, resn6, _, _ = timed(q6_step, devn, nsteps, 1)
        nil_run = {"q1_ms_per_step": msn1 / nsteps, "q6_ms_per_step": msn6 / nsteps, "steps": nsteps,
                   "q1_rows_per_s": total * nsteps / (msn1 * 1e-3), "q6_rows_per_s": total * nsteps / (msn6 * 1e-3)}
        for b in devn.values():
            b.free()
        info.clear()

    if rank != 0:
        if world > 1:
            dist.barrier()
            G.comm_destroy()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_hbm_peak()
    by1 = pipelines.q1_bytes(total, nsel_tot) if world == 1 else pipelines.q1_bytes_sharded(total, nsel_tot)
    by6 = sum(pipelines.q6_bytes(total, n1, n2, n3).values())
    value = total * args.steps / (ms1 * 1e-3)
    # dominant kernel of the Q1 step = the timed kernel with the largest share of the step (per-kernel CUDA events)
    dom = max(k1.items(), key=lambda kv: kv[1]["ms_per_step"]) if k1 else (None, None)
    dram1, dram_src = load_dram_table("q1", n)
    dram6, _ = load_dram_table("q6", n)
    roof = {"bound": "hbm", "kernel": None, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": None,
            "peak_source": peak_src}
    if dom[0]:
        name, d = dom
        ach = d["algorithmic_bytes_per_launch"] / (d["ms_per_launch"] * 1e-3) / 1e9
        roof.update(kernel=name, achieved=ach, frac=ach / peak, launch_ms=d["ms_per_launch"],
                    algorithmic_bytes_per_launch=d["algorithmic_bytes_per_launch"],
                    share_of_step=d["ms_per_step"] / (ms1 / args.steps),
                    traffic=(dram1 or {}).get(name), traffic_source=dram_src if dram1 and name in dram1 else None,
                    note="kernel = the timed kernel with the largest share of the Q1 step; achieved = the operator's "
                         "API-faithful algorithmic bytes (SURVEY.md §8d) / launch time; traffic = ncu dram bytes of the "
                         "same launch where a capture at this row count is committed")
    pipe_gbs = by1 * args.steps / (ms1 * 1e-3) / 1e9
    dram_step = sum(v * k1[k]["launches_per_step"] for k, v in (dram1 or {}).items() if k in k1) if dram1 else None
    line = {
        "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms1 / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": DTYPE, "data": "synthetic",
        "config": {
            "workload": workload_label(total), "rows": total, "rows_per_gpu": n,
            "plan": "SURVEY §8(d) config 3 plan" if world == 1 else
                    "two-phase (mergetable): phase 1 per rank, b200_comm_pack (one ncclAllGather), phase 2 on the packed BATs",
            "selected": nsel_tot, "l2": f"inputs ({table_bytes(n) / 1e9:.1f} GB per GPU) larger than L2; no flush needed",
            "pipeline_algorithmic_bytes": by1, "pipeline_gbs": pipe_gbs,
            "pipeline_frac_of_hbm_peak_algorithmic": pipe_gbs / (peak * world),
            "pipeline_dram_bytes_per_gpu": dram_step,
            "pipeline_frac_of_hbm_peak_dram": (dram_step / (ms1 / args.steps * 1e-3) / 1e9 / peak) if dram_step else None,
            "kernels": {k: {kk: (round(vv, 5) if isinstance(vv, float) else vv) for kk, vv in v.items()} for k, v in k1.items()},
            "timed_kernel_ms_per_step": round(sum(v["ms_per_step"] for v in k1.values()), 4),
            "setup_s": round(setup_s, 1),
            "q6": {
                "value": total * args.steps / (ms6 * 1e-3), "unit": "rows/s", "ms_per_step": ms6 / args.steps,
                "selected": {"n1": n1, "n2": n2, "n3": n3}, "pipeline_algorithmic_bytes": by6,
                "pipeline_gbs": by6 * args.steps / (ms6 * 1e-3) / 1e9,
                "pipeline_frac_of_hbm_peak_algorithmic": by6 * args.steps / (ms6 * 1e-3) / 1e9 / (peak * world),
                "gpu_launches": int(launches6), "result_hge": str(res6),
                "kernels": {k: {kk: (round(vv, 5) if isinstance(vv, float) else vv) for kk, vv in v.items()} for k, v in k6.items()},
                "dram_bytes_per_launch": dram6,
                "e2e": {"value": total * e2e_steps / (ms6e * 1e-3), "unit": "rows/s", "ms_per_step": ms6e / e2e_steps},
            },
            "nil_run": nil_run,
        },
        "gpu_launches": int(launches1),
        "result": {"sum_charge": [str(x) for x in q1_res["sum_charge"]], "count_order": q1_res["count_order"],
                   "extents": q1_res["extents"], "avg_disc": q1_res["avg_disc"]},
        "clocks": clocks,
        "e2e": {"value": total * e2e_steps / (ms1e * 1e-3), "unit": "rows/s", "h2d_bytes_per_step": table_bytes(n) * world,
                "d2h_bytes_per_step": 8 * 16 * 8 * world, "ms_per_step": ms1e / e2e_steps, "steps": e2e_steps},
        "roofline": roof,
    }
    # golden answers of the full SF100 table (written from the reference's own run, tests/golden/)
    gold = ROOT / "tests" / "golden" / "lineitem_answers.json"
    if gold.exists():
        g = json.loads(gold.read_text()).get(str(total))
        if g:
            ok = ([str(x) for x in q1_res["sum_charge"]] == g["q1_sum_charge"] and q1_res["count_order"] == g["q1_count_order"]
                  and str(res6) == g["q6_hge"])
            assert ok, "results differ from the golden answers of the reference GDK"
            line["golden_checked"] = True
    if world == 1 and not args.no_cpu:
        from concurrent.futures import ThreadPoolExecutor
        cthr = host_threads()
        rows = total
        while not host_mem_ok(table_bytes(rows)) and rows > 10_000_000:
            rows //= 2
        tab = ReferenceTable(rows, cthr)
        with ThreadPoolExecutor(max_workers=cthr) as ex:
            rq1, rq6, ta, tb = tab.run(ex)
        if rows == total:
            gq1, gq6 = q1_res, res6
        else:
            # the reference could only hold a prefix of the table: run the GPU plans on the same prefix
            workload.lineitem_into(lo, hi, pinned, False, thr)
            sub = {k: G.BATfromhost(pinned[k][:rows], types[k], hseq=0, nonil=True) for k in types}
            r = q1_step(sub)
            gq1 = q1_unsharded_host(G, r, info["cand"])
            gq6 = q6_step(sub)
        assert same_q1(gq1, rq1), "Q1: the B200 result differs from the reference GDK"
        assert gq6 == rq6, "Q6: the B200 result differs from the reference GDK"
        line["parity_checked"] = True
        line["parity_rows"] = rows
        line["cpu_baseline"] = {
            "value": rows / ta, "unit": "rows/s", "cores": cthr, "kind": "reference",
            "sample": f"rows [0, {rows}) of the same table ({'the whole table' if rows == total else 'prefix: host memory'}), "
                      f"{tab.nsl} mitosis slices on {cthr} threads, one pass; Q6 on the same rows: {rows / tb:.4g} rows/s",
            "q6_value": rows / tb}
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        G.comm_destroy()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------- other BASELINE configs
def other_workload(args):
    """--workload window | config1 | join: the other BASELINE configs on the B200 path (same timing rules as
    the contract line: warm-up, CUDA events on the library's stream, max over ranks).  They are parity-test
    cases first; these lines record their throughput for profiles/.  --rows is PER GPU here (weak scaling)."""
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
        G.comm_init_torch(dist)
    else:
        G.comm_init(0, 1, None)
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

    if w == "config1":
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
                r1, r2 = shard.join_repartitioned(G, dev["l"], dev["r"], phases if args.phases else None)
            info["pairs_rank0"] = len(r1)
            return len(r1)
        metric = "GDK hash join rows/s (lineitem JOIN orders shape, probe rows)"
        # per rank: partition pass 2*(4+8) per tuple of both sides + exchange + build 4+8+8 per build row + probe
        bytes_of = lambda: world * ((4 + 8 + 8) * nr + (4 + 2 * 8 + 16) * nl + (0 if world == 1 else 2 * (4 + 8) * (nl + nr)))
        label = (f"config 4 shape: {nl} int32 probe keys x {nr} unique build keys per GPU (BAThash + BATjoin)"
                 + ("" if world == 1 else "; hash repartition of both sides + grouped ncclSend/ncclRecv of (key, oid) tuples "
                                          "inside the library"))
        info["phases_ms_last_step"] = phases
    elif w == "window":
        blk = workload.window_layout()
        n = max(blk, n // blk * blk)
        lo, hi = rank * n, (rank + 1) * n
        types = {"pkey": T.lng, "ord": T.lng, "val": T.lng}
        dev = upload(lambda a, b: workload.window_fast(a, b, threads=max(1, host_threads() // world)), types, lo, hi, chunk=blk * 400)

        def step():
            # one plan = one step: whatever the library cached about the value column in the previous step (its
            # running totals, shared by the four aggregates of a step) is dropped, so every step pays for it once
            G.lib.b200_BATinvalidate(dev["val"]._p)
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
        dist.barrier()
        G.comm_destroy()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=None,
                    help="lineitem rows of the WHOLE table (default SF100 = 600037902); per GPU for --workload config1/join/window")
    ap.add_argument("--device-only", nargs="?", const="q1", default=None, choices=["q1", "q6"],
                    help="run only the device-resident steps of one plan (for ncu launch lists)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity leg")
    ap.add_argument("--no-nil-run", action="store_true", help="skip the 1 %%-nil second run")
    ap.add_argument("--phases", action="store_true", help="join: time partition / exchange / join separately (adds syncs)")
    ap.add_argument("--workload", default="q1", choices=["q1", "config1", "join", "window"],
                    help="q1 (default) is the contract line (Q1 + Q6 on SF100); the others record the remaining BASELINE configs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.rows is None:
        args.rows = N_SF100 if args.workload == "q1" else N_SF10
    if args.impl == "reference":
        reference_arm(args)
    elif args.workload != "q1":
        other_workload(args)
    else:
        if args.warmup < 3 and not args.device_only:
            args.warmup = 3
        b200_arm(args)


if __name__ == "__main__":
    main()
