#!/usr/bin/env python
"""Turn an ncu launch list (csv, `--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum`)
of `bench.py --device-only --steps 1` into
  * a markdown summary (kernel, launches per step, time per launch, share of the step, DRAM bytes per launch,
    DRAM GB/s) for profiles/, and
  * an entry of profiles/r2_dram_bytes.json that bench.py reads for `roofline.traffic`.

    python scripts/ncu_dram_table.py launches.csv q1 ROWS_PER_GPU STEPS_IN_CAPTURE out.md

Only the launches of the LAST step of the capture are used (the warm-up steps come first)."""
import csv
import json
import re
import sys
from collections import OrderedDict
from pathlib import Path

# per-kernel CUDA-event names used by bench.py (KTimer, csrc/*.cu) -> kernel function names in the launch list
KTIMER = OrderedDict([
    ("calc", r"calc_vec_kernel|calc_gen_kernel"),
    ("aggr", r"aggr_reg_fast_kernel|aggr_reg_kernel|aggr_smem_kernel|aggr_glob_kernel"),
    ("project", r"project_kernel|compact_gather_kernel|project_void_kernel"),
    ("group_assign_g", r"grp_byte_assign_kernel<.*true|grp_assign_kernel<[^,]+, *\(bool\)1|grp_assign_kernel<[^,]+, *true"),
    ("group_assign", r"grp_byte_assign_kernel|grp_assign_kernel"),
    ("group_first_g", r"grp_byte_first_kernel<.*true|grp_first_kernel<[^,]+, *\(bool\)1|grp_first_kernel<[^,]+, *true"),
    ("group_first", r"grp_byte_first_kernel|grp_first_kernel|grp_direct_first_kernel"),
    ("gid_count", r"gid_count_kernel|gid8_count_kernel"),
    ("compact_expand", r"compact_expand_kernel"),
    ("compact", r"compact_mask2_kernel|compact_mask_kernel"),
    ("select_dense", r"select_dense32_kernel|select_dense_kernel"),
])


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name)
    return name


def main():
    src, workload, rows, steps, out_md = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    launches = OrderedDict()           # id -> {name, ns, rd, wr}
    with open(src, newline="") as f:
        rows_iter = (ln for ln in f if ln.startswith('"'))
        for r in csv.DictReader(rows_iter):
            e = launches.setdefault(r["ID"], {"name": short(r["Kernel Name"]), "ns": 0.0, "rd": 0.0, "wr": 0.0})
            v = float(r["Metric Value"].replace(",", ""))
            unit = r["Metric Unit"]
            mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6}.get(unit, 1)
            if r["Metric Name"] == "gpu__time_duration.sum":
                e["ns"] = v * mul
            elif r["Metric Name"] == "dram__bytes_read.sum":
                e["rd"] = v * mul
            elif r["Metric Name"] == "dram__bytes_write.sum":
                e["wr"] = v * mul
    ours = [e for e in launches.values() if not e["name"].startswith("at::") and "nccl" not in e["name"].lower()]
    per_step = len(ours) // steps
    last = ours[len(ours) - per_step:]
    tot_ns = sum(e["ns"] for e in last)
    agg = OrderedDict()
    for e in last:
        base = e["name"].split("(")[0]
        a = agg.setdefault(base, {"n": 0, "ns": 0.0, "dram": 0.0})
        a["n"] += 1
        a["ns"] += e["ns"]
        a["dram"] += e["rd"] + e["wr"]
    kt = {}
    for base, a in agg.items():
        for k, rx in KTIMER.items():
            if re.search(rx, base):
                t = kt.setdefault(k, {"n": 0, "ns": 0.0, "dram": 0.0})
                t["n"] += a["n"]
                t["ns"] += a["ns"]
                t["dram"] += a["dram"]
                break
    lines = [f"# ncu launch list, {workload}, {rows} rows per GPU — last of {steps} captured steps ({per_step} launches, "
             f"{tot_ns / 1e6:.3f} ms of kernel time; per-launch times under ncu are serialised and cold-cache: compare shares, not absolutes)",
             "", "| kernel | launches | us / launch | share of step | DRAM MB / launch (read+write) | DRAM GB/s |", "|---|---|---|---|---|---|"]
    for base, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        lines.append(f"| `{base[:150]}` | {a['n']} | {a['ns'] / a['n'] / 1e3:.1f} | {100 * a['ns'] / tot_ns:.1f} % | "
                     f"{a['dram'] / a['n'] / 1e6:.1f} | {a['dram'] / a['ns']:.0f} |")
    lines += ["", "| bench.py kernel group | launches | ms / step | DRAM MB / launch | DRAM GB/s |", "|---|---|---|---|---|"]
    for k, t in sorted(kt.items(), key=lambda kv: -kv[1]["ns"]):
        lines.append(f"| {k} | {t['n']} | {t['ns'] / 1e6:.3f} | {t['dram'] / t['n'] / 1e6:.1f} | {t['dram'] / t['ns']:.0f} |")
    lines.append(f"\nwhole step: {sum(a['dram'] for a in agg.values()) / 1e9:.3f} GB of DRAM traffic")
    Path(out_md).write_text("\n".join(lines) + "\n")
    table = Path(__file__).resolve().parent.parent / "profiles" / "r2_dram_bytes.json"
    d = json.loads(table.read_text()) if table.exists() else {}
    ent = {"rows_per_gpu": rows, "source": f"ncu launch list {Path(out_md).name} (dram__bytes_read.sum + dram__bytes_write.sum per launch)",
           "step_dram_bytes": sum(a["dram"] for a in agg.values()),
           "kernels": {k: t["dram"] / t["n"] for k, t in kt.items()}}
    d[workload] = [e for e in d.get(workload, []) if e.get("rows_per_gpu") != rows] + [ent]
    table.write_text(json.dumps(d, indent=1) + "\n")
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()
