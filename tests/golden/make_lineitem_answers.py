#!/usr/bin/env python
"""Golden answers of the Q1 / Q6 plans over the full synthetic lineitem tables, computed by the REAL reference
GDK (oracle/_ref) with bench.py's mitosis + mergetable driver:

    python tests/golden/make_lineitem_answers.py 600037902 [59986052 ...]

writes tests/golden/lineitem_answers.json {rows: {q1_sum_charge, q1_count_order, q6_hge}}.  bench.py compares
its result with these at every GPU count (`golden_checked`), in addition to the full in-run comparison with
the reference at N = 1 (`parity_checked`).  SF100 needs about 60 GB of host memory."""
import json
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

out = Path(__file__).with_name("lineitem_answers.json")
d = json.loads(out.read_text()) if out.exists() else {}
for rows in map(int, sys.argv[1:]):
    thr = bench.host_threads()
    tab = bench.ReferenceTable(rows, thr)
    with ThreadPoolExecutor(max_workers=thr) as ex:
        q1, q6, _, _ = tab.run(ex)
    d[str(rows)] = {"q1_sum_charge": [str(x) for x in q1["sum_charge"]], "q1_count_order": q1["count_order"], "q6_hge": str(q6),
                    "q1_extents": q1["extents"], "q1_sum_qty": [str(x) for x in q1["sum_qty"]]}
    del tab
out.write_text(json.dumps(d, indent=1) + "\n")
print(out.read_text())
