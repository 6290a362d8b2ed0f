#!/usr/bin/env python
"""Extract the golden vectors the reference's own tests hold for the hot path into small JSON
fixtures (run in the build container, where /root/reference exists; the fixtures are committed).

  select_maltest.json  monetdb5/modules/kernel/Tests/select.maltest: the full li/hi/anti matrix of
                       algebra.select on {7,6,8,nil,9,5,0,3,2,1,4}, its sorted and reverse-sorted
                       copies (192 cases) -> expected sorted value lists
  known_answers.json   bigsum.maltest (aggr.sum lng->dbl = 1.000000001e+16) and the TPC-H SF0.01 /
                       SF1 Q6 answers quoted by sql/benchmarks/tpch
"""
import json
import re
import sys
from pathlib import Path

REF = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
OUT = Path(__file__).resolve().parent


def select_cases():
    lines = (REF / "monetdb5/modules/kernel/Tests/select.maltest").read_text().splitlines()
    pat = re.compile(r"x := algebra\.select\((\w), nil:bat\[:oid\], ([^,]+), ([^,]+), (\w+), (\w+), (\w+)\)")
    cases, i = [], 0
    while i < len(lines):
        m = pat.match(lines[i].strip())
        if not m:
            i += 1
            continue
        col, lo, hi, li, hi_incl, anti = m.groups()
        # the next "query II rowsort / io.print(z) / ----" block holds (row, value) pairs
        j = i
        while not lines[j].startswith("io.print(z)"):
            j += 1
        j += 2
        vals = []
        while j < len(lines) and lines[j].strip() != "":
            vals.append(lines[j].strip())
            j += 1
        vals = vals[1::2]
        cases.append({
            "col": col,
            "lo": None if lo.startswith("nil") else int(lo), "hi": None if hi.startswith("nil") else int(hi),
            "li": li == "true", "hi_incl": hi_incl == "true", "anti": anti == "true",
            "values": sorted([None if v == "NULL" else int(v) for v in vals], key=lambda v: (v is not None, v)),
        })
        i = j
    return cases


def main():
    cases = select_cases()
    (OUT / "select_maltest.json").write_text(json.dumps({
        "source": "monetdb5/modules/kernel/Tests/select.maltest",
        "b": [7, 6, 8, None, 9, 5, 0, 3, 2, 1, 4], "cases": cases}, indent=0))
    q6 = (REF / "sql/benchmarks/tpch/Tests/one.test.in").read_text()
    m = re.search(r"1193053\.225\d*", q6)
    (OUT / "known_answers.json").write_text(json.dumps({
        "bigsum": {"source": "monetdb5/modules/mal/Tests/bigsum.maltest", "first": 10000000000000000, "ones": 10000000,
                   "sum_dbl": 1.000000001e+16},
        "tpch_sf001_q6": {"source": "sql/benchmarks/tpch/Tests/one.test.in", "revenue": m.group(0) if m else None},
    }, indent=1))
    print(len(cases), "select cases")


if __name__ == "__main__":
    main()
