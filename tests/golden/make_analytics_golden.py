"""Golden vectors for the window path (sql/test/analytics/Tests/analytics02.test).

Reads the reference's own sqllogictest file where it lies under /root/reference, keeps the framed aggregates the
device path covers — sum / avg / min / max / count over `ROWS BETWEEN n PRECEDING AND m FOLLOWING`, with and without
`PARTITION BY x ORDER BY x` — and writes tests/golden/analytics02.json: the table's rows and, per query, the expected
column exactly as the test file lists it (strings, NULL for nil, reals with three decimals) together with its sort mode.
The SQL text is kept beside every case so a reader can find it in the reference.  Run in the build container only (the
GPU box has no /root/reference); the JSON is committed.
"""
import json
import re
from pathlib import Path

SRC = Path("/root/reference/sql/test/analytics/Tests/analytics02.test")
OUT = Path(__file__).parent / "analytics02.json"

INSERT = re.compile(r"insert into analytics values (.*)$")
QUERY = re.compile(
    r"^select (?:cast\()?(sum|avg|min|max|count)\((\*|aa|bb|cc)\) over \("
    r"(?:partition by (\w+) order by (\w+) )?rows between (\d+) preceding and (\d+) following\)"
    r"(?: as bigint\))? from analytics$")


def main():
    lines = SRC.read_text().splitlines()
    table, cases = None, []
    i = 0
    while i < len(lines):
        ln = lines[i].strip()
        m = INSERT.match(ln)
        if m and table is None:
            rows = re.findall(r"\(([^)]*)\)", m.group(1))
            table = [[None if v.strip() == "NULL" else int(v) for v in r.split(",")] for r in rows]
        if ln.startswith("query "):
            kind, mode = ln.split()[1], ln.split()[2]
            sql = lines[i + 1].strip()
            j = i + 2
            assert lines[j].strip() == "----", (i, lines[j])
            j += 1
            exp = []
            while j < len(lines) and lines[j].strip() != "":
                exp.append(lines[j].strip())
                j += 1
            q = QUERY.match(sql)
            if q and len(kind) == 1 and kind in "IR":
                fn, col, part, order, lo, hi = q.groups()
                if part == order:      # both absent, or the same column
                    cases.append({"sql": sql, "fn": fn, "col": col, "partition": part, "preceding": int(lo),
                                  "following": int(hi), "type": kind, "sort": mode, "expected": exp})
            i = j
            continue
        i += 1
    assert table is not None and cases
    OUT.write_text(json.dumps({"source": "sql/test/analytics/Tests/analytics02.test", "columns": ["aa", "bb", "cc"],
                               "types": ["int", "int", "lng"], "rows": table, "cases": cases}, indent=1) + "\n")
    print(len(cases), "cases ->", OUT)


if __name__ == "__main__":
    main()
