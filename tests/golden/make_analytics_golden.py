"""Golden vectors for the window path (sql/test/analytics/Tests/analytics02.test).

Reads the reference's own sqllogictest files (analytics00 and analytics02) where they lie under /root/reference, keeps the
aggregates the device path covers — sum / avg / min / max / count over the default frames (whole partition; unbounded
preceding to the current row's peers with ORDER BY) and over `ROWS BETWEEN n PRECEDING AND m FOLLOWING` — and writes
tests/golden/analytics0N.json: the table's rows and, per query, the expected
column exactly as the test file lists it (strings, NULL for nil, reals with three decimals) together with its sort mode.
The SQL text is kept beside every case so a reader can find it in the reference.  Run in the build container only (the
GPU box has no /root/reference); the JSON is committed.
"""
import json
import re
from pathlib import Path

SRCDIR = Path("/root/reference/sql/test/analytics/Tests")
OUTDIR = Path(__file__).parent

INSERT = re.compile(r"insert into analytics values (.*)$")
# fn(col) OVER ([PARTITION BY x ORDER BY x] ROWS BETWEEN n PRECEDING AND m FOLLOWING)
ROWS = re.compile(
    r"^select (?:cast\()?(sum|avg|min|max|count)\((\*|aa|bb|cc)\) over \("
    r"(?:partition by (\w+) order by (\w+) )?rows between (\d+) preceding and (\d+) following\)"
    r"(?: as bigint\))? from analytics$")
# fn(col) OVER ([PARTITION BY x] [ORDER BY x [ASC|DESC]]) — the default frames: whole partition without ORDER BY,
# RANGE UNBOUNDED PRECEDING .. CURRENT ROW (peers included) with it
DEFAULT = re.compile(
    r"^select (?:cast\()?(sum|avg|min|max|count)\((\*|aa|bb|cc)\) over \("
    r"(?:partition by (\w+))? ?(?:order by (\w+)(?: (asc|desc))?)?\)"
    r"(?: as bigint\))? from analytics$")


def convert(name):
    lines = (SRCDIR / (name + ".test")).read_text().splitlines()
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
            j = i + 1
            while lines[j].strip() != "----":      # a query may span several lines
                j += 1
            sql = " ".join(x.strip() for x in lines[i + 1:j])
            j += 1
            exp = []
            while j < len(lines) and lines[j].strip() != "":
                exp.append(lines[j].strip())
                j += 1
            if len(kind) == 1 and kind in "IR":
                q, d = ROWS.match(sql), DEFAULT.match(sql)
                if q:
                    fn, col, part, order, lo, hi = q.groups()
                    if part == order:      # both absent, or the same column
                        cases.append({"sql": sql, "frame": "rows", "fn": fn, "col": col, "partition": part, "order": order,
                                      "desc": False, "preceding": int(lo), "following": int(hi), "type": kind, "sort": mode,
                                      "expected": exp})
                elif d:
                    fn, col, part, order, direction = d.groups()
                    if part is None or order is None or part == order:
                        cases.append({"sql": sql, "frame": "default", "fn": fn, "col": col, "partition": part, "order": order,
                                      "desc": direction == "desc", "type": kind, "sort": mode, "expected": exp})
            i = j
            continue
        i += 1
    assert table is not None and cases, name
    out = OUTDIR / (name + ".json")
    out.write_text(json.dumps({"source": "sql/test/analytics/Tests/%s.test" % name, "columns": ["aa", "bb", "cc"],
                               "types": ["int", "int", "lng"], "rows": table, "cases": cases}, indent=1) + "\n")
    print(len(cases), "cases ->", out)


if __name__ == "__main__":
    for name in ("analytics00", "analytics02"):     # analytics01 holds ranking functions only
        convert(name)
