"""Golden vectors for BATgroup refinement (monetdb5/modules/mal/Tests/groupby00.maltest).

Reads the reference's own MAL test where it lies under /root/reference: three int columns filled by bat.append and ten
group.multicolumn(...) calls whose group-id columns the test prints.  Writes tests/golden/groupby00.json: the columns
and, per call, the argument list and the expected group id of every row.  Run in the build container only; the JSON is
committed."""
import json
import re
from pathlib import Path

SRC = Path("/root/reference/monetdb5/modules/mal/Tests/groupby00.maltest")
OUT = Path(__file__).parent / "groupby00.json"


def main():
    lines = [ln.strip() for ln in SRC.read_text().splitlines()]
    cols, cases, pending = {}, [], {}
    i = 0
    while i < len(lines):
        ln = lines[i]
        m = re.match(r"bat\.append\((\w+),(-?\d+)\)", ln)
        if m:
            cols.setdefault(m.group(1), []).append(int(m.group(2)))
        m = re.match(r"\((\w+),grp,ext\):= group\.multicolumn\(([\w,]+)\)", ln)
        if m:
            pending[m.group(1)] = m.group(2).split(",")
        m = re.match(r"io\.print\((\w+)\)", ln)
        if m and m.group(1) in pending and lines[i - 1].startswith("query II"):
            assert lines[i + 1] == "----"
            vals, j = [], i + 2
            while j < len(lines) and lines[j] != "":
                vals.append(int(lines[j]))
                j += 1
            pairs = sorted(zip(vals[0::2], vals[1::2]))
            assert [p[0] for p in pairs] == list(range(len(pairs)))
            cases.append({"mal": "group.multicolumn(%s)" % ",".join(pending[m.group(1)]), "columns": pending[m.group(1)],
                          "gids": [p[1] for p in pairs]})
            i = j
            continue
        i += 1
    assert len(cols) == 3 and cases
    OUT.write_text(json.dumps({"source": "monetdb5/modules/mal/Tests/groupby00.maltest", "columns": cols, "cases": cases},
                              indent=1) + "\n")
    print(len(cases), "cases ->", OUT)


if __name__ == "__main__":
    main()
