"""The drop-in boundary: integration/gdk_shim.c exports the reference's own entry points
(BATselect, BATproject, BATgroup, BATgroupsum, BATjoin, GDKanalytical* ... with the prototypes of gdk.h /
gdk_calc.h / gdk_analytic.h — the file is compiled against those headers) and is interposed in front of the
real libbat.  The same driver script runs config 1, Q6, Q1, a join, error cases and a window chain on BATs
created by the real GDK, once against plain libbat_ref and once with the shim loaded first; tails, counts,
nil properties and GDKerrbuf texts must be identical, and the shim must actually have run the calls on the
device with intermediates served from its device-twin cache."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
SHIM = ROOT / "integration" / "_build" / "libbat_b200shim.so"


def _run(n, shim):
    env = dict(os.environ)
    env.pop("REFGDK_SHIM", None)
    if shim:
        env["REFGDK_SHIM"] = str(SHIM)
    r = subprocess.run([sys.executable, str(ROOT / "tests" / "shim_driver.py"), str(n)], env=env, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_shim_is_a_drop_in(ref, gdk):
    if not SHIM.exists():
        pytest.skip("integration/_build/libbat_b200shim.so not built (needs /root/reference at build time)")
    n = 400_000
    plain = _run(n, False)
    shim = _run(n, True)
    assert plain["shim_stats"] is None
    dev, fwd, hits, staged = shim["shim_stats"]
    assert dev >= 40, shim["shim_stats"]           # the plans ran on the device ...
    assert hits >= 25, shim["shim_stats"]          # ... with intermediates taken from the device-twin cache
    for k in ("config1", "q1", "window", "order", "calc"):
        for name, d in plain[k].items():
            assert shim[k][name] == d, (k, name, d, shim[k][name])
    assert shim["q6"] == plain["q6"]
    # dictionary strings: BATgroup on str flags, (anti-)equality selects incl. the nil forms
    assert shim["str"] == plain["str"], (shim["str"], plain["str"])
    assert dev >= 80, shim["shim_stats"]           # + the two str BATgroup calls, 12 str selects, the ORDER BY / join block
    assert shim["join"] == plain["join"]
    assert shim["errors"] == plain["errors"], (shim["errors"], plain["errors"])
    assert "overflow" in plain["errors"]["mul_overflow"] and "division by zero" in plain["errors"]["div_zero"]


def test_shim_exports_reference_entry_points():
    """nm: the symbols north_star names are defined (T) by the shim"""
    if not SHIM.exists():
        pytest.skip("shim not built")
    out = subprocess.run(["nm", "-D", "--defined-only", str(SHIM)], capture_output=True, text=True).stdout
    have = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    for s in ("BATselect", "BATthetaselect", "BATproject", "BATproject2", "BATgroup", "BATgroupsum", "BATgroupprod", "BATgroupcount", "BATgroupmin",
              "BATgroupmax", "BATgroupavg", "BATsum", "BATjoin", "BATleftjoin", "BATouterjoin", "BATsemijoin", "BATintersect", "BATdiff",
              "BATunique", "BATconvert", "BATsort", "BATfirstn", "BATcalcmod", "BATcalcand", "BATcalcxorcst", "BATcalcnot", "BATcalcisnil",
              "BATcalcmin", "BATcalccstmax_no_nil", "BATcalclsh", "BATcalcbetween", "BATcalcbetweencstcst", "BATcalcifthenelse",
              "BATcalcifthencstelsecst", "BATgroupprod", "BATcalcadd", "BATcalcsub", "BATcalcmul", "BATcalcdiv",
              "BATcalcmulcst", "BATcalccstsub", "BATcalclt", "BATcalceq", "GDKanalyticaldiff", "GDKanalyticalwindowbounds",
              "GDKanalyticalsum", "GDKanalyticalavg", "GDKanalyticalcount", "GDKanalyticalmin", "GDKanalyticalmax"):
        assert s in have, s
