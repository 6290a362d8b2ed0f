"""Build the B200 GDK library in-tree with nvcc (sm_100a only).

    python -m monetdb_b200.build            # incremental
    python -m monetdb_b200.build --force

Produces monetdb_b200/libgdk_b200.so (C-ABI declared in include/gdk_b200.h).
nvcc cross-compiles without a GPU, so this runs on the CPU build box; the .so
travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
OBJ = CSRC / "_obj"
LIB = PKG / "libgdk_b200.so"
WLIB = PKG / "libworkload.so"          # host-side synthetic data generator (plain C, no CUDA)

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"),
]


def _deps_mtime() -> float:
    hdrs = list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return max(p.stat().st_mtime for p in hdrs)


def _compile(src: Path, force: bool, hdr_m: float, verbose: bool) -> Path:
    obj = OBJ / (src.stem + ".o")
    if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_m):
        return obj
    cmd = [NVCC, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose or r.stderr.strip():
        sys.stderr.write(r.stderr)
    return obj


def build_workload(force: bool = False) -> Path:
    src = PKG / "hostsrc" / "workload_host.c"
    if force or not WLIB.exists() or WLIB.stat().st_mtime < src.stat().st_mtime:
        tmp = WLIB.with_suffix(".so.tmp")
        r = subprocess.run([os.environ.get("CC", "gcc"), "-O2", "-fPIC", "-shared", "-Wall", "-o", str(tmp), str(src), "-lpthread"],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"gcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, WLIB)
    return WLIB


def build_library(force: bool = False, verbose: bool = False) -> Path:
    build_workload(force)
    OBJ.mkdir(exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    if not srcs:
        raise RuntimeError("no CUDA sources found")
    hdr_m = _deps_mtime()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, force, hdr_m, verbose), srcs))
    if force or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        tmp = LIB.with_suffix(".so.tmp")
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
               "-o", str(tmp), *map(str, objs), "-cudart", "static", "-lpthread", "-ldl", "-lrt",
               "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, LIB)  # atomic: a snapshot taken mid-build never sees a half-written library
    return LIB


if __name__ == "__main__":
    p = build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
