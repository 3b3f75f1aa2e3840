#!/usr/bin/env python
"""Build the native libraries in-tree.

  libllmctl_b200.so : every sm_100a CUDA kernel behind a C ABI (include/llmctl_b200.h)
  libllmctl_host.so : C++ host runtime (KV block manager, batch scheduler, ...), no CUDA needed

Sources are compiled one translation unit per nvcc/g++ process (in parallel) and only when the
source or a header is newer than the object, so incremental rebuilds take seconds.  The reference
has no build system at all — nothing to mirror here.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "distributed-llm-training-and-inference-system_b200"
LIBDIR = PKG / "lib"
OBJDIR = ROOT / "build" / "obj"
CSRC = ROOT / "csrc"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-I", str(ROOT / "include")]


def find_nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def _newest_header_mtime() -> float:
    m = 0.0
    for pat in ("**/*.cuh", "**/*.h", "**/*.hpp"):
        for d in (CSRC, ROOT / "include"):
            for p in d.glob(pat):
                m = max(m, p.stat().st_mtime)
    return m


def _compile(cmd: list[str], src: Path, obj: Path, verbose: bool) -> tuple[Path, str]:
    obj.parent.mkdir(parents=True, exist_ok=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {src}\n{' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    return obj, (r.stdout + r.stderr) if verbose else ""


def build(force: bool = False, verbose: bool = False, ptxas_info: bool = False, jobs: int | None = None) -> dict:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    hdr_m = _newest_header_mtime()
    nvcc = find_nvcc()
    out: dict = {}

    # ---- CUDA kernels ----
    cu_srcs = sorted(p for p in CSRC.rglob("*.cu"))
    if cu_srcs:
        if nvcc is None:
            raise RuntimeError("nvcc not found; cannot build libllmctl_b200.so")
        tasks = []
        objs = []
        for src in cu_srcs:
            obj = OBJDIR / src.relative_to(CSRC).with_suffix(".o")
            objs.append(obj)
            if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_m):
                continue
            cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
            if ptxas_info:
                cmd[1:1] = ["-Xptxas", "-v"]
            tasks.append((cmd, src, obj))
        with cf.ThreadPoolExecutor(max_workers=jobs or os.cpu_count() or 4) as ex:
            for obj, log in ex.map(lambda t: _compile(*t, verbose or ptxas_info), tasks):
                if log.strip():
                    print(f"--- {obj.name}\n{log}")
        lib = LIBDIR / "libllmctl_b200.so"
        if tasks or not lib.exists():
            cmd = [nvcc, "-shared", "-o", str(lib), *map(str, objs), "-gencode",
                   "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-cudart", "shared"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        out["kernels"] = str(lib)

    # ---- host C++ ----
    cpp_srcs = sorted(p for p in (CSRC / "host").glob("*.cpp"))
    if cpp_srcs:
        cxx = os.environ.get("CXX", "g++")
        tasks = []
        objs = []
        for src in cpp_srcs:
            obj = OBJDIR / "host" / (src.stem + ".o")
            objs.append(obj)
            if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_m):
                continue
            tasks.append(([cxx, *CXX_FLAGS, "-c", str(src), "-o", str(obj)], src, obj))
        with cf.ThreadPoolExecutor(max_workers=jobs or os.cpu_count() or 4) as ex:
            list(ex.map(lambda t: _compile(*t, verbose), tasks))
        lib = LIBDIR / "libllmctl_host.so"
        if tasks or not lib.exists():
            r = subprocess.run([cxx, "-shared", "-o", str(lib), *map(str, objs), "-lpthread"],
                               capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        out["host"] = str(lib)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--ptxas-info", action="store_true", help="pass -Xptxas -v (registers/spills/smem)")
    ap.add_argument("-j", type=int, default=None)
    a = ap.parse_args()
    res = build(force=a.force, verbose=a.verbose, ptxas_info=a.ptxas_info, jobs=a.j)
    for k, v in res.items():
        print(f"built {k}: {v}")
    sys.exit(0)
