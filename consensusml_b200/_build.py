"""In-tree native build: nvcc (sm_100a) for ``csrc/*.cu`` and g++ for ``csrc/host/*.cpp``.

Two shared libraries are produced next to the package so they travel with the source tree:

* ``consensusml_b200/_lib/libcml_b200.so``  - every CUDA kernel + its C-ABI launcher
  (declared in ``include/cml_b200.h``), compiled with
  ``-gencode arch=compute_100a,code=sm_100a -lineinfo``.
* ``consensusml_b200/_lib/libcml_host.so``  - host-side C++ runtime (float64 oracles that double
  as the timed CPU baseline, count-file parser).

The reference has no build system at all; this file is the whole of ours.
nvcc cross-compiles without a GPU, so ``build()`` works on CPU-only machines.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import List

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = Path(__file__).resolve().parent / "_lib"
OBJDIR = ROOT / "build" / "obj"

CUDA_LIB = LIBDIR / "libcml_b200.so"
HOST_LIB = LIBDIR / "libcml_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "--expt-extended-lambda",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
    "-I", str(INCLUDE), "-I", str(CSRC),
]
GXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-fno-fast-math", "-I", str(INCLUDE), "-I", str(CSRC)]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")


def _digest(paths: List[Path], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _headers() -> List[Path]:
    return sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(INCLUDE.glob("*.h"))
                  + list((CSRC / "host").glob("*.h")))


def _run(cmd: List[str], log: Path) -> None:
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.write_text(" ".join(cmd) + "\n" + proc.stdout)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout)
        raise RuntimeError(f"native build failed: {' '.join(cmd[:3])} ... (see {log})")


def _compile_objects(sources: List[Path], compiler: List[str], flags: List[str], tag: str, force: bool) -> List[Path]:
    OBJDIR.mkdir(parents=True, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src in sources:
        obj = OBJDIR / f"{tag}_{src.stem}.o"
        stamp = OBJDIR / f"{tag}_{src.stem}.sha"
        dig = _digest([src] + hdrs, " ".join(flags))
        objs.append(obj)
        if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
            continue
        jobs.append((src, obj, stamp, dig))

    def one(job):
        src, obj, stamp, dig = job
        _run(compiler + flags + ["-c", str(src), "-o", str(obj)], OBJDIR / f"{tag}_{src.stem}.log")
        stamp.write_text(dig)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(one, jobs))
    return objs


def build_cuda(force: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libcml_b200.so``."""
    sources = sorted(CSRC.glob("*.cu"))
    if not sources:
        raise RuntimeError("no CUDA sources under csrc/")
    nvcc = _nvcc()
    objs = _compile_objects(sources, [nvcc], NVCC_FLAGS, "cu", force)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime for o in objs)
    if force or not CUDA_LIB.exists() or CUDA_LIB.stat().st_mtime < newest:
        _run([nvcc, "-shared", "-o", str(CUDA_LIB)] + [str(o) for o in objs]
             + ["-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-cudart", "static"],
             OBJDIR / "link_cuda.log")
    return CUDA_LIB


def build_host(force: bool = False) -> Path:
    """Compile ``csrc/host/*.cpp`` with g++ and link ``libcml_host.so``."""
    sources = sorted((CSRC / "host").glob("*.cpp"))
    gxx = shutil.which("g++") or "g++"
    objs = _compile_objects(sources, [gxx], GXX_FLAGS, "host", force)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime for o in objs)
    if force or not HOST_LIB.exists() or HOST_LIB.stat().st_mtime < newest:
        _run([gxx, "-shared", "-o", str(HOST_LIB)] + [str(o) for o in objs] + ["-lpthread"], OBJDIR / "link_host.log")
    return HOST_LIB


def build_all(force: bool = False, cuda: bool = True) -> None:
    build_host(force)
    if cuda:
        build_cuda(force)


def ptxas_report() -> str:
    """Concatenated ``-Xptxas -v`` output of the last CUDA build (registers / spills / smem per kernel)."""
    out = []
    for log in sorted(OBJDIR.glob("cu_*.log")):
        out.append(f"==== {log.stem[3:]}.cu ====")
        for line in log.read_text().splitlines():
            if "ptxas info" in line and ("Used" in line or "Compiling" in line or "spill" in line):
                out.append(line)
    return "\n".join(out)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, cuda="--host-only" not in sys.argv)
    print(ptxas_report())
