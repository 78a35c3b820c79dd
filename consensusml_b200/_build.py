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
EMU_LIB = LIBDIR / "libcml_emu.so"
EMUDIR = ROOT / "build" / "emu"

# every kernel source is also built for the host execution model (csrc/emu/): warp / block primitives in cuda_emu.h, the
# asynchronous units the tcgen05 / TMA GEMMs drive (mbarrier, TMA, TMEM, tcgen05.mma) in async_emu.cpp
EMU_EXCLUDE: set = set()

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
                  + list((CSRC / "host").glob("*.h")) + list((CSRC / "emu").glob("*.h")))


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


def _match_paren(text: str, i: int) -> int:
    """Index just past the parenthesis that closes the one at ``text[i]``."""
    depth = 0
    for j in range(i, len(text)):
        c = text[j]
        if c == "(":
            depth += 1
        elif c == ")":
            depth -= 1
            if depth == 0:
                return j + 1
    raise ValueError("unbalanced parentheses in a kernel launch")


def _split_top(text: str) -> List[str]:
    parts, depth, cur = [], 0, []
    for c in text:
        if c in "([{":
            depth += 1
        elif c in ")]}":
            depth -= 1
        if c == "," and depth == 0:
            parts.append("".join(cur).strip())
            cur = []
        else:
            cur.append(c)
    parts.append("".join(cur).strip())
    return parts


def emu_translate(src: str) -> str:
    """Rewrite the two CUDA constructs g++ cannot parse: ``kernel<<<grid, block, smem, stream>>>(args);`` becomes a call
    of ``cml_emu::launch`` with the kernel invocation as its body, and ``extern __shared__ T name[];`` becomes a pointer
    to the emulated dynamic shared memory.  Everything else is handled by macros and inline functions in cuda_emu.h."""
    import re
    out = []
    pos = 0
    while True:
        i = src.find("<<<", pos)
        if i < 0:
            out.append(src[pos:])
            break
        # kernel name (with optional template arguments) ends at i
        j = i
        if src[j - 1] == ">":
            depth = 0
            while True:
                j -= 1
                if src[j] == ">":
                    depth += 1
                elif src[j] == "<":
                    depth -= 1
                    if depth == 0:
                        break
        while j > 0 and (src[j - 1].isalnum() or src[j - 1] in "_:"):
            j -= 1
        name = src[j:i]
        k = src.index(">>>", i)
        cfg = _split_top(src[i + 3:k])
        while len(cfg) < 3:
            cfg.append("0")
        a0 = src.index("(", k)
        a1 = _match_paren(src, a0)
        args = src[a0:a1]
        out.append(src[pos:j])
        label = name.replace('"', "").replace("\\", "")
        out.append(f"(::cml_emu::launch_name(\"{label}\"), ::cml_emu::launch(dim3({cfg[0]}), dim3({cfg[1]}), (size_t)({cfg[2]}), [&]() {{ {name}{args}; }}))")
        pos = a1
    text = "".join(out)
    text = re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?([\w:]+(?:\s+[\w:]+)*?)\s+(\w+)\[\];",
                  r"\1* \2 = reinterpret_cast<\1*>(::cml_emu::dyn_smem());", text)
    return text


def build_emu(force: bool = False) -> Path:
    """Build ``libcml_emu.so``: the kernel sources compiled by g++ against the host execution model."""
    EMUDIR.mkdir(parents=True, exist_ok=True)
    gen = []
    for src in sorted(CSRC.glob("*.cu")):
        if src.name in EMU_EXCLUDE:
            continue
        dst = EMUDIR / f"{src.stem}.emu.cpp"
        text = f'#line 1 "{src}"\n' + emu_translate(src.read_text())
        if not dst.exists() or dst.read_text() != text:
            dst.write_text(text)
        gen.append(dst)
    gen.append(CSRC / "emu" / "cuda_emu.cpp")
    gen.append(CSRC / "emu" / "async_emu.cpp")
    gxx = shutil.which("g++") or "g++"
    flags = ["-O1", "-g", "-std=c++17", "-fPIC", "-DCML_EMU=1", "-fno-fast-math", "-ffp-contract=off", "-Wno-unknown-pragmas",
             "-Wno-attributes", "-I", str(INCLUDE), "-I", str(CSRC), "-I", str(CSRC / "emu"), "-include", str(CSRC / "emu" / "cuda_emu.h")]
    san = os.environ.get("CML_EMU_SANITIZE")          # "1" / "address" = ASan + UBSan (needs LD_PRELOAD of libasan), "undefined" = UBSan only
    san_flags = ["-fsanitize=undefined", "-fno-sanitize-recover=undefined"] if san == "undefined" else (["-fsanitize=address,undefined"] if san else [])
    if san:
        flags += san_flags + ["-fno-omit-frame-pointer"]
    objs = _compile_objects(gen, [gxx], flags, "emu", force)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime for o in objs)
    if force or not EMU_LIB.exists() or EMU_LIB.stat().st_mtime < newest:
        link = [gxx, "-shared", "-o", str(EMU_LIB)] + [str(o) for o in objs] + ["-lpthread"]
        if san:
            link += san_flags
        _run(link, OBJDIR / "link_emu.log")
    return EMU_LIB


def build_all(force: bool = False, cuda: bool = True) -> None:
    build_host(force)
    build_emu(force)
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
