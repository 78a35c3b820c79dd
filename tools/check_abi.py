#!/usr/bin/env python
"""Python <-> C ABI check: every signature the op modules register with ``_native.register_signature`` is compared with the
prototype of the same name in include/cml_b200.h (argument count, pointer / integer / floating class of every argument, return
class), and every name must be exported by the built libraries.  ctypes would silently pass a mis-declared call.

    python tools/check_abi.py"""
import ctypes
import importlib
import os
import pkgutil
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def c_class(decl: str) -> str:
    d = decl.strip()
    if "*" in d or d.endswith("[]"):
        return "ptr"
    base = re.sub(r"\b(const|unsigned|signed|struct)\b", " ", d)
    toks = base.split()
    ty = " ".join(toks[:-1]) if len(toks) > 1 else toks[0]
    if "double" in ty:
        return "f64"
    if "float" in ty:
        return "f32"
    if "size_t" in ty or "long long" in ty or "uint64" in ty or "int64" in ty:
        return "i64"
    return "i32"


def py_class(t) -> str:
    if t in (ctypes.c_void_p, ctypes.c_char_p) or (isinstance(t, type) and issubclass(t, ctypes._Pointer)):
        return "ptr"
    if t is ctypes.c_double:
        return "f64"
    if t is ctypes.c_float:
        return "f32"
    if t in (ctypes.c_longlong, ctypes.c_ulonglong, ctypes.c_size_t, ctypes.c_int64, ctypes.c_uint64):
        return "i64"
    if t in (ctypes.c_int, ctypes.c_uint, ctypes.c_int32, ctypes.c_uint32, ctypes.c_short, ctypes.c_ushort, ctypes.c_bool):
        return "i32"
    return "?"


def header_prototypes():
    text = open(os.path.join(ROOT, "include", "cml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(cml_\w+)\s*\(([^;{]*?)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        arglist = [] if args in ("", "void") else [a for a in args.split(",")]
        protos[name] = (c_class(ret + " x") if "*" not in ret else "ptr", [c_class(a) for a in arglist])
    return protos


def main():
    import consensusml_b200
    for m in pkgutil.walk_packages(consensusml_b200.__path__, "consensusml_b200."):
        if m.name.endswith("__main__"):
            continue
        try:
            importlib.import_module(m.name)
        except Exception as exc:                               # noqa: BLE001
            print(f"import of {m.name} failed: {exc!r}")
    from consensusml_b200 import _build, _native
    protos = header_prototypes()
    problems = []
    for name, (args, res) in sorted(_native._SIGNATURES.items()):
        if name not in protos:
            problems.append(f"{name}: registered in Python, no prototype in include/cml_b200.h")
            continue
        ret_c, args_c = protos[name]
        args_p = [py_class(a) for a in args]
        if len(args_p) != len(args_c):
            problems.append(f"{name}: {len(args_p)} arguments registered, {len(args_c)} in the header")
            continue
        for i, (a, b) in enumerate(zip(args_p, args_c)):
            if a != b:
                problems.append(f"{name}: argument {i} is {a} in Python, {b} in the header")
        rp = "i32" if res is ctypes.c_int else py_class(res)
        if rp != ret_c and not (ret_c == "i64" and rp == "i64"):
            problems.append(f"{name}: return {rp} in Python, {ret_c} in the header")
    for lib_path in (_build.CUDA_LIB, _build.EMU_LIB):
        if not os.path.exists(lib_path):
            print(f"(not built: {lib_path})")
            continue
        import subprocess
        syms = set(l.split()[-1] for l in subprocess.run(["nm", "-D", "--defined-only", str(lib_path)], capture_output=True, text=True).stdout.splitlines() if l.strip())
        for name in sorted(_native._SIGNATURES):
            if name not in syms and not (name.startswith("cml_host_") or name in ("cml_lognet_cpu",)):
                problems.append(f"{name}: not exported by {os.path.basename(str(lib_path))}")
    print("\n".join(problems) if problems else f"{len(_native._SIGNATURES)} registered entry points agree with the header and are exported")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
