#!/usr/bin/env python
"""Undefined-global check without third-party linters: every LOAD_GLOBAL / LOAD_NAME of every code object of a module must
be a name the module binds at top level (or a builtin).  Catches the typo that only a GPU-only branch would execute.

    python tools/check_names.py [paths...]        default: the package, bench*, tests, tools"""
import builtins
import dis
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def module_bindings(code):
    names = set()
    for ins in dis.get_instructions(code):
        if ins.opname in ("STORE_NAME", "STORE_GLOBAL", "IMPORT_NAME"):
            names.add(str(ins.argval).split(".")[0])
    return names


def walk(code):
    yield code
    for c in code.co_consts:
        if isinstance(c, types.CodeType):
            yield from walk(c)


def check(path):
    src = open(path).read()
    try:
        code = compile(src, path, "exec")
    except SyntaxError as exc:
        return [f"{path}: syntax error {exc}"]
    bound = module_bindings(code) | set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__builtins__", "__spec__", "__package__", "__class__", "__annotations__", "__module__", "__qualname__"}
    for c in walk(code):                          # global statements inside functions bind module names too
        for ins in dis.get_instructions(c):
            if ins.opname == "STORE_GLOBAL":
                bound.add(ins.argval)
    out = []
    for c in walk(code):
        for ins in dis.get_instructions(c):
            if ins.opname in ("LOAD_GLOBAL", "LOAD_NAME") and ins.argval not in bound:
                if c is code and ins.opname == "LOAD_NAME":
                    continue                      # module level: order-dependent, the import itself would have failed
                if ins.opname == "LOAD_NAME" and ins.argval in c.co_names and c.co_name != "<module>" and "<" not in c.co_name:
                    # class bodies load their own attributes by name
                    cls_bound = {i.argval for i in dis.get_instructions(c) if i.opname == "STORE_NAME"}
                    if ins.argval in cls_bound:
                        continue
                out.append(f"{path}:{ins.positions.lineno if ins.positions else '?'}: undefined name '{ins.argval}' in {c.co_name}")
    return out


def main():
    paths = sys.argv[1:] or [os.path.join(ROOT, d) for d in ("consensusml_b200", "bench", "tests", "tools")] + [os.path.join(ROOT, f) for f in ("bench.py", "__graft_entry__.py")]
    files = []
    for p in paths:
        if os.path.isdir(p):
            for d, _, fs in os.walk(p):
                files += [os.path.join(d, f) for f in fs if f.endswith(".py")]
        else:
            files.append(p)
    problems = []
    for f in sorted(files):
        problems += check(f)
    print("\n".join(problems) if problems else f"{len(files)} files: no undefined names")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
