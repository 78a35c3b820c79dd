"""Static checks that stand in for a first execution of the GPU-only branches: no undefined global names anywhere in the Python
sources, and every ctypes signature agrees with its prototype in include/cml_b200.h (count and class of the arguments) and is
exported by the libraries that are built."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(tool):
    return subprocess.run([sys.executable, os.path.join(ROOT, "tools", tool)], capture_output=True, text=True, cwd=ROOT, timeout=600)


def test_no_undefined_names():
    r = _run("check_names.py")
    assert r.returncode == 0, r.stdout[-3000:]


def test_python_signatures_match_the_c_header():
    r = _run("check_abi.py")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
