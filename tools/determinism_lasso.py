"""Run the same batched path problem several times through both entry points and report bitwise differences."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_zz_no_alloc as T
lib = T._lib()
B, n, p, L = [int(a) for a in sys.argv[1:5]] if len(sys.argv) > 4 else (300, 70, 3000, 12)
x, y, w, lam = T._problem(100, B, n, p, L)
runs = []
for i, static in enumerate([False, False, True, True, False]):
    o = T._run(lib, static, x, y, w, lam)
    runs.append(o)
    if i:
        a = runs[0]
        d = {k: int((a[k] != o[k]).sum()) for k in ("npasses", "kin", "a0", "dev", "freq")}
        bad = torch.nonzero(a["npasses"] != o["npasses"]).flatten().tolist()[:10]
        print(f"run {i} static={static}: mismatching entries vs run 0: {d}; reps with different npasses: {bad}", flush=True)
