import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from consensusml_b200 import datasets
from consensusml_b200.models import lasso as L
co = datasets.load_target_aml(); dev = torch.device("cuda")
L.run_lasso(co, device=dev)
for n in (1, 15):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    reps = L.exclusion_reps(co, nreps=n, device=dev)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{n} exclusion reps (LOOCV each): {dt:.3f} s  -> {dt/n*1e3:.0f} ms per rep; nnz {[len(r.nonzero_coef) for r in reps]}", flush=True)
