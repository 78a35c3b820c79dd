import os, sys, ctypes
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import torch
import test_zz_no_alloc as T
lib = T._lib()
B, n, p, L = 64, 70, 3000, 12
x, y, w, lam = T._problem(100, B, n, p, L)
n_pad = lib.cml_lognet_n_pad(n); nx = 128
xt = torch.zeros(p, n_pad, device="cuda"); xt[:, :n] = x.t().cuda()
yp = torch.zeros(n_pad, device="cuda"); yp[:n] = y.cuda()
wp = torch.zeros(B, n_pad, device="cuda"); wp[:, :n] = w.cuda()
lam_d = lam.cuda()
ptr = lambda t: ctypes.c_void_p(t.data_ptr())
nbytes = lib.cml_lognet_scratch_bytes(B, n_pad, p)
def run(fill):
    o = T._outputs(B, L, nx, n_pad, p)
    scratch = torch.empty(nbytes // 4, dtype=torch.float32, device="cuda")
    if fill == "zero": scratch.zero_()
    elif fill == "nan": scratch.fill_(float("nan"))
    elif fill == "big": scratch.fill_(1e30)
    head = [B, n, n_pad, p, L, nx, ptr(xt), ptr(yp), 0, ptr(wp), None, 0, None, ptr(lam_d), 1.0, 1e-7, 0, 0, 100000]
    outs = [ptr(o[k]) for k in ("m", "ca", "kin", "a0", "dev", "eta", "freq", "nulldev", "npasses", "lmu", "jerr")]
    rc = lib.cml_lognet_path(*head, ptr(scratch), nbytes, *outs, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    return o
base = run("zero")
for fill in ["zero", "zero", "nan", "nan", "big", "zero"]:
    o = run(fill)
    bad = torch.nonzero(base["npasses"] != o["npasses"]).flatten().tolist()
    print(fill, "npasses diffs at reps", bad, "a0 diff", int((base["a0"] != o["a0"]).sum()), "eta diff", int((base["eta"] != o["eta"]).sum()), flush=True)
