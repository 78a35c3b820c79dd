"""Summarise an .ncu-rep by CUDA source line (SASS metrics from the report + line table from nvdisasm).
usage: python tools/ncu_hot.py report.ncu-rep <cubin-name-substring> <kernel-name-substring> [top_n]"""
import csv, io, os, re, subprocess, sys, collections, glob, tempfile
rep, cub_pat, kern_pat = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "consensusml_b200/_lib/libcml_b200.so")], cwd=tmp, capture_output=True)
cub = [f for f in glob.glob(tmp + "/*.cubin") if cub_pat in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
line_of = {}
cur = None; inside = False
for l in dis:
    if l.startswith(".text.") and l.rstrip().endswith(":"):
        inside = kern_pat in l
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/', l)
    if m: line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
h = rows[hi]; ix = {n: i for i, n in enumerate(h)}
base = None
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot_i = tot_s = 0
for r in rows[hi + 1:]:
    if len(r) != len(h): continue
    addr = int(r[ix["Address"]], 16)
    if base is None: base = addr
    inst = int(r[ix["Instructions Executed"]]); samp = int(r[ix["# Samples"]])
    key = line_of.get(addr - base, ("?", 0))
    a = agg[key]; a[0] += inst; a[1] += samp
    for k in h:
        if k.startswith("stall_") and "Not Issued" not in k and r[ix[k]].isdigit(): a[2][k[6:]] += int(r[ix[k]])
    tot_i += inst; tot_s += samp
src_cache = {}
def src(key):
    f, n = key
    p = os.path.join(root, "csrc", f)
    if p not in src_cache and os.path.exists(p): src_cache[p] = open(p).read().splitlines()
    L = src_cache.get(p)
    return L[n - 1].strip()[:90] if L and 0 < n <= len(L) else ""
print(f"total warp-instructions {tot_i:,}  samples {tot_s:,}")
print("== by samples")
for key, (inst, samp, st) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    s3 = ", ".join(f"{k}:{100*v//max(samp,1)}%" for k, v in st.most_common(3))
    print(f"{100*samp/tot_s:5.1f}% samp {100*inst/tot_i:5.1f}% inst  {key[0]}:{key[1]:<4} {src(key)}  [{s3}]")
