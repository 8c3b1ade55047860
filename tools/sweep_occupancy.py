"""A/B of the hop kernels' register cap: libsgc_b200.so (launch bounds 128 x 5: 72 registers, 7 CTAs per SM) against
builds with -DSGC_HOP_MIN_CTAS=8 (56 registers) and =10 (48 registers), no spills in any of them.  Each build runs
in its own process (SGC_B200_LIB).  Reddit shape (full width, an 80-column block), products shape; results
checked bit-identical across builds.
    python tools/sweep_occupancy.py > profiles/r02_sweep_occupancy.log"""
import hashlib, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import os, sys, hashlib
import numpy as np, torch
sys.path.insert(0, %r)
import sgc_b200
from sgc_b200 import synth
dev = torch.device("cuda", 0)
def t_ms(fn, reps=9):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
out = []
for name in ("reddit", "products"):
    s = synth.make_shape(name, device=dev)
    g = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    x = sgc_b200.to_padded(s["x"])
    y = sgc_b200.features.padded_rows(s["n"], s["f"], dev)
    ms = t_ms(lambda: sgc_b200.hop(g, x, out=y[:, :s["f"]], pad_owned=True))
    sha = hashlib.sha256(y[:4096].cpu().numpy().tobytes()).hexdigest()[:12]
    out.append("%%s full width %%.3f ms (sha %%s)" %% (name, ms, sha))
    if name == "reddit":
        xb = sgc_b200.to_padded(s["x"][:, :80].contiguous())
        yb = sgc_b200.features.padded_rows(s["n"], 80, dev)
        out.append("reddit 80-column block %%.3f ms" %% t_ms(lambda: sgc_b200.hop(g, xb, out=yb[:, :80], pad_owned=True)))
        k2 = t_ms(lambda: sgc_b200.propagate.propagate(g, x, 2), reps=7)
        out.append("reddit K=2 step %%.3f ms" %% k2)
    del s, g, x, y
    torch.cuda.empty_cache()
print(" | ".join(out))
''' % ROOT

for lib, what in (("libsgc_b200.so", "bounds 128x5 (72 regs, 7 CTAs/SM)"), ("libsgc_b200_occ8.so", "bounds 128x8 (56 regs)"),
                  ("libsgc_b200_occ10.so", "bounds 128x10 (48 regs)"), ("libsgc_b200.so", "bounds 128x5 again")):
    if not os.path.exists(os.path.join(ROOT, "sgc_b200", lib)):
        print("%s: not built" % lib)
        continue
    env = dict(os.environ, SGC_B200_LIB=lib)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print("%-34s %s" % (what, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "FAILED: " + r.stderr[-400:]), flush=True)
