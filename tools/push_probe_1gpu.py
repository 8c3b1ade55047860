"""Single-GPU emulation of one rank of the fused exchange: 1/P of the rows of a workload, P-1
local stand-in buffers in place of the peers' operands.  Separates the cost of the peer-variant
kernel, of issuing the pushes (bulk / per-lane stores) and of the extra HBM writes from NVLink.
    python tools/push_probe_1gpu.py [workload=products] [P=8]"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
from sgc_b200.graph import SgcGraph
name = sys.argv[1] if len(sys.argv) > 1 else "products"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda", 0)
s = synth.make_shape(name, device=dev)
n, f = s["n"], s["f"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = sgc_b200.to_padded(s["x"])
del s
ld = x.stride(0)
n_pad = (n + P - 1) // P
a, b = 0, n_pad
sub = SgcGraph(b - a, n, g.row_ptr[a:b + 1], g.col, None, None, sval=g.values32())
own = torch.zeros((n, ld), device=dev)
fake = [torch.zeros((n_pad, ld), device=dev) for _ in range(P - 1)]
out = own[a:b, :f]
peers = [t.data_ptr() for t in fake]
def timeit(fn, reps=15):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))
H = sgc_b200.propagate.default_hints()
print("workload %s rows %d of %d, f %d, stand-ins %d" % (name, b - a, n, f, P - 1), flush=True)
for rnd in range(2):
    print("plain variant, no peers            %.3f ms" % timeit(lambda: sgc_b200.hop(sub, x, out=out)), flush=True)
    print("peer variant, pushes skipped       %.3f ms" % timeit(lambda: sgc_b200.hop(sub, x, out=out, peers=peers, hints=H | 8)), flush=True)
    print("peer variant, bulk pushes (local)  %.3f ms" % timeit(lambda: sgc_b200.hop(sub, x, out=out, peers=peers, hints=H)), flush=True)
    print("peer variant, lane stores (local)  %.3f ms" % timeit(lambda: sgc_b200.hop(sub, x, out=out, peers=peers, hints=H | 16)), flush=True)
    for k in (1, 3):
        print("peer variant, bulk pushes to %d     %.3f ms" % (k, timeit(lambda: sgc_b200.hop(sub, x, out=out, peers=peers[:k], hints=H))), flush=True)
