"""Hop time on a column block of the Reddit-shaped X (feature-sharded mode): block widths x slabs.
    python tools/sweep_block.py [ncu_width]   (with an argument: 3 hops of that width only, for ncu)"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
dev = torch.device("cuda", 0)
s = synth.make_shape("reddit", device=dev)
n = s["n"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
xfull = s["x"]
def timeit(fn, reps=9):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
if len(sys.argv) > 1:
    w = int(sys.argv[1])
    x = sgc_b200.to_padded(xfull[:, :w].contiguous()); out = sgc_b200.features.padded_rows(n, w, dev)
    for _ in range(3):
        sgc_b200.hop(g, x, out=out)
    torch.cuda.synchronize(); print("done", w); sys.exit(0)
full = timeit(lambda: sgc_b200.hop(g, sgc_b200.to_padded(xfull), out=sgc_b200.features.padded_rows(n, 602, dev)))
print("full 602 columns: %.3f ms -> ideal per 1/8: %.3f ms" % (full, full / 8), flush=True)
for w in (64, 76, 80, 96, 128, 152, 192, 304, 602):
    x = sgc_b200.to_padded(xfull[:, :w].contiguous())
    out = sgc_b200.features.padded_rows(n, w, dev)
    row = ["w=%-4d default %.3f ms" % (w, timeit(lambda: sgc_b200.hop(g, x, out=out)))]
    for slab in (32, 64, 128, 256):
        if slab <= 2 * w:
            row.append("slab%d %.3f" % (slab, timeit(lambda: sgc_b200.hop(g, x, out=out, slab=slab))))
    print("  ".join(row), "  eff vs full/602*w: %.2f" % (full / 602 * w / timeit(lambda: sgc_b200.hop(g, x, out=out))), flush=True)
