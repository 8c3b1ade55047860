"""Column-slab width lists for sgc_b200.precompute_host on the Reddit shape: where the narrow slab goes
(first: the download stream starts earlier; last: shorter tail).  python tools/e2e_widths.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200  # noqa: E402
from sgc_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
s = synth.make_shape("reddit", device=dev)
n, f, k = s["n"], s["f"], s["k"]
host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val", "x")}
out = torch.empty((n, f), dtype=torch.float32).pin_memory()
del s
torch.cuda.empty_cache()
csr = (n, host["row_ptr"], host["col"], host["val"])


def timeit(fn, reps=6):
    fn()
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e3


ref = None
cands = [
    [128, 128, 128, 128, 90],
    [92, 128, 128, 128, 126],
    [64, 128, 128, 128, 128, 26],
    [64, 128, 128, 128, 88, 66],
    [32, 128, 128, 128, 128, 58],
    [64, 64, 128, 128, 128, 90],
    [32, 96, 128, 128, 128, 90],
    [48, 128, 128, 128, 128, 42],
    [128, 128, 128, 128, 56, 34],
    [256, 256, 90],
    [64] * 9 + [26],
]
best = {}
for rnd in range(2):
    for w in cands:
        ms = timeit(lambda: sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, col_slab=w))
        if ref is None:
            ref = out.clone()
        same = torch.equal(out, ref)
        key = "+".join(map(str, w))
        best[key] = min(ms, best.get(key, 1e9))
        print("round %d  %-34s %.2f ms  identical=%s" % (rnd, key, ms, same), flush=True)
print("best:", sorted(best.items(), key=lambda kv: kv[1])[:4])
w = [int(x) for x in min(best, key=best.get).split("+")]
tr = []
sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, col_slab=w, trace=tr)
print("timeline of %s (device ms / host ms):" % w)
for label, d_ms, h_ms in tr:
    print("   %-34s %8.2f %8.2f" % (label, d_ms, h_ms))
