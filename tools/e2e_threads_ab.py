"""Host-to-host step (Reddit shape) against the number of host threads of the value scan (SGC_B200_HOST_THREADS):
the scan's verdict gates the normalisation, which gates the first slab's hops."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200
from sgc_b200 import synth

dev = torch.device("cuda", 0)
s = synth.make_shape("reddit", device=dev)
n, f, k = s["n"], s["f"], s["k"]
host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val", "x")}
out = torch.empty((n, f), dtype=torch.float32).pin_memory()
del s
csr = (n, host["row_ptr"], host["col"], host["val"])
print("host cpus: %d" % os.cpu_count())
for rep in range(2):
    for threads in ("4", "8", "12", "16", "24", "32"):
        if int(threads) > (os.cpu_count() or 1):
            continue
        os.environ["SGC_B200_HOST_THREADS"] = threads
        sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out)
        t0 = time.perf_counter()
        for _ in range(5):
            sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out)
        ms = (time.perf_counter() - t0) / 5 * 1e3
        tr = []
        sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, trace=tr)
        marks = {l: (d, h) for l, d, h in tr}
        print("scan threads %2s: %.2f ms per step   values scanned at host %.2f ms, normalised at device %.2f ms, first slab computed %.2f ms"
              % (threads, ms, marks["values scanned"][1], marks["normalised"][0],
                 [d for l, d, h in tr if l.endswith("computed")][0]), flush=True)
