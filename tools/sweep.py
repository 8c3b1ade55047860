"""Parameter sweep of the hop kernel on the Reddit-shaped (or other) workload; prints ms per hop.
Run on the GPU box:  python tools/sweep.py [workload]"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "reddit"
dev = torch.device("cuda", 0)
s = synth.make_shape(name, device=dev)
n, f = s["n"], s["f"]
graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = s["x"]
mean, std = sgc_b200.col_stats(x)
ld = (f + 7) // 8 * 8
xpad = torch.zeros((n, ld), dtype=torch.float32, device=dev)
xpad[:, :f] = x
xv = xpad[:, :f]
bytes_hop = graph.nnz * (8 + 4 * f) + 4 * f * n + 8 * (n + 1)
print("workload", name, "n", n, "nnz", graph.nnz, "f", f, "GB/hop", bytes_hop / 1e9, flush=True)

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))

out = torch.empty((n, ld), dtype=torch.float32, device=dev)[:, :f]
outc = torch.empty((n, f), dtype=torch.float32, device=dev)
rows = []
def run(tag, xin, o, **kw):
    ms = timeit(lambda: sgc_b200.hop(graph, xin, out=o, **kw))
    rows.append((tag, kw.get("slab"), kw.get("hints"), kw.get("chunk"), ms, bytes_hop / ms / 1e6))
    print("%-28s slab=%-4s hints=%s chunk=%-5s %8.3f ms  %7.0f GB/s" % (tag, kw.get("slab"), kw.get("hints"), kw.get("chunk"), ms, bytes_hop / ms / 1e6), flush=True)

slabs = [32, 64, 128, 256]
for slab in slabs:
    for hints in (0, 1, 2, 3):
        run("plain pad->pad", xv, out, slab=slab, hints=hints, chunk=512)
for chunk in (0, 128, 256, 1024, 2048):
    run("plain pad->pad", xv, out, slab=128, hints=3, chunk=chunk)
    run("plain pad->pad", xv, out, slab=64, hints=3, chunk=chunk)
for slab in slabs:
    run("plain pad->contig", xv, outc, slab=slab, hints=3, chunk=512)
for slab in (32, 64, 128):
    run("plain contig(v2)->pad", x, out, slab=slab, hints=3, chunk=512)
for mode in ("div", "rcp"):
    for slab in slabs:
        run("std-%s pad->pad" % mode, xv, out, slab=slab, hints=3, chunk=512, mean=mean, std=std, std_mode=mode)
    for slab in (32, 64, 128):
        run("std-%s contig(v2)->pad" % mode, x, out, slab=slab, hints=3, chunk=512, mean=mean, std=std, std_mode=mode)
# materialised S (non-fused weights)
st = sgc_b200.SgcGraph(n, n, graph.row_ptr, graph.col, None, None, sval=graph.values32())
for slab in (64, 128):
    ms = timeit(lambda: sgc_b200.hop(st, xv, out=out, slab=slab, hints=3, chunk=512))
    print("materialised-S pad->pad      slab=%-4s %8.3f ms  %7.0f GB/s" % (slab, ms, bytes_hop / ms / 1e6), flush=True)
# incumbent: torch.spmm (cuSPARSE) on the same S
tsp = graph.to_torch_sparse().coalesce()
ms = timeit(lambda: torch.spmm(tsp, x), reps=3)
print("torch.spmm (cuSPARSE, COO)   %8.3f ms  %7.0f GB/s" % (ms, bytes_hop / ms / 1e6), flush=True)
tcsr = tsp.to_sparse_csr()
ms = timeit(lambda: torch.mm(tcsr, x), reps=3)
print("torch.mm (cuSPARSE, CSR)     %8.3f ms  %7.0f GB/s" % (ms, bytes_hop / ms / 1e6), flush=True)
