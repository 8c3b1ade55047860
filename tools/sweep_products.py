"""Single-GPU hop timings on the products-shaped workload (F=100): slab widths, chunking, and the
slab-major layout used by the sharded path."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
name = sys.argv[1] if len(sys.argv) > 1 else "products"
dev = torch.device("cuda", 0)
s = synth.make_shape(name, device=dev)
n, f = s["n"], s["f"]
graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = sgc_b200.to_padded(s["x"])
bytes_hop = graph.nnz * (8 + 4 * f) + 4 * f * n + 8 * (n + 1)
d = (graph.row_ptr[1:] - graph.row_ptr[:-1])
print("workload", name, "n", n, "nnz", graph.nnz, "f", f, "GB/hop", bytes_hop / 1e9, "max row", int(d.max()), flush=True)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
out = sgc_b200.features.padded_rows(n, f, dev)
for slab in (32, 64, 128):
    for chunk in (256, 512, 2048):
        for hints in (0, 3):
            ms = timeit(lambda: sgc_b200.hop(graph, x, out=out, slab=slab, chunk=chunk, hints=hints))
            print("row-major   slab=%-4d chunk=%-5d hints=%d %8.3f ms %7.0f GB/s" % (slab, chunk, hints, ms, bytes_hop / ms / 1e6), flush=True)
# slab-major (what the sharded path uses): one narrow matrix per slab
for w in (32, 64):
    ns = (f + w - 1) // w
    xs = [torch.zeros((n, w), device=dev) for _ in range(ns)]
    ys = [torch.zeros((n, w), device=dev) for _ in range(ns)]
    for i in range(ns):
        fw = min(w, f - i * w)
        xs[i][:, :fw] = x[:, i * w:i * w + fw]
    def run():
        for i in range(ns):
            fw = min(w, f - i * w)
            sgc_b200.hop(graph, xs[i][:, :fw], out=ys[i][:, :fw])
    ms = timeit(run)
    print("slab-major  w=%-4d (%d launches) %8.3f ms %7.0f GB/s" % (w, ns, ms, bytes_hop / ms / 1e6), flush=True)
tsp = graph.to_torch_sparse().coalesce().to_sparse_csr()
xc = s["x"]
ms = timeit(lambda: torch.mm(tsp, xc), reps=3)
print("torch.mm (cuSPARSE, CSR)     %8.3f ms  %7.0f GB/s" % (ms, bytes_hop / ms / 1e6), flush=True)
