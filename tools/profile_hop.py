"""Short program for ncu: a few hops of the Reddit-shaped workload (plain, fused-std rcp, fused-std div).
    python tools/profile_hop.py [workload] [slab]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "reddit"
slab = int(sys.argv[2]) if len(sys.argv) > 2 else 0
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
out = torch.empty((n, ld), dtype=torch.float32, device=dev)[:, :f]
for rep in range(2):
    sgc_b200.hop(graph, xv, out=out, slab=slab)                                   # plain
    sgc_b200.hop(graph, xv, out=out, slab=slab, mean=mean, std=std, std_mode="rcp")
    sgc_b200.hop(graph, xv, out=out, slab=slab, mean=mean, std=std, std_mode="div")
torch.cuda.synchronize()
print("done", n, graph.nnz, f)
