"""Short program for ncu: 3 plain hops of a named workload with the default settings.
    python tools/profile_plain.py [workload] [scale]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
name = sys.argv[1] if len(sys.argv) > 1 else "products"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
dev = torch.device("cuda", 0)
s = synth.make_shape(name, device=dev, scale=scale)
n, f = s["n"], s["f"]
graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = sgc_b200.to_padded(s["x"])
out = sgc_b200.features.padded_rows(n, f, dev)
for rep in range(3):
    sgc_b200.hop(graph, x, out=out)
torch.cuda.synchronize()
plan = graph.plan(sgc_b200.propagate.default_chunk())
print("done", n, graph.nnz, f, "n_hot", plan.n_hot)
