"""Short program for ncu: the adjacency normalisation (K1, K2, S values) of the Reddit-shaped graph,
fast path, twice (the second pass is the warm one).   python tools/profile_norm.py [workload]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "reddit"
dev = torch.device("cuda", 0)
s = synth.make_shape(name, device=dev, with_features=False)
for rep in range(2):
    g = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    torch.cuda.synchronize()
print("done", s["n"], g.nnz)
