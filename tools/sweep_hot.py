"""Hot-column budget sweep (L2 pinning of hub rows) on a workload whose X exceeds L2.
    python tools/sweep_hot.py [workload=products] [scale=1.0]"""
import os, sys
import numpy as np
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
del s
bytes_hop = graph.nnz * (8 + 4 * f) + 4 * f * n + 8 * (n + 1)
print("workload", name, "n", n, "nnz", graph.nnz, "f", f, "GB/hop", bytes_hop / 1e9, flush=True)
def timeit(fn, reps=7):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
out = sgc_b200.features.padded_rows(n, f, dev)
plan = graph.plan(sgc_b200.propagate.default_chunk())
from sgc_b200 import _ffi
import ctypes
for persist_mb in (0, 32, 64, 96, 1000):
  granted = ctypes.c_int64(0)
  _ffi.call("sgc_l2_set_persisting_bytes", persist_mb << 20, ctypes.byref(granted))
  print("== L2 persisting set-aside requested %d MB granted %.1f MB" % (persist_mb, granted.value / 2**20), flush=True)
  for rnd in range(1):
    for mb in (0, 16, 32, 48, 64, 80, 96, 112):
        if persist_mb == 0 and mb not in (0, 64): continue
        os.environ["SGC_B200_HOT_MB"] = str(mb)
        ms = timeit(lambda: sgc_b200.hop(graph, x, out=out))
        print("round %d hot_mb=%-4d n_hot=%-8d %8.3f ms %7.0f GB/s" % (rnd, mb, plan.n_hot, ms, bytes_hop / ms / 1e6), flush=True)
  for hints in (0, 1, 2, 3):
    os.environ["SGC_B200_HOT_MB"] = "0"
    ms = timeit(lambda: sgc_b200.hop(graph, x, out=out, hints=hints))
    print("no tags hints=%d %8.3f ms" % (hints, ms), flush=True)
