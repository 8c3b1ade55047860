"""How much of the hop's gather traffic could on-SM staging of hub rows remove?  

For the Reddit-shaped graph: the fraction of stored entries of S whose column is among the H most frequently
gathered ones -- i.e. the share of the gathers a per-SM shared-memory copy of the H hottest rows of the current
feature slab would serve.  One 128-float slab piece of a row is 512 B, so ~200 KB of shared memory hold H ~ 400
rows per SM; a cluster of 8 CTAs sharing their copies over DSMEM holds ~3200.  Runs on the CPU (the synthetic
graph is a counter-based hash: identical on any device).

    python tools/hub_coverage.py [workload] > profiles/r02_hub_coverage_<workload>.log
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import importlib.util
spec = importlib.util.spec_from_file_location("_synth", os.path.join(ROOT, "sgc_b200", "synth.py"))
synth = importlib.util.module_from_spec(spec)
spec.loader.exec_module(synth)

name = sys.argv[1] if len(sys.argv) > 1 else "reddit"
s = synth.make_shape(name, device="cpu", with_features=False)
col, n = s["col"].long(), s["n"]
nnz = col.numel() + n                                  # + the self-loops A + I adds
cnt = torch.bincount(col, minlength=n) + 1
srt, _ = torch.sort(cnt, descending=True)
cum = torch.cumsum(srt, 0).double() / nnz
print("workload %s: N = %d, nnz(S) = %d, max gather count = %d, mean = %.1f" % (name, n, nnz, int(srt[0]), nnz / n))
print("%8s %10s %14s %s" % ("H rows", "coverage", "min count", "smem for a 128-float slab piece (512 B/row)"))
# L2 budgets: how many whole padded rows of X fit, and the share of all gathers they would serve if exactly the
# most-gathered rows stayed resident (the ceiling of any hot-set / residency scheme)
f = synth.SHAPES[name]["f"]
row_bytes = (f + 7) // 8 * 8 * 4
print("L2 residency ceiling, whole rows of %d bytes:" % row_bytes)
for mb in (16, 32, 48, 64, 96, 126):
    h = min(n, mb * (1 << 20) // row_bytes)
    print("  %4d MB = %8d rows (%5.2f %% of N): %5.1f %% of the gathers, min count %d" % (
        mb, h, 100.0 * h / n, 100 * cum[h - 1].item(), int(srt[h - 1])))
for h in (128, 256, 384, 400, 512, 768, 1024, 1600, 2048, 3200, 4096, 8192, 16384, 32768):
    if h <= n:
        print("%8d %9.1f%% %14d %8.0f KB" % (h, 100 * cum[h - 1].item(), int(srt[h - 1]), h * 512 / 1024))
