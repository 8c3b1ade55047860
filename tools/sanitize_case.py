"""Small cases for compute-sanitizer (memcheck / racecheck, ONE tool per gpurun call -- B200_PROFILING.md):
the Cora golden smoke (K1, K2, S values, plan, K=2 hops incl. the sparse-X first hop), a Reddit-shaped toy with
long-row splitting + fix-up, the fused standardisation variants, and the multi-GPU hop's peer epilogue on one
GPU (two local buffers stand in for the peers' mapped operands: staging in shared memory, cp.async.bulk tile
and row stores, halo masks, the per-lane fallback).  Every result is compared with the plain hop.

    compute-sanitizer --tool memcheck  python tools/sanitize_case.py
    compute-sanitizer --tool racecheck python tools/sanitize_case.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth

dev = torch.device("cuda", 0)
g = np.load(os.path.join(ROOT, "tests", "golden", "cora.npz"), allow_pickle=False)
n = g["adj_indptr"].size - 1
graph = sgc_b200.normalize_csr(n, torch.from_numpy(g["adj_indptr"]).to(dev), torch.from_numpy(g["adj_indices"]).to(dev),
                               torch.from_numpy(g["adj_data"]).to(dev))
x = torch.randn(n, 1433, device=dev)
y, _ = sgc_b200.sgc_precompute(x, graph, 2)
print("cora K=2 ok", float(y.abs().sum()))

s = synth.make_shape("reddit", device=dev, scale=0.004, seed=5)
n = s["n"]
gr = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
for f in (100, 130, 602, 7):
    ld = (f + 7) // 8 * 8
    xs = s["x"][:, :f].contiguous()
    want = sgc_b200.hop(gr, xs, chunk=64)
    for rows_per_tile in (0, 4):
        os.environ["SGC_B200_PUSH_ROWS"] = str(rows_per_tile)
        bufs = [torch.full((n + 3, ld), -7.0, device=dev) for _ in range(3)]
        sgc_b200.hop(gr, xs, chunk=64, out=bufs[0][2:2 + n, :f], peers=[p[2:].data_ptr() for p in bufs[1:]],
                     pad_owned=True, pitch_owned=True)
        assert all(torch.equal(b[2:2 + n, :f], want) for b in bufs), (f, rows_per_tile)
        mask = torch.randint(0, 4, (n,), dtype=torch.int32, device=dev)
        bufs = [torch.full((n + 3, ld), -7.0, device=dev) for _ in range(3)]
        sgc_b200.hop(gr, xs, chunk=64, out=bufs[0][2:2 + n, :f], peers=[p[2:].data_ptr() for p in bufs[1:]],
                     peer_mask=mask, pad_owned=True, pitch_owned=True)
        for slot, b in enumerate(bufs[1:]):
            sent = ((mask >> slot) & 1).bool()
            assert torch.equal(b[2:2 + n, :f][sent], want[sent]) and bool((b[2:2 + n][~sent] == -7.0).all())
    ldu = f + 1 if (f + 1) % 2 else f + 2
    bufs = [torch.full((n + 3, ldu), -7.0, device=dev) for _ in range(3)]
    sgc_b200.hop(gr, xs, chunk=64, out=bufs[0][1:1 + n, :f], peers=[p[1:].data_ptr() for p in bufs[1:]])
    assert all(torch.equal(b[1:1 + n, :f], want) for b in bufs), f
    print("peer epilogue f=%d ok" % f)
mean, std = sgc_b200.col_stats(s["x"])
for mode in ("rcp", "div"):
    sgc_b200.hop(gr, s["x"], mean=mean, std=std, std_mode=mode)
xp = sgc_b200.to_padded(s["x"].clone())
sgc_b200.standardize_(xp, mean, std)
y, _ = sgc_b200.sgc_precompute(xp, gr, 2)
torch.cuda.synchronize()
print("sanitize_case done")
