"""Column-blocked hop (SGC_HOP_ACCUMULATE, propagate.hop_col_blocked) against the single-pass hop on the
products shape (X = 1 GB, 8x the L2): per-hop time for 1..24 column blocks, bit-identity on an exact-mode
plan, and K = 3 through the public entry point.
    python tools/sweep_col_blocks.py [products|reddit] > profiles/r02_sweep_col_blocks.log"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
from sgc_b200.propagate import hop_col_blocked

name = sys.argv[1] if len(sys.argv) > 1 else "products"
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)


def t_ms(fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


s = synth.make_shape(name, device=dev)
n, f, k = s["n"], s["f"], s["k"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = sgc_b200.to_padded(s["x"])
del s
out = sgc_b200.features.padded_rows(n, f, dev)
os.environ["SGC_B200_COL_BLOCKS"] = "0"
print("%s shape: N = %d, nnz(S) = %d, F = %d, X = %.0f MB" % (name, n, g.nnz, f, n * x.stride(0) * 4 / 1e6), flush=True)
base = t_ms(lambda: sgc_b200.hop(g, x, out=out[:, :f], pad_owned=True))
print("single-pass hop (hot-column tags, ticket grid):  %.3f ms" % base, flush=True)
exact = sgc_b200.hop(g, x, chunk=0)
for nb in (2, 4, 6, 8, 10, 12, 16, 24):
    got = hop_col_blocked(g, x, nb, chunk=0)
    same = bool(torch.equal(got, exact))
    del got
    ms = t_ms(lambda: hop_col_blocked(g, x, nb, out=out[:, :f], pad_owned=True))
    blocks = g.col_blocks(nb)
    print("column-blocked hop, %2d blocks of %7d columns (%5.1f MB of X each): %.3f ms   exact mode bit-identical: %s"
          % (len(blocks), blocks[0][1] - blocks[0][0], (blocks[0][1] - blocks[0][0]) * x.stride(0) * 4 / 1e6, ms, same), flush=True)
    g._plans.pop(("col_blocks", nb), None)          # one block set at a time
    torch.cuda.empty_cache()
for env in ("0", "auto"):
    os.environ["SGC_B200_COL_BLOCKS"] = env
    ms = t_ms(lambda: sgc_b200.propagate.propagate(g, x, k), reps=5)
    print("K = %d through propagate(), SGC_B200_COL_BLOCKS=%s (%d blocks): %.3f ms" %
          (k, env, sgc_b200.propagate.col_blocks_for(g, x.stride(0) * 4), ms), flush=True)
