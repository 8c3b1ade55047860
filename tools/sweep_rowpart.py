"""What does one rank of a (column blocks x row parts) grid pay for working on a PART of the Reddit-shaped
problem?  One GPU: the hop on the first 1/P_r of the rows of S and a column block of X, one-shot grid vs ticket
grid, plain kernel vs the peer variant (pushes skipped / pushed into local stand-in buffers), tile sizes.
    python tools/sweep_rowpart.py > profiles/r02_sweep_rowpart.log"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
import sgc_b200.dist
from sgc_b200 import synth
from sgc_b200.graph import SgcGraph
from sgc_b200.propagate import default_hints

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)


def t_ms(fn, reps=9):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


s = synth.make_shape("reddit", device=dev)
n, f = s["n"], s["f"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
g.values32()
x_full = s["x"]
del s
H = default_hints()
full = t_ms(lambda: sgc_b200.hop(g, sgc_b200.to_padded(x_full)))
print("whole problem, 602 columns: %.3f ms per hop" % full, flush=True)
for p_cols, p_rows in ((1, 2), (1, 4), (1, 8), (2, 2), (2, 4), (4, 2), (8, 1), (2, 1), (4, 1)):
    w = sgc_b200.dist.partition_features(f, p_cols)[0]
    w = w[1] - w[0]
    x = sgc_b200.to_padded(x_full[:, :w].contiguous())
    n_pad = (n + p_rows - 1) // p_rows
    r0, r1 = 0, n_pad
    b, e = int(g.row_ptr[r0]), int(g.row_ptr[r1])
    sub = SgcGraph(r1 - r0, n, (g.row_ptr[r0:r1 + 1] - b).contiguous(), g.col[b:e].contiguous(), None, None,
                   sval=g.sval[b:e].contiguous())
    out = sgc_b200.features.padded_rows(n, w, dev)
    own = out[r0:r1, :w]
    peers = [sgc_b200.features.padded_rows(n, w, dev) for _ in range(p_rows - 1)]
    pp = [p[r0:].data_ptr() for p in peers]
    ideal = full * (w / f) / p_rows
    line = "grid %dx%d: %3d columns, %6d rows, %8d entries | ideal %.3f" % (p_cols, p_rows, w, r1 - r0, sub.nnz, ideal)
    line += " | plain auto %.3f  one-shot %.3f  tickets %.3f" % (
        t_ms(lambda: sgc_b200.hop(sub, x, out=own, pad_owned=True)),
        t_ms(lambda: sgc_b200.hop(sub, x, out=own, pad_owned=True, hints=H | 128)),
        t_ms(lambda: sgc_b200.hop(sub, x, out=own, pad_owned=True, hints=H | 256)))
    if pp:
        for rows in ("1", "2", "8"):
            os.environ["SGC_B200_PUSH_ROWS"] = rows
            line += " | peer variant tile %s: no push %.3f  push to %d local buffers %.3f" % (
                rows, t_ms(lambda: sgc_b200.hop(sub, x, out=own, peers=pp, pad_owned=True, hints=H | 8)), len(pp),
                t_ms(lambda: sgc_b200.hop(sub, x, out=own, peers=pp, pad_owned=True)))
        os.environ.pop("SGC_B200_PUSH_ROWS", None)
    print(line, flush=True)
    del peers, out, x, sub
    torch.cuda.empty_cache()
