"""A/B: the single-GPU hop on the persistent ticket-scheduled grid (default) vs one warp per task in a one-shot
grid (hints bit 7), for 1..16 tasks per ticket.
    python tools/sweep_persistent.py [workload]"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
from sgc_b200.propagate import default_hints

def t_hop(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(min(ts))

dev = torch.device("cuda", 0)
for name in (sys.argv[1:] or ["reddit", "products"]):
    s = synth.make_shape(name, device=dev)
    n, f = s["n"], s["f"]
    graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    x = sgc_b200.to_padded(s["x"])
    out = sgc_b200.features.padded_rows(n, f, dev)
    out2 = sgc_b200.features.padded_rows(n, f, dev)
    h = default_hints()
    base = t_hop(lambda: sgc_b200.hop(graph, x, out=out, hints=h | 128, pad_owned=True))
    print("%-9s one-shot grid                      median %.3f ms  best %.3f ms" % (name, *base), flush=True)
    for rows in (1, 2, 4, 8, 16):
        os.environ["SGC_B200_PUSH_ROWS"] = str(rows)
        t = t_hop(lambda: sgc_b200.hop(graph, x, out=out2, pad_owned=True))
        same = bool(torch.equal(out, out2))
        print("%-9s persistent, %2d tasks per ticket     median %.3f ms  best %.3f ms  bit-identical %s" % (name, rows, *t, same), flush=True)
    os.environ.pop("SGC_B200_PUSH_ROWS")
    t = t_hop(lambda: sgc_b200.hop(graph, x, out=out2, pad_owned=True))
    print("%-9s persistent, default tasks per ticket median %.3f ms  best %.3f ms  bit-identical %s"
          % (name, *t, bool(torch.equal(out, out2))), flush=True)
    del graph, x, out, out2, s
    torch.cuda.empty_cache()
