"""Peer variant of the hop on ONE GPU (stand-in local peer buffers), products shape: tile size (tasks per ticket and\nbulk store) against compute time -- the shared memory a tile stages in is taken from L1.\n    python tools/sweep_peer_tiles_1gpu.py"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200
from sgc_b200 import synth
from sgc_b200.graph import SgcGraph
from sgc_b200.propagate import default_hints
dev = torch.device("cuda", 0); H = default_hints()
def t_ms(fn, reps=9):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))
s = synth.make_shape("products", device=dev); n, f = s["n"], s["f"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"]); g.values32(); x = sgc_b200.to_padded(s["x"]); del s
for p_rows in (2, 4, 8):
    n_pad = (n + p_rows - 1) // p_rows
    b, e = int(g.row_ptr[0]), int(g.row_ptr[n_pad])
    sub = SgcGraph(n_pad, n, (g.row_ptr[:n_pad + 1] - b).contiguous(), g.col[b:e].contiguous(), None, None, sval=g.sval[b:e].contiguous())
    own = sgc_b200.features.padded_rows(n, f, dev); peers = [sgc_b200.features.padded_rows(n, f, dev) for _ in range(p_rows - 1)]
    pp = [p.data_ptr() for p in peers]; o = own[:n_pad, :f]
    line = "products 1/%d rows, %d stand-in peers: plain %.3f" % (p_rows, p_rows - 1, t_ms(lambda: sgc_b200.hop(sub, x, out=o, pad_owned=True)))
    for rows in ("1", "2", "4", "8"):
        os.environ["SGC_B200_PUSH_ROWS"] = rows
        line += " | tile %s: no push %.3f pushing %.3f" % (rows, t_ms(lambda: sgc_b200.hop(sub, x, out=o, peers=pp, pad_owned=True, pitch_owned=True, hints=H | 8)),
                                                            t_ms(lambda: sgc_b200.hop(sub, x, out=o, peers=pp, pad_owned=True, pitch_owned=True)))
    print(line, flush=True)
    del peers, own, sub; torch.cuda.empty_cache()
