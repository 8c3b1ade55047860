"""A/B of the peer variant's register cap on ONE GPU (pushes into local stand-in buffers, as
tools/push_probe_1gpu.py does): libsgc_b200.so against builds with -DSGC_PEERS_MIN_CTAS=7 / 8.  A row part of the
products shape (1/8 of the rows, 7 stand-in peers; 1/2 of the rows, 1 peer) and of the Reddit shape (1/4 of the
rows, 3 peers); each build in its own process (SGC_B200_LIB); results compared across builds by a hash.
    python tools/sweep_peers_bounds.py > profiles/r02_sweep_peers_bounds.log"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import os, sys, hashlib
import numpy as np, torch
sys.path.insert(0, %r)
import sgc_b200
from sgc_b200 import synth
from sgc_b200.graph import SgcGraph
from sgc_b200.propagate import default_hints
dev = torch.device("cuda", 0)
H = default_hints()
def t_ms(fn, reps=9):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
out = []
for name, parts in (("products", (8, 2)), ("reddit", (4,))):
    s = synth.make_shape(name, device=dev)
    n, f = s["n"], s["f"]
    g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    g.values32()
    x = sgc_b200.to_padded(s["x"])
    del s
    for p_rows in parts:
        n_pad = (n + p_rows - 1) // p_rows
        b, e = int(g.row_ptr[0]), int(g.row_ptr[n_pad])
        sub = SgcGraph(n_pad, n, (g.row_ptr[:n_pad + 1] - b).contiguous(), g.col[b:e].contiguous(), None, None,
                       sval=g.sval[b:e].contiguous())
        own = sgc_b200.features.padded_rows(n, f, dev)
        peers = [sgc_b200.features.padded_rows(n, f, dev) for _ in range(p_rows - 1)]
        pp = [p.data_ptr() for p in peers]
        o = own[:n_pad, :f]
        plain = t_ms(lambda: sgc_b200.hop(sub, x, out=o, pad_owned=True))
        nopush = t_ms(lambda: sgc_b200.hop(sub, x, out=o, peers=pp, pad_owned=True, hints=H | 8))
        push = t_ms(lambda: sgc_b200.hop(sub, x, out=o, peers=pp, pad_owned=True))
        sha = hashlib.sha256(peers[-1][:2048].cpu().numpy().tobytes()).hexdigest()[:10]
        out.append("%%s 1/%%d rows, %%d peers: plain %%.3f  peer variant no push %%.3f  pushing %%.3f (sha %%s)"
                   %% (name, p_rows, p_rows - 1, plain, nopush, push, sha))
        del peers, own, sub
        torch.cuda.empty_cache()
    del g, x
    torch.cuda.empty_cache()
print(" | ".join(out))
''' % ROOT

for lib, what in (("libsgc_b200.so", "peer bounds 128x6 (80-reg cap, spills)"), ("libsgc_b200_peers7.so", "peer bounds 128x7 (72 regs)"),
                  ("libsgc_b200_peers8.so", "peer bounds 128x8 (64 regs)"), ("libsgc_b200.so", "peer bounds 128x6 again")):
    if not os.path.exists(os.path.join(ROOT, "sgc_b200", lib)):
        print("%s: not built" % lib)
        continue
    r = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, SGC_B200_LIB=lib), capture_output=True, text=True)
    print("%-40s %s" % (what, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "FAILED: " + r.stderr[-600:]), flush=True)
