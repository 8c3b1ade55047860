"""EXPERIMENT (VERDICT r1 item 6c / SURVEY section 7 step 5): gathers of X staged through shared memory by the
TMA unit (cp.async.bulk global -> shared per neighbour row piece, mbarrier completion, 1-3 batches of 8 pieces in
flight per warp; sgc_exp_spmm_hop_bulk_gather -- and its tensor form, cp.async.bulk.tensor.2d tile::gather4, four rows
per instruction through a CUtensorMap; sgc_exp_spmm_hop_gather4) against the product hop kernel (per-lane LDG.128 gathers).
Checks bit-identity on a chunk = 0 plan first, then times both on the Reddit shape.
    python tools/sweep_tma_gather.py > profiles/r02_tma_gather_experiment.log"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth, _ffi
from sgc_b200.propagate import _hop_args, _partial, default_hints
from sgc_b200.graph import stream_ptr

GATHER4 = os.environ.get("SGC_SWEEP_GATHER4", "1") != "0"
GATHER4_OK = True
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)


def exp_hop(graph, x, out, chunk, depth, ctas, entry="sgc_exp_spmm_hop_bulk_gather"):
    plan = graph.plan(chunk)
    f = x.shape[1]
    partial = _partial(plan, f, dev)
    a = _hop_args(graph, plan, x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), f, partial, None, None,
                  "inplace", 0, default_hints(), pad_owned=True)
    _ffi.call(entry, plan.handle, _ffi.ctypes.byref(a), depth, ctas, stream_ptr())
    return partial


def t_ms(fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


# 1. correctness at small scale, no row splitting: bit-identical to the product kernel
s = synth.make_shape("reddit", device=dev, scale=0.02, seed=3)
g = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
for f in (602, 128, 100):
    x = sgc_b200.to_padded(s["x"][:, :f].contiguous())
    want = sgc_b200.hop(g, x, chunk=0)
    for depth in (1, 2, 3):
        out = sgc_b200.features.padded_rows(s["n"], f, dev)
        out.fill_(-1.0)
        exp_hop(g, x, out, 0, depth, 4)
        torch.cuda.synchronize()
        print("check f=%d depth=%d bit-identical to the product hop: %s" % (f, depth, bool(torch.equal(out, want))), flush=True)
        assert torch.equal(out, want)
        if GATHER4:
            out.fill_(-1.0)
            exp_hop(g, x, out, 0, depth, 4, "sgc_exp_spmm_hop_gather4")
            torch.cuda.synchronize()
            same = bool(torch.equal(out, want))
            print("check f=%d depth=%d tile::gather4 bit-identical to the product hop: %s" % (f, depth, same), flush=True)
            if not same:
                bad = (out != want).nonzero()
                print("   first mismatches (row, col):", bad[:5].tolist(), "of", int(bad.shape[0]), flush=True)
                GATHER4_OK = False
del s, g
# 2. timing at the Reddit shape
s = synth.make_shape("reddit", device=dev)
n, f = s["n"], s["f"]
g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
x = sgc_b200.to_padded(s["x"])
out = sgc_b200.features.padded_rows(n, f, dev)
print("product hop (LDG.128 gathers, ticket grid):      %.3f ms" % t_ms(lambda: sgc_b200.hop(g, x, out=out, pad_owned=True)), flush=True)
for depth in (1, 2, 3):
    for ctas in (3, 4, 5, 6, 7):
        if depth * 16 * ctas > 220:
            continue
        ms = t_ms(lambda: exp_hop(g, x, out, 512, depth, ctas))
        print("TMA-staged gathers depth=%d (%2d KB smem/CTA) %d CTAs/SM: %.3f ms" % (depth, depth * 16, ctas, ms), flush=True)
if GATHER4 and GATHER4_OK:
    for depth in (1, 2, 3):
        for ctas in (3, 4, 5, 6, 7):
            if depth * 16 * ctas > 220:
                continue
            ms = t_ms(lambda: exp_hop(g, x, out, 512, depth, ctas, "sgc_exp_spmm_hop_gather4"))
            print("tile::gather4 (4 rows per TMA instruction) depth=%d (%2d KB smem/CTA) %d CTAs/SM: %.3f ms" % (depth, depth * 16, ctas, ms), flush=True)
elif GATHER4:
    print("tile::gather4 results differ from the product hop: not timed", flush=True)
