"""Host-to-host precompute on the ogbn-products shape (X far larger than L2): one contiguous upload
against column slabs (pipelined upload / hops / download), value scan on and off; checks every result
against sgc_precompute on device-resident inputs.  python tools/e2e_products.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200  # noqa: E402
from sgc_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
s = synth.make_shape("products", device=dev)
n, f, k = s["n"], s["f"], s["k"]
host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
xh = s["x"].cpu().pin_memory()
graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(s["x"]), graph, k)
want = want.cpu()
del s, graph
torch.cuda.empty_cache()
out = torch.empty((n, f), dtype=torch.float32).pin_memory()


def run(label, reps=3, **kw):
    st = {}
    sgc_b200.precompute_host(host, xh, k, out=out, stats=st, **kw)
    t0 = time.perf_counter()
    for _ in range(reps):
        sgc_b200.precompute_host(host, xh, k, out=out, stats=st, **kw)
    ms = (time.perf_counter() - t0) / reps * 1e3
    print("products precompute_host %-34s %.1f ms  h2d %.0f MB  uniform=%s  identical=%s"
          % (label, ms, st["h2d_bytes"] / 1e6, st["uniform_value"], torch.equal(out, want)), flush=True)
    return ms


run("one slab, values uploaded", detect_uniform_values=False)
run("one slab")
for widths in ([52, 48], [36, 32, 32], [28, 24, 24, 24]):
    run("slabs %s" % "+".join(map(str, widths)), col_slab=widths)
tr = []
sgc_b200.precompute_host(host, xh, k, out=out, col_slab=[52, 48], trace=tr)
print("timeline, slabs 52+48 (device ms / host ms):")
for label, d_ms, h_ms in tr:
    print("   %-34s %8.2f %8.2f" % (label, d_ms, h_ms))
tr = []
sgc_b200.precompute_host(host, xh, k, out=out, trace=tr)
print("timeline, one slab (device ms / host ms):")
for label, d_ms, h_ms in tr:
    print("   %-34s %8.2f %8.2f" % (label, d_ms, h_ms))
