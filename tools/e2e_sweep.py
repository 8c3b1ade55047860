"""Times sgc_b200.precompute_host (host CSR(A) + X -> host result) on the Reddit shape for several
last-hop block counts, next to the serial schedule it replaces.  python tools/e2e_sweep.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200  # noqa: E402
from sgc_b200 import synth  # noqa: E402
from sgc_b200.features import StandardizedFeatures  # noqa: E402

dev = torch.device("cuda", 0)
s = synth.make_shape("reddit", device=dev)
n, f, k = s["n"], s["f"], s["k"]
host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val", "x")}
out = torch.empty((n, f), dtype=torch.float32).pin_memory()
del s
torch.cuda.empty_cache()


def serial():
    rp = host["row_ptr"].to(dev, non_blocking=True)
    col = host["col"].to(dev, non_blocking=True)
    val = host["val"].to(dev, non_blocking=True)
    x = sgc_b200.to_padded(host["x"], dev)
    graph = sgc_b200.normalize_csr(n, rp, col, val)
    mean, std = sgc_b200.col_stats(x)
    y, _ = sgc_b200.sgc_precompute(StandardizedFeatures(x, mean, std), graph, k)
    out.copy_(y, non_blocking=True)
    torch.cuda.synchronize()


def timeit(fn, reps=5):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e3


def copy_only():
    a = host["row_ptr"].to(dev, non_blocking=True)
    b = host["col"].to(dev, non_blocking=True)
    c = host["val"].to(dev, non_blocking=True)
    x = sgc_b200.to_padded(host["x"], dev)
    out.copy_(x[:, :f], non_blocking=True)
    torch.cuda.synchronize()


print("copies only (H2D 841 MB + D2H 561 MB, serial): %.2f ms" % timeit(copy_only))
print("serial schedule: %.2f ms" % timeit(serial))
ref = out.clone()
csr = (n, host["row_ptr"], host["col"], host["val"])
for blocks in (1, 8, 16):
    ms = timeit(lambda: sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, out_blocks=blocks,
                                                 col_slab=0))
    print("row-block schedule, out_blocks=%-2d: %.2f ms   identical to serial: %s" % (blocks, ms, torch.equal(out, ref)))
for slab in (64, 128, 256):
    out.zero_()
    ms = timeit(lambda: sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, col_slab=slab))
    print("column-slab schedule, slab=%-3d: %.2f ms   identical to serial: %s" % (slab, ms, torch.equal(out, ref)))


for ws in ([128] * 4 + [90], [128] * 4 + [64, 26], [64] + [128] * 4 + [26]):
    ms = timeit(lambda: sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, col_slab=ws))
    print("column-slab schedule, widths=%s: %.2f ms   identical: %s" % (ws, ms, torch.equal(out, ref)))
tr = []
sgc_b200.precompute_host(csr, host["x"], k, standardize=True, out=out, trace=tr)
print("device-side timeline of the default schedule (ms since start):")
for label, ms, host_ms in tr:
    print("   %8.2f  (enqueued at %6.2f)  %s" % (ms, host_ms, label))


def copy2d_only():
    from sgc_b200 import _ffi
    x = sgc_b200.features.padded_rows(n, f, dev)
    ld = x.stride(0)
    st = torch.cuda.current_stream().cuda_stream
    for c0 in range(0, f, 128):
        w = min(128, f - c0)
        _ffi.call("sgc_copy_2d", x.data_ptr() + 4 * c0, 4 * ld, host["x"].data_ptr() + 4 * c0, 4 * f, 4 * w, n, st)
    for c0 in range(0, f, 128):
        w = min(128, f - c0)
        _ffi.call("sgc_copy_2d", out.data_ptr() + 4 * c0, 4 * f, x.data_ptr() + 4 * c0, 4 * ld, 4 * w, n, st)
    torch.cuda.synchronize()


print("X up + down as 128-column 2-D slab copies, serial: %.2f ms (contiguous: see 'copies only' minus CSR)"
      % timeit(copy2d_only))
# stage breakdown of the adjacency side (device-resident inputs)
rp, col, val = (host[kk].to(dev) for kk in ("row_ptr", "col", "val"))


def norm_side():
    g = sgc_b200.normalize_csr(n, rp, col, val)
    g.values32()
    g.plan(512)
    torch.cuda.synchronize()


print("normalize_csr + values32 + plan on resident CSR: %.2f ms" % timeit(norm_side))
