"""L2 -> SM read bandwidth of the box: sgc_l2_read_probe over buffer sizes from
L2-resident to DRAM-resident, both access orders, several CTA counts.  bench.py runs the 32/64 MB points in
every N=1 line (`roofline.l2.peak`); this tool prints the whole curve.

    python tools/l2_peak.py > profiles/r02_l2_peak.log
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sgc_b200 import _ffi

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
sink = torch.zeros(1, device=dev)
print("%8s %8s %6s %10s %10s" % ("MB", "order", "CTAs", "ms", "GB/s"))
for mb in (8, 16, 32, 48, 64, 96, 128, 256, 1024):
    buf = torch.ones(mb * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    reps = max(2, 2048 // mb)
    for pattern in (0, 1, 2, 3):
        for ctas in (4, 8):
            call = lambda: _ffi.call("sgc_l2_read_probe", buf.data_ptr(), buf.numel() * 4, 0, reps, pattern, ctas,
                                     sink.data_ptr(), torch.cuda.current_stream().cuda_stream)
            call()
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); call(); b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            ms = min(ts)
            print("%8d %8s %6d %10.3f %10.0f" % (mb, ("hashed" if pattern & 1 else "address") + ("+hint" if pattern & 2 else ""),
                                                 ctas, ms, mb * 1024 * 1024 * reps / (ms * 1e-3) / 1e9), flush=True)
    del buf

# the Reddit-shaped hop's footprint: 512-byte pieces (one 128-float slab) of 232,965 rows at a 2432-byte pitch
n, pitch = 232965, 2432
buf = torch.ones(n * pitch // 4, dtype=torch.float32, device=dev)
print("strided: %d pieces of 512 B at a %d-byte pitch (%.0f MB touched in %.0f MB)" % (n, pitch, n * 512 / 1e6, n * pitch / 1e6))
for pattern in (0, 1, 2, 3):
    for ctas in (4, 7, 8):
        reps = 20
        call = lambda: _ffi.call("sgc_l2_read_probe", buf.data_ptr(), buf.numel() * 4, pitch, reps, pattern, ctas,
                                 sink.data_ptr(), torch.cuda.current_stream().cuda_stream)
        call()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); call(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = min(ts)
        print("%8s %8s %6d %10.3f %10.0f" % ("slab", ("hashed" if pattern & 1 else "address") + ("+hint" if pattern & 2 else ""), ctas, ms,
                                             n * 512 * reps / (ms * 1e-3) / 1e9), flush=True)
