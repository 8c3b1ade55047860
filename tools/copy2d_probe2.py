"""Strided slab copies of the host pipeline (column slabs of a row-major pinned [N, 602] matrix <-> padded device
rows) -- how much of the contiguous duplex rate (50 + 50 GB/s, tools/pcie_duplex_probe.py) do they reach, as a
function of the piece width and of how many copy streams share one direction?  One GPU.
    python tools/copy2d_probe2.py > profiles/r02_copy2d_probe2.log"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sgc_b200 import _ffi

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
rows, f, ld = 232965, 602, 608
hx = torch.ones((rows, f), dtype=torch.float32).pin_memory()
hy = torch.empty((rows, f), dtype=torch.float32).pin_memory()
dx = torch.empty((rows, ld), dtype=torch.float32, device=dev)
dy = torch.ones((rows, ld), dtype=torch.float32, device=dev)
up_streams = [torch.cuda.Stream() for _ in range(4)]
dn_streams = [torch.cuda.Stream() for _ in range(4)]


def copy2d(dst, dpitch, src, spitch, width_bytes, n_rows, stream):
    _ffi.call("sgc_copy_2d", dst, dpitch, src, spitch, width_bytes, n_rows, stream.cuda_stream)


def enqueue(direction, widths, n_streams, row_split):
    """All column slabs of `widths` (floats) once, each slab split into `row_split` row ranges dealt round-robin
    to `n_streams` streams.  Returns bytes moved."""
    streams = (up_streams if direction == "up" else dn_streams)[:n_streams]
    c0, i, total = 0, 0, 0
    per = (rows + row_split - 1) // row_split
    for w in widths:
        for r0 in range(0, rows, per):
            r1 = min(rows, r0 + per)
            s = streams[i % n_streams]
            i += 1
            if direction == "up":
                copy2d(dx.data_ptr() + 4 * (r0 * ld + c0), 4 * ld, hx.data_ptr() + 4 * (r0 * f + c0), 4 * f, 4 * w, r1 - r0, s)
            else:
                copy2d(hy.data_ptr() + 4 * (r0 * f + c0), 4 * f, dy.data_ptr() + 4 * (r0 * ld + c0), 4 * ld, 4 * w, r1 - r0, s)
            total += 4 * w * (r1 - r0)
        c0 += w
    return total


def measure(label, widths, n_streams=1, row_split=1, reps=4):
    out = {}
    for mode in ("up", "down", "both"):
        best = None
        for _ in range(reps):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            cur = torch.cuda.current_stream()
            for s in up_streams + dn_streams:
                s.wait_stream(cur)
            a.record()
            for s in up_streams + dn_streams:
                s.wait_event(a)
            nb = {}
            if mode in ("up", "both"):
                nb["up"] = enqueue("up", widths, n_streams, row_split)
            if mode in ("down", "both"):
                nb["down"] = enqueue("down", widths, n_streams, row_split)
            for s in up_streams + dn_streams:
                cur.wait_stream(s)
            b.record()
            torch.cuda.synchronize()
            t = a.elapsed_time(b)
            if best is None or t < best[0]:
                best = (t, nb)
        out[mode] = best
    line = "%-58s" % label
    for mode in ("up", "down", "both"):
        t, nb = out[mode]
        line += " | %s %6.2f ms: " % (mode, t) + " ".join("%s %5.1f GB/s" % (k, v / t / 1e6) for k, v in nb.items())
    print(line, flush=True)


print("column slabs of a pinned row-major [%d, %d] fp32 matrix (row pitch %d B) <-> device rows of %d B" % (rows, f, 4 * f, 4 * ld))
measure("slabs 128+128+128+128+90, 1 stream per direction", [128, 128, 128, 128, 90])
measure("slabs 128 x4 + 90, 2 streams per direction, rows in 2", [128, 128, 128, 128, 90], 2, 2)
measure("slabs 128 x4 + 90, 4 streams per direction, rows in 4", [128, 128, 128, 128, 90], 4, 4)
measure("slabs 128 x4 + 90, 1 stream, rows in 8 (more, smaller copies)", [128, 128, 128, 128, 90], 1, 8)
measure("slabs 256+256+90, 1 stream", [256, 256, 90])
measure("slabs 256+256+90, 2 streams, rows in 2", [256, 256, 90], 2, 2)
measure("slabs 304+298, 1 stream", [304, 298])
measure("slabs 304+298, 2 streams, rows in 2", [304, 298], 2, 2)
measure("whole rows (602 of 602 columns; device side pitched), 1 stream", [602])
measure("whole rows, 2 streams, rows in 2", [602], 2, 2)
measure("slabs 64 x9 + 26, 1 stream", [64] * 9 + [26])
# host matrix with rows padded to 64-byte multiples (pitch 2432 B): are the unaligned piece edges the cost?
hx2 = torch.ones((rows, ld), dtype=torch.float32).pin_memory()
hy2 = torch.empty((rows, ld), dtype=torch.float32).pin_memory()
hx, hy, f_save = hx2, hy2, f
f = ld
measure("slabs 128 x4 + 96 of a 64-byte-aligned host matrix (pitch 2432)", [128, 128, 128, 128, 96])
