"""Rate of strided 2-D DMA copies (sgc_copy_2d) between a pinned row-major host matrix and padded
device rows, as a function of the host row pitch and the slab width.   python tools/copy2d_probe.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sgc_b200 import _ffi
dev = torch.device("cuda", 0)
n = 232965
st = torch.cuda.current_stream().cuda_stream
def rate(width_f, pitch_f, d2h, dev_ld=None, off_f=0):
    host = torch.empty(n * pitch_f + 64, dtype=torch.float32).pin_memory()
    ld = dev_ld or (width_f + 7) // 8 * 8
    d = torch.empty((n, ld), dtype=torch.float32, device=dev)
    def go():
        if d2h:
            _ffi.call("sgc_copy_2d", host.data_ptr() + 4 * off_f, 4 * pitch_f, d.data_ptr(), 4 * ld, 4 * width_f, n, st)
        else:
            _ffi.call("sgc_copy_2d", d.data_ptr(), 4 * ld, host.data_ptr() + 4 * off_f, 4 * pitch_f, 4 * width_f, n, st)
    go(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); go(); go(); b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 2
    return ms, n * width_f * 4 / ms / 1e6
for width in (128, 90, 32, 16):
    for pitch in (128, 130, 132, 136, 138, 140, 144, 150, 160, 192, 256, 301, 304, 602, 608):
        if pitch < width:
            continue
        h = rate(width, pitch, False)
        d = rate(width, pitch, True)
        print("width %3d floats  host pitch %3d floats (%4d B): H2D %6.2f ms %5.1f GB/s   D2H %6.2f ms %5.1f GB/s"
              % (width, pitch, pitch * 4, h[0], h[1], d[0], d[1]), flush=True)
