"""NVLink store-bandwidth probe for the fused exchange (run under torchrun, N >= 2):
  1. sgc_push_rows: pure remote stores of [rows, f] pieces to every peer (GB/s egress per GPU)
  2. the sharded hop on this rank's rows: no push / push to peers
    python -m torch.distributed.run --nproc-per-node 2 tools/nvlink_probe.py [workload] [scale]"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth, _ffi, dist as sd
from sgc_b200.propagate import hop

name = sys.argv[1] if len(sys.argv) > 1 else "products"
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
shape = synth.SHAPES[name]
n, f, k = shape["n"], shape["f"], shape["k"]
n_pad, parts = sd.partition_rows(n, world)
row0, row1 = parts[rank]
s = synth.make_shape(name, device=dev, with_features=False)
rp, col, val = sd.local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
del s
sg = sd.shard_normalize(n, rp, col, val)
x_local = synth.gaussian_features(n, f, seed=1, scaled=shape["scaled"], row_range=(row0, row1), device=dev)
prop = sd.ShardedPropagator(sg, f, mode="push")
prop.load(x_local)
torch.cuda.synchronize(); dist.barrier()

def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(reps):
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([float(np.median(ts))], device=dev)
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    return float(t.item()), float(tmax.item())

def report(tag, ms, ms_max, nbytes=None):
    extra = "" if nbytes is None else "  %.0f GB/s egress per GPU" % (nbytes / ms_max / 1e6)
    if rank == 0:
        print("%-44s rank0 %.3f ms  max %.3f ms%s" % (tag, ms, ms_max, extra), flush=True)

ld = prop.ld
# 1. pure push of this rank's rows to all peers / to self only
dsts_all = prop.symm[1].row_ptrs(row0, peers_only=True)
for npeers in sorted(set([1, len(dsts_all)])):
    arr = (_ffi.c_p * npeers)(*dsts_all[:npeers])
    fn = lambda: _ffi.call("sgc_push_rows", x_local.data_ptr(), x_local.stride(0), sg.n_local, f, arr, npeers, ld,
                           torch.cuda.current_stream().cuda_stream)
    ms, mx = timeit(fn)
    report("push_rows to %d peer(s)" % npeers, ms, mx, sg.n_local * f * 4 * npeers)
own = (_ffi.c_p * 1)(prop.symm[1].row_ptrs(row0, peers_only=False)[rank])
ms, mx = timeit(lambda: _ffi.call("sgc_push_rows", x_local.data_ptr(), x_local.stride(0), sg.n_local, f, own, 1, ld,
                                  torch.cuda.current_stream().cuda_stream))
report("push_rows to own buffer (local copy)", ms, mx, sg.n_local * f * 4)
# 1b. TMA bulk variant (contiguous padded rows of buf[0] -> peers' buf[1])
srcb = prop.buf[0][row0:row1]
nbytes = sg.n_local * ld * 4
for tile, ctas in ((416, 8), (512, 8), (1664, 8), (8192, 4), (32768, 2)):
    arr = (_ffi.c_p * len(dsts_all))(*dsts_all)
    fn = lambda: _ffi.call("sgc_push_bytes_bulk", srcb.data_ptr(), nbytes, arr, len(dsts_all), tile, ctas,
                           torch.cuda.current_stream().cuda_stream)
    ms, mx = timeit(fn)
    report("bulk push to %d peers tile=%d ctas/SM=%d" % (len(dsts_all), tile, ctas), ms, mx, nbytes * len(dsts_all))
torch.cuda.synchronize(); dist.barrier()
for p_i, ptr in enumerate(dsts_all[:1]):
    pass
# 2. hop variants
src, dstb = prop.buf[0], prop.symm[1]
out = dstb.tensor[row0:row1, :f]
ms, mx = timeit(lambda: hop(sg.graph, src[:, :f], out=out))
report("hop, no push (compute only, 1/%d of rows)" % world, ms, mx)
peers = dstb.row_ptrs(row0, peers_only=True)
for npeers in sorted(set([1, len(peers)])):
    ms, mx = timeit(lambda: hop(sg.graph, src[:, :f], out=out, peers=peers[:npeers]))
    report("hop + push to %d peer(s)" % npeers, ms, mx, sg.n_local * f * 4 * npeers)
# local stand-ins for the peers: store-issue cost without NVLink
fake = [torch.empty((sg.n_local, ld), device=dev) for _ in range(len(peers))]
ms, mx = timeit(lambda: hop(sg.graph, src[:, :f], out=out, peers=[t.data_ptr() for t in fake]))
report("hop + push to %d LOCAL stand-in buffers" % len(fake), ms, mx)
flag = torch.zeros(1, device=dev)
ms, mx = timeit(lambda: dist.all_reduce(flag))
report("barrier (1-element all-reduce)", ms, mx)
ms, mx = timeit(lambda: prop.propagate_resident(k))
report("propagate_resident K=%d (push)" % k, ms, mx)
dist.barrier()
dist.destroy_process_group()
