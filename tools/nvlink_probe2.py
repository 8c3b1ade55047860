"""Round-2 probe of the fused exchange (run under torchrun, N >= 2): the peer variant of the hop on this rank's
rows of the products-shaped graph -- without pushes, with tile stores of 1..32 rows, with row stores only, with
the end-of-kernel wait on source reads only -- and the whole K-hop step.  Device-timed, max over ranks.
    python -m torch.distributed.run --nproc-per-node 8 tools/nvlink_probe2.py [workload]"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth, _ffi, dist as sd
from sgc_b200.propagate import hop, default_hints

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

def timeit(fn, reps=7):
    fn(); torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(reps):
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([float(np.median(ts))], device=dev)
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    return float(tmin.item()), float(tmax.item())

nbytes = sg.n_local * prop.ld * 4 * (world - 1) * prop.push_fraction
def report(tag, mn, mx, push=True):
    if rank == 0:
        extra = "  %.0f GB/s egress per GPU" % (nbytes / mx / 1e6) if push else ""
        print("%-58s min %.3f ms  max %.3f ms%s" % (tag, mn, mx, extra), flush=True)

src, dstb = prop.buf[0], prop.symm[1]
out = dstb.tensor[row0:row1, :f]
peers = dstb.row_ptrs(row0, peers_only=True)
h = default_hints()
kw = dict(out=out, peers=peers, peer_mask=prop.peer_mask, pad_owned=True, pitch_owned=True)
report("plain kernel, no peers (last hop)", *timeit(lambda: hop(sg.graph, src[:, :f], out=out, pad_owned=True)), push=False)
report("peer variant, pushes skipped (hints 8)", *timeit(lambda: hop(sg.graph, src[:, :f], hints=h | 8, **kw)), push=False)
for rows in (1, 2, 4, 8, 16, 32):
    os.environ["SGC_B200_PUSH_ROWS"] = str(rows)
    report("fused hop, tiles of %2d rows" % rows, *timeit(lambda: hop(sg.graph, src[:, :f], **kw)))
os.environ.pop("SGC_B200_PUSH_ROWS")
report("fused hop, row stores only (hints 32)", *timeit(lambda: hop(sg.graph, src[:, :f], hints=h | 32, **kw)))
report("fused hop, final wait on reads only (hints 64)", *timeit(lambda: hop(sg.graph, src[:, :f], hints=h | 64, **kw)))
report("fused hop, per-lane stores (hints 16)", *timeit(lambda: hop(sg.graph, src[:, :f], hints=h | 16, **kw)))
srcb = prop.buf[0][row0:row1]
arr = (_ffi.c_p * len(peers))(*peers)
for tile, ctas in ((1664, 8), (3328, 8), (8192, 4), (32768, 2)):
    fn = lambda: _ffi.call("sgc_push_bytes_bulk", srcb.data_ptr(), sg.n_local * prop.ld * 4, arr, len(peers), tile, ctas,
                           torch.cuda.current_stream().cuda_stream)
    report("exchange alone: bulk push tile=%d ctas/SM=%d" % (tile, ctas), *timeit(fn))
report("barrier (1-element all-reduce)", *timeit(prop._barrier), push=False)
for rows in (4, 8, 16):
    os.environ["SGC_B200_PUSH_ROWS"] = str(rows)
    report("propagate_resident K=%d, tiles of %d rows" % (k, rows), *timeit(lambda: prop.propagate_resident(k)), push=False)
dist.barrier()
prop.close()
dist.destroy_process_group()
