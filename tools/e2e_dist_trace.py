"""Per-rank, per-call wall times of back-to-back feature-sharded host-to-host steps under torchrun."""
import os, sys, time
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sgc_b200
from sgc_b200 import synth
from sgc_b200.dist import partition_features

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
s = synth.make_shape("reddit", device=dev)
n, f, k = s["n"], s["f"], s["k"]
c0, c1 = partition_features(f, world)[rank]
host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
xh = s["x"][:, c0:c1].contiguous().cpu().pin_memory()
out = torch.empty((n, c1 - c0), dtype=torch.float32).pin_memory()
del s
torch.cuda.empty_cache()
for label, kw in (("cols+shared", dict(group=dist.group.WORLD)), ("cols+local csr", dict()),
                  ("rows+shared", dict(group=dist.group.WORLD, col_slab=0))):
    for _ in range(2):
        sgc_b200.precompute_host(host, xh, k, standardize=True, out=out, **kw)
    dist.barrier()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        sgc_b200.precompute_host(host, xh, k, standardize=True, out=out, **kw)
        ts.append((time.perf_counter() - t0) * 1e3)
    dist.barrier()
    tr = []
    sgc_b200.precompute_host(host, xh, k, standardize=True, out=out, trace=tr, **kw)
    dist.barrier()
    print("rank %d cols %d %-14s calls: %s   traced: %s" % (rank, c1 - c0, label, " ".join("%.1f" % t for t in ts),
          " | ".join("%s %.1f/%.1f" % (l[:14], d, h) for l, d, h in tr if "up" in l or l in ("normalised", "returned"))), flush=True)
dist.destroy_process_group()
