"""Row-sharded propagation on real GPUs: world 1 always (exercises shard_normalize + the slab
pipeline through NCCL on one device), world 2 when two GPUs are visible.  The sharded result must
be bit-identical to the single-GPU path on the same graph."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def worker(rank, world, port, scale, f, k, slab_w, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import sgc_b200
        from sgc_b200 import dist as sd, synth
        s = synth.make_shape("reddit", device=dev, scale=scale, seed=11)
        n = s["n"]
        x = s["x"][:, :f].contiguous()
        n_pad, parts = sd.partition_rows(n, world)
        row0, row1 = parts[rank]
        rp, col, val = sd.local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
        sg = sd.shard_normalize(n, rp, col, val)
        y = sd.propagate_sharded(sg, x[row0:row1].contiguous(), k, row_blocks=slab_w)
        # single-GPU result on the same device
        full = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
        want, _ = sgc_b200.sgc_precompute(x, full, k)
        same_struct = torch.equal(sg.graph.col, full.col[int(full.row_ptr[row0]):int(full.row_ptr[row1])])
        same_d = torch.equal(sg.graph.dinv[:n], full.dinv)
        ret[rank] = (bool(torch.equal(y, want[row0:row1])), bool(same_struct), bool(same_d), int(y.shape[0]))
    finally:
        dist.destroy_process_group()


def run(world, **kw):
    ret = mp.Manager().dict()
    args = (world, free_port(), kw.get("scale", 0.02), kw.get("f", 100), kw.get("k", 3), kw.get("slab_w", 4), ret)
    mp.spawn(worker, args=args, nprocs=world, join=True)
    for rank in range(world):
        ok, same_struct, same_d, rows = ret[rank]
        assert same_struct and same_d, rank
        assert ok, "rank %d: sharded result differs from the single-GPU result" % rank


def test_sharded_world1_bit_identical():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    run(1, f=100, k=3, slab_w=4)
    run(1, f=602, k=2, slab_w=1)


def test_sharded_world2_bit_identical():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run(2, f=100, k=3, slab_w=4)
    run(2, f=130, k=2, slab_w=3)
