"""K5: row-sharded propagation on the GPUs of one box (one process per GPU, torch.distributed).

The reference is single-process; this is the north_star's capability (3): A is
row-partitioned, rank g owns rows [g*n_pad, (g+1)*n_pad) of S (CSR slice with global column ids),
of D^-1/2 and of X, and every hop exchanges X over NVLink.

Exchange and compute form one pipeline over FEATURE SLABS rather than two phases:

    X is kept slab-major on every rank: xfull[s] is [world*n_pad, w] (all rows, w feature columns,
    contiguous), so (a) the all-gather of slab s is one contiguous NCCL collective and (b) the hop
    over slab s is the ordinary single-GPU kernel on a narrow matrix -- identical summation order,
    hence results are bit-identical to the single-GPU path and independent of the GPU count.

    for hop k, slab s:   wait(all-gather[k][s]) -> SpMM(xfull[s] -> ysl[s]) -> start all-gather[k+1][s]

While slab s is being multiplied, the NCCL stream is already moving slab s+1 (and slab s of the
next hop as soon as it has been produced), so NVLink transfers overlap the gathers from HBM.
No data-path collective other than that all-gather exists; D^-1/2 is all-gathered once at set-up.
"""
import os
import time

import torch
import torch.distributed as dist

from .graph import SgcGraph, augment_csr, finish_normalization


def partition_rows(n, world):
    """Equal contiguous row blocks of n_pad = ceil(n / world) rows (the last one may be short):
    a global row id is also its index in the gathered operand."""
    n_pad = (int(n) + world - 1) // world
    return n_pad, [(min(r * n_pad, n), min((r + 1) * n_pad, n)) for r in range(world)]


def _all_gather(out, inp, group, async_op):
    """out [world*rows, w] <- every rank's inp [rows, w]."""
    if out.is_cuda:
        return dist.all_gather_into_tensor(out, inp, group=group, async_op=async_op)
    world = dist.get_world_size(group)
    return dist.all_gather(list(out.chunk(world, dim=0)), inp, group=group, async_op=async_op)


class ShardedGraph(object):
    """This rank's slice of S plus the partition description."""

    def __init__(self, graph, rank, world, n_global, n_pad, row0, row1):
        self.graph, self.rank, self.world = graph, rank, world
        self.n_global, self.n_pad, self.row0, self.row1 = int(n_global), int(n_pad), int(row0), int(row1)
        self.n_local = self.row1 - self.row0


def shard_normalize(n_global, row_ptr_local, col, val, group=None):
    """aug_normalized_adjacency (normalization.py:5-12) for a row-sharded adjacency: every rank
    passes the CSR of ITS rows (global column ids).  K1/K2 run locally; the D^-1/2 entries the other
    ranks own arrive through one all-gather."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    n_pad, parts = partition_rows(n_global, world)
    row0, row1 = parts[rank]
    n_local = row1 - row0
    assert row_ptr_local.numel() == n_local + 1, "row_ptr_local must describe this rank's rows"
    rp, cl, av, rowsum, dinv_rows = augment_csr(n_local, n_global, row0, row_ptr_local, col, val)
    dev = rp.device
    mine = torch.zeros(n_pad, dtype=torch.float64, device=dev)
    mine[:n_local] = dinv_rows
    dinv_all = torch.empty(world * n_pad, dtype=torch.float64, device=dev)
    _all_gather(dinv_all.view(world * n_pad, 1), mine.view(n_pad, 1), group, False)
    graph = finish_normalization(n_local, world * n_pad, rp, cl, av, rowsum, dinv_rows, dinv_all)
    return ShardedGraph(graph, rank, world, n_global, n_pad, row0, row1)


def default_slab_width(f):
    """Feature columns per pipeline stage (SGC_B200_DIST_SLAB): at least 3 stages when F allows."""
    env = int(os.environ.get("SGC_B200_DIST_SLAB", "0"))
    if env > 0:
        return env
    for w in (128, 64, 32):
        if (f + w - 1) // w >= 3:
            return w
    return 32


def _cuda_hop(graph, x, out):
    from .propagate import hop
    hop(graph, x, out=out)


class ShardedPropagator(object):
    """Buffers + pipeline for K-hop propagation on a ShardedGraph (see module docstring)."""

    def __init__(self, sg, f, slab_w=None, group=None, hop_fn=None):
        self.sg, self.f, self.group = sg, int(f), group
        self.w = int(slab_w) if slab_w else default_slab_width(self.f)
        self.n_slabs = (self.f + self.w - 1) // self.w
        self.hop_fn = hop_fn or _cuda_hop
        dev = sg.graph.device
        n_all = sg.world * sg.n_pad
        # slab-major operands; pad rows / columns stay zero
        self.xfull = [torch.zeros((n_all, self.w), dtype=torch.float32, device=dev) for _ in range(self.n_slabs)]
        self.ysl = [torch.zeros((sg.n_pad, self.w), dtype=torch.float32, device=dev) for _ in range(self.n_slabs)]

    def slab_cols(self, s):
        c0 = s * self.w
        return c0, min(self.w, self.f - c0)

    def propagate(self, x_local, degree):
        """x_local: this rank's rows of X, fp32 [n_local, F] -> this rank's rows of S^K X."""
        sg = self.sg
        assert x_local.shape == (sg.n_local, self.f)
        if degree <= 0:
            return x_local
        out = torch.empty((sg.n_local, self.f), dtype=torch.float32, device=x_local.device)
        works = []
        for s in range(self.n_slabs):                      # distribute the initial features
            c0, fw = self.slab_cols(s)
            self.ysl[s][:sg.n_local, :fw].copy_(x_local[:, c0:c0 + fw])
            works.append(_all_gather(self.xfull[s], self.ysl[s], self.group, True))
        for k in range(degree):
            last = (k == degree - 1)
            for s in range(self.n_slabs):
                c0, fw = self.slab_cols(s)
                works[s].wait()                             # slab s of every rank has landed
                dst = out[:, c0:c0 + fw] if last else self.ysl[s][:sg.n_local, :fw]
                self.hop_fn(sg.graph, self.xfull[s][:, :fw], dst)
                if not last:                                # exchange for the next hop starts now
                    works[s] = _all_gather(self.xfull[s], self.ysl[s], self.group, True)
        return out


def propagate_sharded(sg, x_local, degree, slab_w=None, group=None, hop_fn=None):
    return ShardedPropagator(sg, x_local.shape[1], slab_w, group, hop_fn).propagate(x_local, degree)


# --------------------------------------------------------------------------------------------------
def local_rows_of(graph_csr, row0, row1):
    """Slice [row0, row1) of a CSR given as (row_ptr, col, val) tensors."""
    row_ptr, col, val = graph_csr
    b, e = int(row_ptr[row0].item()), int(row_ptr[row1].item())
    return (row_ptr[row0:row1 + 1] - b).contiguous(), col[b:e].contiguous(), val[b:e].contiguous()


def bench(args, rank, world, dev, B):
    """bench.py --gpus N (N > 1): row-sharded K-hop propagation of the products-shaped graph
    (BASELINE configs[3]; strong scaling: the graph is fixed, rows are split over the ranks).
    B: bench.py's module globals (helpers)."""
    import json
    import numpy as np
    from . import _ffi, synth
    from .features import padded_rows  # noqa: F401
    name = args.workload if args.workload_given else "products"
    shape = synth.SHAPES[name]
    n, f, k = shape["n"], shape["f"], shape["k"]
    n_pad, parts = partition_rows(n, world)
    row0, row1 = parts[rank]
    if shape["symmetric"]:
        s = synth.make_shape(name, device=dev, with_features=False)     # every rank builds the graph, keeps its rows
        rp, col, val = local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
        del s
    else:
        s = synth.make_shape(name, device=dev, with_features=False, row_range=(row0, row1))
        rp, col, val = s["row_ptr"], s["col"], s["val"]
    torch.cuda.empty_cache()
    x_local = synth.gaussian_features(n, f, seed=1, scaled=shape["scaled"], row_range=(row0, row1), device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sg = shard_normalize(n, rp, col, val)
    torch.cuda.synchronize()
    normalise_ms = (time.perf_counter() - t0) * 1e3
    nnz_local = torch.tensor([sg.graph.nnz], dtype=torch.int64, device=dev)
    nnz_all = [torch.zeros_like(nnz_local) for _ in range(world)]
    dist.all_gather(nnz_all, nnz_local)
    nnz = int(sum(int(t.item()) for t in nnz_all))
    from .propagate import default_chunk
    sg.graph.plan(default_chunk())
    prop = ShardedPropagator(sg, f)

    ev = lambda: torch.cuda.Event(enable_timing=True)
    for _ in range(args.warmup):
        prop.propagate(x_local, k)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = B["ClockSampler"](dev.index)
    sampler.start()
    launches0 = _ffi.launch_count()
    torch.cuda.synchronize()
    start, stop = ev(), ev()
    start.record()
    for _ in range(args.steps):
        y = prop.propagate(x_local, k)
    stop.record()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([start.elapsed_time(stop)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)            # max over ranks, device-timed
    launches = _ffi.launch_count() - launches0
    clocks = sampler.stop()
    ms_per_step = float(ms.item()) / args.steps
    value = k * nnz * f / (ms_per_step * 1e-3)

    # end to end: this rank's shard from pinned host memory and back
    host = {"rp": rp.cpu().pin_memory(), "col": col.cpu().pin_memory(), "val": val.cpu().pin_memory(),
            "x": x_local.cpu().pin_memory()}
    out_host = torch.empty((sg.n_local, f), dtype=torch.float32).pin_memory()

    def e2e_step():
        sgl = shard_normalize(n, host["rp"].to(dev, non_blocking=True), host["col"].to(dev, non_blocking=True),
                              host["val"].to(dev, non_blocking=True))
        xl = host["x"].to(dev, non_blocking=True)
        yl = ShardedPropagator(sgl, f).propagate(xl, k)
        out_host.copy_(yl, non_blocking=True)
        torch.cuda.synchronize()

    e2e_step()
    dist.barrier()
    e2e_steps = max(2, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    dist.barrier()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=dev)
    dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    h2d = torch.tensor([sum(t.numel() * t.element_size() for t in host.values())], dtype=torch.int64, device=dev)
    dist.all_reduce(h2d)
    if rank == 0:
        peak, peak_src = B["peaks"]()
        b_hop = B["hop_bytes"](n, nnz, f)
        achieved = k * b_hop / world / (ms_per_step * 1e-3) / 1e9      # per-GPU algorithmic GB/s
        line = {
            "metric": B["METRIC"], "value": value, "unit": B["UNIT"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": B["workload_config"](name, n, nnz, f, k, parallelism="row-sharded x%d, per-hop NCCL all-gather "
                                           "pipelined over %d feature slabs of %d" % (world, prop.n_slabs, prop.w),
                                           normalise_ms=normalise_ms, rows_per_rank=n_pad),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "spmm_hop_kernel (per GPU; includes time waiting for the exchange)"},
            "cpu_baseline": None,
            "e2e": {"value": k * nnz * f / float(e2e_s.item()), "unit": B["UNIT"],
                    "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(n * f * 4),
                    "ms_per_step": float(e2e_s.item()) * 1e3},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
