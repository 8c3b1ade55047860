"""bench.py --gpus N (N > 1): the multi-GPU legs.  One process per GPU (torchrun), NCCL for rendezvous.

Headline: the SAME workload as N = 1 (Reddit-shaped, BASELINE configs[2]) sharded over the N GPUs, so that
value(N) / value(1) is a strong-scaling ratio (Reddit takes feature sharding: dist.choose_sharding).
First-class extras in the same line (`other_workloads`):
  * products-shaped (BASELINE configs[3]) row-sharded with the exchange fused into the hop kernel (the series
    north_star's >= 0.7 target names), partition locality lambda = 0 and 0.8, with per-hop breakdown;
  * papers100M-shaped (BASELINE configs[4]) at N = 8 only, lambda = 0 and 0.8.
EVERY leg checks its output after the timed steps (dist.verify_sharded / verify_against_single): against the
single-GPU result of the same workload bit for bit where one GPU can hold it, against the reference's
torch.spmm on sampled rows otherwise, plus owner-vs-copy comparison of the pushed rows.  A mismatch makes the
run exit non-zero after printing the line with "parity": false.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

from sgc_b200 import _ffi
from sgc_b200 import dist as sd


def _device_ms(fn, reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def _single_gpu_reference(s, x, n, f, k):
    """K hops of the whole graph on one GPU (rank 0), device-timed; returns the timing and the result (kept
    for the bit-for-bit comparison with the sharded outputs)."""
    from sgc_b200.graph import normalize_csr
    from sgc_b200.propagate import propagate
    from sgc_b200.features import to_padded
    g = normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    xp = to_padded(x)
    for _ in range(3):
        propagate(g, xp, k)
    torch.cuda.synchronize()
    ms = _device_ms(lambda: propagate(g, xp, k), 5)
    y = propagate(g, xp, k)
    torch.cuda.synchronize()
    out = {"n_gpus": 1, "ms_per_step": ms, "value": k * g.nnz * f / (ms * 1e-3)}
    del g, xp
    torch.cuda.empty_cache()
    return out, y


def _timed_steps(args, dev, B, step, before_step=None):
    """Contract timing: W warm-ups, then K steps bracketed by barrier + synchronize, CUDA events, max over
    ranks.  `before_step` (untimed) restores the step's starting state."""
    for _ in range(args.warmup):
        if before_step:
            before_step()
        step()
    torch.cuda.synchronize()
    dist.barrier()
    sampler = B["ClockSampler"](dev.index)
    sampler.start()
    launches0 = _ffi.launch_count()
    total = 0.0
    if before_step is None:
        torch.cuda.synchronize()
        dist.barrier()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        for _ in range(args.steps):
            step()
        stop.record()
        torch.cuda.synchronize()
        total = start.elapsed_time(stop)
    else:
        for _ in range(args.steps):
            before_step()
            torch.cuda.synchronize()
            dist.barrier()
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
            step()
            stop.record()
            torch.cuda.synchronize()
            total += start.elapsed_time(stop)
    dist.barrier()
    ms = torch.tensor([total], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)            # max over ranks, device-timed
    launches = _ffi.launch_count() - launches0
    return float(ms.item()) / args.steps, int(launches), sampler.stop()


# --------------------------------------------------------------------------------------------------
def run_feature_sharded(name, args, rank, world, dev, B, lam=0.0, with_e2e=True, e2e_only=False):
    """S replicated, X split by columns; no exchange.  Checked against rank 0's single-GPU result.
    e2e_only: just the host-to-host figure of this sharding (the device-timed leg ran on another grid)."""
    synth = B["load_synth"]()
    from sgc_b200.features import col_stats, standardize_, to_padded
    from sgc_b200.graph import normalize_csr
    from sgc_b200.propagate import default_chunk, propagate
    shape = synth.SHAPES[name]
    n, f, k = shape["n"], shape["f"], shape["k"]
    c0, c1 = sd.partition_features(f, world)[rank]
    s = synth.make_shape(name, device=dev, with_features=False, lam=lam)
    x = synth.gaussian_features(n, f, seed=1, scaled=shape["scaled"], device=dev)
    x_raw_local = x[:, c0:c1].contiguous()
    if e2e_only:
        del x
        torch.cuda.empty_cache()
        nnz = int(s["col"].numel()) + n          # A has no self-loops: S = A + I stores nnz(A) + N entries
        e2e = None
        if c1 > c0:
            e2e = _feature_sharded_e2e(args, s, x_raw_local, shape, n, f, k, nnz, c0, c1, world, dev, B)
        else:
            dist.barrier(); dist.barrier()
            z = torch.zeros(1, dtype=torch.float64, device=dev); dist.all_reduce(z, op=dist.ReduceOp.MAX)
            z = torch.zeros(1, dtype=torch.int64, device=dev); dist.all_reduce(z)
        return {"e2e": e2e}
    if shape["scaled"]:
        mean, std = col_stats(x)
        standardize_(x, mean, std)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    graph = normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    torch.cuda.synchronize()
    normalise_ms = (time.perf_counter() - t0) * 1e3
    single = y_single = None
    if rank == 0:
        single, y_single = _single_gpu_reference(s, x, n, f, k)
    x_local = to_padded(x[:, c0:c1]) if c1 > c0 else None
    del x
    torch.cuda.empty_cache()
    nnz = graph.nnz
    graph.plan(default_chunk())
    last = {}

    def step():
        last["y"] = propagate(graph, x_local, k) if x_local is not None else None

    ms_per_step, launches, clocks = _timed_steps(args, dev, B, step)
    y_local = last["y"] if last.get("y") is not None else torch.empty((n, 0), dtype=torch.float32, device=dev)
    parity = sd.verify_against_single(y_local, y_single, 0, c0, f)
    del y_single, y_local
    last.clear()
    out = {"name": name, "n": n, "f": f, "k": k, "nnz": nnz, "n_pad": n, "mode": "features", "lam": lam,
           "push_fraction": 0.0, "n_blocks": 1, "normalise_ms": normalise_ms, "ms_per_step": ms_per_step,
           "value": k * nnz * f / (ms_per_step * 1e-3), "launches": launches, "clocks": clocks,
           "single": single, "e2e": None, "parity": parity, "breakdown": None,
           "cols_per_rank": max(b - a for a, b in sd.partition_features(f, world))}
    if with_e2e and x_local is not None:
        out["e2e"] = _feature_sharded_e2e(args, s, x_raw_local, shape, n, f, k, nnz, c0, c1, world, dev, B)
    elif with_e2e:
        dist.barrier(); dist.barrier()
        z = torch.zeros(1, dtype=torch.float64, device=dev); dist.all_reduce(z, op=dist.ReduceOp.MAX)
        z = torch.zeros(1, dtype=torch.int64, device=dev); dist.all_reduce(z)
    del graph, x_local, s
    torch.cuda.empty_cache()
    return out


def _feature_sharded_e2e(args, s, x_raw_local, shape, n, f, k, nnz, c0, c1, world, dev, B):
    from sgc_b200.pipeline import precompute_host
    host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val")}
    host["x"] = x_raw_local.cpu().pin_memory()
    out_host = torch.empty((n, c1 - c0), dtype=torch.float32).pin_memory()
    host_csr = (n, host["row_ptr"], host["col"], host["val"])
    # one node, one copy of the adjacency in host memory: every rank uploads 1/P of it and NVLink
    # replicates it (SGC_B200_SHARED_CSR=0: every rank uploads all of it)
    shared_csr = os.environ.get("SGC_B200_SHARED_CSR", "1") != "0" and \
        all(b > a for a, b in sd.partition_features(f, world))       # the all-gather needs every rank
    st = {}

    def e2e_step():
        # this rank's column block through the pipelined host-to-host path (pipeline.py): CSR up,
        # K1/K2 under the upload of X, per slab K3 + hops + strided D2H
        precompute_host(host_csr, host["x"], k, standardize=bool(shape["scaled"]), out=out_host,
                        group=dist.group.WORLD if shared_csr else None, stats=st)

    e2e_step()
    dist.barrier()
    e2e_steps = max(2, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    dist.barrier()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=dev)
    dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    # bytes this rank actually copied (col / val: 1/P per rank when shared, the rest over NVLink; a
    # constant value array is filled on the GPU, not uploaded), summed over ranks
    h2d = torch.tensor([st["h2d_bytes"]], dtype=torch.int64, device=dev)
    dist.all_reduce(h2d)
    return {"value": k * nnz * f / float(e2e_s.item()), "unit": B["UNIT"],
            "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(n * f * 4),
            "ms_per_step": float(e2e_s.item()) * 1e3,
            "path": "per rank: sgc_b200.precompute_host on its column block (pipeline.py); adjacency "
                    + ("uploaded once per node (1/P per rank over its own PCIe link, NCCL all-gather "
                       "over NVLink)" if shared_csr else "uploaded by every rank")
                    + ("" if st.get("uniform_value") is None else
                       "; its stored values are all %g (exact host scan every step, 1/P per rank + "
                       "one-word all-reduce): filled on the GPU instead of uploaded" % st["uniform_value"])}


# --------------------------------------------------------------------------------------------------
def _workload_inputs(name, rank, world, dev, lam, want_single, B, grid=None):
    """This rank's rows of the named synthetic workload: raw CSR slice (global column ids) and the
    feature rows the hops start from (Reddit shape: standardised as utils.py:119 does, with the
    statistics of ALL rows); on a grid (dist.Grid) the rows of this rank's row part and the columns of its
    column block.  Rank 0 also runs the whole workload on its one GPU (timing + result)."""
    synth = B["load_synth"]()
    from sgc_b200.features import col_stats, standardize_
    shape = synth.SHAPES[name]
    n, f, k = shape["n"], shape["f"], shape["k"]
    part_rank, part_world = (grid.ri, grid.p_rows) if grid is not None else (rank, world)
    n_pad, parts = sd.partition_rows(n, part_world)
    row0, row1 = parts[part_rank]
    single = y_single = None
    if shape["symmetric"]:
        s = synth.make_shape(name, device=dev, with_features=False, lam=lam)     # every rank builds the graph, keeps its rows
        rp, col, val = sd.local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
        if shape["scaled"]:
            x = synth.gaussian_features(n, f, seed=1, scaled=True, device=dev)
            mean, std = col_stats(x)
            standardize_(x, mean, std)
            x_local = x[row0:row1].clone()
        else:
            x = None
            x_local = synth.gaussian_features(n, f, seed=1, scaled=False, row_range=(row0, row1), device=dev)
        if rank == 0 and want_single:
            if x is None:
                x = synth.gaussian_features(n, f, seed=1, scaled=False, device=dev)
            single, y_single = _single_gpu_reference(s, x, n, f, k)
        del s, x
    else:
        s = synth.make_shape(name, device=dev, with_features=False, row_range=(row0, row1), lam=lam)
        rp, col, val = s["row_ptr"], s["col"], s["val"]
        x_local = synth.gaussian_features(n, f, seed=1, scaled=False, row_range=(row0, row1), device=dev)
        del s
    if grid is not None and grid.p_cols > 1:
        x_local = x_local[:, grid.cols[0]:grid.cols[1]].contiguous()
    torch.cuda.empty_cache()
    return n, f, k, n_pad, rp, col, val, x_local, single, y_single


def _push_breakdown(prop, k_hops, dev, reps=5):
    """Where an exchanging hop's time goes (device-timed, max over ranks): the hop without pushes (peer
    variant, hints bit 3), the fused hop, the last hop (no exchange), the barrier."""
    from sgc_b200.propagate import default_hints, hop
    sg, f = prop.sg, prop.f
    if prop.mode != "push" or sg.world == 1:
        return None
    src, dst = prop.buf[0], prop.symm[1]
    out = dst.tensor[sg.row0:sg.row1, :f]
    peers = dst.row_ptrs(sg.row0, peers_only=True)

    def timed(fn):
        fn()
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        t = torch.tensor([sorted(ts)[len(ts) // 2]], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    h = default_hints()
    res = {
        "hop_without_pushes_ms": timed(lambda: hop(sg.graph, src[:, :f], out=out, peers=peers, peer_mask=prop.peer_mask,
                                                   hints=h | 8, pad_owned=True, pitch_owned=True)),
        "fused_hop_ms": timed(lambda: hop(sg.graph, src[:, :f], out=out, peers=peers, peer_mask=prop.peer_mask,
                                          pad_owned=True, pitch_owned=True)),
        "fused_hop_row_stores_ms": timed(lambda: hop(sg.graph, src[:, :f], out=out, peers=peers, peer_mask=prop.peer_mask,
                                                     hints=h | 32, pad_owned=True, pitch_owned=True)),
        "last_hop_ms": timed(lambda: hop(sg.graph, src[:, :f], out=out, pad_owned=True)),
        "barrier_ms": timed(prop._barrier),
    }
    bytes_out = sg.n_local * prop.ld * 4 * (sg.world - 1) * prop.push_fraction
    res["egress_gbs_per_gpu_in_fused_hop"] = bytes_out / (res["fused_hop_ms"] * 1e-3) / 1e9
    res["exchange_bytes_per_gpu_per_hop"] = int(bytes_out)
    res["model"] = "step ~ (K-1) x (fused_hop + barrier) + last_hop + barrier"
    return res


def run_row_sharded(name, args, rank, world, dev, B, lam=0.0, with_e2e=True, want_single=True, steps=None, grid=None):
    """Times K-hop row-sharded propagation of one named workload on `world` GPUs (strong scaling: the
    graph is fixed, rows are split over the ranks), then checks the result (dist.verify_sharded).  With a
    dist.Grid of more than one column block, every row group (the ranks sharing a column block) runs the
    row-sharded propagation on its block: rows are split over grid.p_rows ranks only."""
    from sgc_b200.propagate import default_chunk
    n, f, k, n_pad, rp, col, val, x_local, single, y_single = _workload_inputs(name, rank, world, dev, lam, want_single, B, grid)
    group = grid.row_group if grid is not None else None
    c0, c1 = grid.cols if grid is not None else (0, f)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sg = sd.shard_normalize(n, rp, col, val, group=group)
    torch.cuda.synchronize()
    normalise_ms = (time.perf_counter() - t0) * 1e3
    nnz_local = torch.tensor([sg.graph.nnz], dtype=torch.int64, device=dev)
    dist.all_reduce(nnz_local, group=group)
    nnz = int(nnz_local.item())
    prop = sd.ShardedPropagator(sg, c1 - c0, group=group)
    mode = prop.mode
    for blk in prop.blocks:
        if blk is not None:
            blk.plan(default_chunk())
    if steps is not None:
        class _A(object):
            pass
        a2 = _A()
        a2.steps, a2.warmup = steps, min(args.warmup, 3)
    else:
        a2 = args
    # load (untimed): X resident on every rank, the reference's starting state of sgc_precompute
    ms_per_step, launches, clocks = _timed_steps(a2, dev, B, lambda: prop.propagate_resident(k),
                                                 before_step=lambda: prop.load(x_local))
    out = {"name": name, "n": n, "f": f, "k": k, "nnz": nnz, "n_pad": n_pad, "mode": mode, "lam": lam,
           "push_fraction": getattr(prop, "push_fraction", 1.0),
           "n_blocks": prop.n_blocks, "normalise_ms": normalise_ms, "ms_per_step": ms_per_step,
           "value": k * nnz * f / (ms_per_step * 1e-3), "launches": launches, "clocks": clocks,
           "single": single, "e2e": None, "steps": a2.steps,
           "grid": [grid.p_cols, grid.p_rows] if grid is not None else [1, world], "cols_per_rank": c1 - c0}
    if mode != "push":
        out["parity"] = None
    elif grid is None or grid.p_cols == 1:
        out["parity"] = sd.verify_sharded(prop, x_local, k, B, y_single=y_single)
    else:
        # inside the row group: pushed rows owner-vs-copy and sampled rows against the reference's torch.spmm;
        # over all ranks: every rank's [rows, columns] block against the single-GPU result, bit for bit
        rep = sd.verify_sharded(prop, x_local, k, B, y_single=None, group=group)
        y_local = prop.buf[k % 2][sg.row0:sg.row1, :c1 - c0]
        rep["vs_single_gpu"] = sd.verify_against_single(y_local, y_single, sg.row0, c0, f)
        ok = torch.tensor([1 if (rep["ok"] and rep["vs_single_gpu"]["ok"]) else 0], dtype=torch.int64, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        rep["ok"] = bool(int(ok.item()))
        rep["scope"] = "hop-by-hop figures are those of rank 0's row group; ok and vs_single_gpu cover all ranks"
        out["parity"] = rep
    del y_single
    out["breakdown"] = _push_breakdown(prop, k, dev)
    if with_e2e:
        out["e2e"] = _row_sharded_e2e(args, prop, sg, rp, col, val, x_local, n, f, k, nnz, mode, dev, B)
    prop.close()
    del prop, sg
    torch.cuda.empty_cache()
    return out


def _row_sharded_e2e(args, prop, sg, rp, col, val, x_local, n, f, k, nnz, mode, dev, B):
    # end to end: this rank's shard from pinned host memory and back
    host = {"rp": rp.cpu().pin_memory(), "col": col.cpu().pin_memory(), "val": val.cpu().pin_memory(),
            "x": x_local.cpu().pin_memory()}
    out_host = torch.empty((sg.n_local, f), dtype=torch.float32).pin_memory()

    def e2e_step():
        sgl = sd.shard_normalize(n, host["rp"].to(dev, non_blocking=True), host["col"].to(dev, non_blocking=True),
                                 host["val"].to(dev, non_blocking=True))
        xl = host["x"].to(dev, non_blocking=True)
        if mode == "push":
            yl = prop.rebind(sgl).propagate(xl, k)      # symmetric buffers + peer mappings are reused
        else:
            yl = sd.ShardedPropagator(sgl, f).propagate(xl, k)
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
    return {"value": k * nnz * f / float(e2e_s.item()), "unit": B["UNIT"],
            "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(n * f * 4),
            "ms_per_step": float(e2e_s.item()) * 1e3}


def run_workload(name, args, rank, world, dev, B, lam=0.0, with_e2e=True, want_single=True, steps=None):
    """One leg: the device-timed K hops on the grid dist.choose_grid picks (column blocks x row parts), its
    result check, and -- when asked -- the host-to-host figure: through the pipelined feature-sharded host path
    wherever the shape allows feature sharding (dist.choose_sharding; PCIe-bound, the hops are a tenth of it),
    else through the row-sharded one."""
    synth = B["load_synth"]()
    shape = synth.SHAPES[name]
    approx_nnz = shape["pairs"] * (2 if shape["symmetric"] else 1) + shape["n"]
    p_cols, p_rows = sd.choose_grid(shape["n"], approx_nnz, shape["f"], shape["k"], world)
    feature_e2e = sd.choose_sharding(shape["n"], approx_nnz, shape["f"], world) == "features"
    if p_rows == 1:
        return run_feature_sharded(name, args, rank, world, dev, B, lam, with_e2e)
    grid = sd.make_grid(shape["f"], p_cols, p_rows) if p_cols > 1 else None
    r = run_row_sharded(name, args, rank, world, dev, B, lam, with_e2e and not feature_e2e, want_single, steps, grid)
    r["grid_model_ms"] = {"%dx%d" % (pc, world // pc): round(sd.grid_cost_ms(shape["n"], approx_nnz, shape["f"], shape["k"],
                                                                              pc, world // pc), 3)
                          for pc in range(1, world + 1) if world % pc == 0 and shape["f"] // pc >= 64}
    if with_e2e and feature_e2e:
        r["e2e"] = run_feature_sharded(name, args, rank, world, dev, B, lam, True, e2e_only=True)["e2e"]
        if r["e2e"] is not None:
            r["e2e"]["value"] = shape["k"] * r["nnz"] * shape["f"] / (r["e2e"]["ms_per_step"] * 1e-3)
            r["e2e"]["sharding"] = "feature-sharded x%d host pipeline (the device-timed value above uses the %dx%d grid)" \
                                   % (world, p_cols, p_rows)
    return r


def parallelism_text(r, world):
    if r["mode"] == "features":
        return ("feature-sharded x%d: S replicated on every GPU, X split into contiguous column blocks of <= %d "
                "columns; column c of S^K X depends on column c of X only, so the K hops (and the per-column "
                "standardisation) need no exchange; output column-sharded" % (world, r.get("cols_per_rank", 0)))
    if r["mode"] == "push" and r.get("grid", [1, world])[0] > 1:
        pc, pr = r["grid"]
        return ("grid-sharded %d column blocks x %d row parts: the %d GPUs that share a column block (<= %d columns) "
                "split the rows of S among them and run the row-sharded hop on that block -- exchange fused into the "
                "SpMM kernel (output rows staged in shared memory, cp.async.bulk to the %d peers' next-hop operands "
                "over NVLink, CUDA-IPC mapped), one 1-element all-reduce inside the row group as barrier per hop; "
                "%.1f%% of rows x peers pushed; column blocks never communicate (column c of S^K X depends on column "
                "c of X only); output sharded by rows and columns"
                % (pc, pr, pr, r.get("cols_per_rank", 0), pr - 1, 100.0 * r["push_fraction"]))
    if r["mode"] == "push":
        return ("row-sharded x%d; exchange fused into the SpMM kernel: output rows are staged in shared memory and "
                "written by cp.async.bulk (tiles of consecutive rows) to all peers' next-hop operands over NVLink "
                "(CUDA-IPC mapped), one 1-element all-reduce as barrier per hop; a row goes only to the peers whose "
                "shard gathers it (%.1f%% of rows x peers here); X replicated at start, output sharded"
                % (world, 100.0 * r["push_fraction"]))
    return ("row-sharded x%d; per-hop NVLink exchange (NCCL send/recv groups) pipelined with the SpMM over %d row "
            "blocks; X replicated at start, output sharded" % (world, r["n_blocks"]))


def extra_record(x, world, B):
    """`other_workloads` entry of one measured leg."""
    n, f, k, nnz = x["n"], x["f"], x["k"], x["nnz"]
    hbm, hbm_src = B["peaks"]()
    per_gpu_hop_ms = x["ms_per_step"] / k
    rec = {"workload": B["workload_config"](x["name"], n, nnz, f, k)["workload"], "n_gpus": world,
           "value": x["value"], "unit": B["UNIT"], "ms_per_step": x["ms_per_step"], "steps": x.get("steps"),
           "nnz_S": nnz, "dist_mode": x["mode"], "partition_locality_lambda": x["lam"],
           "pushed_rows_x_peers_fraction": x["push_fraction"], "parallelism": parallelism_text(x, world),
           "single_gpu_same_workload": x["single"],
           "strong_scaling_efficiency": (x["value"] / (world * x["single"]["value"]) if x["single"] else None),
           "parity": x["parity"], "per_hop_breakdown": x["breakdown"],
           "roofline": {"bound": "hbm (gather) / nvlink (exchange)", "unit": "GB/s", "peak": hbm, "peak_source": hbm_src,
                        "gather_model_gbs_per_gpu": B["hop_bytes"](n, nnz, f) / world / (per_gpu_hop_ms * 1e-3) / 1e9,
                        "compulsory_bytes_per_gpu_per_hop": B["compulsory_bytes"](n, nnz, f) / world,
                        "traffic": None,
                        "note": "multi-GPU kernels are never run under ncu (profiling recipe): no dram counters; the "
                                "single-GPU capture of the same kernel is other_workloads.products of the N=1 line"}}
    rec["roofline"]["frac_of_hbm_gather_model"] = rec["roofline"]["gather_model_gbs_per_gpu"] / hbm
    return rec


def weak_scaling_reference(name, world, lam, dev, B):
    """SURVEY 8(e)'s yardstick for a workload no single GPU can hold (papers100M shape): the SAME generator at
    1/world of the size -- nodes and pairs scaled together, same degree law, same feature width and K -- run on ONE
    GPU of this box with the single-GPU hop kernel, device-timed like every other single-GPU leg.  The caller
    reports  weak_scaling_efficiency = rate(world GPUs, full size) / (world * rate(1 GPU, 1/world size)),
    rate = K * nnz(S) * F / step time.  Rank 0 only; no collective inside."""
    synth = B["load_synth"]()
    shape = synth.SHAPES[name]
    t0 = time.perf_counter()
    s = synth.make_shape(name, device=dev, scale=1.0 / world, lam=lam, with_features=False)
    n1, f, k = s["n"], shape["f"], shape["k"]
    x = synth.gaussian_features(n1, f, seed=1, scaled=False, device=dev)
    single, y = _single_gpu_reference(s, x, n1, f, k)
    del y, s, x
    torch.cuda.empty_cache()
    single.update({"n": n1, "f": f, "k": k, "nnz_S": int(round(single["value"] * single["ms_per_step"] * 1e-3 / (k * f))),
                   "generator_scale": 1.0 / world, "partition_locality_lambda": lam,
                   "wall_s": time.perf_counter() - t0,
                   "what": "the same generator at 1/%d size (nodes and pairs scaled together) on ONE GPU, K hops on "
                           "device-resident inputs, CUDA events, median-free mean of 5 steps after 3 warm-ups" % world})
    return single


def _r3(v):
    return None if v is None else float("%.3g" % v)


def summary_of(name, line, extras):
    """Compact digest of every leg of this run -- {leg: {ms, eff, par}} -- kept under 600 characters and placed at
    the very end of the JSON line, so that a reader who only has the last 1500 characters of the output still
    sees every leg's step time, efficiency (strong scaling against one GPU of this box, null where no single-GPU
    run of the leg exists) and whether its result check passed."""
    out = {name: {"ms": _r3(line["ms_per_step"]), "eff": _r3(line["strong_scaling_efficiency"]),
                  "par": line["parity"].get(name)}}
    for key, v in extras.items():
        if not isinstance(v, dict):
            continue
        if "ms_per_step" in v:
            eff = v.get("strong_scaling_efficiency")
            if eff is None:
                eff = v.get("weak_scaling_efficiency")
            out[key] = {"ms": _r3(v["ms_per_step"]), "eff": _r3(eff), "par": (v.get("parity") or {}).get("ok")}
        else:
            out[key] = {"skip": str(v.get("skipped") or v.get("error") or "?")[:40]}
    return out


def bench(args, rank, world, dev, B):
    name = args.workload
    failed = []
    r = run_workload(name, args, rank, world, dev, B, lam=0.0,
                     with_e2e=(not getattr(args, "skip_e2e", False))
                     and (name != "papers100m" or os.environ.get("SGC_BENCH_PAPERS_E2E") == "1"),
                     want_single=(name != "papers100m"))
    if r["parity"] is not None and not r["parity"]["ok"]:
        failed.append(name)
    extras = {}
    extra_names = os.environ.get("SGC_BENCH_EXTRA", None)
    if extra_names is None:
        extra_names = "" if args.workload_given else ("products,papers100m" if world >= 8 else "products")
    lams = [float(v) for v in os.environ.get("SGC_BENCH_LAMBDAS", "0,0.8").split(",") if v != ""]
    papers_budget = float(os.environ.get("SGC_BENCH_PAPERS_BUDGET_S", "300"))
    for other in [e for e in extra_names.split(",") if e and e != name]:
        for lam in lams:
            key = other if lam == 0 else "%s_lambda%g" % (other, lam)
            t0 = time.perf_counter()
            try:
                if other == "papers100m":
                    spent = torch.tensor([extras.get("_papers_s", 0.0)], dtype=torch.float64, device=dev)
                    dist.all_reduce(spent, op=dist.ReduceOp.MAX)
                    if float(spent.item()) > papers_budget:
                        extras[key] = {"skipped": "time guard: the papers100M legs already took %.0f s (budget %.0f s)"
                                                  % (float(spent.item()), papers_budget)}
                        continue
                x = run_workload(other, args, rank, world, dev, B, lam=lam, with_e2e=False,
                                 want_single=(other != "papers100m"), steps=(3 if other == "papers100m" else None))
                extras[key] = extra_record(x, world, B)
                if x["parity"] is not None and not x["parity"]["ok"]:
                    failed.append(key)
                if x["single"] is None and rank == 0 and os.environ.get("SGC_BENCH_WEAK_REF", "1") == "1":
                    # no one-GPU run of this size exists: weak scaling against the 1/world-size generator (8e)
                    del x
                    torch.cuda.empty_cache()
                    try:
                        ref = weak_scaling_reference(other, world, lam, dev, B)
                        extras[key]["weak_scaling_reference"] = ref
                        extras[key]["weak_scaling_efficiency"] = extras[key]["value"] / (world * ref["value"])
                        extras[key]["weak_scaling_efficiency_is"] = (
                            "rate(%d GPUs, full size) / (%d x rate(1 GPU, 1/%d-size graph of the same generator)), "
                            "rate = K*nnz(S)*F / step time (SURVEY 8e)" % (world, world, world))
                    except Exception as exc:
                        extras[key]["weak_scaling_reference"] = {"error": repr(exc)[:300]}
            except Exception as exc:            # an optional extra must never cost the primary line
                extras[key] = {"error": repr(exc)[:400]}
                torch.cuda.empty_cache()
            if other == "papers100m":
                extras["_papers_s"] = extras.get("_papers_s", 0.0) + (time.perf_counter() - t0)
                if isinstance(extras.get(key), dict):
                    extras[key]["leg_wall_s"] = time.perf_counter() - t0
    extras.pop("_papers_s", None)
    if rank == 0:
        n, f, k, nnz = r["n"], r["f"], r["k"], r["nnz"]
        peak, peak_src = B["peaks"]()
        b_hop = B["hop_bytes"](n, nnz, f)
        achieved = k * b_hop / world / (r["ms_per_step"] * 1e-3) / 1e9      # per-GPU algorithmic GB/s
        line = {
            "metric": B["METRIC"], "value": r["value"], "unit": B["UNIT"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": B["workload_config"](name, n, nnz, f, k),
            # one yardstick, stated: SURVEY 8(d)'s gather-model bytes against the measured HBM copy figure.  A
            # fraction above 1 is not a faster-than-HBM kernel: it is the share of the gathers that L2 served
            # (Reddit slabs are L2-resident; the N=1 line carries the L2 -> SM probe and the ncu DRAM bytes)
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "achieved_is": "gather-model (algorithmic) bytes per GPU / step time; multi-GPU kernels are "
                                        "not run under ncu, so there is no DRAM counter here.  frac > 1 = re-reads of "
                                        "L2-resident rows served by L2 instead of HBM, not skipped work (every leg is "
                                        "verified); the hardware-counter view of this kernel is the N=1 line's "
                                        "roofline.l2 / roofline.dram",
                         "l2_resident_operand": bool(r["mode"] == "features" or n * 128 * 4 <= 128 * 1024 * 1024),
                         "kernel": "spmm_hop_kernel (per GPU)"},
            "cpu_baseline": None, "e2e": r["e2e"], "gpu_launches": r["launches"], "clocks": r["clocks"],
            "parity": {name: (r["parity"]["ok"] if r["parity"] else None),
                       **{kk: (v.get("parity") or {}).get("ok") for kk, v in extras.items() if isinstance(v, dict) and "parity" in v},
                       "detail": r["parity"]},
            "details": {"parallelism": parallelism_text(r, world), "dist_mode": r["mode"],
                        "grid_cols_x_rows": r.get("grid", [world, 1]), "grid_model_ms": r.get("grid_model_ms"),
                        "normalise_ms": r["normalise_ms"], "partition_locality_lambda": r["lam"],
                        "pushed_rows_x_peers_fraction": r["push_fraction"], "rows_per_rank": r["n_pad"],
                        "per_hop_breakdown": r["breakdown"]},
            # the same workload on ONE GPU of this box, measured by rank 0 in this run, so that the
            # strong-scaling efficiency value / (n_gpus * single.value) can be read off this line
            "single_gpu_same_workload": r["single"],
            "strong_scaling_efficiency": (r["value"] / (world * r["single"]["value"]) if r["single"] else None),
            "other_workloads": extras,
        }
        line["summary"] = summary_of(name, line, extras)      # LAST key: survives a tail-truncated log
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if failed:
        if rank == 0:
            print("bench: multi-GPU output differs from the single-GPU / reference result: %s" % failed, file=sys.stderr)
        sys.exit(4)
