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


def worker(rank, world, port, scale, f, k, slab_w, mode, lam, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import sgc_b200
        from sgc_b200 import dist as sd, synth
        s = synth.make_shape("reddit", device=dev, scale=scale, seed=11, lam=lam)
        n = s["n"]
        x = s["x"][:, :f].contiguous()
        n_pad, parts = sd.partition_rows(n, world)
        row0, row1 = parts[rank]
        rp, col, val = sd.local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
        sg = sd.shard_normalize(n, rp, col, val)
        y = sd.propagate_sharded(sg, x[row0:row1].contiguous(), k, row_blocks=slab_w, mode=mode)
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
    args = (world, free_port(), kw.get("scale", 0.02), kw.get("f", 100), kw.get("k", 3), kw.get("slab_w", 4), kw.get("mode", "push"),
            kw.get("lam", 0.0), ret)
    mp.spawn(worker, args=args, nprocs=world, join=True)
    for rank in range(world):
        ok, same_struct, same_d, rows = ret[rank]
        assert same_struct and same_d, rank
        assert ok, "rank %d: sharded result differs from the single-GPU result" % rank


def test_sharded_world1_bit_identical():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    run(1, f=100, k=3, slab_w=4, mode="exchange")
    run(1, f=602, k=2, slab_w=1, mode="exchange")
    run(1, f=100, k=3, mode="push")
    run(1, f=602, k=2, mode="push")


def test_sharded_world2_bit_identical():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run(2, f=100, k=3, slab_w=4, mode="exchange")
    run(2, f=130, k=2, slab_w=3, mode="exchange")
    run(2, f=100, k=3, mode="push")
    run(2, f=130, k=2, mode="push")
    run(2, f=602, k=2, mode="push")
    run(2, f=100, k=3, mode="push", lam=0.95)      # partition locality: many rows no peer needs


def test_hop_pushes_rows_to_peer_buffers():
    """The fused exchange on ONE GPU: two extra local buffers stand in for the peers' mapped
    operands; after the hop they must hold exactly the rows the kernel wrote locally (long rows
    go through the fix-up kernel, which pushes too), and nothing else."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import sgc_b200
    from sgc_b200 import synth
    dev = torch.device("cuda", 0)
    s = synth.make_shape("reddit", device=dev, scale=0.02, seed=5)
    n = s["n"]
    g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    for f, tile_rows in ((100, "0"), (100, "1"), (100, "4"), (100, "16"), (130, "0"), (602, "0"), (7, "0"), (7, "3"), (64, "0")):
        monkey_env("SGC_B200_PUSH_ROWS", tile_rows)
        ld = (f + 7) // 8 * 8
        x = s["x"][:, :f].contiguous()
        want = sgc_b200.hop(g, x, chunk=64)
        bufs = [torch.full((n + 3, ld), -7.0, device=dev) for _ in range(3)]
        own, peers = bufs[0], bufs[1:]
        sgc_b200.hop(g, x, chunk=64, out=own[2:2 + n, :f], peers=[p[2:].data_ptr() for p in peers], pad_owned=True)
        r4 = (f + 3) // 4 * 4
        for b in bufs:
            assert torch.equal(b[2:2 + n, :f], want), (f, tile_rows)
            assert bool((b[:2] == -7.0).all()) and bool((b[2 + n:] == -7.0).all())
            # pad_owned hands over [f, round4(f)) only: everything beyond keeps its value, here and on the peers
            assert bool((b[2:2 + n, r4:] == -7.0).all()), (f, tile_rows)
            pad = b[2:2 + n, f:r4]
            assert bool(((pad == -7.0) | (pad == 0.0)).all())
        # pitch_owned: whole padded rows -- tiles of consecutive rows may leave as one store, zero-filling [f, ld)
        bufs = [torch.full((n + 3, ld), -7.0, device=dev) for _ in range(3)]
        sgc_b200.hop(g, x, chunk=64, out=bufs[0][2:2 + n, :f], peers=[p[2:].data_ptr() for p in bufs[1:]],
                     pitch_owned=True)
        for b in bufs:
            assert torch.equal(b[2:2 + n, :f], want), (f, tile_rows)
            assert bool((b[:2] == -7.0).all()) and bool((b[2 + n:] == -7.0).all())
            pad = b[2:2 + n, f:]
            assert bool(((pad == -7.0) | (pad == 0.0)).all())
        # halo mask: a row reaches peer p only if bit p of its mask is set
        mask = torch.randint(0, 4, (n,), dtype=torch.int32, device=dev)
        mask[: n // 3] = 3                        # long runs of rows every peer needs: the tile-store path
        bufs = [torch.full((n + 3, ld), -7.0, device=dev) for _ in range(3)]
        sgc_b200.hop(g, x, chunk=64, out=bufs[0][2:2 + n, :f], peers=[p[2:].data_ptr() for p in bufs[1:]],
                     peer_mask=mask, pitch_owned=True)
        assert torch.equal(bufs[0][2:2 + n, :f], want)
        for slot, b in enumerate(bufs[1:]):
            sent = ((mask >> slot) & 1).bool()
            assert torch.equal(b[2:2 + n, :f][sent], want[sent]), (f, tile_rows, slot)
            assert bool((b[2:2 + n][~sent] == -7.0).all()), (f, tile_rows, slot)
        # unaligned output (odd leading dimension): the per-lane store fallback
        ldu = f + 1 if (f + 1) % 2 else f + 2
        bufs = [torch.full((n + 3, ldu), -7.0, device=dev) for _ in range(3)]
        sgc_b200.hop(g, x, chunk=64, out=bufs[0][1:1 + n, :f], peers=[p[1:].data_ptr() for p in bufs[1:]])
        for b in bufs:
            assert torch.equal(b[1:1 + n, :f], want)
            assert bool((b[:1] == -7.0).all()) and bool((b[1 + n:] == -7.0).all()) and bool((b[:, f:] == -7.0).all())
    monkey_env("SGC_B200_PUSH_ROWS", None)
    # widths that are not a multiple of 4 floats cannot be bulk-pushed into operands whose padding is not owned
    x = s["x"][:, :130].contiguous()
    bufs = [torch.zeros((n, 136), device=dev) for _ in range(2)]
    with pytest.raises(sgc_b200._ffi.SgcError):
        sgc_b200.hop(g, x, out=bufs[0][:, :130], peers=[bufs[1].data_ptr()])


def test_peer_push_into_column_views_leaves_neighbours_alone():
    """Without pitch_owned the peers' buffers may be column views of wider live matrices, exactly like `out`:
    a narrow width on a pitch that fits one feature slab (the shape where tiles of rows would otherwise travel
    as one store of rows*ld floats) must leave every neighbouring column on every peer untouched."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import sgc_b200
    from sgc_b200 import synth
    dev = torch.device("cuda", 0)
    s = synth.make_shape("reddit", device=dev, scale=0.02, seed=7)
    n = s["n"]
    g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    for f, ld, c0, pad_owned in ((64, 128, 8, False), (100, 104, 0, False), (100, 104, 0, True), (30, 64, 4, True)):
        x = s["x"][:, :f].contiguous()
        want = sgc_b200.hop(g, x, chunk=64)
        bufs = [torch.full((n + 2, ld), 9.0, device=dev) for _ in range(3)]
        views = [b[1:1 + n, c0:c0 + f] for b in bufs]
        sgc_b200.hop(g, x, chunk=64, out=views[0], peers=[v.data_ptr() for v in views[1:]], pad_owned=pad_owned)
        keep = c0 + ((f + 3) // 4 * 4 if pad_owned else f)
        for b in bufs:
            assert torch.equal(b[1:1 + n, c0:c0 + f], want), (f, ld, c0)
            assert bool((b[:, :c0] == 9.0).all()) and bool((b[:, keep:] == 9.0).all()), (f, ld, c0)
            assert bool((b[0] == 9.0).all()) and bool((b[1 + n] == 9.0).all())


def monkey_env(key, value):
    if value is None:
        os.environ.pop(key, None)
    else:
        os.environ[key] = value


def test_hop_out_column_view_leaves_neighbours_alone():
    """hop(out=big[:, :f]) writes columns [0, f) only: the columns of a wider live matrix next to the view
    keep their values unless pad_owned=True hands the pad columns to the kernel."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import sgc_b200
    from sgc_b200 import synth
    dev = torch.device("cuda", 0)
    s = synth.make_shape("reddit", device=dev, scale=0.01, seed=3)
    n = s["n"]
    g = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    for f, width in ((30, 32), (30, 40), (101, 104), (7, 8), (601, 608)):
        x = s["x"][:, :f].contiguous()
        want = sgc_b200.hop(g, x, chunk=64)
        big = torch.full((n, width), 5.0, device=dev)
        sgc_b200.hop(g, x, chunk=64, out=big[:, :f])
        assert torch.equal(big[:, :f], want) and bool((big[:, f:] == 5.0).all()), (f, width)
        big = torch.full((n, width), 5.0, device=dev)
        sgc_b200.hop(g, x, chunk=64, out=big[:, :f], pad_owned=True)
        r4 = (f + 3) // 4 * 4
        assert torch.equal(big[:, :f], want) and bool((big[:, r4:] == 5.0).all())
    with pytest.raises(ValueError):
        sgc_b200.hop(g, x, out=torch.empty((n, 5), device=dev))
    with pytest.raises(ValueError):
        sgc_b200.hop(g, x, out=torch.empty((n, x.shape[1]), device=dev, dtype=torch.float64))


def test_push_rows_kernel():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ctypes
    from sgc_b200 import _ffi
    dev = torch.device("cuda", 0)
    for f, ld_src in ((100, 100), (602, 602), (602, 608), (5, 8)):
        rows, ld = 37, (f + 7) // 8 * 8
        src = torch.randn((rows, ld_src), device=dev)
        dst = [torch.full((rows + 2, ld), 3.0, device=dev) for _ in range(3)]
        arr = (_ffi.c_p * 3)(*[d[1:].data_ptr() for d in dst])
        _ffi.call("sgc_push_rows", src.data_ptr(), ld_src, rows, f, arr, 3, ld, torch.cuda.current_stream().cuda_stream)
        for d in dst:
            assert torch.equal(d[1:1 + rows, :f], src[:, :f])
            assert bool((d[0] == 3.0).all()) and bool((d[-1] == 3.0).all())
            if ld_src == f:
                assert bool((d[:, f:] == 3.0).all())


def test_feature_sharded_columns_bit_identical():
    """Column blocks of X propagated independently (S replicated, no exchange) equal the same
    columns of the single-GPU result bit for bit, for every block of an 8-way split."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import sgc_b200
    from sgc_b200 import synth
    from sgc_b200.dist import partition_features, propagate_feature_sharded
    dev = torch.device("cuda", 0)
    s = synth.make_shape("reddit", device=dev, scale=0.02, seed=21)
    g = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    x = s["x"]
    want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(x), g, 2)
    for c0, c1 in partition_features(x.shape[1], 8):
        got = propagate_feature_sharded(g, sgc_b200.to_padded(x[:, c0:c1]), 2)
        assert torch.equal(got, want[:, c0:c1])


def host_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import sgc_b200
        from sgc_b200 import synth
        from sgc_b200.dist import partition_features
        s = synth.make_shape("reddit", device=dev, scale=0.01, seed=13)
        n, k = s["n"], 2
        x = s["x"][:, :200].contiguous()
        c0, c1 = partition_features(200, world)[rank]
        host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
        xh = x[:, c0:c1].contiguous().cpu().pin_memory()
        st = {}
        got = sgc_b200.precompute_host(host, xh, k, standardize=True, group=dist.group.WORLD, col_slab=64, stats=st,
                                       detect_uniform_values=True)
        alone = sgc_b200.precompute_host(host, xh, k, standardize=True, col_slab=64)
        graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
        mean, std = sgc_b200.col_stats(x)
        want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded((x - mean) / std), graph, k)
        ok = [bool(torch.equal(got, alone)), bool(torch.equal(got, want[:, c0:c1].cpu()))]
        # the constant value array (2.0: adj + adj.T) is recognised piecewise, by consensus, and filled
        ok.append(st["uniform_value"] == float(host[3][0]))
        os.environ["SGC_B200_UNIFORM_VALUES"] = "0"           # always upload + all-gather: same bits
        st0 = {}
        got0 = sgc_b200.precompute_host(host, xh, k, standardize=True, group=dist.group.WORLD, col_slab=64, stats=st0,
                                        detect_uniform_values=True)
        os.environ["SGC_B200_UNIFORM_VALUES"] = "1"
        ok.append(st0["uniform_value"] is None and bool(torch.equal(got0, got)))
        ok.append(st0["h2d_bytes"] > st["h2d_bytes"])
        # a deviation that only the LAST rank's piece holds: every rank must upload (common verdict)
        val2 = s["val"].clone()
        val2[-1] = 3.5
        g2 = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], val2)
        want2, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded((x - mean) / std), g2, k)
        st2 = {}
        got2 = sgc_b200.precompute_host((n, host[1], host[2], val2.cpu().pin_memory()), xh, k, standardize=True,
                                        group=dist.group.WORLD, col_slab=64, stats=st2, detect_uniform_values=True)
        ok.append(st2["uniform_value"] is None and bool(torch.equal(got2, want2[:, c0:c1].cpu())))
        st3 = {}
        got3 = sgc_b200.precompute_host(host, xh, k, standardize=True, group=dist.group.WORLD, col_slab=64, stats=st3,
                                        detect_uniform_values=False)
        ok.append(st3["uniform_value"] is None and bool(torch.equal(got3, got)))      # switched off per call
        st4 = {}
        got4 = sgc_b200.precompute_host(host, xh, k, standardize=True, group=dist.group.WORLD, col_slab=64, stats=st4)
        ok.append(st4["uniform_value"] is None and bool(torch.equal(got4, got)))      # opt-in under `group`
        ret[rank] = tuple(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2])
def test_host_pipeline_shared_adjacency_upload(world):
    """precompute_host(group=...): every rank uploads 1/P of the adjacency and an NCCL all-gather
    replicates it; the result equals the rank-local upload and the single-GPU columns bit for bit; a
    constant value array is recognised piecewise (one-word all-reduce) and filled instead."""
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    ret = mp.Manager().dict()
    mp.spawn(host_worker, args=(world, free_port(), ret), nprocs=world, join=True)
    for rank in range(world):
        assert ret[rank] == (True,) * 8, (rank, ret[rank])
