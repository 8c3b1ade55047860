"""Host-side logic of the row-sharded propagation (sgc_b200/dist.py) on CPU: world_size 2 and 3
over gloo.  The CUDA hop is replaced by the oracle's CPU hop (tests may use the oracle), so what
is exercised is the partition, the ping-pong gathered operands, the per-row-block send/receive
pipeline and its ordering across hops.  Results must equal the single-process oracle bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def make_problem(n=53, f=70, seed=0):
    from oracle import normalization_oracle as no
    rng = np.random.RandomState(seed)
    dense = (rng.rand(n, n) < 0.15).astype(np.float64)
    dense = np.maximum(dense, dense.T)
    np.fill_diagonal(dense, 0)
    dense[7, :] = 1.0                                  # a hub row
    dense[:, 7] = 1.0
    indptr = np.zeros(n + 1, dtype=np.int64)
    rows, cols = np.nonzero(dense)
    np.cumsum(np.bincount(rows, minlength=n), out=indptr[1:])
    r = no.aug_normalized_adjacency_oracle(n, indptr, cols.astype(np.int32), dense[rows, cols])
    x = rng.standard_normal((n, f)).astype(np.float32)
    return r["indptr"], r["indices"].astype(np.int32), r["data"].astype(np.float32), x


def worker(rank, world, port, k, blocks, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import spmm_oracle as so
        from sgc_b200 import dist as sd
        from sgc_b200.graph import SgcGraph
        indptr, indices, data, x = make_problem()
        n, f = x.shape
        n_pad, parts = sd.partition_rows(n, world)
        row0, row1 = parts[rank]
        b, e = indptr[row0], indptr[row1]
        g = SgcGraph(row1 - row0, world * n_pad, torch.from_numpy(indptr[row0:row1 + 1] - b),
                     torch.from_numpy(indices[b:e]), None, None, sval=torch.from_numpy(data[b:e]))
        sg = sd.ShardedGraph(g, rank, world, n, n_pad, row0, row1)
        calls = []

        def cpu_hop(graph, xin, out):
            calls.append(tuple(xin.shape))
            rp = graph.row_ptr.numpy()             # a row block: offsets into the shared col/val arrays
            y = so.spmm_hop(rp - rp[0], graph.col.numpy()[rp[0]:rp[-1]], graph.sval.numpy()[rp[0]:rp[-1]],
                            np.ascontiguousarray(xin.numpy()))
            out.copy_(torch.from_numpy(y))

        y = sd.propagate_sharded(sg, torch.from_numpy(x[row0:row1].copy()), k, row_blocks=blocks, hop_fn=cpu_hop)
        assert 1 <= len(calls) <= k * blocks
        assert all(c == (world * n_pad, f) for c in calls)
        ret[rank] = (row0, row1, y.numpy().copy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,k,blocks", [(2, 2, 1), (2, 3, 4), (3, 1, 2), (3, 3, 5)])
def test_sharded_pipeline_matches_single_process(world, k, blocks):
    from oracle import spmm_oracle as so
    indptr, indices, data, x = make_problem()
    want, _ = so.sgc_precompute_oracle(x, indptr, indices, data, k)
    ret = mp.Manager().dict()
    mp.spawn(worker, args=(world, free_port(), k, blocks, ret), nprocs=world, join=True)
    got = np.zeros_like(want)
    covered = 0
    for rank in range(world):
        row0, row1, y = ret[rank]
        got[row0:row1] = y
        covered += row1 - row0
    assert covered == x.shape[0]
    assert np.array_equal(got, want)


def test_partition_rows():
    from sgc_b200.dist import partition_rows
    for n, world in ((10, 4), (5, 4), (8, 8), (2449029, 8), (1, 2)):
        n_pad, parts = partition_rows(n, world)
        assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        assert all(0 <= r1 - r0 <= n_pad for r0, r1 in parts)
        # a global row id is its own index in the gathered operand
        for rank, (r0, r1) in enumerate(parts):
            if r1 > r0:
                assert r0 == rank * n_pad


def test_partition_features_and_sharding_choice():
    from sgc_b200.dist import FEATURE_UNIT, choose_sharding, partition_features
    for f, world in ((602, 8), (602, 2), (100, 8), (128, 8), (1433, 4), (7, 3), (16, 1)):
        parts = partition_features(f, world)
        assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == f
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        widths = [b - a for a, b in parts]
        assert all(a % FEATURE_UNIT == 0 for a, _ in parts if a < f)       # blocks start sector-aligned
        assert max(widths) - min(w for w in widths if w or True) <= 2 * FEATURE_UNIT
    # Reddit shape: S fits everywhere and blocks stay wide -> columns; narrow products/papers -> rows
    assert choose_sharding(232965, 23.4e6, 602, 8) == "features"
    assert choose_sharding(2449029, 126e6, 100, 8) == "rows"
    assert choose_sharding(111059956, 1.75e9, 128, 8) == "rows"
    assert choose_sharding(232965, 23.4e6, 602, 8, mem_bytes=1e8) == "rows"   # S replica does not fit


def shared_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgc_b200 import pipeline
        cpu = torch.device("cpu")
        group = dist.group.WORLD
        pieces = {}
        for nnz in (0, 1, 2, 3, 7, 64, 1001, 23163630):
            lo, hi, per, w = pipeline._shared_piece(nnz, group)
            pieces[nnz] = (lo, hi, per, w)
        verdicts = []
        os.environ["SGC_B200_UNIFORM_VALUES"] = "1"
        val = torch.full((200001,), 2.0, dtype=torch.float64)
        verdicts.append(pipeline._uniform_value_shared(val, cpu, group))            # constant -> 2.0 everywhere
        val[-1] = 3.0                                                                 # only the last rank sees it
        verdicts.append(pipeline._uniform_value_shared(val, cpu, group))
        val[-1] = 2.0
        val[0] = 5.0                                                                  # only the first rank sees it
        verdicts.append(pipeline._uniform_value_shared(val, cpu, group))
        val[0] = float("inf")
        verdicts.append(pipeline._uniform_value_shared(val, cpu, group))
        val[0] = 2.0
        verdicts.append(pipeline._uniform_value_shared(val.float(), cpu, group))    # not fp64: never scanned
        verdicts.append(pipeline._uniform_value_shared(val[:1], cpu, group))        # one entry, empty pieces elsewhere
        os.environ["SGC_B200_UNIFORM_VALUES"] = "0"
        verdicts.append(pipeline._uniform_value_shared(val, cpu, group))
        ret[rank] = (pieces, verdicts)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_shared_adjacency_pieces_and_common_verdict(world):
    """pipeline._shared_piece tiles the stored entries (even piece length: 16-byte aligned fp64 / 8-byte
    int32 pieces); pipeline._uniform_value_shared gives every rank the same verdict whichever rank's
    piece holds the deviation (gloo, verdict tensor on the CPU; the scan itself is host code)."""
    import __graft_entry__ as entry
    entry.build()
    ret = mp.Manager().dict()
    mp.spawn(shared_worker, args=(world, free_port(), ret), nprocs=world, join=True)
    for nnz in ret[0][0]:
        spans = [ret[r][0][nnz] for r in range(world)]
        per = spans[0][2]
        assert all(s[2] == per and s[3] == world for s in spans) and per % 2 == 0 and per * world >= nnz
        assert spans[0][0] == 0 and spans[-1][1] == nnz
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert all(0 <= hi - lo <= per for lo, hi, _, _ in spans)
    for r in range(world):
        assert ret[r][1] == [2.0, None, None, None, None, 2.0, None], (r, ret[r][1])


def test_block_checksum_is_partition_independent_and_bit_sensitive():
    """dist.block_checksum: the sum over the blocks of any partition equals the checksum of the whole matrix
    (wrap-around int64), and any single-bit change anywhere -- including +0.0 vs -0.0 and element swaps --
    changes it."""
    from sgc_b200.dist import block_checksum
    torch.manual_seed(0)
    y = torch.randn(57, 23)
    f = y.shape[1]
    whole = int(block_checksum(y, 0, 0, f))
    M = (1 << 64)
    rows = sum(int(block_checksum(y[a:b], a, 0, f)) for a, b in ((0, 10), (10, 11), (11, 57))) % M
    cols = sum(int(block_checksum(y[:, a:b], 0, a, f, chunk_elems=64)) for a, b in ((0, 16), (16, 23))) % M
    assert rows == whole % M and cols == whole % M
    z = y.clone()
    z[31, 7] = torch.nextafter(z[31, 7], torch.tensor(10.0))
    assert int(block_checksum(z, 0, 0, f)) != whole
    z = y.clone()
    z[3, 4], z[3, 5] = y[3, 5], y[3, 4]
    assert int(block_checksum(z, 0, 0, f)) != whole
    zero = torch.zeros(4, 4)
    neg = zero.clone()
    neg[2, 2] = -0.0
    assert int(block_checksum(zero, 0, 0, 4)) != int(block_checksum(neg, 0, 0, 4))
    assert int(block_checksum(y[:0], 0, 0, f)) == 0


def verify_worker(rank, world, port, corrupt, by_cols, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgc_b200 import dist as sd
        torch.manual_seed(1)
        y = torch.randn(41, 19)
        if by_cols:
            c0, c1 = sd.partition_features(19, world, unit=4)[rank]
            mine, r0 = y[:, c0:c1].clone(), 0
        else:
            _, parts = sd.partition_rows(41, world)
            r0, r1 = parts[rank]
            mine, c0 = y[r0:r1].clone(), 0
        if corrupt == rank and mine.numel():
            mine.view(-1)[mine.numel() // 2] *= (1 + 2 ** -23)           # one ulp, one element, one rank
        rep = sd.verify_against_single(mine, y if rank == 0 else None, r0, c0, 19)
        ret[rank] = (rep["ok"], rep["per_rank"], rep["checksum"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,by_cols", [(2, False), (3, False), (2, True)])
def test_verify_against_single_flags_the_rank_that_differs(world, by_cols):
    """dist.verify_against_single (what every bench.py --gpus N leg runs after its timed steps): all ranks
    agree on the verdict; a one-ulp deviation of one element on one rank is found and attributed."""
    for corrupt in (-1, world - 1):
        ret = mp.Manager().dict()
        mp.spawn(verify_worker, args=(world, free_port(), corrupt, by_cols, ret), nprocs=world, join=True)
        want = [r != corrupt for r in range(world)]
        for rank in range(world):
            ok, per_rank, checksum = ret[rank]
            assert per_rank == want and ok == all(want), (corrupt, rank, per_rank)
            assert checksum == ret[0][2]


def replication_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgc_b200 import dist as sd
        from sgc_b200.graph import SgcGraph
        indptr, indices, data, x = make_problem()
        n, f = x.shape
        n_pad, parts = sd.partition_rows(n, world)
        row0, row1 = parts[rank]
        b, e = indptr[row0], indptr[row1]
        g = SgcGraph(row1 - row0, world * n_pad, torch.from_numpy(indptr[row0:row1 + 1] - b),
                     torch.from_numpy(indices[b:e]), None, None, sval=torch.from_numpy(data[b:e]))
        sg = sd.ShardedGraph(g, rank, world, n, n_pad, row0, row1)
        prop = sd.ShardedPropagator(sg, f, row_blocks=2, hop_fn=lambda *a: None)
        prop.load(torch.from_numpy(x[row0:row1].copy()))
        clean = sd._replication_mismatches(prop, prop.buf[0], False, 1000, None)
        masked = sd._replication_mismatches(prop, prop.buf[0], True, 1000, None)
        if rank == 1:                                  # a stale copy of one of rank 0's rows, on rank 1 only
            victim = int(indices[indptr[row0]:indptr[row1]][indices[indptr[row0]:indptr[row1]] < parts[0][1]][0])
            prop.buf[0][victim, 3] += 1.0
        dirty = sd._replication_mismatches(prop, prop.buf[0], True, 1000, None)
        ret[rank] = (clean, masked, dirty)
    finally:
        dist.destroy_process_group()


def test_replication_check_finds_a_stale_pushed_row():
    """dist._replication_mismatches (part of verify_sharded): owner-vs-copy comparison of sampled rows of the
    full-height operand, restricted to the rows a rank's shard gathers when halo masks are in use."""
    world = 2
    ret = mp.Manager().dict()
    mp.spawn(replication_worker, args=(world, free_port(), ret), nprocs=world, join=True)
    for rank in range(world):
        clean, masked, dirty = ret[rank]
        assert clean[0] == 0 and clean[1] > 0 and masked[0] == 0 and 0 < masked[1] <= clean[1]
        assert dirty[0] == 1 and dirty[1] == masked[1]


# --------------------------------------------------------------------------------------------------
# 2-D grid sharding (dist.make_grid / choose_grid): column blocks x row parts
def grid_worker(rank, world, port, p_cols, p_rows, k, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import spmm_oracle as so
        from sgc_b200 import dist as sd
        from sgc_b200.graph import SgcGraph
        indptr, indices, data, x = make_problem(n=61, f=150)
        n, f = x.shape
        grid = sd.make_grid(f, p_cols, p_rows)
        assert (grid.ci, grid.ri) == (rank // p_rows, rank % p_rows)
        assert dist.get_world_size(grid.row_group) == p_rows and dist.get_rank(grid.row_group) == grid.ri
        c0, c1 = grid.cols
        n_pad, parts = sd.partition_rows(n, p_rows)
        row0, row1 = parts[grid.ri]
        b, e = indptr[row0], indptr[row1]
        g = SgcGraph(row1 - row0, p_rows * n_pad, torch.from_numpy(indptr[row0:row1 + 1] - b),
                     torch.from_numpy(indices[b:e]), None, None, sval=torch.from_numpy(data[b:e]))
        sg = sd.ShardedGraph(g, grid.ri, p_rows, n, n_pad, row0, row1)

        def cpu_hop(graph, xin, out):
            rp = graph.row_ptr.numpy()
            y = so.spmm_hop(rp - rp[0], graph.col.numpy()[rp[0]:rp[-1]], graph.sval.numpy()[rp[0]:rp[-1]],
                            np.ascontiguousarray(xin.numpy()))
            out.copy_(torch.from_numpy(y))

        y = sd.propagate_sharded(sg, torch.from_numpy(x[row0:row1, c0:c1].copy()), k, row_blocks=2,
                                 group=grid.row_group, hop_fn=cpu_hop)
        # every rank's [rows, columns] block against the single-process result held by rank 0 only
        want = None
        if rank == 0:
            want = torch.from_numpy(so.sgc_precompute_oracle(x, indptr, indices, data, k)[0])
        rep = sd.verify_against_single(y, want, row0, c0, f)
        ret[rank] = (row0, row1, c0, c1, y.numpy().copy(), rep["ok"], rep["per_rank"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,p_cols,p_rows,k", [(4, 2, 2, 2), (2, 1, 2, 3), (2, 2, 1, 2), (4, 1, 4, 2)])
def test_grid_sharded_matches_single_process(world, p_cols, p_rows, k):
    from oracle import spmm_oracle as so
    indptr, indices, data, x = make_problem(n=61, f=150)
    want, _ = so.sgc_precompute_oracle(x, indptr, indices, data, k)
    ret = mp.Manager().dict()
    mp.spawn(grid_worker, args=(world, free_port(), p_cols, p_rows, k, ret), nprocs=world, join=True)
    got = np.full_like(want, np.nan)
    for rank in range(world):
        row0, row1, c0, c1, y, ok, per_rank = ret[rank]
        assert ok and all(per_rank), (rank, per_rank)
        got[row0:row1, c0:c1] = y
    assert np.array_equal(got, want)             # the blocks tile the whole matrix


def test_choose_grid(monkeypatch):
    from sgc_b200 import dist as sd
    monkeypatch.delenv("SGC_B200_GRID", raising=False)
    monkeypatch.delenv("SGC_B200_SHARDING", raising=False)
    monkeypatch.setenv("SGC_B200_GRID_AUTO", "1")
    for world in (1, 2, 4, 8):
        for shape in ((232965, 23.4e6, 602, 2), (2449029, 126e6, 100, 3), (111059956, 1.75e9, 128, 3), (19717, 1.1e5, 500, 2)):
            pc, pr = sd.choose_grid(*shape, world)
            assert pc * pr == world and pc >= 1 and pr >= 1
            if pc > 1:
                assert shape[2] // pc >= 64                      # blocks stay at least 64 columns wide
    # operands far larger than L2 and narrow feature matrices: pure row sharding
    assert sd.choose_grid(2449029, 126e6, 100, 3, 8) == (1, 8)
    assert sd.choose_grid(111059956, 1.75e9, 128, 3, 8) == (1, 8)
    # Reddit shape on 8 GPUs: neither 8 narrow column blocks nor 7 peers per row
    pc, pr = sd.choose_grid(232965, 23.4e6, 602, 2, 8)
    assert 1 < pc < 8 and 1 < pr < 8
    monkeypatch.setenv("SGC_B200_GRID", "4x2")
    assert sd.choose_grid(232965, 23.4e6, 602, 2, 8) == (4, 2)
    with pytest.raises(ValueError):
        sd.choose_grid(232965, 23.4e6, 602, 2, 4)
    monkeypatch.delenv("SGC_B200_GRID")
    monkeypatch.setenv("SGC_B200_SHARDING", "rows")
    assert sd.choose_grid(232965, 23.4e6, 602, 2, 8) == (1, 8)
    monkeypatch.setenv("SGC_B200_SHARDING", "features")
    assert sd.choose_grid(232965, 23.4e6, 602, 2, 8) == (8, 1)
    # the cost model: more row parts -> less compute per rank, more egress per exchanging hop
    assert sd.grid_cost_ms(232965, 23.4e6, 602, 2, 1, 8) < sd.grid_cost_ms(232965, 23.4e6, 602, 2, 1, 2)
    assert sd.grid_cost_ms(232965, 23.4e6, 602, 1, 2, 4) < sd.grid_cost_ms(232965, 23.4e6, 602, 2, 2, 4)
