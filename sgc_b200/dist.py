"""K5: row-sharded propagation on the GPUs of one box (one process per GPU, torch.distributed).

The reference is single-process; this is the north_star's capability (3): A is
row-partitioned, rank g owns rows [g*n_pad, (g+1)*n_pad) of S (CSR slice with global column ids)
and of D^-1/2, produces those rows of every hop's output, and the hops exchange X over NVLink.

Data layout.  Every rank keeps two full-height gathered operands ``buf[0|1]`` of shape
[world*n_pad, ld] (row-major, ld = F rounded up to 8 floats).  Hop k gathers from buf[k%2] and
writes this rank's rows straight into its own segment of buf[(k+1)%2]; the other segments of
that buffer are filled by the peers.  A global row id is its own index in the buffer, so the
single-GPU kernel runs unchanged and every output element is accumulated in the same order as on
one GPU: sharded results are bit-identical to single-GPU results for any GPU count.

Two ways of filling the peers' segments:

``push`` (default on CUDA)  the exchange is fused into the SpMM kernel.  The operands are
    "symmetric" buffers (sgc_symm_alloc) whose CUDA IPC handles the ranks swap once; every rank
    maps every peer's buffers into its own address space.  spmm_hop_kernel then stores each
    output row from its epilogue registers to the local buffer AND to the same offset of every
    peer's buffer over NVLink (sgc_hop_args.peer_y_d) while other warps are still gathering: one
    kernel per hop, no all-gather, and the ranks only meet at a tiny barrier between hops.

``exchange``  one NCCL group of sends/receives per ROW BLOCK, placed directly at their final
    position, pipelined with the SpMM of the following blocks (only the last block's transfer is
    exposed).  Library-collective baseline for the fused path; also the mode the CPU (gloo) tests
    drive.
"""
import ctypes
import os
import time
import weakref

import torch
import torch.distributed as dist

from . import _ffi
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


def default_row_blocks():
    """Pipeline stages per hop in ``exchange`` mode (SGC_B200_DIST_BLOCKS)."""
    return max(1, int(os.environ.get("SGC_B200_DIST_BLOCKS", "4")))


def default_mode(device):
    """SGC_B200_DIST_MODE: 'push' (fused into the hop kernel; CUDA only) or 'exchange'."""
    mode = os.environ.get("SGC_B200_DIST_MODE", "push" if torch.device(device).type == "cuda" else "exchange")
    if mode not in ("push", "exchange"):
        raise ValueError("SGC_B200_DIST_MODE must be 'push' or 'exchange'")
    return mode


def _cuda_hop(graph, x, out, peers=None, peer_mask=None):
    from .propagate import hop
    hop(graph, x, out=out, peers=peers, peer_mask=peer_mask)


def halo_enabled():
    """SGC_B200_DIST_HALO=0 pushes every row to every peer (plain all-gather semantics); default 1:
    a row is pushed only to the peers whose shard actually gathers it."""
    return os.environ.get("SGC_B200_DIST_HALO", "1") != "0"


def needed_by_peers(sg, group=None):
    """int32 [n_local]: bit p set = the p-th peer (ranks in ascending order, this rank left out)
    stores at least one entry in column row0+i, i.e. gathers this rank's row i in every hop.
    One pass over the local column indices plus one all-to-all of world*n_pad bytes."""
    g, world, rank, n_pad = sg.graph, sg.world, sg.rank, sg.n_pad
    dev = g.device
    touch = torch.zeros(world * n_pad, dtype=torch.uint8, device=dev)
    if g.nnz:
        touch[g.col.long()] = 1
    recv = torch.empty_like(touch)
    dist.all_to_all_single(recv, touch, group=group)         # recv[q*n_pad + i] = rank q touches my row i
    recv = recv.view(world, n_pad)[:, :sg.n_local].to(torch.int32)
    mask = torch.zeros(sg.n_local, dtype=torch.int32, device=dev)
    slot = 0
    for q in range(world):
        if q == rank:
            continue
        mask |= recv[q] << slot
        slot += 1
    return mask


def _pad8(f):          # historical name: the padded leading dimension, see features.pad_ld
    from .features import pad_ld
    return pad_ld(f)


class _DeviceMemory(object):
    """Lets torch.as_tensor view a raw device allocation (CUDA array interface, no copy)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def _release_symm(local_ptr, peer_ptrs):
    lib = _ffi.lib()
    for ptr in peer_ptrs:
        lib.sgc_ipc_close(ptr)
    lib.sgc_symm_free(local_ptr)


class SymmetricBuffer(object):
    """fp32 [rows, ld] buffer that exists on every rank of the group at the same logical layout,
    each rank holding device addresses of all of them: ``ptrs[r]`` is rank r's copy as seen from
    this process (its own allocation for r == rank, a CUDA-IPC mapping reached over NVLink
    otherwise).  Collective: every rank of the group must construct it at the same time."""

    def __init__(self, rows, ld, device, group=None):
        lib = _ffi.lib()
        self.rows, self.ld, self.group = int(rows), int(ld), group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        nbytes = self.rows * self.ld * 4
        with torch.cuda.device(device):
            ptr = _ffi.c_p()
            _ffi.call("sgc_symm_alloc", nbytes, ctypes.byref(ptr))
            handle = ctypes.create_string_buffer(_ffi.IPC_HANDLE_BYTES)
            _ffi.call("sgc_ipc_export", ptr, handle)
            handles = [None] * self.world
            dist.all_gather_object(handles, handle.raw, group=group)
            self.ptrs, opened = [], []
            for r in range(self.world):
                if r == self.rank:
                    self.ptrs.append(int(ptr.value))
                    continue
                peer = _ffi.c_p()
                _ffi.call("sgc_ipc_open", ctypes.create_string_buffer(handles[r], _ffi.IPC_HANDLE_BYTES),
                          ctypes.byref(peer))
                self.ptrs.append(int(peer.value))
                opened.append(int(peer.value))
            self.tensor = torch.as_tensor(_DeviceMemory(ptr.value, (self.rows, self.ld)), device=device)
        dist.barrier(group=group)          # nobody frees or writes before everyone has mapped
        self._fin = weakref.finalize(self, _release_symm, int(ptr.value), opened)

    def row_ptrs(self, row, peers_only):
        """Device address of row `row` in every rank's copy (optionally without this rank's)."""
        off = int(row) * self.ld * 4
        return [p + off for r, p in enumerate(self.ptrs) if not (peers_only and r == self.rank)]

    def close(self):
        self.tensor = None
        self._fin()


class ShardedPropagator(object):
    """Buffers + schedule for K-hop propagation on a ShardedGraph (see module docstring)."""

    def __init__(self, sg, f, row_blocks=None, group=None, hop_fn=None, mode=None):
        self.sg, self.f, self.group = sg, int(f), group
        self.hop_fn = hop_fn or _cuda_hop
        g = sg.graph
        dev = g.device
        from .propagate import fused_norm
        if g.fused and not fused_norm() and hop_fn is None:
            g.values32()                      # materialise once so that row blocks share it
        self.mode = mode or ("exchange" if hop_fn is not None else default_mode(dev))
        self.ld = _pad8(self.f)
        n_all = sg.world * sg.n_pad
        _, self.parts = partition_rows(sg.n_global, sg.world)
        if self.mode == "push":
            torch.cuda.empty_cache()          # the symmetric buffers come straight from cudaMalloc
            self.symm = [SymmetricBuffer(n_all, self.ld, dev, group) for _ in range(2)]
            self.buf = [b.tensor for b in self.symm]
            self._flag = torch.zeros(1, dtype=torch.float32, device=dev)
            self.n_blocks, self.blk, self.blocks = 1, sg.n_pad, [g]
            self.peer_mask = needed_by_peers(sg, group) if (sg.world > 1 and halo_enabled()) else None
            self.push_fraction = self._push_fraction()
            return
        self.buf = [torch.zeros((n_all, self.ld), dtype=torch.float32, device=dev) for _ in range(2)]
        self.n_blocks = max(1, min(int(row_blocks or default_row_blocks()), max(sg.n_pad, 1)))
        self.blk = (sg.n_pad + self.n_blocks - 1) // self.n_blocks
        # one sub-graph (views into the same CSR arrays) per row block of this rank
        self.blocks = []
        for r in range(self.n_blocks):
            a, b = self.block_rows(sg.rank, r)
            if b <= a:
                self.blocks.append(None)
                continue
            sub = SgcGraph(b - a, g.n_cols, g.row_ptr[a:b + 1], g.col, g.aval, g.dinv, sval=g.sval,
                           dinv_rows=None if g.dinv_rows is None else g.dinv_rows[a:b])
            self.blocks.append(sub)

    def block_rows(self, rank, r):
        """Local row range of block r on `rank` (empty when that rank has fewer rows)."""
        n_local = self.parts[rank][1] - self.parts[rank][0]
        return min(r * self.blk, n_local), min((r + 1) * self.blk, n_local)

    # -- push mode -----------------------------------------------------------------------------
    def _push_fraction(self):
        """Rows x peers actually pushed / (rows x peers), over all ranks (1.0 without halo masks)."""
        sg = self.sg
        if self.peer_mask is None or sg.world == 1:
            return 1.0
        bits = torch.zeros(1, dtype=torch.float64, device=self.peer_mask.device)
        m = self.peer_mask
        for slot in range(sg.world - 1):
            bits += ((m >> slot) & 1).sum()
        tot = torch.tensor([float(sg.n_local * (sg.world - 1))], dtype=torch.float64, device=m.device)
        both = torch.cat([bits, tot])
        dist.all_reduce(both, group=self.group)
        return float(both[0].item() / max(both[1].item(), 1.0))

    def _barrier(self):
        """Stream-ordered rendezvous of all ranks: a one-element all-reduce enqueued behind the hop
        kernel.  When it completes on a rank, every rank's preceding kernel -- and therefore every
        NVLink store it issued into this rank's buffers -- has completed."""
        if self.sg.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def _push_load(self, x_local):
        sg = self.sg
        if sg.n_local == 0:
            self._barrier()
            return
        x_local = x_local if (x_local.dtype == torch.float32 and x_local.stride(1) == 1) else x_local.float().contiguous()
        dsts = self.symm[0].row_ptrs(sg.row0, peers_only=False)
        arr = (_ffi.c_p * len(dsts))(*dsts)
        with torch.cuda.device(x_local.device):
            _ffi.call("sgc_push_rows", x_local.data_ptr(), x_local.stride(0), sg.n_local, self.f, arr, len(dsts),
                      self.ld, torch.cuda.current_stream().cuda_stream)
        self._barrier()

    def _push_propagate(self, degree):
        sg = self.sg
        for k in range(degree):
            src, dst = self.buf[k % 2], self.symm[(k + 1) % 2]
            last = (k == degree - 1)
            if sg.n_local > 0:
                peers = None if (last or sg.world == 1) else dst.row_ptrs(sg.row0, peers_only=True)
                self.hop_fn(sg.graph, src[:, :self.f], dst.tensor[sg.row0:sg.row1, :self.f], peers,
                            None if peers is None else self.peer_mask)
            # after the barrier: all rows of hop k are in place everywhere, and nobody still reads
            # the buffer the next hop (or the next load) overwrites
            self._barrier()
        return self.buf[degree % 2][sg.row0:sg.row1, :self.f]

    # -- exchange mode ---------------------------------------------------------------------------
    def _exchange_block(self, dst, r):
        """Send this rank's rows of block r (already written into dst) to every peer and receive
        theirs, each piece landing at its final position in dst."""
        sg = self.sg
        ops = []
        a, b = self.block_rows(sg.rank, r)
        for q in range(sg.world):
            if q == sg.rank:
                continue
            if b > a:
                ops.append(dist.P2POp(dist.isend, dst[sg.row0 + a:sg.row0 + b], q, self.group))
            qa, qb = self.block_rows(q, r)
            if qb > qa:
                base = q * sg.n_pad
                ops.append(dist.P2POp(dist.irecv, dst[base + qa:base + qb], q, self.group))
        return dist.batch_isend_irecv(ops) if ops else []

    def load(self, x_local):
        """Distribute the initial features: afterwards buf[0] holds all of X on every rank (the
        state the reference starts sgc_precompute from -- X fully resident)."""
        sg = self.sg
        assert tuple(x_local.shape) == (sg.n_local, self.f)
        if self.mode == "push":
            return self._push_load(x_local)
        self.buf[0][sg.row0:sg.row1, :self.f].copy_(x_local)
        works = []
        for r in range(self.n_blocks):
            works += self._exchange_block(self.buf[0], r)
        for w in works:
            w.wait()

    def propagate_resident(self, degree):
        """K hops starting from buf[0]; returns this rank's rows of S^K X as a [n_local, F] view."""
        if self.mode == "push":
            return self._push_propagate(degree)
        sg = self.sg
        pending = []
        for k in range(degree):
            src, dst = self.buf[k % 2], self.buf[(k + 1) % 2]
            for w in pending:                               # every peer's rows of hop k-1 have landed
                w.wait()
            pending = []
            last = (k == degree - 1)
            for r in range(self.n_blocks):
                sub = self.blocks[r]
                a, b = self.block_rows(sg.rank, r)
                if sub is not None:
                    self.hop_fn(sub, src[:, :self.f], dst[sg.row0 + a:sg.row0 + b, :self.f])
                if not last:                                # rows of block r travel while r+1.. compute
                    pending += self._exchange_block(dst, r)
        return self.buf[degree % 2][sg.row0:sg.row1, :self.f]

    def propagate(self, x_local, degree):
        """x_local: this rank's rows of X, fp32 [n_local, F] -> this rank's rows of S^K X."""
        if degree <= 0:
            return x_local
        self.load(x_local)
        return self.propagate_resident(degree)

    def rebind(self, sg):
        """Reuse the buffers (and, in push mode, the peer mappings) for another ShardedGraph of the
        same partition -- they are a per-shape resource like a communicator."""
        assert (sg.world, sg.n_pad, sg.row0, sg.row1) == (self.sg.world, self.sg.n_pad, self.sg.row0, self.sg.row1)
        changed = sg is not self.sg
        self.sg = sg
        if self.mode == "push":
            self.blocks = [sg.graph]
            if changed and self.peer_mask is not None:
                self.peer_mask = needed_by_peers(sg, self.group)     # the structure may differ
        return self

    def close(self):
        if self.mode == "push":
            self.buf = []
            for b in self.symm:
                b.close()


def propagate_sharded(sg, x_local, degree, row_blocks=None, group=None, hop_fn=None, mode=None):
    prop = ShardedPropagator(sg, x_local.shape[1], row_blocks, group, hop_fn, mode)
    out = prop.propagate(x_local, degree).clone()
    if prop.mode == "push":
        torch.cuda.synchronize(out.device)
        prop.close()
    return out


# --------------------------------------------------------------------------------------------------
def local_rows_of(graph_csr, row0, row1):
    """Slice [row0, row1) of a CSR given as (row_ptr, col, val) tensors."""
    row_ptr, col, val = graph_csr
    b, e = int(row_ptr[row0].item()), int(row_ptr[row1].item())
    return (row_ptr[row0:row1 + 1] - b).contiguous(), col[b:e].contiguous(), val[b:e].contiguous()


# --------------------------------------------------------------------------------------------------
# Feature-sharded propagation: S replicated, X split by COLUMNS.  Column c of S^K X depends on column
# c of X only, so the K hops of a column block need no exchange at all -- not even the Reddit
# standardisation (utils.py:119 is per column).  This is the sharding of choice whenever S fits on
# every GPU and the per-GPU block stays wide enough for efficient gathers; row sharding (above) is
# for adjacencies that exceed one GPU or for narrow feature matrices.
FEATURE_UNIT = 16            # floats: blocks are multiples of 64 bytes so that gathers stay sector-aligned


def partition_features(f, world, unit=FEATURE_UNIT):
    """Contiguous column blocks [(c0, c1)] in units of `unit` floats, sizes differing by at most one
    unit; trailing ranks may be empty when f is small."""
    units = (int(f) + unit - 1) // unit
    base, extra = divmod(units, world)
    out, c = [], 0
    for r in range(world):
        w = (base + (1 if r < extra else 0)) * unit
        out.append((min(c, f), min(c + w, f)))
        c += w
    return out


def choose_sharding(n, nnz, f, world, mem_bytes=170e9):
    """'features' when a replica of S (12 B per entry + row pointers) plus two N x F/world buffers fit on
    each GPU and every GPU gets at least 64 columns (256-byte gathers); else 'rows'."""
    forced = os.environ.get("SGC_B200_SHARDING")
    if forced in ("rows", "features"):
        return forced
    widest = max(c1 - c0 for c0, c1 in partition_features(f, world))
    fits = nnz * 20.0 + n * 8.0 + 3.0 * n * _pad8(widest) * 4.0 < mem_bytes
    return "features" if (fits and f // world >= 64) else "rows"


def propagate_feature_sharded(graph, x_cols, degree, **kw):
    """K hops on this rank's column block [N, fc] of X with the full S: no communication.  Returns
    the same columns of S^K X, bit-identical to those columns of the single-GPU result."""
    from .propagate import propagate
    return propagate(graph, x_cols, degree, **kw)


def run_feature_sharded_workload(name, args, rank, world, dev, B, with_e2e=True, want_single=True):
    """bench leg for sharding == 'features' (see run_sharded_workload for the contract)."""
    from . import _ffi, synth
    from .features import col_stats, standardize_, to_padded
    from .graph import normalize_csr
    from .propagate import default_chunk, propagate
    shape = synth.SHAPES[name]
    n, f, k = shape["n"], shape["f"], shape["k"]
    c0, c1 = partition_features(f, world)[rank]
    s = synth.make_shape(name, device=dev, with_features=False, lam=bench_lambda())
    x = synth.gaussian_features(n, f, seed=1, scaled=shape["scaled"], device=dev)
    x_raw_local = x[:, c0:c1].contiguous()
    if shape["scaled"]:
        mean, std = col_stats(x)
        standardize_(x, mean, std)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    graph = normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    torch.cuda.synchronize()
    normalise_ms = (time.perf_counter() - t0) * 1e3
    single = None
    if rank == 0 and want_single:
        xp = to_padded(x)
        for _ in range(3):
            propagate(graph, xp, k)
        torch.cuda.synchronize()
        ms1 = _device_ms(lambda: propagate(graph, xp, k), 5)
        single = {"n_gpus": 1, "ms_per_step": ms1, "value": k * graph.nnz * f / (ms1 * 1e-3)}
        del xp
    x_local = to_padded(x[:, c0:c1]) if c1 > c0 else None
    del x
    torch.cuda.empty_cache()
    nnz = graph.nnz
    graph.plan(default_chunk())

    def step():
        return propagate(graph, x_local, k) if x_local is not None else None

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    sampler = B["ClockSampler"](dev.index)
    sampler.start()
    launches0 = _ffi.launch_count()
    torch.cuda.synchronize()
    dist.barrier()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(args.steps):
        step()
    stop.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([start.elapsed_time(stop)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)            # max over ranks, device-timed
    launches = _ffi.launch_count() - launches0
    clocks = sampler.stop()
    ms_per_step = float(ms.item()) / args.steps
    out = {"name": name, "n": n, "f": f, "k": k, "nnz": nnz, "n_pad": n, "mode": "features",
           "push_fraction": 0.0, "n_blocks": 1, "normalise_ms": normalise_ms, "ms_per_step": ms_per_step,
           "value": k * nnz * f / (ms_per_step * 1e-3), "launches": int(launches), "clocks": clocks,
           "single": single, "e2e": None, "cols_per_rank": max(b - a for a, b in partition_features(f, world))}
    if with_e2e and x_local is not None:
        host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val")}
        host["x"] = x_raw_local.cpu().pin_memory()
        out_host = torch.empty((n, c1 - c0), dtype=torch.float32).pin_memory()

        from .pipeline import precompute_host
        host_csr = (n, host["row_ptr"], host["col"], host["val"])
        # one node, one copy of the adjacency in host memory: every rank uploads 1/P of it and NVLink
        # replicates it (SGC_B200_SHARED_CSR=0: every rank uploads all of it)
        shared_csr = os.environ.get("SGC_B200_SHARED_CSR", "1") != "0" and \
            all(b > a for a, b in partition_features(f, world))       # the all-gather needs every rank

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
        out["e2e"] = {"value": k * nnz * f / float(e2e_s.item()), "unit": B["UNIT"],
                      "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(n * f * 4),
                      "ms_per_step": float(e2e_s.item()) * 1e3,
                      "path": "per rank: sgc_b200.precompute_host on its column block (pipeline.py); adjacency "
                              + ("uploaded once per node (1/P per rank over its own PCIe link, NCCL all-gather "
                                 "over NVLink)" if shared_csr else "uploaded by every rank")
                              + ("" if st.get("uniform_value") is None else
                                 "; its stored values are all %g (exact host scan every step, 1/P per rank + "
                                 "one-word all-reduce): filled on the GPU instead of uploaded" % st["uniform_value"])}
    elif with_e2e:
        dist.barrier(); dist.barrier()
        z = torch.zeros(1, dtype=torch.float64, device=dev); dist.all_reduce(z, op=dist.ReduceOp.MAX)
        z = torch.zeros(1, dtype=torch.int64, device=dev); dist.all_reduce(z)
    del graph, x_local, s
    torch.cuda.empty_cache()
    return out


def _device_ms(fn, reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def _single_gpu_reference(s, x, n, f, k, dev):
    """K hops of the whole graph on one GPU (rank 0), device-timed; for the strong-scaling ratio."""
    from .graph import normalize_csr
    from .propagate import propagate
    from .features import to_padded
    g = normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    xp = to_padded(x)
    for _ in range(3):
        propagate(g, xp, k)
    torch.cuda.synchronize()
    ms = _device_ms(lambda: propagate(g, xp, k), 5)
    out = {"n_gpus": 1, "ms_per_step": ms, "value": k * g.nnz * f / (ms * 1e-3)}
    del g, xp
    torch.cuda.empty_cache()
    return out


def bench_lambda():
    """SGC_BENCH_LAMBDA: fraction of sampled edges forced into the source's own 1/8 block of nodes
    (synth.powerlaw_graph `lam`; 0 = uniformly random neighbours, the default)."""
    return float(os.environ.get("SGC_BENCH_LAMBDA", "0"))


def _workload_inputs(name, rank, world, dev, want_single):
    """This rank's rows of the named synthetic workload: raw CSR slice (global column ids) and the
    feature rows the hops start from (Reddit shape: standardised as utils.py:119 does, with the
    statistics of ALL rows).  Rank 0 also times the whole workload on its one GPU."""
    from . import synth
    from .features import col_stats, standardize_
    shape = synth.SHAPES[name]
    n, f, k = shape["n"], shape["f"], shape["k"]
    n_pad, parts = partition_rows(n, world)
    row0, row1 = parts[rank]
    lam = bench_lambda()
    single = None
    if shape["symmetric"]:
        s = synth.make_shape(name, device=dev, with_features=False, lam=lam)     # every rank builds the graph, keeps its rows
        rp, col, val = local_rows_of((s["row_ptr"], s["col"], s["val"]), row0, row1)
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
            single = _single_gpu_reference(s, x, n, f, k, dev)
        del s, x
    else:
        s = synth.make_shape(name, device=dev, with_features=False, row_range=(row0, row1), lam=lam)
        rp, col, val = s["row_ptr"], s["col"], s["val"]
        x_local = synth.gaussian_features(n, f, seed=1, scaled=False, row_range=(row0, row1), device=dev)
        del s
    torch.cuda.empty_cache()
    return n, f, k, n_pad, rp, col, val, x_local, single


def run_sharded_workload(name, args, rank, world, dev, B, with_e2e=True, want_single=True):
    """Times K-hop sharded propagation of one named workload on `world` GPUs (strong scaling: the
    graph is fixed; rows or feature columns are split over the ranks, see choose_sharding).
    Returns the measurements on rank 0."""
    from . import _ffi, synth
    from .propagate import default_chunk
    shape = synth.SHAPES[name]
    approx_nnz = shape["pairs"] * (2 if shape["symmetric"] else 1) + shape["n"]
    if choose_sharding(shape["n"], approx_nnz, shape["f"], world) == "features":
        return run_feature_sharded_workload(name, args, rank, world, dev, B, with_e2e, want_single)
    n, f, k, n_pad, rp, col, val, x_local, single = _workload_inputs(name, rank, world, dev, want_single)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sg = shard_normalize(n, rp, col, val)
    torch.cuda.synchronize()
    normalise_ms = (time.perf_counter() - t0) * 1e3
    nnz_local = torch.tensor([sg.graph.nnz], dtype=torch.int64, device=dev)
    dist.all_reduce(nnz_local)
    nnz = int(nnz_local.item())
    prop = ShardedPropagator(sg, f)
    mode = prop.mode
    for blk in prop.blocks:
        if blk is not None:
            blk.plan(default_chunk())

    for _ in range(args.warmup):
        prop.load(x_local)
        prop.propagate_resident(k)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = B["ClockSampler"](dev.index)
    sampler.start()
    launches0 = _ffi.launch_count()
    total = 0.0
    for _ in range(args.steps):
        prop.load(x_local)         # untimed: X resident on every rank, the reference's starting state
        torch.cuda.synchronize()
        dist.barrier()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        prop.propagate_resident(k)
        stop.record()
        torch.cuda.synchronize()
        total += start.elapsed_time(stop)
    dist.barrier()
    ms = torch.tensor([total], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)            # max over ranks, device-timed
    launches = _ffi.launch_count() - launches0
    clocks = sampler.stop()
    ms_per_step = float(ms.item()) / args.steps
    out = {"name": name, "n": n, "f": f, "k": k, "nnz": nnz, "n_pad": n_pad, "mode": mode,
           "push_fraction": getattr(prop, "push_fraction", 1.0),
           "n_blocks": prop.n_blocks, "normalise_ms": normalise_ms, "ms_per_step": ms_per_step,
           "value": k * nnz * f / (ms_per_step * 1e-3), "launches": int(launches), "clocks": clocks,
           "single": single, "e2e": None}
    if with_e2e:
        # end to end: this rank's shard from pinned host memory and back
        host = {"rp": rp.cpu().pin_memory(), "col": col.cpu().pin_memory(), "val": val.cpu().pin_memory(),
                "x": x_local.cpu().pin_memory()}
        out_host = torch.empty((sg.n_local, f), dtype=torch.float32).pin_memory()

        def e2e_step():
            sgl = shard_normalize(n, host["rp"].to(dev, non_blocking=True), host["col"].to(dev, non_blocking=True),
                                  host["val"].to(dev, non_blocking=True))
            xl = host["x"].to(dev, non_blocking=True)
            if mode == "push":
                yl = prop.rebind(sgl).propagate(xl, k)      # symmetric buffers + peer mappings are reused
            else:
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
        out["e2e"] = {"value": k * nnz * f / float(e2e_s.item()), "unit": B["UNIT"],
                      "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(n * f * 4),
                      "ms_per_step": float(e2e_s.item()) * 1e3}
        del host, out_host
    prop.close()
    del prop, sg
    torch.cuda.empty_cache()
    return out


def _parallelism_text(r, world):
    if r["mode"] == "features":
        return ("feature-sharded x%d: S replicated on every GPU, X split into contiguous column blocks of <= %d "
                "columns; column c of S^K X depends on column c of X only, so the K hops (and the per-column "
                "standardisation) need no exchange; output column-sharded" % (world, r.get("cols_per_rank", 0)))
    if r["mode"] == "push":
        return ("row-sharded x%d; exchange fused into the SpMM kernel: every output row is staged in shared memory "
                "and written by cp.async.bulk to all peers' next-hop operands over NVLink (CUDA-IPC mapped), one "
                "1-element all-reduce as barrier per hop; a row goes only to the peers whose shard gathers it "
                "(%.1f%% of rows x peers here); X replicated at start, output sharded"
                % (world, 100.0 * r["push_fraction"]))
    return ("row-sharded x%d; per-hop NVLink exchange (NCCL send/recv groups) pipelined with the SpMM over %d row "
            "blocks; X replicated at start, output sharded" % (world, r["n_blocks"]))


def bench(args, rank, world, dev, B):
    """bench.py --gpus N (N > 1): the SAME workload as N = 1 (Reddit-shaped, BASELINE configs[2]) row-sharded
    over the N GPUs, so that value(N) / value(1) is a strong-scaling ratio; by default the line also
    carries the products-shaped workload (BASELINE configs[3], the config north_star names for the
    1/2/4/8-GPU series) with its own single-GPU reference (SGC_BENCH_EXTRA="" to skip).
    B: bench.py's module globals (helpers)."""
    import json
    name = args.workload
    r = run_sharded_workload(name, args, rank, world, dev, B,
                             with_e2e=(name != "papers100m" or os.environ.get("SGC_BENCH_PAPERS_E2E") == "1"))
    extras = {}
    extra_names = os.environ.get("SGC_BENCH_EXTRA", "products" if not args.workload_given else "")
    for other in [e for e in extra_names.split(",") if e and e != name]:
        try:
            x = run_sharded_workload(other, args, rank, world, dev, B, with_e2e=False)
            extras[other] = {"workload": B["workload_config"](other, x["n"], x["nnz"], x["f"], x["k"])["workload"],
                             "n_gpus": world, "value": x["value"], "unit": B["UNIT"], "ms_per_step": x["ms_per_step"],
                             "nnz_S": x["nnz"], "dist_mode": x["mode"], "single_gpu_same_workload": x["single"],
                             "strong_scaling_efficiency": (x["value"] / (world * x["single"]["value"])
                                                           if x["single"] else None)}
        except Exception as exc:            # an optional extra must never cost the primary line
            extras[other] = {"error": repr(exc)[:300]}
    if rank == 0:
        n, f, k, nnz = r["n"], r["f"], r["k"], r["nnz"]
        peak, peak_src = B["peaks"]()
        b_hop = B["hop_bytes"](n, nnz, f)
        achieved = k * b_hop / world / (r["ms_per_step"] * 1e-3) / 1e9      # per-GPU algorithmic GB/s
        line = {
            "metric": B["METRIC"], "value": r["value"], "unit": B["UNIT"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": B["workload_config"](name, n, nnz, f, k, parallelism=_parallelism_text(r, world),
                                           dist_mode=r["mode"], normalise_ms=r["normalise_ms"],
                                           partition_locality_lambda=bench_lambda(),
                                           pushed_rows_x_peers_fraction=r["push_fraction"],
                                           rows_per_rank=r["n_pad"]),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "spmm_hop_kernel (per GPU; includes the fused NVLink pushes and the barrier)"},
            "cpu_baseline": None, "e2e": r["e2e"], "gpu_launches": r["launches"], "clocks": r["clocks"],
            # the same workload on ONE GPU of this box, measured by rank 0 in this run, so that the
            # strong-scaling efficiency value / (n_gpus * single.value) can be read off this line
            "single_gpu_same_workload": r["single"],
            "other_workloads": extras,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
