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
    # `out` is always a row range of a full-height padded operand (ld = pad_ld(f), pad columns zero)
    hop(graph, x, out=out, peers=peers, peer_mask=peer_mask, pad_owned=True, pitch_owned=bool(peers))


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


def _release_symm(local_ptr, peer_ptrs, state):
    """Finaliser of last resort (interpreter exit / a rank that died): unmaps and frees without a rendezvous.
    The orderly path is SymmetricBuffer.close()."""
    if state["closed"]:
        return
    state["closed"] = True
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
        self.device = torch.device(device)
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
        self._local, self._opened, self._state = int(ptr.value), opened, {"closed": False}
        self._fin = weakref.finalize(self, _release_symm, self._local, opened, self._state)

    def row_ptrs(self, row, peers_only):
        """Device address of row `row` in every rank's copy (optionally without this rank's)."""
        off = int(row) * self.ld * 4
        return [p + off for r, p in enumerate(self.ptrs) if not (peers_only and r == self.rank)]

    def close(self):
        """Collective: every rank unmaps its peers' buffers, the ranks meet, and only then is the local
        allocation freed -- cudaFree of exported memory while another process still has it mapped
        (cudaIpcOpenMemHandle without the matching close) is undefined behaviour."""
        if self._state["closed"]:
            return
        self.tensor = None
        lib = _ffi.lib()
        try:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize(self.device)           # this buffer's GPU, whatever the current device is
                while self._opened:                           # emptied as it goes: the finaliser shares this list
                    lib.sgc_ipc_close(self._opened.pop())
                dist.barrier(group=self.group)
        finally:
            # also when the rendezvous failed (a peer died or timed out): the local allocation is still freed,
            # and only now does the finaliser of last resort have nothing left to do
            with torch.cuda.device(self.device):
                lib.sgc_symm_free(self._local)
            self._state["closed"] = True


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

    def _push_hop(self, k, degree):
        """Hop k of `degree` in push mode: gathers from buf[k % 2], writes this rank's rows into its own
        segment of buf[(k+1) % 2] and (unless it is the last hop) into every peer's, then the barrier."""
        sg = self.sg
        src, dst = self.buf[k % 2], self.symm[(k + 1) % 2]
        last = (k == degree - 1)
        if sg.n_local > 0:
            peers = None if (last or sg.world == 1) else dst.row_ptrs(sg.row0, peers_only=True)
            self.hop_fn(sg.graph, src[:, :self.f], dst.tensor[sg.row0:sg.row1, :self.f], peers,
                        None if peers is None else self.peer_mask)
        # after the barrier: all rows of hop k are in place everywhere, and nobody still reads
        # the buffer the next hop (or the next load) overwrites
        self._barrier()

    def _push_propagate(self, degree):
        for k in range(degree):
            self._push_hop(k, degree)
        return self.buf[degree % 2][self.sg.row0:self.sg.row1, :self.f]

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
            # P2POp takes the peer's GLOBAL rank; q counts inside the group (a row group of a grid)
            peer = q if self.group is None else dist.get_global_rank(self.group, q)
            if b > a:
                ops.append(dist.P2POp(dist.isend, dst[sg.row0 + a:sg.row0 + b], peer, self.group))
            qa, qb = self.block_rows(q, r)
            if qb > qa:
                base = q * sg.n_pad
                ops.append(dist.P2POp(dist.irecv, dst[base + qa:base + qb], peer, self.group))
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
# Result checks of the multi-GPU paths (bench.py --gpus N runs them after the timed steps and fails on a
# mismatch; tests/test_gpu_dist.py drives them too).  The contract is the reference's: "same numbers as one
# process" (utils.py:92-97 has no notion of ranks) -- here bit for bit.
def block_checksum(t, row0, col0, f_total, chunk_elems=1 << 23):
    """Position-sensitive 64-bit checksum of the BITS of an fp32 block [r, c] whose element (i, j) is element
    (row0 + i, col0 + j) of a logical [*, f_total] matrix: sum over elements of bits * (hash(global index) | 1)
    in wrap-around int64.  Two ranks (or one GPU and many) hold the same numbers iff the sums agree, whatever
    the partition.  Device-side, no host copy."""
    from . import synth
    r, c = int(t.shape[0]), int(t.shape[1])
    dev = t.device
    total = torch.zeros((), dtype=torch.int64, device=dev)
    if r == 0 or c == 0:
        return total
    rows_per = max(1, chunk_elems // c)
    cols = torch.arange(c, dtype=torch.int64, device=dev) + int(col0)
    for a in range(0, r, rows_per):
        b = min(a + rows_per, r)
        bits = t[a:b].contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        idx = (torch.arange(a, b, dtype=torch.int64, device=dev) + int(row0))[:, None] * int(f_total) + cols[None, :]
        total = total + (bits * (synth.mix64(idx) | 1)).sum()
    return total


def verify_against_single(y_local, y_single, row0, col0, f_total, group=None):
    """Collective.  Every rank's output block [rows row0.., columns col0..] against the same block of the
    single-GPU result `y_single` (held by rank 0 only; None elsewhere), by block_checksum.  Returns
    {"ok", "per_rank", "checksum"} on every rank."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = y_local.device
    r, c = int(y_local.shape[0]), int(y_local.shape[1])
    mine = torch.stack([torch.tensor(v, dtype=torch.int64, device=dev) for v in (row0, r, col0, c)]
                       + [block_checksum(y_local, row0, col0, f_total)])
    allv = torch.empty((world * 5, 1), dtype=torch.int64, device=dev)
    _all_gather(allv, mine.view(5, 1), group, False)
    allv = allv.view(world, 5).cpu().tolist()
    ok = torch.ones(world, dtype=torch.int64, device=dev)
    if rank == 0:
        for q, (r0, rr, c0, cc, got) in enumerate(allv):
            want = int(block_checksum(y_single[r0:r0 + rr, c0:c0 + cc], r0, c0, f_total).item())
            ok[q] = 1 if want == got else 0
    dist.broadcast(ok, 0, group=group)
    per_rank = [bool(v) for v in ok.cpu().tolist()]
    total = 0
    for row in allv:
        total = (total + row[4]) & 0xFFFFFFFFFFFFFFFF
    return {"ok": all(per_rank), "per_rank": per_rank, "checksum": "%016x" % total,
            "what": "every rank's output block vs the same block of the single-GPU result (rank 0), bit-level "
                    "position-sensitive 64-bit checksum of all %d elements" % sum(row[1] * row[3] for row in allv)}


def _bits_differ(a, b):
    return (a.contiguous().view(torch.int32) != b.contiguous().view(torch.int32)).any(dim=1)


def _replication_mismatches(prop, buf, masked, rows_per_rank, group):
    """Collective.  Sampled rows of every rank's segment of the full-height operand `buf`: the owner's own
    copy is the truth (all-gathered), every other rank compares its copy bit for bit -- only where its shard
    gathers that row when `masked` (halo masks: other rows are never pushed to it).  Returns the global
    number of differing (row, rank) pairs and of compared pairs."""
    sg = prop.sg
    dev = buf.device
    m = max(1, min(int(rows_per_rank), min(b - a for a, b in prop.parts)))
    gens = []
    for q in range(sg.world):
        g = torch.Generator(device="cpu")
        g.manual_seed(1000 + q)
        nq = prop.parts[q][1] - prop.parts[q][0]
        gens.append(torch.sort(torch.randperm(nq, generator=g)[:m])[0].to(dev) + q * sg.n_pad)
    truth = torch.empty((sg.world * m, buf.shape[1]), dtype=torch.float32, device=dev)
    _all_gather(truth, buf[gens[sg.rank]].contiguous(), group, False)
    touch = None
    if masked:
        touch = torch.zeros(sg.world * sg.n_pad, dtype=torch.bool, device=dev)
        if sg.graph.nnz:
            touch[sg.graph.col.long()] = True
    bad = torch.zeros(2, dtype=torch.int64, device=dev)
    for q in range(sg.world):
        if q == sg.rank:
            continue
        diff = _bits_differ(buf[gens[q]], truth[q * m:(q + 1) * m])
        need = touch[gens[q]] if masked else torch.ones_like(diff)
        bad[0] += (diff & need).sum()
        bad[1] += need.sum()
    dist.all_reduce(bad, group=group)
    return int(bad[0].item()), int(bad[1].item())


def verify_sharded(prop, x_local, degree, B, y_single=None, rows_per_rank=256, group=None):
    """Collective check of the row-sharded (push) propagation, hop by hop, on the live buffers:
      * after the load and after every exchanging hop: sampled rows of every rank's segment are bit-identical
        on the owner and on every rank that gathers them (the NVLink pushes landed, complete and in place);
      * every hop: sampled output rows of this rank recomputed by the reference's torch.spmm on the CPU from
        this rank's full-height operand (B["sampled_rows_check"]), north_star tolerance;
      * with y_single (rank 0's single-GPU result of the same workload): every rank's final rows against it,
        bit for bit (block_checksum).
    B: bench.py's globals (the CPU checker lives there, not in this package).  Returns the report (same on
    every rank); "ok" False on any mismatch."""
    sg, f = prop.sg, prop.f
    dev = sg.graph.device
    assert prop.mode == "push", "verify_sharded drives the fused (push) path"
    ref = B["Reference"]()
    rep = {"hops": [], "rows_per_rank": rows_per_rank}
    prop.load(x_local)
    torch.cuda.synchronize(dev)
    bad, cmp_ = _replication_mismatches(prop, prop.buf[0], False, rows_per_rank, group)
    rep["load"] = {"replicated_rows_compared": cmp_, "mismatches": bad}
    ok = bad == 0
    rows = B["sample_rows"](sg.n_local, rows_per_rank, dev, seed=11 + sg.rank) if sg.n_local else None
    for k in range(degree):
        prop._push_hop(k, degree)
        torch.cuda.synchronize(dev)
        src, out = prop.buf[k % 2][:, :f], prop.buf[(k + 1) % 2][sg.row0:sg.row1, :f]
        viol = torch.zeros(3, dtype=torch.int64, device=dev)
        maxd = torch.zeros(1, dtype=torch.float64, device=dev)
        if rows is not None:
            r = B["sampled_rows_check"](ref, sg.graph, src, out, rows)
            viol += torch.tensor([r["violations"], r["elements"], r["bit_identical_elements"]], device=dev)
            maxd[0] = r["max_abs_diff"]
        dist.all_reduce(viol, group=group)
        dist.all_reduce(maxd, op=dist.ReduceOp.MAX, group=group)
        hop_rep = {"hop": k + 1, "vs_reference_spmm": {"elements": int(viol[1]), "violations": int(viol[0]),
                                                       "bit_identical_elements": int(viol[2]),
                                                       "max_abs_diff": float(maxd[0])}}
        ok = ok and int(viol[0]) == 0
        if k < degree - 1 and sg.world > 1:
            bad, cmp_ = _replication_mismatches(prop, prop.buf[(k + 1) % 2], prop.peer_mask is not None,
                                                rows_per_rank, group)
            hop_rep["pushed_rows"] = {"compared": cmp_, "mismatches": bad}
            ok = ok and bad == 0
        rep["hops"].append(hop_rep)
    y_local = prop.buf[degree % 2][sg.row0:sg.row1, :f]
    single_flag = torch.tensor([1 if y_single is not None else 0], dtype=torch.int64, device=dev)
    dist.all_reduce(single_flag, group=group)
    if int(single_flag.item()) > 0:
        rep["vs_single_gpu"] = verify_against_single(y_local, y_single, sg.row0, 0, f, group)
        ok = ok and rep["vs_single_gpu"]["ok"]
    rep["ok"] = bool(ok)
    rep["checker"] = ("the reference's torch.spmm on CPU (%s) on %d sampled rows per rank and hop; owner-vs-copy "
                      "bit compare of sampled pushed rows; tolerance %s" % (ref.kind, rows_per_rank, B["TOLERANCE"]))
    return rep


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


# --------------------------------------------------------------------------------------------------
# 2-D ("grid") sharding: P_c column blocks x P_r row parts, world = P_c * P_r.  The P_r ranks that share a
# column block form a ROW GROUP and run the row-sharded propagation above on that block (exchange fused into
# the hop, barrier inside the group only); different column blocks never talk to each other.  (P_c, P_r) =
# (world, 1) is feature sharding, (1, world) is row sharding.  Why both axes: splitting columns costs lane
# efficiency once a block is narrower than a few 128-float slabs (an 80-column block costs 0.85 ms where 1/8 of
# 602 columns would cost 0.62 ms, profiles/r01_sweep_block.log), splitting rows costs NVLink egress
# (rows/P_r x (P_r - 1) row pushes per exchanging hop) that hides behind the gathers only while it is shorter
# than them.  Reddit shape on 8 GPUs: 2 x 4.
GRID_AUTO = False          # the cost model below picks the grid (else: the degenerate grid choose_sharding names)
# Constants fitted to tools/sweep_rowpart.py on one B200 (profiles/r02_sweep_rowpart.log: the hop on a column block
# and a row part of the Reddit shape, plain and peer variants) and to the 8-GPU exchange probes:
SLAB_PASS_MS_PER_ENTRY = 3.5e-8           # one 128-float slab pass over all stored entries (Reddit shape: 0.82 ms)
HOP_FIXED_MS = 0.07                       # launches (main + narrow remainder + fix-ups) and ramp of one hop
PEER_VARIANT_FACTOR = 1.15                # peer variant vs plain kernel, pushes skipped
PUSH_MS_PER_BYTE = 0.33e-9                # issuing the pushes + the HBM writes they cause at the receivers
NVLINK_PUSH_BYTES_PER_MS = 5.5e8          # egress per GPU inside the fused hop (profiles/r02_nvlink_probe_n8_products.log)
ROWS_MUST_WIN_BY = 0.08                   # an exchanging grid is taken only when the model prefers it by this margin


def _block_cost_slabs(w):
    """Cost of one hop over a w-column block in units of a full 128-float slab pass: full slabs, plus the
    remainder -- at most 64 columns run as their own launch with narrower lane groups at ~0.7 of a full pass
    (304 columns = 2.7, 160 = 1.67, 80 = 0.92 passes measured), wider ones as one more (partly idle) slab."""
    full, r = divmod(int(w), 128)
    rem = 0.0 if r == 0 else (0.7 if r <= 64 else (0.92 if r <= 96 else 1.0))
    return full + rem


def grid_cost_ms(n, nnz, f, k, p_cols, p_rows):
    """Modelled time of K hops on a p_cols x p_rows grid (ranks the grids; not a prediction)."""
    w = max(c1 - c0 for c0, c1 in partition_features(f, p_cols))
    compute = SLAB_PASS_MS_PER_ENTRY * nnz * _block_cost_slabs(w) / p_rows + HOP_FIXED_MS
    if p_rows == 1:
        return k * compute
    rows = (n + p_rows - 1) // p_rows
    pushed = rows * _pad8(w) * 4.0 * (p_rows - 1)
    fused = max(PEER_VARIANT_FACTOR * compute + PUSH_MS_PER_BYTE * pushed, 1.1 * pushed / NVLINK_PUSH_BYTES_PER_MS)
    barrier = 0.04
    return (k - 1) * (fused + barrier) + compute + barrier


def choose_grid(n, nnz, f, k, world, mem_bytes=170e9):
    """(p_cols, p_rows) for `world` GPUs.  SGC_B200_GRID=CxR forces a grid, SGC_B200_SHARDING=rows|features the
    two degenerate ones.  Column splits are considered only where `choose_sharding` would allow them (every
    block at least 64 columns) and a 128-float slab of the gathered operand stays L2-resident (the regime the
    cost model was measured in); operands far larger than L2 (products, papers100M) take pure row sharding."""
    from .graph import L2_SLAB_BUDGET
    forced = os.environ.get("SGC_B200_GRID")
    if forced:
        pc, pr = (int(v) for v in forced.lower().split("x"))
        if pc * pr != world or pc < 1 or pr < 1:
            raise ValueError("SGC_B200_GRID=%s does not describe %d GPUs" % (forced, world))
        return pc, pr
    forced = os.environ.get("SGC_B200_SHARDING")
    if forced == "rows":
        return 1, world
    if forced == "features":
        return world, 1
    if not GRID_AUTO and os.environ.get("SGC_B200_GRID_AUTO", "0") != "1":
        return (world, 1) if choose_sharding(n, nnz, f, world, mem_bytes) == "features" else (1, world)
    best = None
    for pc in range(1, world + 1):
        if world % pc:
            continue
        pr = world // pc
        w = max(c1 - c0 for c0, c1 in partition_features(f, pc))
        if pc > 1 and (f // pc < 64 or n * 128 * 4.0 > L2_SLAB_BUDGET):
            continue
        if nnz / pr * 20.0 + n * 8.0 + 3.0 * n * _pad8(w) * 4.0 >= mem_bytes:
            continue
        cost = grid_cost_ms(n, nnz, f, k, pc, pr)
        if pr > 1:
            cost *= 1.0 + ROWS_MUST_WIN_BY      # feature sharding needs no exchange, no peer mappings, no barrier
        if best is None or cost < best[0] - 1e-9:
            best = (cost, pc, pr)
    if best is None:
        return 1, world
    return best[1], best[2]


class Grid(object):
    """This rank's place in a p_cols x p_rows grid: ranks [ci*p_rows, (ci+1)*p_rows) share column block ci."""

    def __init__(self, rank, world, p_cols, p_rows, row_group, cols):
        self.rank, self.world, self.p_cols, self.p_rows = rank, world, p_cols, p_rows
        self.ci, self.ri = rank // p_rows, rank % p_rows
        self.row_group, self.cols = row_group, cols


def make_grid(f, p_cols, p_rows):
    """Collective over the default group: builds the row groups (every rank must take part in the creation
    of every group, torch.distributed.new_group) and returns this rank's Grid."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if p_cols * p_rows != world:
        raise ValueError("grid %dx%d does not describe %d ranks" % (p_cols, p_rows, world))
    mine = None
    for ci in range(p_cols):
        ranks = list(range(ci * p_rows, (ci + 1) * p_rows))
        g = dist.group.WORLD if p_cols == 1 else dist.new_group(ranks)
        if rank in ranks:
            mine = g
    return Grid(rank, world, p_cols, p_rows, mine, partition_features(f, p_cols)[rank // p_rows])


def propagate_feature_sharded(graph, x_cols, degree, **kw):
    """K hops on this rank's column block [N, fc] of X with the full S: no communication.  Returns
    the same columns of S^K X, bit-identical to those columns of the single-GPU result."""
    from .propagate import propagate
    return propagate(graph, x_cols, degree, **kw)
