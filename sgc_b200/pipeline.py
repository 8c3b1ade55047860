"""Host-to-host precompute: the whole path of the reference's loaders + `sgc_precompute` for data
that lives in host memory, with the PCIe copies overlapped with the GPU work.

What it replaces, in reference order (paths relative to /root/reference):
    utils.py:118-119   feature standardisation        (optional, `standardize=True`)
    utils.py:120-122   aug_normalized_adjacency + sparse_mx_to_torch_sparse_tensor
    utils.py:126-130   .cuda() of adjacency and features            (H2D)
    utils.py:92-97     sgc_precompute                               (K hops)
    reddit.py:71 / citation.py: the propagated features going back to the host consumer (D2H)

Two schedules, both bit-identical to `sgc_precompute` on device-resident inputs
(tests/test_gpu_parity.py); `in`/`out` are copy streams (one DMA engine per direction), `main`
carries every kernel.  The adjacency always goes up first, so its normalisation (K1, K2, fp32 S
values, split plans) runs while the features are still crossing PCIe.

(1) Column-slab pipeline -- feature matrices wider than one L2-sized slab (Reddit shape: 602
    columns = 4 x 128 + 90).  Column c of S^K X depends on column c of X only, and the
    standardisation of utils.py:119 is per column, so every slab is an independent job:

    in:    H2D CSR(A) | H2D X[:,s0] | H2D X[:,s1] | H2D X[:,s2] | ...
    main:             | K1 K2 S plan | K3+std+K hops s0 | K3+std+K hops s1 | ...
    out:                                               | D2H Y[:,s0]      | D2H Y[:,s1] | ...

    The slab copies are strided 2-D DMA transfers (sgc_copy_2d) between the caller's row-major
    pinned [N, F] matrix and padded device rows.  The slab width is the one the hop kernel tiles
    with anyway (one slab of the gathered operand stays resident in L2).

(2) Row-block pipeline -- narrow feature matrices (one slab): uploads as above, then the LAST hop is
    launched per block of output rows (row-slice views of the same CSR arrays, the same kernel) and
    block b is copied back while block b+1 is computed.
"""
import contextlib
import os
import time

import torch

from . import _ffi
from .features import col_stats, padded_rows, standardize_
from .graph import L2_SLAB_BUDGET, SgcGraph, normalize_csr, require_cuda
from .propagate import default_chunk, hop

_copy_streams = {}


def _copy_streams_for(dev):
    s = _copy_streams.get(dev.index)
    if s is None:
        s = _copy_streams[dev.index] = (torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev))
    return s


def pinned_rows(n, f, align_bytes=64):
    """Pinned fp32 [n, f] host matrix whose row pitch is a multiple of `align_bytes` (a column view of a wider
    pinned buffer): a staging layout for `precompute_host`'s `features` / `out`.  A column slab of a row-major
    matrix is a strided transfer of 512-byte pieces; with a 2408-byte pitch (602 floats) every piece starts and
    ends inside a 64-byte line, with 2432 bytes none does: H2D 47.9 -> 54.2 GB/s, D2H 46.1 -> 52.3 GB/s, both
    at once 34.6 -> 40.8 GB/s each way (profiles/r02_copy2d_probe2.log)."""
    per = max(1, align_bytes // 4)
    ld = (int(f) + per - 1) // per * per
    return torch.empty((int(n), ld), dtype=torch.float32).pin_memory()[:, :int(f)]


def default_out_blocks():
    """Row blocks of the last hop (SGC_B200_OUT_BLOCKS; 1 = one launch, then one D2H)."""
    return max(1, int(os.environ.get("SGC_B200_OUT_BLOCKS", "8")))


def row_blocks(graph, n_blocks):
    """Row-slice views [(r0, r1, SgcGraph)] of `graph` with roughly equal numbers of stored entries
    (not rows: power-law graphs), sharing its CSR arrays.  One small D2H read of the block
    boundaries."""
    n = graph.n_rows
    n_blocks = max(1, min(int(n_blocks), n))
    if n_blocks == 1:
        return [(0, n, graph)]
    targets = torch.arange(1, n_blocks, device=graph.device, dtype=torch.int64) * (graph.nnz // n_blocks)
    cuts = torch.searchsorted(graph.row_ptr, targets).clamp_(0, n).cpu().tolist()
    # 128-row alignment keeps each block a whole number of warps' worth of tasks
    bounds = [0] + [min(n, (c + 127) // 128 * 128) for c in cuts] + [n]
    out = []
    for r0, r1 in zip(bounds[:-1], bounds[1:]):
        if r1 <= r0:
            continue
        sub = SgcGraph(r1 - r0, graph.n_cols, graph.row_ptr[r0:r1 + 1], graph.col, graph.aval, graph.dinv,
                       sval=graph.sval, dinv_rows=None if graph.dinv_rows is None else graph.dinv_rows[r0:r1])
        out.append((r0, r1, sub))
    return out


def default_col_slab(n, f):
    """Feature-slab width (floats) of the column pipeline: the widest slab the hop kernel is built
    for whose N x width x 4-byte working set fits the L2 budget (spmm.cu: default_slab); 0 = the
    matrix is one slab (or no slab fits L2) -> row-block schedule.  SGC_B200_COL_SLAB overrides."""
    env = os.environ.get("SGC_B200_COL_SLAB")
    if env is not None:
        w = int(env)
        return w if 0 < w < f else 0
    for w in (256, 128, 64, 32):
        if n * w * 4 <= L2_SLAB_BUDGET:
            # at least two full slabs: with fewer the pipeline has nothing to overlap, and strided H2D
            # copies whose host pitch is only slightly larger than the slab (512-byte pieces at a
            # 520..600-byte pitch) run at 5-13 GB/s instead of 35-46 (tools/copy2d_probe.py,
            # profiles/r01_copy2d_probe.log) -- such blocks take the contiguous row-block schedule
            return w if 2 * w <= f else 0
    return 0


def slab_widths(f, width):
    """Column slabs of the pipeline, in processing order: full slabs of `width` floats, then the
    remainder.  (Cutting the last slab further to shorten the un-overlapped tail was measured and
    does not pay: narrow strided copies run at 2/3 of the PCIe rate, 24.9 vs 24.0 ms end to end.)"""
    if not width or width >= f:
        return [f]
    widths = [width] * (f // width)
    if f % width:
        widths.append(f % width)
    return widths


def _host_csr(adj):
    if isinstance(adj, tuple):
        return adj
    import numpy as np
    import scipy.sparse as sp
    m = adj.tocsr() if sp.issparse(adj) else sp.csr_matrix(adj)
    return (m.shape[0], torch.from_numpy(np.ascontiguousarray(m.indptr, dtype=np.int64)),
            torch.from_numpy(np.ascontiguousarray(m.indices, dtype=np.int32)),
            torch.from_numpy(np.ascontiguousarray(m.data, dtype=np.float64)))


def _shared_piece(nnz, group):
    """[lo, hi) of the stored entries this rank uploads, and the padded piece length."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = max(2, ((nnz + world - 1) // world + 1) // 2 * 2)
    return min(nnz, rank * per), min(nnz, (rank + 1) * per), per, world


def _upload_shared(host, dtype, dev, group):
    """One array of the shared adjacency: this rank's 1/P over its own PCIe link, then an NCCL
    all-gather over NVLink assembles the whole array on every GPU.  Returns (array, bytes uploaded)."""
    import torch.distributed as dist
    nnz = int(host.numel())
    lo, hi, per, world = _shared_piece(nnz, group)
    piece = torch.zeros(per, dtype=dtype, device=dev)
    if hi > lo:
        piece[:hi - lo].copy_(host[lo:hi], non_blocking=True)
    full = torch.empty(per * world, dtype=dtype, device=dev)
    dist.all_gather_into_tensor(full, piece, group=group)
    return full[:nnz], (hi - lo) * host.element_size()


def _uniform_value_shared(val_h, dev, group, stream=None):
    """`uniform_value` for an adjacency every rank of `group` holds in host memory: each rank compares
    its own 1/P of the values with the first one, and a one-word all-reduce (MIN) makes the verdict
    common -- every rank fills, or every rank uploads and all-gathers (a split decision would leave
    the all-gather short of participants).  Collective.  No time budget: the piece is small."""
    import math
    import torch.distributed as dist
    if os.environ.get("SGC_B200_UNIFORM_VALUES", "1") == "0":
        return None
    nnz = int(val_h.numel())
    if val_h.dtype != torch.float64 or not val_h.is_contiguous() or val_h.is_cuda or nnz == 0:
        return None
    first = float(val_h[0])
    if not math.isfinite(first):
        return None
    lo, hi, _, world = _shared_piece(nnz, group)
    threads = max(1, host_threads() // world)
    same = hi <= lo or _ffi.lib().sgc_host_all_equal_f64(val_h.data_ptr() + 8 * lo, hi - lo, first, threads, 0) == 1
    ctx = torch.cuda.stream(stream) if (stream is not None and dev.type == "cuda") else contextlib.nullcontext()
    with ctx:
        verdict = torch.tensor([1 if same else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(verdict, op=dist.ReduceOp.MIN, group=group)
        agreed = int(verdict.item()) == 1
    return first if agreed else None


def host_threads():
    """Workers of the host-side scan of the adjacency values (SGC_B200_HOST_THREADS)."""
    return max(1, int(os.environ.get("SGC_B200_HOST_THREADS", min(12, os.cpu_count() or 1))))


PCIE_BYTES_PER_US = 40e3        # ~40 GB/s: sustained pinned H2D rate of the strided / contiguous copies


def uniform_value(val_h, hidden_bytes=0, report=None):
    """The common value when every stored value of the host CSR holds the same bits, else None.  The
    reference's adjacency matrices are constant-valued: 1.0 for the networkx adjacency of the citation
    graphs (utils.py:62), 2.0 for Reddit (`adj + adj.T` of a symmetric 0/1 matrix, utils.py:116); the
    device array is then filled in place of an nnz x 8-byte upload.  Exact -- a threaded bit compare over
    the whole array (sgc_host_all_equal_f64), run while the DMA engine is busy with the `hidden_bytes`
    enqueued ahead of it.  The scan gets the time those copies take plus the time the upload it replaces
    would take; a host that cannot finish within that (it would lose time, not save it) gives up and the
    array is uploaded.  Nothing is remembered between calls: every call decides for itself, and a call
    that gave up says so (`report["scan_gave_up"]`, surfaced in precompute_host's `stats`) so that a
    caller on a slow host can pass detect_uniform_values=False.  SGC_B200_UNIFORM_VALUES=0 switches the
    scan off, =force removes the time budget."""
    if report is not None:
        report["scan_gave_up"] = False
    mode = os.environ.get("SGC_B200_UNIFORM_VALUES", "1")
    if mode == "0":
        return None
    if val_h.dtype != torch.float64 or not val_h.is_contiguous() or val_h.is_cuda or val_h.numel() == 0:
        return None
    first = float(val_h[0])
    if first != first or first in (float("inf"), float("-inf")):       # NaN payloads would not survive the fill
        return None
    nbytes = int(val_h.numel()) * 8
    budget_us = 0 if mode == "force" else max(500, int((hidden_bytes + nbytes) / PCIE_BYTES_PER_US))
    r = _ffi.lib().sgc_host_all_equal_f64(val_h.data_ptr(), int(val_h.numel()), first, host_threads(), budget_us)
    if r == 2 and report is not None:
        report["scan_gave_up"] = True
    return first if r == 1 else None


def precompute_host(adj, features, degree, standardize=False, out=None, out_blocks=None, chunk=None,
                    col_slab=None, trace=None, group=None, detect_uniform_values=None, stats=None):
    """`sgc_precompute(features, aug_normalized_adjacency(adj), degree)` for HOST inputs, result
    in host memory.

    adj       scipy sparse matrix, or a tuple (n, row_ptr, col, val) of host tensors holding the
              CSR of A (int64 / int32 / float64; pinned memory makes the copies asynchronous)
    features  host fp32 tensor [N, F] with unit column stride (pinned for asynchronous copies).  Rows may be
              padded: the strided slab copies run 13-18 % faster when the host row pitch is a multiple of 64 bytes
              (`pinned_rows`; profiles/r02_copy2d_probe2.log) -- the same holds for `out`
    degree    K >= 1
    standardize   apply utils.py:119 ((x - mean) / std per column, unbiased std) before the hops
    out       optional pinned fp32 [N, F] host tensor receiving the result
    out_blocks / col_slab   schedule knobs (defaults: SGC_B200_OUT_BLOCKS, default_col_slab)
    group     torch.distributed process group whose ranks all hold the same `adj` in host memory and each
              call with their own column block of the features (feature-sharded propagation): the
              adjacency is then uploaded once per node -- 1/P per rank -- and all-gathered over NVLink
    detect_uniform_values   scan the stored values on the host and, when they are all the same, fill the
              device array instead of uploading it (`uniform_value`; under `group` every rank scans its
              own 1/P and a one-word all-reduce makes the verdict common); default: on, except under
              `group`, where it measured neutral to slightly negative and is opt-in
    stats     optional dict; receives h2d_bytes / d2h_bytes actually copied by this call,
              uniform_value (the fill value, or None when the values were uploaded) and scan_gave_up
              (True when the host scan ran out of its time budget in THIS call; no state is kept)
    trace     optional list; receives (label, device ms since the start mark, host ms since the call) for
              the milestones of the schedule (CUDA events), for tools/e2e_sweep.py
    Returns `out` (complete when the call returns)."""
    t_call = time.perf_counter()
    dev = require_cuda()
    degree = int(degree)
    if degree < 1:
        raise ValueError("degree must be >= 1")
    n, rp_h, col_h, val_h = _host_csr(adj)
    n = int(n)
    if features.dim() != 2 or features.shape[0] != n or features.dtype != torch.float32 or features.is_cuda:
        raise ValueError("features must be a host float32 [N, F] tensor with N = %d rows" % n)
    if features.stride(1) != 1 or (n > 1 and features.stride(0) < features.shape[1]):
        features = features.contiguous()
    f = int(features.shape[1])
    fs = int(features.stride(0)) if n > 1 else f          # host row pitch of X in floats (pinned_rows: 64-byte multiples)
    if out is None:
        out = torch.empty((n, f), dtype=torch.float32).pin_memory()
    if tuple(out.shape) != (n, f) or out.dtype != torch.float32 or out.is_cuda or (f > 1 and out.stride(1) != 1) \
            or (n > 1 and out.stride(0) < f):
        raise ValueError("out must be a float32 [N, F] host tensor with unit column stride (rows may be padded)")
    os_ = int(out.stride(0)) if n > 1 else f              # host row pitch of the result
    chunk = default_chunk() if chunk is None else chunk
    if isinstance(col_slab, (list, tuple)):            # explicit slab widths, in processing order
        widths = [int(w) for w in col_slab]
        if sum(widths) != f or any(w <= 0 or w % 4 for w in widths[:-1]):
            raise ValueError("col_slab widths must be multiples of 4 floats summing to F")
    else:
        width = default_col_slab(n, f) if col_slab is None else (int(col_slab) if 0 < int(col_slab) < f else 0)
        if width % 4:
            raise ValueError("col_slab must be a multiple of 4 floats")
        widths = slab_widths(f, width)

    with torch.cuda.device(dev):
        main = torch.cuda.current_stream()
        c_in, c_out = _copy_streams_for(dev)
        marks = []

        def mark(label, stream):
            if trace is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record(stream)
                marks.append((label, e, (time.perf_counter() - t_call) * 1e3))

        x = padded_rows(n, f, dev)
        ld = x.stride(0)
        c_in.wait_stream(main)
        c_out.wait_stream(main)
        mark("start", main)
        starts = [sum(widths[:i]) for i in range(len(widths))]
        slabs = [(c0, c0 + w) for c0, w in zip(starts, widths)]
        by_cols = len(slabs) > 1
        with torch.cuda.stream(c_in):
            x_up = []

            def upload_slab(c0, c1):
                if by_cols:
                    _ffi.call("sgc_copy_2d", x.data_ptr() + 4 * c0, 4 * ld, features.data_ptr() + 4 * c0, 4 * fs,
                              4 * (c1 - c0), n, c_in.cuda_stream)
                elif fs == f:
                    x.copy_(features, non_blocking=True)       # whole rows: one contiguous transfer
                else:
                    _ffi.call("sgc_copy_2d", x.data_ptr(), 4 * x.stride(0), features.data_ptr(), 4 * fs, 4 * f, n,
                              c_in.cuda_stream)
                e = torch.cuda.Event()
                e.record(c_in)
                x_up.append(e)
                mark("x[:, %d:%d] up" % (c0, c1), c_in)

            nbytes = lambda t: t.numel() * t.element_size()
            h2d = nbytes(rp_h) + n * f * 4
            fill = None
            scan_report = {"scan_gave_up": False}
            # under `group` the scan is opt-in: measured neutral to slightly negative (N = 2: 16.3 vs 15.7 ms, N = 8:
            # 12.1 vs 11.2 ms) -- the piece each rank would upload is small, and the verdict costs a rendezvous
            detect = (group is None) if detect_uniform_values is None else bool(detect_uniform_values)
            rp = rp_h.to(dev, non_blocking=True)
            if group is not None:
                # all ranks hold the same adjacency in host memory (feature-sharded propagation): 1/P of
                # `col` / `val` per rank over its own PCIe link + an NCCL all-gather over NVLink, instead of
                # P times the bytes over PCIe (Reddit shape: 280 MB per GPU, 5 ms).  Collective.
                col, up = _upload_shared(col_h, torch.int32, dev, group)
                h2d += up
            else:
                col = col_h.to(dev, non_blocking=True)
                h2d += nbytes(col_h)
            # structure is up: when the values turn out to be constant the normalisation waits for this event
            # only, and runs while the first feature slab is still crossing PCIe
            csr_up = torch.cuda.Event()
            csr_up.record(c_in)
            if detect:
                ahead = nbytes(col_h)
                if by_cols:
                    upload_slab(*slabs[0])                 # keeps the DMA engine fed during the host scan
                    ahead += n * 4 * (slabs[0][1] - slabs[0][0])
                if group is not None:
                    # the verdict's all-reduce goes through the download stream (idle so far), so that it queues
                    # behind the all-gather of `col` in NCCL's order but not behind the slab upload above
                    fill = _uniform_value_shared(val_h, dev, group, c_out)
                else:
                    fill = uniform_value(val_h, ahead, scan_report)
                mark("values scanned", c_in)
            val = None
            if fill is None:
                if group is not None:
                    val, up = _upload_shared(val_h, torch.float64, dev, group)
                    h2d += up
                else:
                    val = val_h.to(dev, non_blocking=True)
                    h2d += nbytes(val_h)
                csr_up = torch.cuda.Event()
                csr_up.record(c_in)
            mark("csr up", c_in)
            for c0, c1 in slabs[len(x_up):]:
                upload_slab(c0, c1)
            if stats is not None:
                stats.update(h2d_bytes=int(h2d), d2h_bytes=int(n * f * 4), uniform_value=fill,
                             scan_gave_up=scan_report["scan_gave_up"])
        for t in (rp, col) + (() if val is None else (val,)):
            t.record_stream(main)
        x.record_stream(c_in)

        # adjacency side, while the features are still on their way
        main.wait_event(csr_up)
        if val is None:                                    # constant-valued adjacency: filled, not uploaded
            val = torch.full((val_h.numel(),), fill, dtype=torch.float64, device=dev)
        graph = normalize_csr(n, rp, col, val)
        graph.values32()
        if by_cols:
            graph.plan(chunk)
            blocks = None
        else:
            # row blocks pay when the gathered operand is L2-resident (Reddit-sized column blocks: 35 ->
            # 31 ms); when X is far larger than L2 every block launch re-streams it and re-tags the hub
            # columns (products shape: 89 ms with one launch, 96 ms with eight) -> one launch
            big = n * 32 * 4 > L2_SLAB_BUDGET
            blocks = row_blocks(graph, (1 if big else default_out_blocks()) if out_blocks is None else int(out_blocks))
            if degree > 1:
                graph.plan(chunk)
            for _, _, sub in blocks:
                sub.plan(chunk)
        mark("normalised", main)
        bufs = [padded_rows(n, f, dev) for _ in range(min(degree - 1, 2))]
        # column slabs land in padded rows (aligned vector stores; the strided D2H drops the padding);
        # whole-row blocks go back as contiguous pieces
        y = padded_rows(n, f, dev) if by_cols else torch.empty((n, f), dtype=torch.float32, device=dev)
        y.record_stream(c_out)

        def back(r0, r1, c0, c1):
            done = torch.cuda.Event()
            done.record(main)
            c_out.wait_event(done)
            mark("y[%d:%d, %d:%d] computed" % (r0, r1, c0, c1), main)
            if by_cols:
                _ffi.call("sgc_copy_2d", out.data_ptr() + 4 * (r0 * os_ + c0), 4 * os_,
                          y.data_ptr() + 4 * (r0 * ld + c0), 4 * ld, 4 * (c1 - c0), r1 - r0, c_out.cuda_stream)
            elif os_ == f:
                with torch.cuda.stream(c_out):
                    out[r0:r1].copy_(y[r0:r1], non_blocking=True)
            else:
                _ffi.call("sgc_copy_2d", out.data_ptr() + 4 * r0 * os_, 4 * os_, y.data_ptr() + 4 * r0 * y.stride(0),
                          4 * y.stride(0), 4 * f, r1 - r0, c_out.cuda_stream)
            mark("y[%d:%d, %d:%d] down" % (r0, r1, c0, c1), c_out)

        for (c0, c1), up in zip(slabs, x_up):
            main.wait_event(up)
            cur = x[:, c0:c1]
            if standardize:
                mean, std = col_stats(cur)
                standardize_(cur, mean, std)
            for k in range(degree - 1):
                nxt = bufs[k & 1][:, c0:c1]
                hop(graph, cur, chunk=chunk, out=nxt, pad_owned=True)
                cur = nxt
            if blocks is None:
                hop(graph, cur, chunk=chunk, out=y[:, c0:c1], pad_owned=by_cols and c1 == f)
                back(0, n, c0, c1)
            else:
                # last hop by row blocks, each block copied back while the next one is computed
                for r0, r1, sub in blocks:
                    hop(sub, cur, chunk=chunk, out=y[r0:r1, c0:c1])
                    back(r0, r1, c0, c1)
        main.wait_stream(c_out)
        c_out.synchronize()
        if trace is not None:
            main.synchronize()
            # (label, device ms since the start mark, host ms since the call when the mark was enqueued)
            trace.extend((label, marks[0][1].elapsed_time(e), h) for label, e, h in marks)
            trace.append(("returned", marks[0][1].elapsed_time(marks[-1][1]), (time.perf_counter() - t_call) * 1e3))
    return out
