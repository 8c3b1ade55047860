"""K-hop feature propagation  X <- S X  on the GPU (the reference's utils.sgc_precompute,
/root/reference/utils.py:92-97, whose inner call is torch.spmm).

Data layout in HBM: the caller's feature matrix is row-major fp32 ``[N, F]``.  Intermediate hops
are written to internal ping-pong buffers with a leading dimension padded to a multiple of 8
floats (32-byte sectors; e.g. 602 -> 608) so that every gather of a neighbour row is a run of whole
16-byte vectors; the last hop writes a contiguous ``[N, F]`` result, as the reference returns.
Only two such buffers exist at any time: no normalised adjacency and no standardised copy of the
features is materialised.
"""
import os
from time import perf_counter

import torch

from . import _ffi
from .graph import SgcGraph, from_torch_sparse, require_cuda, stream_ptr

_ENV = os.environ.get


def default_chunk():
    """Rows with more stored entries than this are split (SGC_B200_CHUNK; 0 = never split, i.e.
    bit-exact sequential accumulation for every row)."""
    return int(_ENV("SGC_B200_CHUNK", "512"))


def default_hints():
    """bit0 evict_last on gathered rows (when the slab fits L2), bit1 evict_first on the CSR
    stream and the output, bit2 use hot-column tags when X does not fit L2."""
    return int(_ENV("SGC_B200_HINTS", "7"))


def default_slab():
    return int(_ENV("SGC_B200_SLAB", "0"))


def default_std_mode():
    """How `sgc_precompute` applies the Reddit feature standardisation (utils.py:119) to a
    `StandardizedFeatures` input (SGC_B200_STD):
      'inplace' (default) one streaming pass standardises the feature storage in place, exactly,
                before the first hop (0.2 ms at the Reddit shape);
      'div' / 'rcp'       fused into the first hop's gather: exact division, or multiplication by
                1/std (<= 1 ulp per element).  Measured on B200 the gather touches every row of X
                ~100 times, so per-element work inside it costs 3-13 ms -- see DESIGN.md."""
    return _ENV("SGC_B200_STD", "inplace")


def fused_norm():
    """SGC_B200_FUSED_NORM=1: the hop forms fp32((d_i*a_ij)*d_j) on the fly from the fp64 values of
    A+I and D^-1/2, so S is never materialised (8 B + a D^-1/2 gather per stored entry and slab
    pass).  Default 0: the same fp32 values are written once by sgc_norm_values at normalisation time (4 B per entry,
    cached on the graph) and streamed by every hop -- bit-identical results, and measured 8-13 %
    faster per hop on the Reddit shape (4.65 ms vs 5.1-5.4 ms), see DESIGN.md."""
    return _ENV("SGC_B200_FUSED_NORM", "0") == "1"


def _pad8(f):          # historical name: the padded leading dimension, see features.pad_ld
    from .features import pad_ld
    return pad_ld(f)


def as_graph(adj):
    if isinstance(adj, SgcGraph):
        return adj
    if isinstance(adj, torch.Tensor) and adj.layout in (torch.sparse_coo, torch.sparse_csr):
        g = getattr(adj, "_sgc_graph", None)
        if g is None:
            g = from_torch_sparse(adj)
            try:
                adj._sgc_graph = g          # reuse the CSR + plan on the next call
            except Exception:
                pass
        return g
    raise TypeError("adj must be an SgcGraph or a CUDA torch sparse tensor, got %r" % type(adj))


def push_tile_tasks():
    """Rows of a peer tile per group slot (SGC_B200_PUSH_ROWS; 0 = the library default: 4 where a tile leaves as one bulk store, else 1): consecutive
    output rows that travel to a peer as ONE cp.async.bulk store when they are contiguous in the operand."""
    return int(_ENV("SGC_B200_PUSH_ROWS", "0"))


def _hop_args(g, plan, x, ldx, y, ldy, f, partial, mean, std, std_mode, slab, hints, peers=None, peer_mask=None,
              pad_owned=False, pitch_owned=False):
    a = _ffi.HopArgs()
    use_fused = g.fused and fused_norm()
    if g.fused and not use_fused:
        g.values32()                 # materialise S (fp32) once; cached on the graph
    if peers:
        if len(peers) > _ffi.MAX_PEERS:
            raise ValueError("at most %d peers" % _ffi.MAX_PEERS)
        a.n_peers = len(peers)
        for i, ptr in enumerate(peers):
            a.peer_y_d[i] = int(ptr)
        if peer_mask is not None:
            a.peer_mask_d = _ffi.ptr(peer_mask)
    a.push_tile_tasks = push_tile_tasks()
    flags = _ffi.HOP_PAD_OWNED if pad_owned else 0
    if pitch_owned:
        flags |= _ffi.HOP_PITCH_OWNED
    if use_fused:
        flags |= _ffi.HOP_FUSED_NORM
        a.aval_d, a.dinv_d, a.dinv_rows_d = _ffi.ptr(g.aval), _ffi.ptr(g.dinv), _ffi.ptr(g.dinv_rows)
    else:
        a.sval_d = _ffi.ptr(g.sval)
    if mean is not None:
        flags |= _ffi.HOP_STD_RCP if std_mode == "rcp" else _ffi.HOP_STD_DIV
        a.mean_d, a.std_d = _ffi.ptr(mean), _ffi.ptr(std)
    a.x_d, a.ldx, a.y_d, a.ldy = x, ldx, y, ldy
    if partial is not None:
        a.partial_d, a.ldp = _ffi.ptr(partial), partial.stride(0)
    a.f, a.slab_floats, a.flags, a.hints = f, slab, flags, hints
    return a


def _check_features(g, x):
    if not isinstance(x, torch.Tensor) or x.dim() != 2:
        raise TypeError("features must be a 2-D torch tensor")
    if x.device != g.device:
        raise ValueError("features are on %s but the graph is on %s" % (x.device, g.device))
    if x.shape[0] != g.n_cols:
        raise ValueError("size mismatch: adj is %s, features are %s" % (g.shape, tuple(x.shape)))
    if x.dtype != torch.float32:
        x = x.float()
    if x.stride(1) != 1 or x.stride(0) < x.shape[1]:
        x = x.contiguous()
    return x


def _partial(plan, f, dev):
    if plan.partial_rows == 0:
        return None
    return torch.empty((plan.partial_rows, (f + 3) // 4 * 4), dtype=torch.float32, device=dev)


def _check_out(g, out, f):
    if not isinstance(out, torch.Tensor) or out.dim() != 2 or tuple(out.shape) != (g.n_rows, f):
        raise ValueError("out must be a [%d, %d] tensor, got %r" % (g.n_rows, f, getattr(out, "shape", type(out))))
    if out.dtype != torch.float32 or out.device != g.device:
        raise ValueError("out must be a float32 tensor on %s (got %s on %s)" % (g.device, out.dtype, out.device))
    if out.stride(1) != 1 or (out.shape[0] > 1 and out.stride(0) < f):
        raise ValueError("out must have unit column stride and a row stride of at least F")


def col_blocks_for(g, row_bytes):
    """How many column blocks a hop over an operand with `row_bytes`-long rows should take (SGC_B200_COL_BLOCKS):
    0 / 1 = the plain single-pass hop (default); N >= 2 = that many; 'auto' = as many as make one block of the
    gathered operand fit the L2 budget, for operands between 2x and 64x that budget whose rows are one slab
    wide (products-like: N = 2.45 M rows of 416 bytes = 1 GB -> 10 blocks); wider operands take feature slabs,
    larger ones would re-read y too often."""
    from .graph import L2_SLAB_BUDGET
    env = _ENV("SGC_B200_COL_BLOCKS", "0")
    if env != "auto":
        return max(0, int(env))
    total = g.n_cols * int(row_bytes)
    if row_bytes > 128 * 4 or total < 2 * L2_SLAB_BUDGET or total > 64 * L2_SLAB_BUDGET:
        return 0
    return int((total + 0.8 * L2_SLAB_BUDGET - 1) // (0.8 * L2_SLAB_BUDGET))


def hop_col_blocked(g, x, n_blocks, chunk=None, hints=None, out=None, pad_owned=False):
    """One hop ``S @ x`` as `n_blocks` launches, one per block of S's columns (SgcGraph.col_blocks), each adding
    its block's contribution onto y (SGC_HOP_ACCUMULATE): the rows of x a launch gathers from are one
    contiguous block small enough to stay in L2, where the single-pass hop on an operand several times larger
    than L2 misses on most gathers (products shape: 45 GB of DRAM traffic per hop against 3 GB of compulsory
    bytes).  The price is that y is read and re-written once per block.  Every row's sum is the same ascending-
    column FMA chain as in the single-pass hop, continued from y: bit-identical with chunk = 0; with row
    splitting the chunk boundaries fall inside the blocks, so split rows agree within north_star's tolerance."""
    dev = require_cuda(g.device)
    x = _check_features(g, x)
    f = int(x.shape[1])
    if out is not None:
        _check_out(g, out, f)
    chunk = default_chunk() if chunk is None else chunk
    hints = default_hints() if hints is None else hints
    with torch.cuda.device(dev):
        y = out if out is not None else torch.empty((g.n_rows, f), dtype=torch.float32, device=dev)
        if f == 0 or g.n_rows == 0:
            return y
        for i, (c0, c1, blk) in enumerate(g.col_blocks(n_blocks)):
            plan = blk.plan(chunk)
            partial = _partial(plan, f, dev)
            a = _hop_args(blk, plan, x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0), f, partial, None, None,
                          "inplace", 0, hints & ~4, pad_owned=pad_owned)
            if i > 0:
                a.flags |= _ffi.HOP_ACCUMULATE
            _ffi.call("sgc_spmm_hop", plan.handle, _ffi.ctypes.byref(a), stream_ptr())
    return y


def hop(adj, x, mean=None, std=None, chunk=None, slab=None, hints=None, std_mode=None, out=None, peers=None,
        peer_mask=None, pad_owned=False, pitch_owned=False):
    """One hop ``S @ x`` (torch.spmm(adj, x), utils.py:95).  Returns a new [n_rows, F] fp32
    tensor.  With ``mean``/``std`` (fp32 [F]) the gathered operand is standardised on the fly.
    ``out``: optional [n_rows, F] fp32 destination (any row stride >= F); only its columns [0, F) are
    written, so it may be a column view of a wider live matrix -- unless ``pad_owned=True`` declares the
    columns up to the next multiple of 4 to be padding of this operand (features.padded_rows storage), which
    lets the kernel finish every row with whole 16-byte stores.
    ``peers``: device addresses of other GPUs' buffers (same row 0 and leading dimension as ``out``)
    that receive every output row from the kernel's epilogue (row-sharded propagation, dist.py);
    ``peer_mask`` (int32/uint32 [n_rows], optional): bit p set = row must go to peers[p].
    ``pitch_owned=True`` (implies ``pad_owned``) declares ALL columns from F to the row stride of ``out`` and of
    every peer buffer to be padding -- whole padded rows, not column views; only then are consecutive rows sent
    to the peers as one store per tile, which zero-fills those columns there (SGC_HOP_PITCH_OWNED)."""
    g = as_graph(adj)
    dev = require_cuda(g.device)
    x = _check_features(g, x)
    f = int(x.shape[1])
    if out is not None:
        _check_out(g, out, f)
    if mean is not None:
        mean, std = mean.to(dev, torch.float32).contiguous(), std.to(dev, torch.float32).contiguous()
    if mean is None and not peers and not slab:
        nb = col_blocks_for(g, _pad8(f) * 4)
        if nb >= 2:
            return hop_col_blocked(g, x, nb, chunk=chunk, hints=hints, out=out, pad_owned=pad_owned)
    with torch.cuda.device(dev):
        y = out if out is not None else torch.empty((g.n_rows, f), dtype=torch.float32, device=dev)
        if f == 0 or g.n_rows == 0:
            return y
        plan = g.plan(default_chunk() if chunk is None else chunk)
        plan.ensure_hot(g, x.stride(0) * 4)
        partial = _partial(plan, f, dev)
        a = _hop_args(g, plan, x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0), f, partial, mean, std,
                      std_mode or default_std_mode(), default_slab() if slab is None else slab,
                      default_hints() if hints is None else hints, peers, peer_mask, pad_owned, pitch_owned)
        _ffi.call("sgc_spmm_hop", plan.handle, _ffi.ctypes.byref(a), stream_ptr())
    return y


def propagate(adj, x, degree, mean=None, std=None, chunk=None, slab=None, hints=None, std_mode=None):
    """K hops, stream-ordered, no synchronisation; returns the [N, F] fp32 result tensor."""
    g = as_graph(adj)
    dev = require_cuda(g.device)
    degree = int(degree)
    if degree <= 0:
        return x
    if g.n_rows != g.n_cols:
        raise ValueError("K-hop propagation needs a square adjacency, got %s" % (g.shape,))
    from .sparse_features import hop_sparse_x, sparse_origin
    sf = sparse_origin(x) if mean is None else None
    x = _check_features(g, x)
    n, f = int(x.shape[0]), int(x.shape[1])
    if mean is not None:
        mean, std = mean.to(dev, torch.float32).contiguous(), std.to(dev, torch.float32).contiguous()
    if sf is not None and f * 4 <= 200 * 1024 and n > 0 and f > 0:
        # citation-style sparse features: the first hop accumulates the stored entries
        # of X only -- same per-element order as the dense kernel's exact mode
        with torch.cuda.device(dev):
            if degree == 1:
                y = torch.empty((n, f), dtype=torch.float32, device=dev)
                return hop_sparse_x(g, sf, out=y)
            first = hop_sparse_x(g, sf)
        return propagate(g, first, degree - 1, chunk=chunk, slab=slab, hints=hints)
    if mean is None and not slab and n > 0 and f > 0 and col_blocks_for(g, _pad8(f) * 4) >= 2:
        # column-blocked hops (operand several times larger than L2): hop by hop through `hop`
        with torch.cuda.device(dev):
            ld = _pad8(f)
            cur = x if (x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0) else None
            if cur is None:
                from .features import to_padded
                cur = to_padded(x)
            bufs = []
            for k in range(degree):
                last = k == degree - 1
                if last:
                    dst = torch.empty((n, f), dtype=torch.float32, device=dev)
                else:
                    if len(bufs) < 2:
                        b = torch.empty((n, ld), dtype=torch.float32, device=dev)
                        if ld != f:
                            b[:, f:].zero_()
                        bufs.append(b)
                    dst = bufs[k & 1][:, :f]
                hop(g, cur, chunk=chunk, hints=hints, out=dst, pad_owned=not last)
                cur = dst
            return cur
    with torch.cuda.device(dev):
        y = torch.empty((n, f), dtype=torch.float32, device=dev)
        if f == 0 or n == 0:
            return y
        plan = g.plan(default_chunk() if chunk is None else chunk)
        plan.ensure_hot(g, _pad8(f) * 4)
        partial = _partial(plan, f, dev)
        ld = _pad8(f)
        bufs = []
        for _ in range(min(degree - 1, 2)):
            b = torch.empty((n, ld), dtype=torch.float32, device=dev)
            if ld != f:
                b[:, f:].zero_()
            bufs.append(b)
        a = _hop_args(g, plan, x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0), f, partial, mean, std,
                      std_mode or default_std_mode(), default_slab() if slab is None else slab,
                      default_hints() if hints is None else hints)
        _ffi.call("sgc_precompute", plan.handle, _ffi.ctypes.byref(a), degree,
                  _ffi.ptr(bufs[0]) if bufs else None, _ffi.ptr(bufs[1]) if len(bufs) > 1 else None, ld,
                  stream_ptr())
    return y


def sgc_precompute(features, adj, degree, **kw):
    """Drop-in for utils.sgc_precompute (utils.py:92-97): returns ``(features, seconds)``.
    Unlike the reference, the clock is read after the stream has drained, so the time is the
    real duration of the K hops."""
    from .features import StandardizedFeatures
    if int(degree) <= 0:
        return features, 0.0            # utils.py:94: the loop does not run; same object back
    mean = std = None
    g = as_graph(adj)
    torch.cuda.synchronize(g.device)
    t = perf_counter()
    if isinstance(features, StandardizedFeatures):
        if (kw.get("std_mode") or default_std_mode()) == "inplace":
            features.apply_()
            kw.pop("std_mode", None)
        features, mean, std = features.unwrap()
    out = propagate(g, features, degree, mean=mean, std=std, **kw)
    torch.cuda.synchronize(g.device)
    return out, perf_counter() - t
