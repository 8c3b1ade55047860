"""Mirror of the propagation part of the reference's downstream/TextSGC/utils.py:
the second consumer of the same SpMM, with a fused min-max epilogue.  Same names, argument order
and return shapes as the reference (paths relative to /root/reference/downstream/TextSGC):

    normalize_adj(adj)                      utils.py:59-66    D^-1/2 A D^-1/2, product order (a*d_j)*d_i
    preprocess_adj(adj)                     utils.py:68-71    normalize_adj(adj + I)
    sparse_to_torch_sparse(sparse_mx)       utils.py:110-124  -> the GPU-resident graph itself
    sparse_to_torch_dense(sparse)           utils.py:126-129  -> a column view (no N x N dense matrix)
    sgc_precompute(adj, features, degree, index_dict)   utils.py:131-152

What changes underneath: the reference builds the dense N x N copy of the normalised adjacency on the
host (20 073^2 floats = 1.6 GB for its covid corpus), slices document columns out of it, copies them
to the GPU and calls torch.spmm; here the selected columns are scattered straight from the CSR
arrays into a dense N x m operand on the GPU (sgc_gather_columns_dense), the hop is the same kernel
as the main path (K4), and the transpose + min-max scaling is one kernel (sgc_transpose_scale).
"""
from time import perf_counter

import numpy as np
import torch

from . import _ffi
from .features import pad_ld
from .graph import SgcGraph, finish_normalization, from_scipy, require_cuda, stream_ptr, upload_csr
from .propagate import hop


def preprocess_adj(adj):
    """``normalize_adj(adj + sp.eye(n))`` (utils.py:68-71) on the GPU: K1 inserts the self-loops,
    K2 the degrees, and the values are formed as (a_ij * d_j) * d_i -- the association of
    ``adj.dot(D).transpose().dot(D).transpose()`` (utils.py:66).  Returns an `SgcGraph`;
    ``.tocoo()`` gives the reference's return value.  The sibling packages (downstream/TextSGC_Bio,
    TextSGC_indexing, TextSGC_ref: utils.py:108-112 there) share normalize_adj / sgc_precompute line for
    line and only append ``.astype(np.float32)`` here -- the rounding `sparse_to_torch_sparse` applies
    anyway (utils.py:112), and the one the graph's fp32 S values already carry -- so this mirror serves them
    unchanged."""
    if isinstance(adj, SgcGraph):
        return adj
    return from_scipy(adj, order=1)


def normalize_adj(adj):
    """Symmetric normalisation WITHOUT self-loop insertion (utils.py:59-66); the input must be a
    canonical scipy matrix without explicit zeros (what build_graph.py writes)."""
    shape, row_ptr, col, val = upload_csr(adj)
    if shape[0] != shape[1]:
        raise ValueError("adjacency must be square, got %s" % (shape,))
    n = shape[0]
    dev = row_ptr.device
    with torch.cuda.device(dev):
        flag = torch.empty(1, dtype=torch.int32, device=dev)
        _ffi.call("sgc_csr_check_canonical", n, n, _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(flag), stream_ptr())
        if int(flag.item()) == 0:
            from .prep import canonicalize
            row_ptr, col, val = canonicalize(n, n, row_ptr, col, val)
        rowsum = torch.empty(n, dtype=torch.float64, device=dev)
        dinv = torch.empty(n, dtype=torch.float64, device=dev)
        _ffi.call("sgc_degree_inv_sqrt", n, _ffi.ptr(row_ptr), _ffi.ptr(val), _ffi.ptr(rowsum), _ffi.ptr(dinv),
                  stream_ptr())
    return finish_normalization(n, n, row_ptr, col, val, rowsum, dinv, dinv, order=1)


def sparse_to_torch_sparse(sparse_mx, device="cuda"):
    """The reference converts scipy -> torch COO fp32 here; an `SgcGraph` already is the
    GPU-resident fp32 operand of the hop and is returned as is (``.to_torch_sparse()`` gives the
    torch tensor)."""
    if isinstance(sparse_mx, SgcGraph):
        return sparse_mx
    coo = sparse_mx.tocoo().astype(np.float32)
    idx = torch.from_numpy(np.vstack((coo.row, coo.col)).astype(np.int64))
    t = torch.sparse_coo_tensor(idx, torch.from_numpy(coo.data), torch.Size(coo.shape))
    return t.cuda() if device == "cuda" else t


class DenseColumns(object):
    """Stand-in for ``torch.from_numpy(sparse.todense().astype(np.float32))`` (utils.py:126-129):
    behaves as the dense N x N matrix for the one thing the precompute does with it --
    ``features[:, idx]`` -- and builds that N x m slice on the GPU from the CSR arrays."""

    def __init__(self, graph):
        self.graph = graph
        self.shape = graph.shape

    def size(self, dim=None):
        return torch.Size(self.shape) if dim is None else self.shape[dim]

    def __getitem__(self, key):
        if not (isinstance(key, tuple) and len(key) == 2 and key[0] == slice(None)):
            raise TypeError("DenseColumns supports features[:, idx] only")
        return gather_columns(self.graph, key[1])


def sparse_to_torch_dense(sparse, device="cuda"):
    if isinstance(sparse, SgcGraph):
        return DenseColumns(sparse)
    dense = torch.from_numpy(sparse.todense().astype(np.float32))
    return dense.to(device=device)


def gather_columns(graph, idx):
    """fp32 [N, m] device matrix equal to ``dense(S)[:, idx]`` (padded row storage)."""
    dev = require_cuda(graph.device)
    idx = torch.as_tensor(np.asarray(idx), dtype=torch.int64, device=dev).contiguous()
    m = int(idx.numel())
    n = graph.n_rows
    out = torch.zeros((n, pad_ld(max(m, 1))), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        s = stream_ptr()
        colmap = torch.empty(graph.n_cols, dtype=torch.int32, device=dev)
        _ffi.call("sgc_subgraph_colmap", graph.n_cols, m, _ffi.ptr(idx), _ffi.ptr(colmap), s)
        _ffi.call("sgc_gather_columns_dense", n, graph.nnz, _ffi.ptr(graph.row_ptr), _ffi.ptr(graph.col),
                  _ffi.ptr(graph.values32()), _ffi.ptr(colmap), out.data_ptr(), out.stride(0), s)
    return out[:, :m]


def row_minmax(y):
    """(min, range) over the columns of a device fp32 [N, m] matrix (utils.py:137-139)."""
    dev = require_cuda(y.device)
    n, m = y.shape
    mn = torch.empty(n, dtype=torch.float32, device=dev)
    rng = torch.empty(n, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _ffi.call("sgc_row_minmax", y.data_ptr(), n, m, y.stride(0), _ffi.ptr(mn), _ffi.ptr(rng), stream_ptr())
    return mn, rng


def transpose_scale(y, keep, mn, rng):
    """[m, k] matrix ``((y[keep].t() - mn[keep]) / rng[keep])`` (utils.py:140-144, 149-150)."""
    dev = require_cuda(y.device)
    m, k = int(y.shape[1]), int(keep.numel())
    out = torch.empty((m, k), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _ffi.call("sgc_transpose_scale", y.data_ptr(), m, y.stride(0), _ffi.ptr(keep), k, _ffi.ptr(mn),
                  _ffi.ptr(rng), out.data_ptr(), max(k, 1), stream_ptr())
    return out


def sgc_precompute(adj, features, degree, index_dict, chunk=None):
    """Drop-in for downstream/TextSGC/utils.py:131-152.  Returns ``(feat_dict, seconds)``:
    feat_dict["train"] on the GPU, "val" / "test" on the host, as the reference leaves them."""
    assert degree == 1, "Only supporting degree 2 now"       # the reference's own guard and message
    if not isinstance(adj, SgcGraph):
        from .propagate import as_graph
        adj = as_graph(adj)
    dev = require_cuda(adj.device)
    feat_dict = {}
    torch.cuda.synchronize(dev)
    start = perf_counter()

    def propagated(phase):
        cols = features[:, index_dict[phase]]
        if not (isinstance(cols, torch.Tensor) and cols.is_cuda):
            cols = torch.as_tensor(cols, dtype=torch.float32).to(dev)
        return hop(adj, cols, chunk=chunk)                    # [N, m]; .t() happens in the epilogue

    y = propagated("train")
    mn, rng = row_minmax(y)
    keep = rng.gt(0).nonzero().squeeze(1)                     # useful_features_dim (utils.py:140)
    feat_dict["train"] = transpose_scale(y, keep, mn, rng)
    for phase in ["test", "val"]:
        feat_dict[phase] = transpose_scale(propagated(phase), keep, mn, rng).cpu()
    torch.cuda.synchronize(dev)
    return feat_dict, perf_counter() - start
