"""Sparse feature matrices on the GPU: the row normalisation of the citation
loaders and a first hop that never densifies X.

Reference behaviour being replaced (paths relative to /root/reference):
    normalization.py:21-28   row_normalize(mx): diag(1 / rowsum) @ mx, inf -> 0       (scipy, host)
    utils.py:74              features = torch.FloatTensor(np.array(features.todense()))  (host)
    utils.py:95              first torch.spmm(adj, features) on that dense matrix

`SparseFeatures` is the GPU-resident CSR (row_ptr int64, col int32, val fp32/fp64).  It answers
``.todense()`` / ``.tocsr()`` / ``.shape`` like the scipy matrix the reference's row_normalize returns,
and ``.to_dense_padded()`` gives the dense device tensor the loaders hand on; that tensor remembers
its sparse origin, so `sgc_precompute` runs its first hop on the stored entries only.
"""
import os

import numpy as np
import torch

from . import _ffi
from .features import pad_ld
from .graph import require_cuda, stream_ptr


def sparse_first_hop_enabled():
    """SGC_B200_SPARSE_FIRST_HOP (default 1): first hop on the CSR of X when the features came from
    `SparseFeatures.to_dense_padded()` and have not been modified since."""
    return os.environ.get("SGC_B200_SPARSE_FIRST_HOP", "1") != "0"


class SparseFeatures(object):
    def __init__(self, shape, row_ptr, col, val):
        self.shape = (int(shape[0]), int(shape[1]))
        self.row_ptr, self.col, self.val = row_ptr, col, val
        self.device = row_ptr.device
        self.nnz = int(col.numel())
        self.dtype = val.dtype
        self._val32 = None

    @classmethod
    def from_scipy(cls, mx, device=None):
        """Upload a scipy sparse matrix (any format; fp32 stays fp32, everything else becomes fp64,
        which is what scipy's own arithmetic would promote to)."""
        import scipy.sparse as sp
        dev = require_cuda(device)
        m = mx.tocsr() if sp.issparse(mx) else sp.csr_matrix(mx)
        m.sum_duplicates()
        m.sort_indices()
        data = np.ascontiguousarray(m.data)
        if data.dtype != np.float32:
            data = data.astype(np.float64)
        return cls(m.shape, torch.from_numpy(np.ascontiguousarray(m.indptr, dtype=np.int64)).to(dev),
                   torch.from_numpy(np.ascontiguousarray(m.indices, dtype=np.int32)).to(dev),
                   torch.from_numpy(data).to(dev))

    def values32(self):
        if self._val32 is None:
            self._val32 = self.val if self.val.dtype == torch.float32 else self.val.to(torch.float32)
        return self._val32

    # -- what the reference's callers do with the scipy matrix ---------------------------------------
    def tocsr(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.val.cpu().numpy(), self.col.cpu().numpy(), self.row_ptr.cpu().numpy()),
                             shape=self.shape)

    def todense(self):
        return self.tocsr().todense()

    def to_dense_padded(self):
        """Dense fp32 [N, F] device tensor in padded row storage (utils.py:74 without the host
        detour), tagged with this CSR for the sparse first hop."""
        n, f = self.shape
        dev = self.device
        base = torch.zeros((n, pad_ld(f)), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            ident = torch.arange(f, dtype=torch.int32, device=dev)
            _ffi.call("sgc_gather_columns_dense", n, self.nnz, _ffi.ptr(self.row_ptr), _ffi.ptr(self.col),
                      _ffi.ptr(self.values32()), _ffi.ptr(ident), base.data_ptr(), base.stride(0), stream_ptr())
        x = base[:, :f]
        x._sgc_sparse = (self, x._version)
        return x


def sparse_origin(x):
    """The `SparseFeatures` a dense tensor was built from, if it is still an exact copy."""
    tag = getattr(x, "_sgc_sparse", None)
    if tag is None or not sparse_first_hop_enabled():
        return None
    sf, version = tag
    if x._version != version or tuple(x.shape) != sf.shape or x.device != sf.device:
        return None
    return sf


def row_normalize(mx, device=None):
    """normalization.py:21-28 on the GPU.  Takes what the reference takes (a scipy sparse matrix) or
    a `SparseFeatures`; returns a `SparseFeatures`."""
    sf = mx if isinstance(mx, SparseFeatures) else SparseFeatures.from_scipy(mx, device)
    dev = sf.device
    n = sf.shape[0]
    out = torch.empty_like(sf.val)
    zero = torch.empty(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _ffi.call("sgc_row_normalize", n, _ffi.ptr(sf.row_ptr), _ffi.ptr(sf.val),
                  1 if sf.val.dtype == torch.float64 else 0, _ffi.ptr(out), None, _ffi.ptr(zero), stream_ptr())
        row_ptr, col = sf.row_ptr, sf.col
        if int(zero.item()) != 0:
            # scipy's CSR product drops the entries that became exactly zero (explicit zeros, rows whose
            # sum is zero): compact with the sort-free reduce of prep.py (fp64 holds fp32 exactly)
            from .prep import coo_to_canonical_csr, expand_rows
            rows = expand_rows(n, row_ptr, col.numel())
            row_ptr, col, v64 = coo_to_canonical_csr(n, sf.shape[1], rows, col, out.to(torch.float64), op="sum",
                                                     drop_zeros=True)
            out = v64.to(sf.val.dtype)
    return SparseFeatures(sf.shape, row_ptr, col, out)


def hop_sparse_x(graph, sf, out=None):
    """One hop ``S @ X`` with X = `sf` (CSR): [n_rows, F] fp32, padded rows unless `out` is given."""
    dev = require_cuda(graph.device)
    n, f = graph.n_rows, sf.shape[1]
    if sf.shape[0] != graph.n_cols:
        raise ValueError("size mismatch: adj is %s, features are %s" % (graph.shape, sf.shape))
    if out is None:
        out = torch.empty((n, pad_ld(f)), dtype=torch.float32, device=dev)[:, :f]
        if out.stride(0) != f:
            out.as_strided((n, out.stride(0) - f), (out.stride(0), 1), out.storage_offset() + f).zero_()
    with torch.cuda.device(dev):
        _ffi.call("sgc_spmm_hop_sparse_x", n, _ffi.ptr(graph.row_ptr), _ffi.ptr(graph.col), _ffi.ptr(graph.values32()),
                  _ffi.ptr(sf.row_ptr), _ffi.ptr(sf.col), _ffi.ptr(sf.values32()), f, out.data_ptr(), out.stride(0),
                  stream_ptr())
    return out
