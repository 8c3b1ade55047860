"""Device-resident normalised adjacency (the `adj` object of the reference's API).

`SgcGraph` is what `normalization.aug_normalized_adjacency` produces here instead of the scipy COO
matrix the reference returns (/root/reference/normalization.py:5-12): the canonical CSR structure
of ``S = D^-1/2 (A+I) D^-1/2`` in HBM together with the fp64 values of ``A+I`` and the fp64
``D^-1/2`` diagonal.  The normalised values themselves are never stored unless a caller asks for
the torch / scipy view -- the propagation kernel forms ``fp32((d_i*a_ij)*d_j)`` on the fly.

HBM layout (all contiguous 1-D arrays, 256-byte aligned by the torch allocator):
    row_ptr  int64 [n+1]     col  int32 [nnz]     aval  float64 [nnz]     dinv  float64 [n]

All arithmetic runs in libsgc_b200.so through sgc_b200._ffi; PyTorch only owns the memory.
"""
import weakref

import numpy as np
import torch

from . import _ffi

_checked_devices = set()


def require_cuda(device=None):
    """Raise unless a B200-class device is usable (no CPU fallback, by design)."""
    if not torch.cuda.is_available():
        raise _ffi.SgcError("sgc_b200 needs a CUDA device (sm_100a); it has no CPU fallback -- run the "
                            "reference implementation for CPU-only (--no-cuda) experiments")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    if dev.index not in _checked_devices:
        _ffi.call("sgc_device_check", dev.index)
        _checked_devices.add(dev.index)
    return dev


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


class SgcGraph(object):
    """Normalised adjacency S on the GPU (see module docstring).  Treat as opaque: pass it to
    ``sgc_precompute`` / ``torch.spmm``; ``.to_torch_sparse()`` and ``.tocoo()`` give the views
    the reference's callers would have had."""

    is_sparse = True

    def __init__(self, n_rows, n_cols, row_ptr, col, aval, dinv, rowsum=None, sval=None, dinv_rows=None, order=0):
        self.n_rows, self.n_cols = int(n_rows), int(n_cols)
        # association of the value product: 0 = (d_i*a)*d_j (normalization.py:12), 1 = (a*d_j)*d_i
        # (TextSGC's normalize_adj, downstream/TextSGC/utils.py:66)
        self.order = int(order)
        self.row_ptr, self.col, self.aval, self.dinv, self.rowsum = row_ptr, col, aval, dinv, rowsum
        self.dinv_rows = dinv if dinv_rows is None else dinv_rows     # differs for a row slice only
        self.sval = sval                      # fp32 values of S when materialised (or given)
        self.nnz = int(col.numel())
        self.device = row_ptr.device
        self._plans = {}
        self._torch_view = None

    # -- shape protocol the reference's callers rely on -----------------------------------
    @property
    def shape(self):
        return (self.n_rows, self.n_cols)

    def size(self, dim=None):
        return torch.Size(self.shape) if dim is None else self.shape[dim]

    @property
    def is_cuda(self):
        return True

    def cuda(self, *a, **k):      # utils.py:84,127-128 call .cuda() on adj: already resident
        return self

    def float(self):              # utils.py:77,122,124 call .float(): fp32 is what the hop computes in
        return self

    def to(self, *a, **k):
        return self

    @property
    def fused(self):
        """True when the hop kernel may form the weights itself from aval and dinv (order 0 only:
        the kernel multiplies in the order of normalization.py:12)."""
        return self.aval is not None and self.dinv is not None and self.order == 0

    def _norm_values(self, s64, s32):
        zero = torch.empty(1, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.call("sgc_norm_values", self.n_rows, self.nnz, _ffi.ptr(self.row_ptr), _ffi.ptr(self.col),
                      _ffi.ptr(self.aval), _ffi.ptr(self.dinv_rows), _ffi.ptr(self.dinv), self.order,
                      _ffi.ptr(s64), _ffi.ptr(s32), _ffi.ptr(zero), stream_ptr())

    # -- plan cache --------------------------------------------------------------------------
    def plan(self, chunk):
        chunk = max(int(chunk), 0)
        hit = self._plans.get(chunk)
        if hit is not None:
            return hit[0]
        lib = _ffi.lib()
        nbytes = lib.sgc_spmm_plan_workspace_bytes(self.n_rows, self.nnz, chunk)
        ws = _ws(nbytes, self.device)
        handle = _ffi.c_p()
        with torch.cuda.device(self.device):
            _ffi.call("sgc_spmm_plan_create", self.n_rows, self.n_cols, self.nnz, _ffi.ptr(self.row_ptr),
                      _ffi.ptr(self.col), chunk, _ffi.ptr(ws), ws.numel(), _ffi.ctypes.byref(handle),
                      stream_ptr())
        plan = _Plan(handle, ws, int(lib.sgc_spmm_plan_partial_rows(handle)))
        self._plans[chunk] = (plan,)
        return plan

    # -- column blocks (operands several times larger than L2) -----------------------------------
    def col_blocks(self, n_blocks):
        """[(c0, c1, SgcGraph)]: S split into `n_blocks` blocks of consecutive COLUMNS, each block a compacted
        CSR of the entries whose column lies in [c0, c1) (global column ids kept; fp32 S values copied, so the
        blocks together hold one more copy of col / sval plus n_blocks row-pointer arrays).  Entries of a row
        keep their ascending order inside every block, which is what lets `propagate.hop_col_blocked` continue
        each row's FMA chain block by block.  Cached per n_blocks.  Each block reports `n_cols` = the number of
        rows of X it gathers from, so that the hop sizes its L2 policy for the block, not for all of X.
        Row-slice views (pipeline.row_blocks, the sharded propagator's sub-graphs) keep absolute offsets in
        `row_ptr` and share the parent's `col` array: only the entries [row_ptr[0], row_ptr[-1]) are theirs, and
        the blocks are built from that range alone (one small D2H read of the two offsets)."""
        n_blocks = max(1, min(int(n_blocks), max(self.n_cols, 1)))
        hit = self._plans.get(("col_blocks", n_blocks))
        if hit is not None:
            return hit
        dev = self.device
        sval = self.values32()
        per = (self.n_cols + n_blocks - 1) // n_blocks
        per = (per + 127) // 128 * 128
        with torch.cuda.device(dev):
            b, e = (int(v) for v in self.row_ptr[[0, -1]].cpu().tolist())
            row_ptr = self.row_ptr if b == 0 else (self.row_ptr - b).contiguous()
            col, sval = self.col[b:e], sval[b:e]
            rows = torch.empty(e - b, dtype=torch.int32, device=dev)
            _ffi.call("sgc_csr_expand_rows", self.n_rows, _ffi.ptr(row_ptr), 0, _ffi.ptr(rows), stream_ptr())
            blk = torch.div(col, per, rounding_mode="floor")
            out = []
            for b in range(n_blocks):
                c0, c1 = b * per, min((b + 1) * per, self.n_cols)
                if c1 <= c0:
                    break
                keep = (blk == b).nonzero().squeeze(1)
                counts = torch.bincount(rows[keep].long(), minlength=self.n_rows)
                rp = torch.zeros(self.n_rows + 1, dtype=torch.int64, device=dev)
                torch.cumsum(counts, 0, out=rp[1:])
                g = SgcGraph(self.n_rows, c1 - c0, rp, col[keep].contiguous(), None, None,
                             sval=sval[keep].contiguous())
                out.append((c0, c1, g))
        self._plans[("col_blocks", n_blocks)] = out
        return out

    # -- views -------------------------------------------------------------------------------
    def values32(self):
        """fp32 values of S: fp32((d_i*a_ij)*d_j), the single rounding of utils.py:25."""
        if self.sval is None and self.order != 0:
            self.sval = torch.empty(self.nnz, dtype=torch.float32, device=self.device)
            self._norm_values(None, self.sval)
        if self.sval is None:
            self.sval = torch.empty(self.nnz, dtype=torch.float32, device=self.device)
            with torch.cuda.device(self.device):
                _ffi.call("sgc_norm_build", self.n_rows, _ffi.ptr(self.row_ptr), _ffi.ptr(self.col),
                          _ffi.ptr(self.aval), _ffi.ptr(self.dinv_rows), _ffi.ptr(self.dinv),
                          _ffi.ptr(self.row_ptr), None, None, None, _ffi.ptr(self.sval), stream_ptr())
        return self.sval

    def values64(self):
        """fp64 values of S as normalization.py:12 returns them."""
        out = torch.empty(self.nnz, dtype=torch.float64, device=self.device)
        if self.order != 0:
            self._norm_values(out, None)
            return out
        with torch.cuda.device(self.device):
            _ffi.call("sgc_norm_build", self.n_rows, _ffi.ptr(self.row_ptr), _ffi.ptr(self.col),
                      _ffi.ptr(self.aval), _ffi.ptr(self.dinv_rows), _ffi.ptr(self.dinv),
                      _ffi.ptr(self.row_ptr), None, None, _ffi.ptr(out), None, stream_ptr())
        return out

    def coo_indices(self):
        """int64 [2, nnz] row-major sorted indices (utils.py:26-27)."""
        idx = torch.empty((2, self.nnz), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.call("sgc_csr_to_coo", self.n_rows, _ffi.ptr(self.row_ptr), _ffi.ptr(self.col), self.nnz,
                      _ffi.ptr(idx), stream_ptr())
        return idx

    def to_torch_sparse(self):
        """torch sparse COO fp32 tensor equal to the reference's
        ``sparse_mx_to_torch_sparse_tensor(aug_normalized_adjacency(adj)).cuda()``."""
        if self._torch_view is None:
            self._torch_view = torch.sparse_coo_tensor(self.coo_indices(), self.values32(),
                                                       torch.Size(self.shape))
        return self._torch_view

    def tocoo(self):
        """scipy COO fp64 on the host: the reference's return value of aug_normalized_adjacency."""
        import scipy.sparse as sp
        idx = self.coo_indices().cpu().numpy()
        return sp.coo_matrix((self.values64().cpu().numpy(), (idx[0], idx[1])), shape=self.shape)

    def tocsr(self):
        return self.tocoo().tocsr()

    # -- torch.spmm(adj, x) keeps working (models.py:38) --------------------------------------
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func in _MATMULS and len(args) == 2 and isinstance(args[0], SgcGraph) \
                and isinstance(args[1], torch.Tensor) and not kwargs:
            from .propagate import hop
            return hop(args[0], args[1])
        raise TypeError("SgcGraph supports torch.spmm / torch.sparse.mm / torch.mm / torch.matmul with a "
                        "dense right operand; use .to_torch_sparse() for anything else (got %r)" % (func,))

    def __matmul__(self, other):
        from .propagate import hop
        return hop(self, other)

    def __repr__(self):
        return "SgcGraph(shape=%s, nnz=%d, device=%s, fused_norm=%s)" % (self.shape, self.nnz, self.device,
                                                                        self.fused)


_MATMULS = (torch.spmm, torch.sparse.mm, torch.mm, torch.matmul)


L2_SLAB_BUDGET = 128 * 1024 * 1024      # same constant as spmm.cu: what one slab may occupy in L2


def hot_budget_bytes():
    """L2 bytes reserved for the most frequently gathered rows of X when X does not fit in L2
    (SGC_B200_HOT_MB; 0 disables hot-column tagging)."""
    import os
    return int(float(os.environ.get("SGC_B200_HOT_MB", "64")) * 1024 * 1024)


class _Plan(object):
    def __init__(self, handle, ws, partial_rows):
        self.handle, self.ws, self.partial_rows = handle, ws, partial_rows
        self._fin = weakref.finalize(self, _ffi.lib().sgc_spmm_plan_destroy, handle)
        self.hot_key, self.col_tagged, self.n_hot = None, None, 0

    def ensure_hot(self, graph, row_bytes):
        """Tag the hub columns for an operand whose rows are `row_bytes` long, if not even the
        narrowest feature slab of it fits in L2 (otherwise slab tiling keeps everything resident
        and tags are switched off).  Cached per (row_bytes, budget)."""
        budget = hot_budget_bytes()
        need = budget > 0 and graph.n_cols * 32 * 4 > L2_SLAB_BUDGET and graph.nnz > 0
        key = (int(row_bytes), budget) if need else None
        if key == self.hot_key:
            return
        lib = _ffi.lib()
        with torch.cuda.device(graph.device):
            if not need:
                _ffi.call("sgc_spmm_plan_tag_hot", self.handle, 0, None, None, 0, None, stream_ptr())
                self.col_tagged, self.n_hot = None, 0
            else:
                ws = _ws(lib.sgc_spmm_hot_workspace_bytes(graph.n_cols), graph.device)
                tagged = torch.empty(graph.nnz, dtype=torch.int32, device=graph.device)
                n_hot = _ffi.c_i64(0)
                _ffi.call("sgc_spmm_plan_tag_hot", self.handle, max(1, budget // int(row_bytes)), _ffi.ptr(tagged),
                          _ffi.ptr(ws), ws.numel(), _ffi.ctypes.byref(n_hot), stream_ptr())
                self.col_tagged, self.n_hot = (tagged if n_hot.value > 0 else None), int(n_hot.value)
        self.hot_key = key


# ---- K1 + K2: normalisation driver ------------------------------------------------------------------
def _augment_general(n, n_cols, row_offset, row_ptr, col, val, nnz, canonical):
    """General path of K1 (csr_build.cu): unsorted / duplicated input is canonicalised first, and
    entries A + I drops (exact zeros, a_ii == -1: scipy's csr_plus_csr prunes them) are compacted
    away by the row-group kernels.  Returns (row_ptr, col, aval) of A + I."""
    dev = row_ptr.device
    lib = _ffi.lib()
    s = stream_ptr()
    scan_ws = _ws(lib.sgc_scan_workspace_bytes(n + 1), dev)
    if not canonical:
        if nnz and (int(col.min().item()) < 0 or int(col.max().item()) >= n_cols):
            raise ValueError("column index out of range")
        rp_host_ok = bool((row_ptr[1:] >= row_ptr[:-1]).all().item()) and int(row_ptr[0].item()) == 0 \
            and int(row_ptr[n].item()) == nnz if n else True
        if not rp_host_ok:
            raise ValueError("row_ptr is not a valid CSR row pointer array")
        if nnz > (1 << 18):
            # large input: radix sort of the COO stream (prep.py); the per-row rank sort below
            # is quadratic in the row length
            from .prep import canonicalize
            row_ptr, col, val = canonicalize(n, n_cols, row_ptr, col, val)
            nnz = int(col.numel())
        else:
            ws = _ws(lib.sgc_csr_canonicalize_workspace_bytes(n, nnz), dev)
            rp2 = torch.empty(n + 1, dtype=torch.int64, device=dev)
            col2 = torch.empty(nnz, dtype=torch.int32, device=dev)
            val2 = torch.empty(nnz, dtype=torch.float64, device=dev)
            _ffi.call("sgc_csr_canonicalize", n, nnz, _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(val),
                      _ffi.ptr(rp2), _ffi.ptr(col2), _ffi.ptr(val2), _ffi.ptr(ws), ws.numel(), s)
            nnz = int(rp2[n].item())
            row_ptr, col, val = rp2, col2[:nnz], val2[:nnz]
    rp_aug = torch.empty(n + 1, dtype=torch.int64, device=dev)
    _ffi.call("sgc_csr_selfloop_count", n, row_offset, _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(val),
              _ffi.ptr(rp_aug), _ffi.ptr(scan_ws), scan_ws.numel(), s)
    nnz_aug = int(rp_aug[n].item()) if n else 0
    col_aug = torch.empty(nnz_aug, dtype=torch.int32, device=dev)
    aval = torch.empty(nnz_aug, dtype=torch.float64, device=dev)
    _ffi.call("sgc_csr_selfloop_build", n, row_offset, _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(val),
              _ffi.ptr(rp_aug), _ffi.ptr(col_aug), _ffi.ptr(aval), s)
    return rp_aug, col_aug, aval


def augment_csr(n_rows, n_cols, row_offset, row_ptr, col, val):
    """K1 + K2 on a row slice [row_offset, row_offset + n_rows) of an n_cols x n_cols adjacency with
    global column ids (row_offset = 0 and n_rows == n_cols for a whole matrix): canonical CSR of
    A + I (normalization.py:6) and the slice's row sums / D^-1/2 (normalization.py:8-10).
    Returns (row_ptr, col, aval, rowsum, dinv_rows)."""
    dev = require_cuda(row_ptr.device)
    lib = _ffi.lib()
    n, n_cols, row_offset = int(n_rows), int(n_cols), int(row_offset)
    with torch.cuda.device(dev):
        s = stream_ptr()
        row_ptr = row_ptr.to(torch.int64).contiguous()
        col = col.to(torch.int32).contiguous()
        val = val.to(torch.float64).contiguous()
        nnz = int(col.numel())
        scan_ws = _ws(lib.sgc_scan_workspace_bytes(n + 1), dev)
        rp_aug = torch.empty(n + 1, dtype=torch.int64, device=dev)
        # fast path (csr_fast.cu): one element-parallel pass proves the input canonical, finds out
        # whether A + I drops anything, and places every self-loop; one host read-back
        info = torch.empty(n, dtype=torch.int32, device=dev)
        summary = torch.empty(2, dtype=torch.int64, device=dev)
        _ffi.call("sgc_csr_augment_count", n, n_cols, row_offset, nnz, _ffi.ptr(row_ptr), _ffi.ptr(col),
                  _ffi.ptr(val), _ffi.ptr(info), _ffi.ptr(rp_aug), _ffi.ptr(summary), _ffi.ptr(scan_ws),
                  scan_ws.numel(), s)
        flags, nnz_aug = summary.cpu().tolist()
        if flags == 0:
            col_aug = torch.empty(nnz_aug, dtype=torch.int32, device=dev)
            aval = torch.empty(nnz_aug, dtype=torch.float64, device=dev)
            _ffi.call("sgc_csr_augment_build", n, row_offset, nnz, _ffi.ptr(row_ptr), _ffi.ptr(col),
                      _ffi.ptr(val), _ffi.ptr(info), _ffi.ptr(rp_aug), _ffi.ptr(col_aug), _ffi.ptr(aval), s)
        else:
            rp_aug, col_aug, aval = _augment_general(n, n_cols, row_offset, row_ptr, col, val, nnz,
                                                     canonical=not (flags & 1))
        # K2: degrees
        rowsum = torch.empty(n, dtype=torch.float64, device=dev)
        dinv = torch.empty(n, dtype=torch.float64, device=dev)
        _ffi.call("sgc_degree_inv_sqrt", n, _ffi.ptr(rp_aug), _ffi.ptr(aval), _ffi.ptr(rowsum),
                  _ffi.ptr(dinv), s)
    return rp_aug, col_aug, aval, rowsum, dinv


def finish_normalization(n_rows, n_cols, rp_aug, col_aug, aval, rowsum, dinv_rows, dinv_cols, materialize=None,
                         order=0):
    """Structure of S after scipy's zero-product pruning (normalization.py:12; it only differs
    from A + I when some D^-1/2 entry is zero) -> `SgcGraph`.  dinv_rows: D^-1/2 of the slice's
    rows; dinv_cols: D^-1/2 of all n_cols columns."""
    dev = require_cuda(rp_aug.device)
    lib = _ffi.lib()
    n = int(n_rows)
    if materialize is None:
        import os
        materialize = os.environ.get("SGC_B200_FUSED_NORM", "0") != "1" or order != 0
    with torch.cuda.device(dev):
        s = stream_ptr()
        # fast path: the fp32 values of S in one element-parallel pass, which also reports whether
        # any product is exactly zero (only then does scipy's CSR product prune the structure)
        nnz = int(col_aug.numel())
        sval = torch.empty(nnz, dtype=torch.float32, device=dev) if materialize else None
        zero = torch.empty(1, dtype=torch.int64, device=dev)
        _ffi.call("sgc_norm_values", n, nnz, _ffi.ptr(rp_aug), _ffi.ptr(col_aug), _ffi.ptr(aval),
                  _ffi.ptr(dinv_rows), _ffi.ptr(dinv_cols), int(order), None, _ffi.ptr(sval), _ffi.ptr(zero), s)
        if int(zero.item()) != 0 and order != 0:
            raise _ffi.SgcError("normalize_adj: a zero degree prunes the structure; only the association of "
                                "normalization.py:12 implements scipy's pruning")
        if int(zero.item()) != 0:
            sval = None
            scan_ws = _ws(lib.sgc_scan_workspace_bytes(n + 1), dev)
            rp_s = torch.empty(n + 1, dtype=torch.int64, device=dev)
            dropped = torch.empty(1, dtype=torch.int64, device=dev)
            _ffi.call("sgc_norm_count", n, _ffi.ptr(rp_aug), _ffi.ptr(col_aug), _ffi.ptr(aval),
                      _ffi.ptr(dinv_rows), _ffi.ptr(dinv_cols), _ffi.ptr(rp_s), _ffi.ptr(dropped),
                      _ffi.ptr(scan_ws), scan_ws.numel(), s)
            if int(dropped.item()) != 0:
                nnz_s = int(rp_s[n].item())
                col_s = torch.empty(nnz_s, dtype=torch.int32, device=dev)
                aval_s = torch.empty(nnz_s, dtype=torch.float64, device=dev)
                _ffi.call("sgc_norm_build", n, _ffi.ptr(rp_aug), _ffi.ptr(col_aug), _ffi.ptr(aval),
                          _ffi.ptr(dinv_rows), _ffi.ptr(dinv_cols), _ffi.ptr(rp_s), _ffi.ptr(col_s),
                          _ffi.ptr(aval_s), None, None, s)
                rp_aug, col_aug, aval = rp_s, col_s, aval_s
    return SgcGraph(n, int(n_cols), rp_aug, col_aug, aval, dinv_cols, rowsum=rowsum, dinv_rows=dinv_rows, sval=sval,
                    order=order)


def normalize_csr(n, row_ptr, col, val, order=0):
    """``aug_normalized_adjacency`` on device CSR arrays (normalization.py:5-12).

    row_ptr int64 [n+1], col int32 [nnz], val float64 [nnz], all on one CUDA device; any stored
    order (non-canonical input is canonicalised first).  Returns an `SgcGraph`.  order=1 multiplies
    in the order of TextSGC's normalize_adj (see `SgcGraph.order`)."""
    rp_aug, col_aug, aval, rowsum, dinv = augment_csr(n, n, 0, row_ptr, col, val)
    return finish_normalization(n, n, rp_aug, col_aug, aval, rowsum, dinv, dinv, order=order)


def upload_csr(adj, device=None):
    """scipy sparse matrix (any format/dtype) -> device CSR arrays (row_ptr int64, col int32,
    val float64).  Host work is limited to scipy's own format conversion; dtype casts run on the
    device."""
    import scipy.sparse as sp
    dev = require_cuda(device)
    m = adj if sp.issparse(adj) else sp.csr_matrix(adj)
    if m.format != "csr":
        m = m.tocsr()
    if m.shape[0] >= 2 ** 31 or m.shape[1] >= 2 ** 31:
        raise ValueError("more than 2^31-1 rows/columns are not supported (int32 column indices)")
    row_ptr = torch.from_numpy(np.ascontiguousarray(m.indptr)).to(dev, non_blocking=False).to(torch.int64)
    col = torch.from_numpy(np.ascontiguousarray(m.indices)).to(dev).to(torch.int32)
    data = np.ascontiguousarray(m.data)
    if data.dtype == np.bool_:
        data = data.astype(np.uint8)
    val = torch.from_numpy(data).to(dev).to(torch.float64)
    return m.shape, row_ptr, col, val


def from_scipy(adj, device=None, order=0):
    """normalization.aug_normalized_adjacency on a scipy matrix: upload, then K1/K2 on the GPU."""
    shape, row_ptr, col, val = upload_csr(adj, device)
    if shape[0] != shape[1]:
        raise ValueError("adjacency must be square, got %s" % (shape,))
    return normalize_csr(shape[0], row_ptr, col, val, order=order)


def from_torch_sparse(t):
    """Wrap an existing torch sparse tensor holding S (e.g. one a caller built with the
    reference's own sparse_mx_to_torch_sparse_tensor and moved to the GPU) without copying its
    values: zero-copy ingestion of device-resident COO/CSR (scope table row f2)."""
    dev = require_cuda(t.device)
    lib = _ffi.lib()
    with torch.cuda.device(dev):
        if t.layout == torch.sparse_csr:
            n_rows, n_cols = t.shape
            return SgcGraph(n_rows, n_cols, t.crow_indices().to(torch.int64).contiguous(),
                            t.col_indices().to(torch.int32).contiguous(), None, None,
                            sval=t.values().to(torch.float32).contiguous())
        if t.layout != torch.sparse_coo:
            raise TypeError("expected a torch sparse COO or CSR tensor")
        n_rows, n_cols = t.shape
        idx, val = t._indices(), t._values()
        nnz = int(val.numel())
        row_ptr = torch.empty(n_rows + 1, dtype=torch.int64, device=dev)
        col = torch.empty(nnz, dtype=torch.int32, device=dev)
        flag = torch.empty(1, dtype=torch.int32, device=dev)
        idx = idx.contiguous()
        _ffi.call("sgc_coo_to_csr", n_rows, nnz, _ffi.ptr(idx[0]), _ffi.ptr(idx[1]), _ffi.ptr(row_ptr),
                  _ffi.ptr(col), _ffi.ptr(flag), stream_ptr())
        if int(flag.item()) == 0:
            # unsorted or duplicated COO: let torch coalesce (sort + merge) once, then retry
            tc = t.coalesce()
            idx, val = tc.indices().contiguous(), tc.values()
            nnz = int(val.numel())
            col = torch.empty(nnz, dtype=torch.int32, device=dev)
            _ffi.call("sgc_coo_to_csr", n_rows, nnz, _ffi.ptr(idx[0]), _ffi.ptr(idx[1]), _ffi.ptr(row_ptr),
                      _ffi.ptr(col), _ffi.ptr(flag), stream_ptr())
            if int(flag.item()) == 0:
                raise _ffi.SgcError("could not convert the sparse tensor to CSR")
        g = SgcGraph(n_rows, n_cols, row_ptr, col, None, None, sval=val.to(torch.float32).contiguous())
        g._torch_view = t
        return g
