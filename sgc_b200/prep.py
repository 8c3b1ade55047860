"""Graph preparation on the GPU: what the reference's loaders do to the raw adjacency right before
they normalise it.

    symmetrize(..., op="sum")    adj + adj.T                          /root/reference/utils.py:116
    symmetrize(..., op="max")    adj + adj.T*(adj.T>adj) - adj*(...)  /root/reference/utils.py:63
    induced_subgraph(..., idx)   adj[idx, :][:, idx]                  /root/reference/utils.py:117
    canonicalize(...)            scipy's implicit sort_indices / sum_duplicates

All of them run in libsgc_b200.so (csrc/graph_prep.cu): a stable radix sort of the COO stream by
column and then by row, followed by one flag + scan + fill pass that merges equal (row, col) pairs.
Inputs and outputs are device CSR triples (row_ptr int64, col int32, val float64).
"""
import torch

from . import _ffi
from .graph import _ws, require_cuda, stream_ptr


def _bits(n):
    return max(1, int(max(int(n) - 1, 1)).bit_length())


def _sort_pairs(key, p32, p64, key_bits):
    """Stable sort of (key, p32, p64) by key; returns new tensors."""
    lib = _ffi.lib()
    n = int(key.numel())
    dev = key.device
    ko, po, do = torch.empty_like(key), torch.empty_like(p32), torch.empty_like(p64)
    ws = _ws(lib.sgc_sort_workspace_bytes(n), dev)
    _ffi.call("sgc_sort_by_key_i32", n, key_bits, _ffi.ptr(key), _ffi.ptr(p32), _ffi.ptr(p64), _ffi.ptr(ko),
              _ffi.ptr(po), _ffi.ptr(do), _ffi.ptr(ws), ws.numel(), stream_ptr())
    return ko, po, do


def expand_rows(n_rows, row_ptr, nnz, row_offset=0):
    """int32 COO row index of every stored entry."""
    rows = torch.empty(int(nnz), dtype=torch.int32, device=row_ptr.device)
    _ffi.call("sgc_csr_expand_rows", int(n_rows), _ffi.ptr(row_ptr), int(row_offset), _ffi.ptr(rows), stream_ptr())
    return rows


def coo_to_canonical_csr(n_rows, n_cols, rows, cols, vals, op="sum", drop_zeros=False):
    """Arbitrary-order COO stream (int32 rows/cols, fp64 vals; duplicates allowed) -> canonical CSR.
    Duplicates are combined in their stored order (op "sum") or by maximum (op "max")."""
    dev = require_cuda(rows.device)
    lib = _ffi.lib()
    n = int(rows.numel())
    with torch.cuda.device(dev):
        if n:
            # LSD: by column first, then (stably) by row -> row-major, ascending columns
            cols, rows, vals = _sort_pairs(cols, rows, vals, _bits(n_cols))
            rows, cols, vals = _sort_pairs(rows, cols, vals, _bits(n_rows))
        row_ptr = torch.empty(int(n_rows) + 1, dtype=torch.int64, device=dev)
        col_out = torch.empty(n, dtype=torch.int32, device=dev)
        val_out = torch.empty(n, dtype=torch.float64, device=dev)
        ws = _ws(lib.sgc_coo_reduce_workspace_bytes(n), dev)
        _ffi.call("sgc_coo_reduce_to_csr", n, int(n_rows), _ffi.ptr(rows), _ffi.ptr(cols), _ffi.ptr(vals),
                  1 if op == "max" else 0, 1 if drop_zeros else 0, _ffi.ptr(row_ptr), _ffi.ptr(col_out),
                  _ffi.ptr(val_out), _ffi.ptr(ws), ws.numel(), stream_ptr())
        nnz = int(row_ptr[int(n_rows)].item())
    return row_ptr, col_out[:nnz], val_out[:nnz]


def canonicalize(n_rows, n_cols, row_ptr, col, val):
    """Sort the columns of every row and sum duplicate entries in stored order (what scipy does
    implicitly before `adj + sp.eye(...)`, normalization.py:6)."""
    row_ptr = row_ptr.to(torch.int64).contiguous()
    rows = expand_rows(n_rows, row_ptr, col.numel())
    return coo_to_canonical_csr(n_rows, n_cols, rows, col.to(torch.int32).contiguous(),
                                val.to(torch.float64).contiguous(), op="sum", drop_zeros=False)


def symmetrize(n, row_ptr, col, val, op="sum"):
    """`A + A.T` (op "sum", utils.py:116) or the element-wise maximum of A and A.T (op "max": the
    reference's `adj + adj.T.multiply(adj.T > adj) - adj.multiply(adj.T > adj)`, utils.py:63, for
    the non-negative adjacencies it is applied to) of a square CSR matrix.  Results that are
    exactly zero are dropped, as scipy's sparse arithmetic does.  Returns canonical CSR."""
    if op not in ("sum", "max"):
        raise ValueError("op must be 'sum' or 'max'")
    dev = require_cuda(row_ptr.device)
    with torch.cuda.device(dev):
        row_ptr = row_ptr.to(torch.int64).contiguous()
        col = col.to(torch.int32).contiguous()
        val = val.to(torch.float64).contiguous()
        rows = expand_rows(n, row_ptr, col.numel())
        rows2 = torch.cat([rows, col])            # entries of A followed by the entries of A^T
        cols2 = torch.cat([col, rows])
        vals2 = torch.cat([val, val])
    return coo_to_canonical_csr(n, n, rows2, cols2, vals2, op=op, drop_zeros=True)


def induced_subgraph(n_rows, n_cols, row_ptr, col, val, idx):
    """`M[idx, :][:, idx]` (utils.py:117) for distinct ids `idx` (any integer tensor / array, any
    order).  Returns canonical CSR of the m x m sub-matrix (m = len(idx))."""
    dev = require_cuda(row_ptr.device)
    lib = _ffi.lib()
    with torch.cuda.device(dev):
        idx = torch.as_tensor(idx).to(dev, torch.int64).contiguous()
        m = int(idx.numel())
        if m and (int(idx.min().item()) < 0 or int(idx.max().item()) >= min(n_rows, n_cols)):
            raise IndexError("index out of range")
        row_ptr = row_ptr.to(torch.int64).contiguous()
        col = col.to(torch.int32).contiguous()
        val = val.to(torch.float64).contiguous()
        colmap = torch.empty(int(n_cols), dtype=torch.int32, device=dev)
        _ffi.call("sgc_subgraph_colmap", int(n_cols), m, _ffi.ptr(idx), _ffi.ptr(colmap), stream_ptr())
        rp = torch.empty(m + 1, dtype=torch.int64, device=dev)
        scan_ws = _ws(lib.sgc_scan_workspace_bytes(m + 1), dev)
        _ffi.call("sgc_subgraph_count", m, _ffi.ptr(idx), _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(colmap),
                  _ffi.ptr(rp), _ffi.ptr(scan_ws), scan_ws.numel(), stream_ptr())
        nnz = int(rp[m].item()) if m else 0
        col_out = torch.empty(nnz, dtype=torch.int32, device=dev)
        val_out = torch.empty(nnz, dtype=torch.float64, device=dev)
        _ffi.call("sgc_subgraph_fill", m, _ffi.ptr(idx), _ffi.ptr(row_ptr), _ffi.ptr(col), _ffi.ptr(val),
                  _ffi.ptr(colmap), _ffi.ptr(rp), _ffi.ptr(col_out), _ffi.ptr(val_out), stream_ptr())
        # ascending idx keeps every row's columns ascending; otherwise sort inside the rows
        flag = torch.empty(1, dtype=torch.int32, device=dev)
        _ffi.call("sgc_csr_check_canonical", m, m, _ffi.ptr(rp), _ffi.ptr(col_out), _ffi.ptr(flag), stream_ptr())
        if int(flag.item()) == 0:
            return canonicalize(m, m, rp, col_out, val_out)
    return rp, col_out, val_out
