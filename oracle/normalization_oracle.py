"""NumPy fp64 restatement of the reference's adjacency normalisation (TEST INFRASTRUCTURE).

Follows /root/reference/normalization.py:5-12 line by line, with the behaviour of the scipy
kernels it reaches written out explicitly (scipy is a third-party dependency of the reference,
pinned ``scipy==1.5.2`` in /root/reference/requirements.txt:7 and absent from /root/reference):

* ``adj + sp.eye(n)``  (normalization.py:6)  -> csr_plus_csr: canonical merge of sorted rows,
  coincident (i,i) entries are summed, results that are exactly 0 are dropped.
* ``adj.sum(1)``       (normalization.py:8)  -> coo @ ones: sequential fp64 accumulation in
  stored (row-major, ascending column) order.
* ``np.power(r,-0.5)`` (normalization.py:9), ``inf -> 0`` (normalization.py:10).
* ``D.dot(adj).dot(D)``(normalization.py:12) -> two csr_matmat products: value
  ``(d_i * a_ij) * d_j`` in fp64, products that are exactly 0 are dropped from the structure.

No scipy import: the whole point is an independent restatement that the golden vectors pin.
"""
import numpy as np


def canonicalize_csr(n_rows, indptr, indices, data):
    """Sort columns inside each row and sum duplicate (row, col) entries (scipy
    ``sum_duplicates`` semantics; explicit zeros are kept)."""
    indptr = np.asarray(indptr, dtype=np.int64)
    indices = np.asarray(indices, dtype=np.int64)
    data = np.asarray(data, dtype=np.float64)
    rows = np.repeat(np.arange(n_rows, dtype=np.int64), np.diff(indptr))
    order = np.lexsort((indices, rows))  # stable: duplicates keep stored order
    rows, cols, vals = rows[order], indices[order], data[order]
    if rows.size:
        new = np.ones(rows.size, dtype=bool)
        new[1:] = (rows[1:] != rows[:-1]) | (cols[1:] != cols[:-1])
        starts = np.flatnonzero(new)
        # sequential sum of each duplicate group (np.add.reduceat is sequential per segment)
        vals = np.add.reduceat(vals, starts)
        rows, cols = rows[starts], cols[starts]
    out_ptr = np.zeros(n_rows + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows, minlength=n_rows), out=out_ptr[1:])
    return out_ptr, cols, vals


def add_self_loops(n, indptr, indices, data):
    """``A + I`` as scipy's csr_plus_csr computes it (normalization.py:6).

    Returns canonical CSR (int64 indptr, int64 indices, fp64 data) of ``A+I`` with exact-zero
    results dropped."""
    indptr, indices, data = canonicalize_csr(n, indptr, indices, data)
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(indptr))
    diag = np.arange(n, dtype=np.int64)
    rows2 = np.concatenate([rows, diag])
    cols2 = np.concatenate([indices, diag])
    vals2 = np.concatenate([data, np.ones(n, dtype=np.float64)])
    order = np.lexsort((cols2, rows2))
    rows2, cols2, vals2 = rows2[order], cols2[order], vals2[order]
    new = np.ones(rows2.size, dtype=bool)
    new[1:] = (rows2[1:] != rows2[:-1]) | (cols2[1:] != cols2[:-1])
    starts = np.flatnonzero(new)
    vals2 = np.add.reduceat(vals2, starts) if rows2.size else vals2
    rows2, cols2 = rows2[starts], cols2[starts]
    keep = vals2 != 0  # csr_plus_csr drops results that compare equal to zero (NaN is kept)
    rows2, cols2, vals2 = rows2[keep], cols2[keep], vals2[keep]
    out_ptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows2, minlength=n), out=out_ptr[1:])
    return out_ptr, cols2, vals2


def row_sums_sequential(n, indptr, data):
    """fp64 row sums accumulated in stored order (normalization.py:8 -> coo_matvec).
    np.bincount with weights adds sequentially in input order."""
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(indptr))
    return np.bincount(rows, weights=data, minlength=n).astype(np.float64)


def aug_normalized_adjacency_oracle(n, indptr, indices, data):
    """Restates normalization.py:5-12.  Input: any CSR (possibly unsorted / with duplicates) of
    the n x n matrix ``A``.  Output: ``(indptr int64, indices int64, data fp64)`` of
    ``S = D^-1/2 (A+I) D^-1/2`` in canonical CSR order, which is also the row-major COO order
    the reference returns, plus ``d`` (the fp64 ``D^-1/2`` diagonal) and the CSR of ``A+I``."""
    ptr, cols, vals = add_self_loops(n, indptr, indices, data)
    r = row_sums_sequential(n, ptr, vals)
    with np.errstate(divide="ignore", invalid="ignore"):
        d = np.power(r, -0.5)            # normalization.py:9
    d[np.isinf(d)] = 0.0                 # normalization.py:10
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(ptr))
    with np.errstate(invalid="ignore"):
        left = d[rows] * vals            # D.dot(adj)      (first csr_matmat)
        keep1 = left != 0
        s = left * d[cols]               # (...).dot(D)    (second csr_matmat)
    keep = keep1 & (s != 0)
    rows_k, cols_k, s_k = rows[keep], cols[keep], s[keep]
    out_ptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows_k, minlength=n), out=out_ptr[1:])
    return {
        "indptr": out_ptr, "indices": cols_k, "data": s_k, "d": d,
        "aug_indptr": ptr, "aug_indices": cols, "aug_data": vals, "row_sum": r,
    }


def row_normalize_oracle(n_rows, indptr, indices, data):
    """Restates normalization.py:21-28: ``diag(1/rowsum) @ mx`` with inf -> 0.

    scipy computes this with csr_matmat, which drops products that are exactly zero; the value
    is ``r_inv_i * x_ij`` in the dtype numpy promotes to (fp32 rowsum of an fp32 matrix stays
    fp32: ``np.power(float32, -1)`` is fp32, ``sp.diags`` keeps it, the product is fp32)."""
    indptr = np.asarray(indptr, dtype=np.int64)
    data = np.asarray(data)
    rows = np.repeat(np.arange(n_rows, dtype=np.int64), np.diff(indptr))
    # mx.sum(1) on a CSR matrix -> csr @ ones, sequential, in the matrix dtype
    acc_dtype = data.dtype if data.dtype.kind == "f" else np.float64
    rowsum = np.zeros(n_rows, dtype=acc_dtype)
    np.add.at(rowsum, rows, data.astype(acc_dtype))
    with np.errstate(divide="ignore"):
        r_inv = np.power(rowsum, -1)
    r_inv[np.isinf(r_inv)] = 0.0
    out = r_inv[rows] * data
    keep = out != 0
    out_ptr = np.zeros(n_rows + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows[keep], minlength=n_rows), out=out_ptr[1:])
    return out_ptr, np.asarray(indices)[keep], out[keep]


def standardize_oracle(x):
    """Restates /root/reference/utils.py:118-119 in fp64 then rounds to fp32 (column mean and
    UNBIASED std over all rows).  The reference does this in fp32 torch; the fp64 result is the
    exact value its fp32 arithmetic approximates, so comparisons use a tolerance."""
    x64 = np.asarray(x, dtype=np.float64)
    mean = x64.mean(axis=0)
    std = x64.std(axis=0, ddof=1)
    return ((x64 - mean) / std).astype(np.float32), mean, std
