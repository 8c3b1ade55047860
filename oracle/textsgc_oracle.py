"""CPU restatement of the propagation part of the reference's TextSGC (TEST INFRASTRUCTURE).

Follows /root/reference/downstream/TextSGC/utils.py:
* ``preprocess_adj`` / ``normalize_adj`` (:59-71): ``adj + eye`` as in the main path, row sums, ``d =
  rowsum^-0.5`` (inf -> 0), then ``adj.dot(D).transpose().dot(D).transpose()``: the value of entry
  (i, j) is ``(a_ij * d_j) * d_i`` in fp64 -- the column factor first, unlike
  normalization.py:12 -- and products that are exactly zero are dropped by scipy's CSR product.
* ``sgc_precompute`` (:131-152): ``feats = spmm(adj, dense_adj[:, idx]).t()``; per-feature max/min
  over the training documents; features with range 0 are dropped; ``(feats - min) / range`` in fp32.
Pinned by tests/golden/textsgc_precompute.npz (generated from the unmodified reference).
"""
import numpy as np

from . import normalization_oracle as no
from . import spmm_oracle as so


def preprocess_adj_oracle(n, indptr, indices, data):
    """CSR (indptr int64, indices int64, data fp64) of normalize_adj(A + I) and d."""
    ptr, cols, vals = no.add_self_loops(n, indptr, indices, data)
    r = no.row_sums_sequential(n, ptr, vals)
    with np.errstate(divide="ignore", invalid="ignore"):
        d = np.power(r, -0.5)
    d[np.isinf(d)] = 0.0
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(ptr))
    left = vals * d[cols]                 # adj.dot(D)
    s = left * d[rows]                    # (...).transpose().dot(D).transpose()
    keep = (left != 0) & (s != 0)
    out_ptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows[keep], minlength=n), out=out_ptr[1:])
    return {"indptr": out_ptr, "indices": cols[keep], "data": s[keep], "d": d}


def dense_columns(n, indptr, indices, data32, idx):
    """``dense(S).astype(float32)[:, idx]`` without the N x N matrix."""
    idx = np.asarray(idx, dtype=np.int64)
    pos = np.full(n, -1, dtype=np.int64)
    pos[idx] = np.arange(idx.size)
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(indptr))
    sel = pos[indices] >= 0
    out = np.zeros((n, idx.size), dtype=np.float32)
    out[rows[sel], pos[indices[sel]]] = data32[sel]
    return out


def sgc_precompute_oracle(n, indptr, indices, data64, index_dict):
    """feat_dict of utils.py:131-152 (degree == 1) as fp32 NumPy arrays, plus the kept feature ids."""
    data32 = np.asarray(data64, dtype=np.float64).astype(np.float32)      # sparse_to_torch_sparse: astype(float32)
    def prop(idx):
        return so.spmm_hop(indptr, indices, data32, dense_columns(n, indptr, indices, data32, idx)).T
    train = prop(index_dict["train"])
    mx, mn = train.max(axis=0, keepdims=True), train.min(axis=0, keepdims=True)
    rng = mx - mn
    keep = np.flatnonzero(rng.squeeze(0) > 0)
    out = {"train": (train[:, keep] - mn[:, keep]) / rng[:, keep]}
    for ph in ("test", "val"):
        out[ph] = (prop(index_dict[ph])[:, keep] - mn[:, keep]) / rng[:, keep]
    return out, keep
