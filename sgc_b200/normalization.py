"""Mirror of the reference's normalization.py (same names, same argument meaning), backed by the
GPU kernels.

    aug_normalized_adjacency(adj)   /root/reference/normalization.py:5-12
    fetch_normalization(type)       /root/reference/normalization.py:14-19
    row_normalize(mx)               /root/reference/normalization.py:21-28
"""
from .graph import SgcGraph, from_scipy


def aug_normalized_adjacency(adj):
    """``D^-1/2 (A + I) D^-1/2`` with D the row sums of A + I.

    Takes what the reference takes (any scipy sparse matrix).  Returns an `SgcGraph` -- the
    normalised adjacency resident on the GPU -- instead of a scipy COO matrix; ``.tocoo()`` on the
    result gives exactly the reference's return value (fp64, row-major sorted), and
    `utils.sparse_mx_to_torch_sparse_tensor` accepts it."""
    if isinstance(adj, SgcGraph):
        return adj
    return from_scipy(adj)


def fetch_normalization(type):
    """Same one-entry table as the reference; an unknown name yields a zero-argument callable, so
    calling it with an adjacency raises TypeError exactly as the reference does."""
    switcher = {"AugNormAdj": aug_normalized_adjacency}
    return switcher.get(type, lambda: "Invalid normalization technique.")


def row_normalize(mx):
    """Feature rows divided by their sum, inf -> 0 (normalization.py:21-28), on the GPU
    (sgc_row_normalize).  Takes the scipy sparse matrix the reference takes; returns a
    `SparseFeatures` (GPU CSR) that answers ``.todense()`` / ``.tocsr()`` like the scipy matrix the
    reference returns."""
    from .sparse_features import row_normalize as _row_normalize
    return _row_normalize(mx)
