"""The reference's own call sequence for the hot path, restated (TEST INFRASTRUCTURE / CPU baseline).

The reference's implementation of this path is a handful of calls into third-party CPU kernels
(scipy.sparse and torch.spmm).  This module makes the same calls in the same order so the CPU
baseline timed by ``bench.py`` exercises the very kernels the reference reaches on the host:

* ``aug_normalized_adjacency``        <- /root/reference/normalization.py:5-12
* ``sparse_mx_to_torch_sparse_tensor``<- /root/reference/utils.py:23-30
* ``sgc_precompute``                  <- /root/reference/utils.py:92-97

It is a port (``cpu_baseline.kind == "port"``): /root/reference does not exist on the GPU box and
the Python reference cannot travel, so the three short functions are re-expressed here; the
golden fixtures under tests/golden/ pin them to the unmodified reference's outputs.
"""
from time import perf_counter

import numpy as np
import scipy.sparse as sp
import torch


def aug_normalized_adjacency(adj):
    n = adj.shape[0]
    aug = sp.coo_matrix(adj + sp.eye(n))
    deg = np.array(aug.sum(1))
    with np.errstate(divide="ignore", invalid="ignore"):
        dis = np.power(deg, -0.5).flatten()
    dis[np.isinf(dis)] = 0.0
    dmat = sp.diags(dis)
    return dmat.dot(aug).dot(dmat).tocoo()


def sparse_mx_to_torch_sparse_tensor(mx):
    coo = mx.tocoo().astype(np.float32)
    idx = torch.from_numpy(np.vstack((coo.row, coo.col)).astype(np.int64))
    val = torch.from_numpy(coo.data)
    return torch.sparse_coo_tensor(idx, val, torch.Size(coo.shape))


def sgc_precompute(features, adj, degree):
    t0 = perf_counter()
    for _ in range(degree):
        features = torch.spmm(adj, features)
    return features, perf_counter() - t0


def standardize(features):
    """utils.py:119."""
    return (features - features.mean(dim=0)) / features.std(dim=0)
