"""Mirror of the reference's utils.py for the propagation path (same public names, argument
order and return shapes), with the heavy steps running in libsgc_b200.so.

    sgc_precompute(features, adj, degree)            /root/reference/utils.py:92-97
    sparse_mx_to_torch_sparse_tensor(sparse_mx)      /root/reference/utils.py:23-30
    preprocess_citation(adj, features, norm)         /root/reference/utils.py:17-21
    load_citation(dataset_str, normalization, cuda)  /root/reference/utils.py:32-90
    load_reddit_data(data_path, normalization, cuda) /root/reference/utils.py:110-131
    loadRedditFromNPZ, parse_index_file, set_seed    /root/reference/utils.py:104-108, 10-15, 99-102

Differences a caller can observe: `adj` is an `SgcGraph` (GPU-resident; pass it back to
`sgc_precompute`, or to `torch.spmm`), Reddit `features` is a `StandardizedFeatures` tensor (raw
values + statistics; behaves as the standardised matrix), and `cuda=False` raises because this
package has no CPU path -- CPU-only runs belong to the reference itself.
"""
import os
import pickle as pkl
import sys

import numpy as np
import scipy.sparse as sp
import torch

from . import _ffi
from .features import StandardizedFeatures, col_stats, to_padded
from .graph import SgcGraph, normalize_csr, require_cuda, upload_csr
from .normalization import aug_normalized_adjacency, fetch_normalization, row_normalize
from .prep import induced_subgraph, symmetrize
from .propagate import sgc_precompute  # noqa: F401  (re-exported: the reference's utils.sgc_precompute)


def parse_index_file(filename):
    """One integer per line (utils.py:10-15)."""
    with open(filename) as fh:
        return [int(line.strip()) for line in fh]


def preprocess_citation(adj, features, normalization="FirstOrderGCN"):
    """Normalise the adjacency (GPU) and row-normalise the features (utils.py:17-21).  As in the
    reference, only 'AugNormAdj' is implemented; the default name is not, and fails the same way."""
    adj = fetch_normalization(normalization)(adj)
    return adj, row_normalize(features)


def sparse_mx_to_torch_sparse_tensor(sparse_mx):
    """fp32 torch sparse COO tensor with int64 [2, nnz] indices (utils.py:23-30).  For an
    `SgcGraph` the tensor is assembled on the GPU from the CSR arrays; a scipy matrix is converted
    on the host exactly as the reference does."""
    if isinstance(sparse_mx, SgcGraph):
        return sparse_mx.to_torch_sparse()
    coo = sparse_mx.tocoo().astype(np.float32)
    index = torch.from_numpy(np.vstack((coo.row, coo.col)).astype(np.int64))
    return torch.sparse_coo_tensor(index, torch.from_numpy(coo.data), torch.Size(coo.shape))


def set_seed(seed, cuda):
    np.random.seed(seed)
    torch.manual_seed(seed)
    if cuda:
        torch.cuda.manual_seed(seed)


def _need_cuda(cuda):
    if not cuda:
        raise _ffi.SgcError("sgc_b200 implements the CUDA path only (cuda=True); use the reference for CPU runs")
    return require_cuda()


def _planetoid_adjacency(graph):
    """Symmetric binary adjacency of the Planetoid dict-of-lists in dict-key order: what
    nx.adjacency_matrix(nx.from_dict_of_lists(graph)) followed by the max-symmetrisation of
    utils.py:62-63 yields (an undirected simple graph: duplicates collapse, self-loops stay)."""
    keys = list(graph.keys())
    pos = {k: i for i, k in enumerate(keys)}
    src = np.fromiter((pos[k] for k in keys for _ in graph[k]), dtype=np.int64)
    dst = np.fromiter((pos[v] for k in keys for v in graph[k]), dtype=np.int64)
    n = len(keys)
    m = sp.coo_matrix((np.ones(src.size, dtype=np.int64), (src, dst)), shape=(n, n)).tocsr()
    m = m + m.T
    m.data[:] = 1
    m.sort_indices()
    return m


def load_citation(dataset_str="cora", normalization="AugNormAdj", cuda=True):
    """Planetoid citation data -> (adj, features, labels, idx_train, idx_val, idx_test), all on
    the GPU (utils.py:32-90).  Reads the same ``data/ind.<name>.*`` files, relative to the cwd."""
    dev = _need_cuda(cuda)
    parts = {}
    for key in ("x", "y", "tx", "ty", "allx", "ally", "graph"):
        with open("data/ind.{}.{}".format(dataset_str.lower(), key), "rb") as fh:
            parts[key] = pkl.load(fh, encoding="latin1") if sys.version_info > (3, 0) else pkl.load(fh)
    test_order = np.asarray(parse_index_file("data/ind.{}.test.index".format(dataset_str)))
    test_sorted = np.sort(test_order)
    tx, ty = parts["tx"], parts["ty"]
    if dataset_str == "citeseer":
        # isolated test nodes are missing from tx/ty: insert zero rows at their positions
        lo, hi = test_order.min(), test_order.max()
        tx_full = sp.lil_matrix((hi - lo + 1, parts["x"].shape[1]))
        tx_full[test_sorted - lo, :] = tx
        ty_full = np.zeros((hi - lo + 1, parts["y"].shape[1]))
        ty_full[test_sorted - lo, :] = ty
        tx, ty = tx_full, ty_full
    features = sp.vstack((parts["allx"], tx)).tolil()
    features[test_order, :] = features[test_sorted, :]
    labels = np.vstack((parts["ally"], ty))
    labels[test_order, :] = labels[test_sorted, :]
    adj = _planetoid_adjacency(parts["graph"])

    adj, features = preprocess_citation(adj, features, normalization)   # adjacency: GPU kernels

    n_train = len(parts["y"])
    # dense fp32 features in padded device storage (row stride a multiple of 8 floats), scattered on
    # the GPU from the row-normalised CSR; the tensor keeps a handle on that CSR so that the first
    # hop of sgc_precompute runs on the stored entries only (sparse_features.py)
    features = features.to_dense_padded()
    labels = torch.from_numpy(labels).long().max(dim=1)[1].to(dev)
    idx_train = torch.arange(n_train, dtype=torch.long, device=dev)
    idx_val = torch.arange(n_train, n_train + 500, dtype=torch.long, device=dev)
    idx_test = torch.as_tensor(test_sorted.tolist(), dtype=torch.long, device=dev)
    return adj, features, labels, idx_train, idx_val, idx_test


def loadRedditFromNPZ(dataset_dir):
    adj = sp.load_npz(dataset_dir + "reddit_adj.npz")
    data = np.load(dataset_dir + "reddit.npz")
    return (adj, data["feats"], data["y_train"], data["y_val"], data["y_test"], data["train_index"],
            data["val_index"], data["test_index"])


def load_reddit_data(data_path="data/", normalization="AugNormAdj", cuda=True):
    """Reddit npz files -> (adj, train_adj, features, labels, train_index, val_index, test_index)
    (utils.py:110-131).  The reference ignores `data_path` and always reads "data/" (reddit.py:38
    even passes the normalisation name in that slot), so a value that is not a directory falls
    back to "data/"."""
    dev = _need_cuda(cuda)
    root = data_path if (isinstance(data_path, str) and os.path.isdir(data_path)) else "data/"
    if not root.endswith(os.sep):
        root += os.sep
    adj, feats, y_train, y_val, y_test, train_index, val_index, test_index = loadRedditFromNPZ(root)
    labels = np.zeros(adj.shape[0])
    labels[train_index] = y_train
    labels[val_index] = y_val
    labels[test_index] = y_test
    raw = to_padded(torch.from_numpy(np.ascontiguousarray(feats, dtype=np.float32)), dev)
    mean, std = col_stats(raw)                                  # utils.py:119 statistics (K3)
    features = StandardizedFeatures(raw, mean, std)             # standardised in place before hop 1
    normalizer = fetch_normalization(normalization)
    if normalizer is not aug_normalized_adjacency:              # unknown name: fail as the reference does
        adj = normalizer(adj + adj.T)
    # graph preparation on the GPU (prep.py): the raw adjacency goes up once
    shape, rp, col, val = upload_csr(adj, dev)
    n = shape[0]
    rp, col, val = symmetrize(n, rp, col, val, op="sum")                        # utils.py:116  adj + adj.T
    trp, tcol, tval = induced_subgraph(n, n, rp, col, val, train_index)         # utils.py:117
    adj = normalize_csr(n, rp, col, val)                                        # K1 + K2
    train_adj = normalize_csr(len(train_index), trp, tcol, tval)
    labels = torch.from_numpy(labels).long().to(dev)
    return adj, train_adj, features, labels, train_index, val_index, test_index
