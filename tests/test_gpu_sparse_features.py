"""SURVEY 8f row 3 on the GPU: row_normalize (normalization.py:21-28) and the first hop on sparse
features, against the golden vectors of the unmodified reference (Cora, Citeseer) and the oracle."""
import hashlib

import numpy as np
import pytest
import scipy.sparse as sp
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sgc_b200 import _ffi
    _ffi.lib()
    return torch.device("cuda", 0)


def raw_features(g):
    n = g["featraw_indptr"].size - 1
    f = {2708: 1433, 3327: 3703}[n]
    return sp.csr_matrix((g["featraw_data"], g["featraw_indices"], g["featraw_indptr"]), shape=(n, f))


@pytest.mark.parametrize("ds", ["cora", "citeseer"])
def test_row_normalize_and_sparse_first_hop(dev, ds, monkeypatch):
    import sgc_b200
    from sgc_b200 import _ffi
    from sgc_b200.sparse_features import SparseFeatures, sparse_origin
    g = load_golden(ds)
    raw = raw_features(g)
    # Cora's features stay fp32 (scipy computes the row sums and products in fp32); Citeseer's are fp64
    # because utils.py:53-55 pads them through an fp64 lil matrix
    assert raw.dtype == (np.float32 if ds == "cora" else np.float64)
    sf = sgc_b200.row_normalize(raw.tolil())                      # utils.py:61 hands a lil matrix over
    assert isinstance(sf, SparseFeatures) and sf.shape == raw.shape
    assert sf.dtype == (torch.float32 if ds == "cora" else torch.float64)
    assert np.array_equal(sf.row_ptr.cpu().numpy(), g["featnorm_indptr"])
    assert np.array_equal(sf.col.cpu().numpy(), g["featnorm_indices"])
    assert sf.val.cpu().numpy().dtype == g["featnorm_data"].dtype
    assert np.array_equal(sf.val.cpu().numpy(), g["featnorm_data"])           # bit for bit, in the matrix dtype
    assert sha(np.asarray(sf.todense(), dtype=np.float32)) == str(g["X_sha"])  # what utils.py:74 densifies
    x = sf.to_dense_padded()
    assert x.is_cuda and x.stride(0) % 8 == 0 and sha(x.cpu().numpy()) == str(g["X_sha"])
    assert sparse_origin(x) is sf
    n = g["adj_indptr"].size - 1
    graph = sgc_b200.normalize_csr(n, *(torch.from_numpy(np.ascontiguousarray(g["adj_" + k])).to(dev)
                                        for k in ("indptr", "indices", "data")))
    before = _ffi.launch_count()
    y1, _ = sgc_b200.sgc_precompute(x, graph, 1)
    assert _ffi.launch_count() - before == 1                      # one sparse-X hop, nothing else
    assert y1.is_contiguous() and sha(y1.cpu().numpy()) == str(g["Y1_sha"])
    y2, _ = sgc_b200.sgc_precompute(x, graph, 2)
    assert sha(y2.cpu().numpy()) == str(g["Y_sha"])               # == the reference's torch.spmm bytes
    # the dense path on the same tensor gives the same bytes; a modified tensor loses its sparse origin
    monkeypatch.setenv("SGC_B200_SPARSE_FIRST_HOP", "0")
    assert sparse_origin(x) is None
    y2d, _ = sgc_b200.sgc_precompute(x, graph, 2)
    assert torch.equal(y2d, y2)
    monkeypatch.setenv("SGC_B200_SPARSE_FIRST_HOP", "1")
    x.mul_(2.0)
    assert sparse_origin(x) is None
    y2m, _ = sgc_b200.sgc_precompute(x, graph, 2, chunk=0)
    assert torch.equal(y2m, y2 * 2.0)


def test_row_normalize_edge_cases(dev):
    """Explicit zeros, rows whose sum is zero (inf -> 0, then scipy drops the zero products), empty
    rows, negative sums, fp64 input -- against the oracle's restatement of normalization.py:21-28."""
    import sgc_b200
    from oracle import normalization_oracle as no
    rng = np.random.RandomState(5)
    m = sp.random(60, 45, density=0.2, random_state=rng, format="lil", dtype=np.float32)
    m[3, :] = 0
    m[5, 1], m[5, 2] = 1.0, -1.0                 # sum 0 -> r_inv = inf -> 0 -> products dropped
    m[7, 4] = -3.0
    c = m.tocsr()
    c.data[::11] = 0.0                           # explicit stored zeros
    for dtype in (np.float32, np.float64):
        cc = c.astype(dtype)
        want_ptr, want_idx, want_val = no.row_normalize_oracle(60, cc.indptr, cc.indices, cc.data)
        got = sgc_b200.row_normalize(cc)
        assert got.val.dtype == (torch.float32 if dtype == np.float32 else torch.float64)
        assert np.array_equal(got.row_ptr.cpu().numpy(), want_ptr)
        assert np.array_equal(got.col.cpu().numpy(), want_idx)
        assert np.array_equal(got.val.cpu().numpy(), want_val.astype(dtype))
    with pytest.raises(ValueError):
        from sgc_b200.sparse_features import hop_sparse_x
        g = sgc_b200.aug_normalized_adjacency(sp.identity(10, format="csr"))
        hop_sparse_x(g, sgc_b200.row_normalize(c))


def test_sparse_hop_wide_and_long_rows(dev):
    """Sparse-X hop on a power-law graph with hub rows and a 5000-column feature matrix == the dense
    hop in exact mode; rows too wide for the shared-memory accumulator are refused loudly."""
    import sgc_b200
    from sgc_b200 import _ffi, synth
    from sgc_b200.sparse_features import SparseFeatures, hop_sparse_x
    s = synth.make_shape("reddit", device=dev, scale=0.01, seed=4)
    graph = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    n = s["n"]
    x = sp.random(n, 5000, density=0.004, random_state=np.random.RandomState(1), format="csr", dtype=np.float32)
    sf = SparseFeatures.from_scipy(x, dev)
    got = hop_sparse_x(graph, sf)
    want = sgc_b200.hop(graph, torch.from_numpy(x.toarray()).to(dev), chunk=0)
    assert torch.equal(got, want)
    wide = SparseFeatures.from_scipy(sp.csr_matrix((n, 60000), dtype=np.float32), dev)
    with pytest.raises(_ffi.SgcError):
        hop_sparse_x(graph, wide)


def test_load_citation_on_synthetic_planetoid_files(dev, tmp_path, monkeypatch):
    """utils.load_citation end to end on files in the Planetoid layout the reference reads
    (utils.py:32-90): adjacency normalised on the GPU, features row-normalised and densified on the GPU
    (fp32 CSR in, as in Cora), labels / splits as the reference builds them, and sgc_precompute on the
    returned tensors (sparse first hop) equal to the oracle."""
    import os
    import pickle
    import sgc_b200
    from sgc_b200 import utils as U
    from sgc_b200.sparse_features import sparse_origin
    from oracle import normalization_oracle as no, spmm_oracle as so
    rng = np.random.RandomState(8)
    n_train, n_all, n_test, f, c = 40, 700, 120, 300, 4
    n = n_all + n_test
    feats = sp.random(n, f, density=0.03, random_state=rng, format="csr", dtype=np.float32)
    feats.data[:] = 1.0                                            # bag-of-words, like Cora
    lab = rng.randint(0, c, n)
    onehot = np.eye(c, dtype=np.int32)[lab]
    test_order = n_all + rng.permutation(n_test)                   # ind.<ds>.test.index is unsorted
    test_sorted = np.sort(test_order)
    src, dst = rng.randint(0, n, 2500), rng.randint(0, n, 2500)
    graph = {i: [] for i in range(n)}
    for a, b in zip(src.tolist(), dst.tolist()):
        graph[a].append(b)
        graph[b].append(a)
    os.makedirs(tmp_path / "data")
    # rows n_all.. of the stored matrices are in test.index order (the loader undoes that, utils.py:60-61)
    stored = feats.tolil()
    stored[test_sorted, :] = feats[test_order, :]
    stored = stored.tocsr()
    lab_stored = onehot.copy()
    lab_stored[test_sorted] = onehot[test_order]
    parts = {"x": stored[:n_train], "y": lab_stored[:n_train], "tx": stored[n_all:], "ty": lab_stored[n_all:],
             "allx": stored[:n_all], "ally": lab_stored[:n_all], "graph": graph}
    for key, obj in parts.items():
        with open(tmp_path / "data" / ("ind.synth.%s" % key), "wb") as fh:
            pickle.dump(obj, fh)
    with open(tmp_path / "data" / "ind.synth.test.index", "w") as fh:
        fh.write("\n".join(str(int(i)) for i in test_order) + "\n")
    monkeypatch.chdir(tmp_path)
    adj, features, labels, idx_train, idx_val, idx_test = U.load_citation("synth", "AugNormAdj", True)
    # expected adjacency: symmetric binary graph of the dict (self-loops stay), then normalisation
    a = sp.coo_matrix((np.ones(2 * src.size), (np.r_[src, dst], np.r_[dst, src])), shape=(n, n)).tocsr()
    a.data[:] = 1.0
    a.sort_indices()
    ref = no.aug_normalized_adjacency_oracle(n, a.indptr, a.indices, a.data)
    assert np.array_equal(adj.row_ptr.cpu().numpy(), ref["indptr"])
    assert np.array_equal(adj.col.cpu().numpy().astype(np.int64), ref["indices"])
    assert np.array_equal(adj.values32().cpu().numpy(), ref["data"].astype(np.float32))
    # expected features: the loader's reordering puts node i's features in row i; then row_normalize
    fc = feats.tocsr()
    fc.sort_indices()
    ptr, idx, val = no.row_normalize_oracle(n, fc.indptr, fc.indices, fc.data)
    want_x = np.zeros((n, f), dtype=np.float32)
    want_x[np.repeat(np.arange(n), np.diff(ptr)), idx] = val
    assert features.is_cuda and features.dtype == torch.float32 and np.array_equal(features.cpu().numpy(), want_x)
    assert sparse_origin(features) is not None
    assert np.array_equal(labels.cpu().numpy(), lab)
    assert idx_train.tolist() == list(range(n_train)) and idx_val.tolist() == list(range(n_train, n_train + 500))
    assert idx_test.tolist() == test_sorted.tolist()
    y, _ = sgc_b200.sgc_precompute(features, adj, 2)              # citation.py:32
    want, _ = so.sgc_precompute_oracle(want_x, ref["indptr"], ref["indices"], ref["data"].astype(np.float32), 2)
    assert np.array_equal(y.cpu().numpy(), want)
