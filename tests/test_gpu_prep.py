"""Graph preparation on the GPU (sgc_b200/prep.py, csrc/graph_prep.cu) against scipy doing what the
reference's loaders do (/root/reference/utils.py:63, 116-117), and the Reddit loader end to end on
synthetic files written in the reference's npz layout (utils.py:104-108).  Integer structure and
the fp64 values of the prepared adjacency are bit-exact; normalised values are within 4 ulp(fp64) of
a host computation because numpy's pow() is itself host-dependent in the last bit."""
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def ulp64(a, b):
    return np.abs(a.view(np.int64) - b.view(np.int64)).max() if a.size else 0


def to_dev(m, dev):
    m = sp.csr_matrix(m)
    return (torch.from_numpy(m.indptr.astype(np.int64)).to(dev), torch.from_numpy(m.indices.astype(np.int32)).to(dev),
            torch.from_numpy(m.data.astype(np.float64)).to(dev))


def canon(m):
    m = sp.csr_matrix(m).copy()
    m.sum_duplicates()
    m.eliminate_zeros()
    m.sort_indices()
    return m


def same(got, want):
    rp, col, val = (t.cpu().numpy() for t in got)
    want = sp.csr_matrix(want)
    assert np.array_equal(rp, want.indptr.astype(np.int64))
    assert np.array_equal(col, want.indices.astype(np.int32))
    assert np.array_equal(val, want.data.astype(np.float64))


def random_graph(n, density, seed, weights="int", hubs=0):
    rng = np.random.RandomState(seed)
    m = sp.random(n, n, density=density, random_state=rng, format="lil")
    for h in range(hubs):                       # long rows and columns
        m[h, rng.rand(n) < 0.6] = 1.0
        m[rng.rand(n) < 0.5, h] = 1.0
    m = sp.csr_matrix(m)
    if weights == "int":
        m.data = rng.randint(1, 6, size=m.nnz).astype(np.float64)
    elif weights == "real":
        m.data = rng.rand(m.nnz) + 0.1
    m.setdiag(rng.randint(0, 3, size=n).astype(np.float64))       # some existing self-loops, some zero
    m = sp.csr_matrix(m)
    m.eliminate_zeros()
    m.sort_indices()
    return m


def test_radix_sort_is_stable(dev):
    from sgc_b200 import prep
    g = torch.Generator(device="cpu").manual_seed(0)
    for n, bits in ((1, 3), (255, 8), (4097, 9), (300_000, 17), (1_000_003, 27)):
        key = torch.randint(0, 1 << bits, (n,), generator=g, dtype=torch.int32).to(dev)
        p32 = torch.arange(n, dtype=torch.int32, device=dev)
        p64 = torch.rand(n, generator=g, dtype=torch.float64).to(dev)
        ko, po, do = prep._sort_pairs(key, p32, p64, bits)
        want_k, order = torch.sort(key.long(), stable=True)
        assert torch.equal(ko.long(), want_k)
        assert torch.equal(po.long(), order)                 # equal keys keep their original order
        assert torch.equal(do, p64[order])


@pytest.mark.parametrize("n,density,hubs", [(1, 1.0, 0), (37, 0.2, 0), (900, 0.02, 3), (5000, 0.004, 2)])
def test_symmetrize_matches_scipy(dev, n, density, hubs):
    from sgc_b200 import prep
    for weights in ("int", "real"):
        a = random_graph(n, density, seed=n, weights=weights, hubs=hubs)
        got = prep.symmetrize(n, *to_dev(a, dev), op="sum")
        same(got, canon(a + a.T))                                            # utils.py:116
    a = random_graph(n, density, seed=n + 1, weights="int", hubs=hubs)
    ref = a + a.T.multiply(a.T > a) - a.multiply(a.T > a)                    # utils.py:63
    got = prep.symmetrize(n, *to_dev(a, dev), op="max")
    same(got, canon(ref))


def test_symmetrize_unsorted_duplicated_input(dev):
    """The stream sort handles input that is not canonical: duplicates are summed first-to-last."""
    from sgc_b200 import prep
    rng = np.random.RandomState(5)
    n, m = 300, 6000
    r, c = rng.randint(0, n, m), rng.randint(0, n, m)
    v = rng.randint(1, 4, m).astype(np.float64)
    order = np.argsort(r, kind="stable")
    r, c, v = r[order], c[order], v[order]
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=n), out=indptr[1:])
    raw = (torch.from_numpy(indptr).to(dev), torch.from_numpy(c.astype(np.int32)).to(dev), torch.from_numpy(v).to(dev))
    a = canon(sp.csr_matrix((v, c, indptr), shape=(n, n)))
    same(prep.canonicalize(n, n, *raw), a)
    same(prep.symmetrize(n, *raw, op="sum"), canon(a + a.T))


@pytest.mark.parametrize("ordered", [True, False])
def test_induced_subgraph_matches_scipy(dev, ordered):
    from sgc_b200 import prep
    n = 3000
    a = random_graph(n, 0.005, seed=9, weights="real", hubs=2)
    a = canon(a + a.T)
    rng = np.random.RandomState(1)
    idx = rng.choice(n, size=1700, replace=False)
    if ordered:
        idx = np.sort(idx)
    got = prep.induced_subgraph(n, n, *to_dev(a, dev), idx)
    same(got, canon(a[idx, :][:, idx]))                                      # utils.py:117
    same(prep.induced_subgraph(n, n, *to_dev(a, dev), np.zeros(0, dtype=np.int64)), sp.csr_matrix((0, 0)))
    with pytest.raises(IndexError):
        prep.induced_subgraph(n, n, *to_dev(a, dev), np.array([0, n]))


def test_large_noncanonical_input_takes_the_radix_path(dev):
    """normalize_csr on > 2^18 unsorted entries with duplicates (scipy canonicalises implicitly in
    `adj + sp.eye`, normalization.py:6): structure and values as the oracle's."""
    import sgc_b200
    from oracle import normalization_oracle as no
    rng = np.random.RandomState(3)
    n, m = 20_000, 400_000
    r, c = np.sort(rng.randint(0, n, m)), rng.randint(0, n, m)
    v = rng.randint(1, 3, m).astype(np.float64)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=n), out=indptr[1:])
    g = sgc_b200.normalize_csr(n, torch.from_numpy(indptr).to(dev), torch.from_numpy(c.astype(np.int32)).to(dev),
                               torch.from_numpy(v).to(dev))
    ref = no.aug_normalized_adjacency_oracle(n, indptr, c, v)
    assert np.array_equal(g.row_ptr.cpu().numpy(), ref["indptr"])
    assert np.array_equal(g.col.cpu().numpy().astype(np.int64), ref["indices"])
    # D^-1/2 comes from numpy's pow() in the oracle, which is host-dependent in the last fp64 bit
    # (glibc vs SVML, tests/test_oracle_golden.py::test_rsqrt_refined_vs_pow): a few ulp on the product
    assert np.array_equal(g.rowsum.cpu().numpy(), ref["row_sum"])
    assert ulp64(g.values64().cpu().numpy(), ref["data"]) <= 4


def test_load_reddit_data_on_synthetic_files(dev, tmp_path, monkeypatch):
    """utils.load_reddit_data on files in the reference's layout: adj / train_adj equal the
    reference's scipy pipeline (symmetrise, sub-graph, normalise) bit for bit, features behave as
    the standardised matrix, and the inductive precompute of reddit.py:43-45 matches the oracle."""
    import sgc_b200
    from sgc_b200 import synth, utils as U
    from oracle import reference_port as rp, spmm_oracle as so
    s = synth.make_shape("reddit", device="cpu", scale=0.01, seed=4)
    n, f = s["n"], 48
    a = sp.csr_matrix((np.ones(s["col"].numel()), s["col"].numpy(), s["row_ptr"].numpy()), shape=(n, n))
    a = sp.triu(a, k=1).tocsr()                      # a directed half: the loader symmetrises it
    feats = s["x"][:, :f].numpy().astype(np.float32)
    rng = np.random.RandomState(0)
    perm = rng.permutation(n)
    tr, va, te = np.sort(perm[:n * 6 // 10]), np.sort(perm[n * 6 // 10:n * 7 // 10]), np.sort(perm[n * 7 // 10:])
    y = rng.randint(0, 5, n)
    os.makedirs(tmp_path / "data")
    sp.save_npz(str(tmp_path / "data" / "reddit_adj.npz"), a)
    np.savez(str(tmp_path / "data" / "reddit.npz"), feats=feats, y_train=y[tr], y_val=y[va], y_test=y[te],
             train_index=tr, val_index=va, test_index=te)
    monkeypatch.chdir(tmp_path)
    adj, train_adj, features, labels, i_tr, i_va, i_te = U.load_reddit_data("AugNormAdj")
    full = a + a.T
    want = rp.aug_normalized_adjacency(full).tocsr()
    want_tr = rp.aug_normalized_adjacency(full[tr, :][:, tr]).tocsr()
    for got, ref in ((adj, want), (train_adj, want_tr)):
        ref.sort_indices()
        assert np.array_equal(got.row_ptr.cpu().numpy(), ref.indptr.astype(np.int64))
        assert np.array_equal(got.col.cpu().numpy(), ref.indices.astype(np.int32))
        assert ulp64(got.values64().cpu().numpy(), ref.data) <= 4       # numpy pow(): last-bit host dependence
    assert np.array_equal(labels.cpu().numpy(), y) and np.array_equal(i_tr, tr)
    xs = torch.from_numpy(feats)
    xs = (xs - xs.mean(dim=0)) / xs.std(dim=0)                                # utils.py:119 on the CPU
    assert torch.allclose(features.standardized().cpu(), xs, rtol=1e-5, atol=1e-6)
    # reddit.py:43-45: transductive and inductive precompute, chunk=0 -> the oracle's summation order
    std_gpu = features.standardized().cpu().numpy()
    yfull, _ = sgc_b200.sgc_precompute(features, adj, 2, chunk=0)
    wfull, _ = so.sgc_precompute_oracle(std_gpu, want.indptr, want.indices, adj.values32().cpu().numpy(), 2)
    assert np.array_equal(yfull.cpu().numpy(), wfull)
    ytr, _ = sgc_b200.sgc_precompute(features[i_tr], train_adj, 2, chunk=0)
    wtr, _ = so.sgc_precompute_oracle(std_gpu[tr], want_tr.indptr, want_tr.indices,
                                      train_adj.values32().cpu().numpy(), 2)
    assert np.array_equal(ytr.cpu().numpy(), wtr)
