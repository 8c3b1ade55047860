"""Generates the golden fixtures under tests/golden/ by importing the UNMODIFIED reference from
/root/reference and running its own functions on CPU.  Run in the build container only
(/root/reference does not exist on the GPU box):

    python tests/golden/make_golden.py

What is recorded (all produced by reference code; nothing here re-implements it):
  * citation graphs (Cora, Citeseer): the symmetrised adjacency handed to the normaliser
    (utils.py:63), the raw feature matrix handed to row_normalize (utils.py:61), and the outputs of
    load_citation (utils.py:32-90): S = aug_normalized_adjacency (fp64 COO) + its fp32 torch view,
    row-normalised X, labels, splits; Y = sgc_precompute(X, S, 2) (utils.py:92-97) as sha256 +
    sampled rows + column sums; the reference's own val/test accuracy for the tuned run.
  * Pubmed: real graph (ind.pubmed.graph; ind.pubmed.allx is missing from the snapshot), S from the
    reference normaliser, synthetic 500-dim features, Y from the reference's sgc_precompute.
  * edge cases: small matrices with existing self-loops, -1 diagonals, explicit zeros, zero and
    negative row sums, unsorted/duplicate entries, asymmetric weighted input -> reference outputs.
  * a weighted asymmetric real-world block (TextSGC covid PMI/TF-IDF graph, induced sub-matrix on
    the 1800 busiest rows + 1200 random rows).
  * the TextSGC precompute (downstream/TextSGC/utils.py: preprocess_adj, sgc_precompute) on that block.
"""
import hashlib
import os
import pickle
import sys
import warnings

import numpy as np
import scipy.sparse as sp
import torch

warnings.filterwarnings("ignore")
REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REF)
sys.path.insert(1, REPO)
os.chdir(REF)  # the reference opens "data/..." relative to cwd

import normalization as ref_norm  # noqa: E402  (the reference's module)
import utils as ref_utils  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def csr_parts(m, prefix):
    m = sp.csr_matrix(m)
    m.sort_indices()
    return {prefix + "_indptr": m.indptr.astype(np.int64), prefix + "_indices": m.indices.astype(np.int32),
            prefix + "_data": m.data}


def sample_rows(n, k=96, seed=0):
    rng = np.random.RandomState(seed)
    return np.sort(rng.choice(n, size=min(k, n), replace=False)).astype(np.int64)


def citation(ds, epochs):
    captured = {}
    orig_fetch, orig_rn = ref_utils.fetch_normalization, ref_utils.row_normalize

    def fetch(kind):
        fn = orig_fetch(kind)

        def wrapped(adj):
            captured["adj_in"] = adj.copy()
            out = fn(adj)
            captured["S"] = out.copy()
            return out
        return wrapped

    def rn(mx):
        captured["feat_in"] = mx.copy()
        out = orig_rn(mx)
        captured["feat_out"] = out.copy()
        return out

    ref_utils.fetch_normalization, ref_utils.row_normalize = fetch, rn
    try:
        ref_utils.set_seed(42, False)
        adj, features, labels, idx_train, idx_val, idx_test = ref_utils.load_citation(ds, "AugNormAdj", False)
    finally:
        ref_utils.fetch_normalization, ref_utils.row_normalize = orig_fetch, orig_rn
    Y, _ = ref_utils.sgc_precompute(features, adj, 2)
    Y1, _ = ref_utils.sgc_precompute(features, adj, 1)
    S = captured["S"].tocoo()
    out = {}
    out.update(csr_parts(captured["adj_in"].astype(np.float64), "adj"))
    out.update(csr_parts(sp.csr_matrix(captured["feat_in"]), "featraw"))
    out.update(csr_parts(sp.csr_matrix(captured["feat_out"]), "featnorm"))
    out["S_row"], out["S_col"], out["S_data64"] = S.row.astype(np.int64), S.col.astype(np.int64), S.data
    adjc = adj  # torch COO as the reference built it (utils.py:30), NOT coalesced
    out["S_idx_sha"] = sha(adjc._indices().numpy())
    out["S_val32"] = adjc._values().numpy()
    out["S_val32_sha"] = sha(adjc._values().numpy())
    X = features.numpy()
    out["X_sha"], out["X_sum"] = sha(X), np.float64(X.astype(np.float64).sum())
    Yn, Y1n = Y.numpy(), Y1.numpy()
    rows = sample_rows(X.shape[0])
    out["Y_sha"], out["Y_sum"], out["Y_absmax"] = sha(Yn), np.float64(Yn.astype(np.float64).sum()), np.float32(np.abs(Yn).max())
    out["Y1_sha"] = sha(Y1n)
    out["rows"], out["Y_rows"], out["Y1_rows"] = rows, Yn[rows], Y1n[rows]
    out["Y_colsum"] = Yn.astype(np.float64).sum(0)
    out["labels"] = labels.numpy().astype(np.int64)
    out["idx_train"], out["idx_val"], out["idx_test"] = idx_train.numpy(), idx_val.numpy(), idx_test.numpy()
    # the reference's own training loop (citation.py:35-62) on its own Y, seed 42, tuned weight decay
    with open("SGC-tuning/{}.txt".format(ds), "rb") as fh:
        wd = pickle.load(fh)["weight_decay"]
    out["weight_decay"], out["epochs"] = np.float64(wd), np.int64(epochs)
    from models import get_model
    from metrics import accuracy
    import torch.nn.functional as F
    import torch.optim as optim
    ref_utils.set_seed(42, False)
    model = get_model("SGC", X.shape[1], int(labels.max()) + 1, 0, 0, False)
    opt = optim.Adam(model.parameters(), lr=0.2, weight_decay=wd)
    for _ in range(epochs):
        model.train(); opt.zero_grad()
        loss = F.cross_entropy(model(Y[idx_train]), labels[idx_train]); loss.backward(); opt.step()
    model.eval()
    with torch.no_grad():
        out["acc_val"] = np.float64(accuracy(model(Y[idx_val]), labels[idx_val]).item())
        out["acc_test"] = np.float64(accuracy(model(Y[idx_test]), labels[idx_test]).item())
    print(ds, "N", X.shape, "nnz(S)", S.nnz, "val/test acc", out["acc_val"], out["acc_test"],
          "Ssha", out["S_val32_sha"][:16], "Ysha", out["Y_sha"][:16])
    np.savez_compressed(os.path.join(HERE, ds + ".npz"), **out)


def pubmed():
    import networkx as nx
    with open("data/ind.pubmed.graph", "rb") as fh:
        graph = pickle.load(fh, encoding="latin1")
    adj = nx.adjacency_matrix(nx.from_dict_of_lists(graph))          # utils.py:62
    adj = adj + adj.T.multiply(adj.T > adj) - adj.multiply(adj.T > adj)  # utils.py:63
    S = ref_norm.aug_normalized_adjacency(adj)
    St = ref_utils.sparse_mx_to_torch_sparse_tensor(S).float()
    from sgc_b200 import synth
    X = synth.tfidf_features(adj.shape[0], 500, density=0.1, seed=1)
    Y, _ = ref_utils.sgc_precompute(X, St, 2)
    Yn = Y.numpy()
    rows = sample_rows(adj.shape[0])
    out = {}
    out.update(csr_parts(sp.csr_matrix(adj).astype(np.float64), "adj"))
    Sc = S.tocoo()
    out["S_row"], out["S_col"], out["S_data64"] = Sc.row.astype(np.int64), Sc.col.astype(np.int64), Sc.data
    out["S_val32"] = St._values().numpy()
    out["S_val32_sha"], out["S_idx_sha"] = sha(St._values().numpy()), sha(St._indices().numpy())
    out["X_sha"] = sha(X.numpy())
    out["Y_sha"], out["Y_sum"] = sha(Yn), np.float64(Yn.astype(np.float64).sum())
    out["rows"], out["Y_rows"], out["Y_colsum"] = rows, Yn[rows], Yn.astype(np.float64).sum(0)
    print("pubmed N", adj.shape, "nnz(S)", Sc.nnz, "Ysha", out["Y_sha"][:16])
    np.savez_compressed(os.path.join(HERE, "pubmed.npz"), **out)


def edge_cases():
    rng = np.random.RandomState(1234)
    cases = {}

    def add(name, m):
        m = sp.csr_matrix(m)
        with np.errstate(all="ignore"):
            S = ref_norm.aug_normalized_adjacency(m).tocoo()
        Sc = sp.csr_matrix(S)  # canonical row-major order
        Sc.sort_indices()
        cases[name] = 1
        out[name + "/n"] = np.int64(m.shape[0])
        out[name + "/in_indptr"], out[name + "/in_indices"], out[name + "/in_data"] = \
            m.indptr.astype(np.int64), m.indices.astype(np.int32), m.data.astype(np.float64)
        out[name + "/S_row"], out[name + "/S_col"], out[name + "/S_data"] = \
            S.row.astype(np.int64), S.col.astype(np.int64), S.data.astype(np.float64)

    out = {}
    n = 40
    base = sp.random(n, n, density=0.12, random_state=rng, format="csr")
    base.data[:] = 1.0
    sym = ((base + base.T) > 0).astype(np.float64)
    add("binary_sym", sym)
    m = sym.tolil(); m[3, 3] = 1.0; m[7, 7] = 5.0; add("self_loops", m.tocsr())
    m = sym.tolil(); m[4, 4] = -1.0; add("diag_minus_one", m.tocsr())
    m = sym.tocsr().copy(); m.data[::7] = 0.0; add("explicit_zeros", m)
    m = sp.lil_matrix((n, n)); m[0, 1] = 1.0; m[0, 2] = -2.0; m[1, 0] = 1.0; m[2, 0] = 3.0; m[5, 6] = 1.0
    add("zero_rowsum", m.tocsr())          # row 0: 1 - 2 + 1(self) = 0  -> d_0 = inf -> 0
    m = sym.tolil(); m[9, 10] = -50.0; add("negative_rowsum", m.tocsr())   # NaN propagation
    w = sp.random(n, n, density=0.15, random_state=rng, format="csr"); add("weighted_asym", w)
    add("empty", sp.csr_matrix((n, n)))
    add("one_by_one", sp.csr_matrix(np.array([[0.0]])))
    # non-canonical: unsorted columns and duplicate entries
    r = rng.randint(0, n, 300); c = rng.randint(0, n, 300); v = rng.rand(300)
    coo = sp.coo_matrix((v, (r, c)), shape=(n, n))
    indptr = np.zeros(n + 1, np.int64); np.cumsum(np.bincount(r, minlength=n), out=indptr[1:])
    order = np.argsort(r, kind="stable")
    nc = sp.csr_matrix((v[order], c[order], indptr), shape=(n, n))  # rows grouped, cols unsorted + dups
    add("non_canonical", nc)
    out["names"] = np.array(sorted(cases))
    np.savez_compressed(os.path.join(HERE, "edge_cases.npz"), **out)
    print("edge cases:", sorted(cases))


def textsgc_block():
    path = os.path.join(REF, "downstream/TextSGC_Bio/data/ind.covid_19_production.B.adj")
    if not os.path.exists(path):
        print("textsgc fixture missing; skipped"); return
    with open(path, "rb") as fh:
        adj = pickle.load(fh, encoding="latin1")
    adj = sp.csr_matrix(adj)
    rng = np.random.RandomState(7)
    rowlen = np.diff(adj.indptr)
    busy = np.argsort(-rowlen, kind="stable")[:1800]
    rest = rng.choice(adj.shape[0], 1200, replace=False)
    sel = np.unique(np.concatenate([busy, rest]))
    blk = adj[sel, :][:, sel].tocsr()   # weighted, asymmetric, with empty rows
    S = ref_norm.aug_normalized_adjacency(blk).tocoo()
    out = {}
    out.update(csr_parts(blk.astype(np.float64), "adj"))
    out["S_row"], out["S_col"], out["S_data64"] = S.row.astype(np.int64), S.col.astype(np.int64), S.data
    print("textsgc block", blk.shape, "nnz", blk.nnz, "empty rows", int((np.diff(blk.indptr) == 0).sum()),
          "asym", (abs(blk - blk.T) > 0).nnz)
    np.savez_compressed(os.path.join(HERE, "textsgc_block.npz"), **out)


def textsgc_precompute():
    """downstream/TextSGC/utils.py (imported unmodified; its dead `eigsh` import no longer exists in
    this scipy and is stubbed, `.cuda()` is made a no-op because this container has no GPU):
    preprocess_adj -> sparse_to_torch_sparse / sparse_to_torch_dense -> sgc_precompute on the same
    covid PMI/TF-IDF block, with random disjoint train/val/test column sets."""
    import importlib.util
    import types
    blk = np.load(os.path.join(HERE, "textsgc_block.npz"))
    n = blk["adj_indptr"].size - 1
    adj = sp.csr_matrix((blk["adj_data"], blk["adj_indices"], blk["adj_indptr"]), shape=(n, n))
    stub = types.ModuleType("scipy.sparse.linalg.eigen.arpack")
    stub.eigsh = None
    sys.modules.setdefault("scipy.sparse.linalg.eigen", types.ModuleType("scipy.sparse.linalg.eigen"))
    sys.modules["scipy.sparse.linalg.eigen.arpack"] = stub
    spec = importlib.util.spec_from_file_location("textsgc_utils", os.path.join(REF, "downstream/TextSGC/utils.py"))
    T = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(T)
    adj32 = adj.astype(np.float32)                                  # load_corpus: utils.py:53
    sp_adj = T.preprocess_adj(adj32)                                # utils.py:54
    t_adj = T.sparse_to_torch_sparse(sp_adj, device="cpu")          # train.py:49
    dense = T.sparse_to_torch_dense(sp_adj, device="cpu")           # train.py:105
    rng = np.random.RandomState(11)
    perm = rng.permutation(n)
    index_dict = {"train": np.sort(perm[:300]).tolist(), "val": np.sort(perm[300:380]).tolist(),
                  "test": np.sort(perm[380:500]).tolist()}
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        feat_dict, _ = T.sgc_precompute(t_adj, dense, 1, index_dict)    # train.py:106
    finally:
        torch.Tensor.cuda = orig_cuda
    S = sp_adj.tocoo()
    out = {}
    out.update(csr_parts(adj32.astype(np.float64), "adj"))
    out["S_row"], out["S_col"], out["S_data64"] = S.row.astype(np.int64), S.col.astype(np.int64), S.data
    for ph in ("train", "val", "test"):
        a = feat_dict[ph].numpy()
        out["idx_" + ph] = np.asarray(index_dict[ph], dtype=np.int64)
        out["feats_%s_shape" % ph] = np.asarray(a.shape, dtype=np.int64)
        out["feats_%s_sha" % ph] = sha(a)
        out["feats_%s_head" % ph] = a[:12].copy()
        out["feats_%s_sum" % ph] = np.float64(a.astype(np.float64).sum())
    print("textsgc precompute", {ph: tuple(feat_dict[ph].shape) for ph in feat_dict})
    np.savez_compressed(os.path.join(HERE, "textsgc_precompute.npz"), **out)


if __name__ == "__main__":
    citation("cora", 100)
    citation("citeseer", 150)
    pubmed()
    edge_cases()
    textsgc_block()
    textsgc_precompute()
