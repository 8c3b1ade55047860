"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the reference's golden
vectors.  Integer / index work must be bit-exact; fp32 propagation is bit-exact for rows that are
not split and within |a-b| <= 1e-6 + 1e-5*|b| (north_star's tolerance) for split rows."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-6


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import sgc_b200  # noqa: F401  (fails loudly if libsgc_b200.so is missing)
    from sgc_b200 import _ffi
    _ffi.lib()
    return torch.device("cuda", 0)


def to_dev(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def gpu_graph(g, dev, prefix="adj"):
    import sgc_b200
    n = g[prefix + "_indptr"].size - 1
    return sgc_b200.normalize_csr(n, to_dev(g[prefix + "_indptr"], dev), to_dev(g[prefix + "_indices"], dev),
                                  to_dev(g[prefix + "_data"], dev))


def coo_rows(graph):
    rp = graph.row_ptr.cpu().numpy()
    return np.repeat(np.arange(rp.size - 1, dtype=np.int64), np.diff(rp))


def ulp64(a, b):
    return np.abs(a.view(np.int64) - b.view(np.int64)).max() if a.size else 0


# ---- K1 / K2 -------------------------------------------------------------------------------------
@pytest.mark.parametrize("ds", ["cora", "citeseer", "pubmed", "textsgc_block"])
def test_normalisation_matches_reference(dev, ds):
    g = load_golden(ds)
    graph = gpu_graph(g, dev)
    assert np.array_equal(coo_rows(graph), g["S_row"])                       # row_ptr bit-exact
    assert np.array_equal(graph.col.cpu().numpy().astype(np.int64), g["S_col"])   # col_idx bit-exact
    s64 = graph.values64().cpu().numpy()
    # fp64 values: d_i and d_j are each within 1 ulp of numpy's pow() (itself host-dependent in the
    # last bit: glibc vs SVML), so the product (d_i*a)*d_j is within a few ulp(fp64); the fp32
    # rounding the reference feeds to torch.spmm is compared bit for bit below
    assert ulp64(s64, g["S_data64"]) <= 4
    if "S_val32_sha" in g.files:
        assert sha(graph.values32().cpu().numpy()) == str(g["S_val32_sha"])   # the fp32 S the reference feeds torch.spmm
        idx = graph.coo_indices().cpu().numpy()
        assert sha(idx) == str(g["S_idx_sha"])                                # int64 [2,nnz] of utils.py:26-27
    # against the oracle restatement as well (same inputs)
    from oracle import normalization_oracle as no
    n = g["adj_indptr"].size - 1
    r = no.aug_normalized_adjacency_oracle(n, g["adj_indptr"], g["adj_indices"], g["adj_data"])
    assert np.array_equal(graph.row_ptr.cpu().numpy(), r["indptr"])
    assert np.array_equal(graph.rowsum.cpu().numpy().view(np.int64), r["row_sum"].view(np.int64))
    assert ulp64(graph.dinv.cpu().numpy(), r["d"]) <= 1


def test_normalisation_edge_cases(dev):
    import sgc_b200
    g = load_golden("edge_cases")
    for name in [str(x) for x in g["names"]]:
        n = int(g[name + "/n"])
        graph = sgc_b200.normalize_csr(n, to_dev(g[name + "/in_indptr"], dev), to_dev(g[name + "/in_indices"], dev),
                                       to_dev(g[name + "/in_data"], dev))
        order = np.lexsort((g[name + "/S_col"], g[name + "/S_row"]))
        assert np.array_equal(coo_rows(graph), g[name + "/S_row"][order]), name
        assert np.array_equal(graph.col.cpu().numpy().astype(np.int64), g[name + "/S_col"][order]), name
        got, ref = graph.values64().cpu().numpy(), g[name + "/S_data"][order]
        assert np.array_equal(np.isnan(got), np.isnan(ref)), name
        ok = ~np.isnan(ref)
        if name == "non_canonical":
            assert np.allclose(got[ok], ref[ok], rtol=1e-15, atol=0), name
        else:
            assert ulp64(got[ok], ref[ok]) <= 4, name


def test_scipy_entry_point(dev):
    """normalization.aug_normalized_adjacency on a scipy matrix == the golden reference output,
    and .tocoo() hands back the reference's return type."""
    import scipy.sparse as sp
    import normalization as drop_in
    g = load_golden("citeseer")
    n = g["adj_indptr"].size - 1
    adj = sp.csr_matrix((g["adj_data"], g["adj_indices"], g["adj_indptr"]), shape=(n, n))
    S = drop_in.fetch_normalization("AugNormAdj")(adj)
    coo = S.tocoo()
    assert sp.isspmatrix_coo(coo) and coo.dtype == np.float64
    assert np.array_equal(coo.row, g["S_row"]) and np.array_equal(coo.col, g["S_col"])
    assert ulp64(coo.data, g["S_data64"]) <= 4
    import utils as drop_utils
    t = drop_utils.sparse_mx_to_torch_sparse_tensor(S)
    assert t.is_sparse and t.dtype == torch.float32 and t._indices().dtype == torch.int64
    assert sha(t._values().cpu().numpy()) == str(g["S_val32_sha"])
    with pytest.raises(TypeError):
        drop_in.fetch_normalization("FirstOrderGCN")(adj)      # same failure mode as the reference


# ---- K4 ------------------------------------------------------------------------------------------
def dense_features(g):
    n = g["featnorm_indptr"].size - 1
    f = {2708: 1433, 3327: 3703}[n]
    x = np.zeros((n, f), dtype=np.float32)
    rows = np.repeat(np.arange(n), np.diff(g["featnorm_indptr"]))
    x[rows, g["featnorm_indices"]] = g["featnorm_data"].astype(np.float32)
    return x


@pytest.mark.parametrize("ds", ["cora", "citeseer"])
def test_precompute_bit_exact_citation(dev, ds):
    import sgc_b200
    g = load_golden(ds)
    graph = gpu_graph(g, dev)
    x = dense_features(g)
    assert sha(x) == str(g["X_sha"])
    xd = to_dev(x, dev)
    y1, _ = sgc_b200.sgc_precompute(xd, graph, 1)
    assert sha(y1.cpu().numpy()) == str(g["Y1_sha"])
    y2, secs = sgc_b200.sgc_precompute(xd, graph, 2)
    assert secs > 0 and y2.shape == xd.shape and y2.is_contiguous()
    assert sha(y2.cpu().numpy()) == str(g["Y_sha"])            # == reference torch.spmm bytes
    y0, _ = sgc_b200.sgc_precompute(xd, graph, 0)
    assert y0 is xd
    # the materialised-S path (torch sparse tensor in, as a reference caller would pass)
    st = graph.to_torch_sparse()
    y2b, _ = sgc_b200.sgc_precompute(xd, st, 2)
    assert sha(y2b.cpu().numpy()) == str(g["Y_sha"])
    # torch.spmm(adj, x) dispatches to the same kernel (models.py:38)
    assert torch.equal(torch.spmm(graph, xd), y1)


def test_precompute_bit_exact_pubmed_all_slabs(dev):
    import sgc_b200
    from sgc_b200 import synth
    g = load_golden("pubmed")
    graph = gpu_graph(g, dev)
    x = synth.tfidf_features(19717, 500, density=0.1, seed=1)
    assert sha(x.numpy()) == str(g["X_sha"])
    xd = x.to(dev)
    for slab in (0, 32, 64, 128, 256):
        for hints in (0, 3):
            y, _ = sgc_b200.sgc_precompute(xd, graph, 2, slab=slab, hints=hints)
            assert sha(y.cpu().numpy()) == str(g["Y_sha"]), (slab, hints)


@pytest.mark.parametrize("f", [1, 3, 30, 100, 128, 257, 602])
@pytest.mark.parametrize("k", [1, 2, 3])
def test_precompute_alignment_cases(dev, f, k):
    """Feature widths with 4-, 8- and 16-byte aligned rows, partial slabs, K up to 3 -- bit-exact
    against the oracle's sequential-FMA loop."""
    import sgc_b200
    from oracle import spmm_oracle as so
    g = load_golden("cora")
    graph = gpu_graph(g, dev)
    rng = np.random.RandomState(f * 10 + k)
    x = rng.standard_normal((2708, f)).astype(np.float32)
    want, _ = so.sgc_precompute_oracle(x, graph.row_ptr.cpu().numpy(), graph.col.cpu().numpy(),
                                       graph.values32().cpu().numpy(), k)
    for slab in (0, 32, 64):
        y, _ = sgc_b200.sgc_precompute(to_dev(x, dev), graph, k, slab=slab)
        assert np.array_equal(y.cpu().numpy(), want), (f, k, slab)


def powerlaw_case(dev, scale=0.02, seed=3):
    import sgc_b200
    from sgc_b200 import synth
    s = synth.make_shape("reddit", device=dev, scale=scale, seed=seed)
    graph = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    return s, graph


def test_long_rows_split_and_exact_mode(dev):
    """Power-law graph with hub rows far longer than the split threshold: exact mode (chunk=0) is
    bit-identical to the oracle; the split path stays within north_star's tolerance and is
    deterministic."""
    import sgc_b200
    from oracle import spmm_oracle as so
    s, graph = powerlaw_case(dev)
    rp, col, val = graph.row_ptr.cpu().numpy(), graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    assert np.diff(rp).max() > 600
    x = s["x"][:, :96].contiguous()
    want, _ = so.sgc_precompute_oracle(x.cpu().numpy(), rp, col, val, 2)
    exact, _ = sgc_b200.sgc_precompute(x, graph, 2, chunk=0)
    assert np.array_equal(exact.cpu().numpy(), want)
    for chunk in (64, 512):
        a, _ = sgc_b200.sgc_precompute(x, graph, 2, chunk=chunk)
        b, _ = sgc_b200.sgc_precompute(x, graph, 2, chunk=chunk)
        assert torch.equal(a, b)                                         # run-to-run deterministic
        assert np.allclose(a.cpu().numpy(), want, rtol=RTOL, atol=ATOL)
        assert graph.plan(chunk).partial_rows > 0


def test_hot_column_tags(dev, monkeypatch):
    """Hot-column tagging (L2 pinning of hub rows for operands larger than L2) changes cache
    policies only: results are bit-identical, the tagged set is exactly the columns whose gather
    count reaches the chosen threshold, and it respects the row budget."""
    import sgc_b200
    from sgc_b200 import graph as graph_mod
    s, graph = powerlaw_case(dev, scale=0.02, seed=13)
    x = s["x"][:, :100].contiguous()
    monkeypatch.setenv("SGC_B200_HOT_MB", "0")
    base, _ = sgc_b200.sgc_precompute(x, graph, 2)
    plan = graph.plan(sgc_b200.propagate.default_chunk())
    assert plan.col_tagged is None
    monkeypatch.setattr(graph_mod, "L2_SLAB_BUDGET", 1024)          # pretend nothing fits in L2
    monkeypatch.setenv("SGC_B200_HOT_MB", "0.25")                   # 256 KiB / 416 B = 630 rows
    got, _ = sgc_b200.sgc_precompute(x, graph, 2)
    assert torch.equal(got, base)
    assert plan.col_tagged is not None and 0 < plan.n_hot <= (256 * 1024) // 416
    tagged = plan.col_tagged.cpu().numpy()
    col = graph.col.cpu().numpy()
    assert np.array_equal(tagged & 0x7fffffff, col)
    counts = np.bincount(col, minlength=graph.n_cols)
    hot_cols = np.unique(col[tagged < 0])
    assert hot_cols.size == plan.n_hot
    thr = counts[hot_cols].min()
    assert np.array_equal(hot_cols, np.flatnonzero(counts >= thr))   # a pure count threshold
    assert np.count_nonzero(counts >= thr - 1) > (256 * 1024) // 416 or thr == 2   # and the lowest that fits
    # the fused-weights and peer-push variants take tags too
    monkeypatch.setenv("SGC_B200_FUSED_NORM", "1")
    got2, _ = sgc_b200.sgc_precompute(x, graph, 2)
    assert torch.equal(got2, base)
    monkeypatch.setenv("SGC_B200_HOT_MB", "0")
    again, _ = sgc_b200.sgc_precompute(x, graph, 2)
    assert plan.col_tagged is None and torch.equal(again, base)


@pytest.mark.parametrize("f", [100, 130, 7])
def test_column_blocked_hop(dev, monkeypatch, f):
    """Column-blocked hop (SGC_HOP_ACCUMULATE; operands several times larger than L2): every block's launch
    continues the rows' FMA chains from y, so exact mode (chunk=0) is bit-identical to the single-pass hop and
    to the oracle for any number of blocks; with row splitting, split rows stay within tolerance and results are
    deterministic; blocks tile the stored entries; the K-hop driver takes the same path hop by hop."""
    import sgc_b200
    from sgc_b200.propagate import hop_col_blocked
    from oracle import spmm_oracle as so
    s, graph = powerlaw_case(dev, scale=0.02, seed=17)
    n = graph.n_rows
    x = s["x"][:, :f].contiguous()
    rp, col, val = graph.row_ptr.cpu().numpy(), graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "0")
    plain = sgc_b200.hop(graph, x, chunk=0)
    assert np.array_equal(plain.cpu().numpy(), so.spmm_hop(rp, col, val, x.cpu().numpy()))
    for nb in (2, 3, 7, 33):
        blocks = graph.col_blocks(nb)
        assert sum(b.nnz for _, _, b in blocks) == graph.nnz
        assert all(int(b.col.min()) >= c0 and int(b.col.max()) < c1 for c0, c1, b in blocks if b.nnz)
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:])) and blocks[0][0] == 0 and blocks[-1][1] == n
        got = hop_col_blocked(graph, x, nb, chunk=0)
        assert torch.equal(got, plain), nb
        # into a column view of a wider live matrix: neighbours untouched
        wide = torch.full((n, f + 9), -3.0, device=dev)
        hop_col_blocked(graph, x, nb, chunk=0, out=wide[:, 4:4 + f])
        assert torch.equal(wide[:, 4:4 + f], plain) and bool((wide[:, :4] == -3.0).all()) and bool((wide[:, 4 + f:] == -3.0).all())
        split_a = hop_col_blocked(graph, x, nb, chunk=64)
        split_b = hop_col_blocked(graph, x, nb, chunk=64)
        assert torch.equal(split_a, split_b)
        assert np.allclose(split_a.cpu().numpy(), plain.cpu().numpy(), rtol=RTOL, atol=ATOL)
    # the K-hop entry points, switched by the environment
    want, _ = sgc_b200.sgc_precompute(x, graph, 3, chunk=0)
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "5")
    got, _ = sgc_b200.sgc_precompute(x, graph, 3, chunk=0)
    assert torch.equal(got, want)
    assert torch.equal(sgc_b200.hop(graph, x, chunk=0), plain)
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "auto")          # small operand: auto keeps the single pass
    assert sgc_b200.propagate.col_blocks_for(graph, 416) == 0


def test_hops_on_one_plan_from_two_streams(dev):
    """Two ticket-scheduled hops on the SAME cached plan, in flight at once on different streams, draw their
    tickets from different counter pairs (the counters used to be shared): both results are the
    single-stream result, repeatedly, and later launches still start from clean counters."""
    import sgc_b200
    from sgc_b200 import _ffi
    s, graph = powerlaw_case(dev, scale=0.05, seed=29)
    h = _ffi_hints_force_tickets()
    xa = s["x"][:, :128].contiguous()
    xb = (s["x"][:, :128] * 0.5 + 1.0).contiguous()
    want_a = sgc_b200.hop(graph, xa, chunk=0, hints=h)
    want_b = sgc_b200.hop(graph, xb, chunk=0, hints=h)
    sa, sb = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    outs = [(torch.empty_like(want_a), torch.empty_like(want_b)) for _ in range(6)]
    torch.cuda.synchronize(dev)
    for ya, yb in outs:
        with torch.cuda.stream(sa):
            sgc_b200.hop(graph, xa, chunk=0, hints=h, out=ya)
        with torch.cuda.stream(sb):
            sgc_b200.hop(graph, xb, chunk=0, hints=h, out=yb)
    torch.cuda.synchronize(dev)
    for ya, yb in outs:
        assert torch.equal(ya, want_a) and torch.equal(yb, want_b)
    assert torch.equal(sgc_b200.hop(graph, xa, chunk=0, hints=h), want_a)


def _ffi_hints_force_tickets():
    from sgc_b200.propagate import default_hints
    return default_hints() | 256                  # hints bit 8: ticket grid whatever the task count


def test_column_blocks_of_row_slice_views(dev, monkeypatch):
    """Row-slice views share the parent's `col` array and keep absolute offsets in `row_ptr` (pipeline.row_blocks,
    the sharded propagator's sub-graphs).  With SGC_B200_COL_BLOCKS set, hop() takes the column-blocked path for
    them too: their blocks must be built from the view's own entries only -- block by block the
    entries tile [row_ptr[0], row_ptr[-1]), the hop equals the matching rows of the single-pass hop bit for bit,
    and precompute_host(out_blocks > 1) gives the same bytes with and without column blocks."""
    import sgc_b200
    from sgc_b200.pipeline import row_blocks
    s, graph = powerlaw_case(dev, scale=0.02, seed=23)
    n = graph.n_rows
    x = s["x"][:, :72].contiguous()
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "0")
    plain = sgc_b200.hop(graph, x, chunk=0)
    views = row_blocks(graph, 5)
    assert len(views) > 2
    for r0, r1, sub in views:
        b, e = int(sub.row_ptr[0]), int(sub.row_ptr[-1])
        for nb in (2, 4):
            blocks = sub.col_blocks(nb)
            assert sum(g.nnz for _, _, g in blocks) == e - b
            assert all(int(g.row_ptr[0]) == 0 and int(g.row_ptr[-1]) == g.nnz for _, _, g in blocks)
        monkeypatch.setenv("SGC_B200_COL_BLOCKS", "3")
        assert torch.equal(sgc_b200.hop(sub, x, chunk=0), plain[r0:r1]), (r0, r1)
        monkeypatch.setenv("SGC_B200_COL_BLOCKS", "0")
    host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
    x_host = x.cpu().pin_memory()
    want = sgc_b200.precompute_host(host, x_host, 2, out_blocks=4, chunk=0, col_slab=0).clone()
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "3")
    got = sgc_b200.precompute_host(host, x_host, 2, out_blocks=4, chunk=0, col_slab=0)
    assert torch.equal(got, want)


def test_fused_norm_mode_is_bit_identical(dev, monkeypatch):
    """SGC_B200_FUSED_NORM=1 (weights formed in the kernel from fp64 A+I values and D^-1/2; S is
    never materialised) and the default (fp32 S values written once by sgc_norm_values at
    normalisation time) give the same bits, split rows included."""
    import sgc_b200
    x = None
    outs = []
    for mode in ("1", "0"):
        monkeypatch.setenv("SGC_B200_FUSED_NORM", mode)
        s, graph = powerlaw_case(dev, scale=0.01, seed=9)
        assert (graph.sval is None) == (mode == "1")
        x = s["x"][:, :100].contiguous()
        y, _ = sgc_b200.sgc_precompute(x, graph, 2)
        assert (graph.sval is None) == (mode == "1")               # fused mode materialises nothing
        outs.append(y)
    assert torch.equal(outs[0], outs[1])
    # the general kernel (sgc_norm_build, used for views and row slices) writes the same fp32 values
    sv = graph.sval
    graph.sval = None
    assert torch.equal(graph.values32(), sv)


def test_fused_standardisation(dev):
    """First-hop standardisation fused into the gather == the oracle's hop on (x-mean)/std with the
    same fp32 statistics (DIV mode bit-exact; RCP mode within tolerance); statistics themselves
    against the oracle's fp64 column statistics."""
    import sgc_b200
    from oracle import spmm_oracle as so
    s, graph = powerlaw_case(dev, scale=0.01, seed=5)
    x = s["x"][:, :130].contiguous()
    mean32, std32, mean64, std64 = sgc_b200.col_stats(x, want64=True)
    om, osd = so.col_stats(x.cpu().numpy())
    assert np.allclose(mean64.cpu().numpy(), om, rtol=1e-12, atol=1e-12)
    assert np.allclose(std64.cpu().numpy(), osd, rtol=1e-12, atol=0)
    # torch's own fp32 statistics (what utils.py:119 computes) agree to fp32 rounding
    assert torch.allclose(mean32, x.mean(0), rtol=1e-5, atol=1e-6)
    assert torch.allclose(std32, x.std(0), rtol=1e-5, atol=0)
    rp, col, val = graph.row_ptr.cpu().numpy(), graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    m, sd = mean32.cpu().numpy(), std32.cpu().numpy()
    want = so.spmm_hop(rp, col, val, x.cpu().numpy(), mean=m, std=sd)
    got = sgc_b200.hop(graph, x, mean=mean32, std=std32, chunk=0, std_mode="div")
    assert np.array_equal(got.cpu().numpy(), want)
    got_rcp = sgc_b200.hop(graph, x, mean=mean32, std=std32, chunk=0, std_mode="rcp")
    assert np.allclose(got_rcp.cpu().numpy(), want, rtol=RTOL, atol=ATOL)
    # through the public API: StandardizedFeatures in, two hops out -- fused and in-place modes
    want2 = so.spmm_hop(rp, col, val, want)
    feats = sgc_b200.StandardizedFeatures(x.clone(), mean32, std32)
    y, _ = sgc_b200.sgc_precompute(feats, graph, 2, chunk=0, std_mode="div")
    assert np.array_equal(y.cpu().numpy(), want2)
    sub = feats[np.arange(0, 50)]                     # reddit.py:45 indexes with an ndarray
    assert isinstance(sub, sgc_b200.StandardizedFeatures)
    assert torch.equal(sub.standardized(), ((x - mean32) / std32)[:50])
    assert sub.unwrap()[0].stride(0) % 8 == 0          # re-padded for the aligned gather path
    # default mode: one exact in-place pass over the feature storage, then plain hops
    feats2 = sgc_b200.StandardizedFeatures(sgc_b200.to_padded(x.clone()), mean32, std32)
    y2, _ = sgc_b200.sgc_precompute(feats2, graph, 2, chunk=0)
    assert np.array_equal(y2.cpu().numpy(), want2)
    raw2, m2, s2 = feats2.unwrap()
    assert m2 is None and torch.equal(raw2, (x - mean32) / std32)      # storage now holds utils.py:119's values
    assert torch.equal(feats2[np.arange(0, 50)], ((x - mean32) / std32)[:50])
    y3, _ = sgc_b200.sgc_precompute(feats2, graph, 2, chunk=0)         # idempotent: not applied twice
    assert torch.equal(y3, y2)


def test_standardised_path_is_no_further_from_fp64_than_the_reference(dev):
    """The path that ships for Reddit (K3 statistics on the GPU -> exact in-place (x-mean)/std -> K hops) against
    the exact answer: utils.py:118-119 and the hops of utils.py:94-95 evaluated in fp64 on the CPU with the same
    fp32 S values.  The reference's own fp32 evaluation (torch mean/std/spmm on the CPU -- its literal
    expressions) differs from the GPU's on a fraction of the elements because its fp32 column statistics carry
    rounding noise; the bar is accuracy: on EVERY element |gpu - fp64| <= |reference - fp64| + 1e-6 + 1e-5*|fp64|
    (north_star's tolerance), through load-time standardisation and through precompute_host alike."""
    import sgc_b200
    s, graph = powerlaw_case(dev, scale=0.03, seed=41)
    n, k = s["n"], 2
    x_raw = s["x"][:, :200].contiguous()
    xh = x_raw.cpu()
    sv = graph.values32().cpu()
    idx = torch.stack([torch.from_numpy(coo_rows(graph)), graph.col.cpu().long()])
    adj32 = torch.sparse_coo_tensor(idx, sv, (n, n))
    ref = (xh - xh.mean(dim=0)) / xh.std(dim=0)                       # utils.py:119, verbatim
    truth = (xh.double() - xh.double().mean(dim=0)) / xh.double().std(dim=0)
    adj64 = adj32.to(torch.float64)
    for _ in range(k):
        ref = torch.spmm(adj32, ref)                                   # utils.py:95
        truth = torch.spmm(adj64, truth)
    mean, std = sgc_b200.col_stats(x_raw)
    feats = sgc_b200.StandardizedFeatures(sgc_b200.to_padded(x_raw.clone()), mean, std)
    got, _ = sgc_b200.sgc_precompute(feats, graph, k)
    host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
    got_host = sgc_b200.precompute_host(host, xh.pin_memory(), k, standardize=True)
    for y in (got.cpu(), got_host):
        e_gpu = (y.double() - truth).abs()
        e_ref = (ref.double() - truth).abs()
        bad = e_gpu > e_ref + ATOL + RTOL * truth.abs()
        assert int(bad.sum()) == 0, (int(bad.sum()), float(e_gpu.max()), float(e_ref.max()))
        assert float(e_gpu.mean()) <= float(e_ref.mean()) * 1.05 + 1e-9   # and not worse on average either


@pytest.mark.parametrize("blocks,col_slab", [(1, None), (3, None), (8, None), (1, 32), (1, 64)])
def test_precompute_host_pipeline(dev, blocks, col_slab):
    """Host-to-host entry point (pipeline.precompute_host: overlapped H2D; column-slab pipeline or
    row-blocked last hop with overlapped D2H) == sgc_precompute on device-resident inputs, bit for
    bit, and == the oracle in exact mode; with and without the utils.py:119 standardisation;
    K = 1, 2, 3."""
    import sgc_b200
    from oracle import spmm_oracle as so
    s, graph = powerlaw_case(dev, scale=0.01, seed=7)
    n = s["n"]
    x = s["x"][:, :77].contiguous()
    host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
    x_host = x.cpu().pin_memory()
    for k in (1, 2, 3):
        want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(x.clone()), graph, k)
        got = sgc_b200.precompute_host(host, x_host, k, out_blocks=blocks, col_slab=col_slab)
        assert not got.is_cuda and torch.equal(got, want.cpu())
    # standardised, exact accumulation order: against the oracle on (x - mean) / std
    mean32, std32 = sgc_b200.col_stats(x)
    xs = ((x - mean32) / std32).cpu().numpy()
    rp, col, val = graph.row_ptr.cpu().numpy(), graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    want2 = so.spmm_hop(rp, col, val, so.spmm_hop(rp, col, val, xs))
    out = torch.empty((n, 77), dtype=torch.float32).pin_memory()
    got2 = sgc_b200.precompute_host(host, x_host, 2, standardize=True, out=out, out_blocks=blocks, chunk=0,
                                    col_slab=col_slab)
    assert got2 is out and np.array_equal(out.numpy(), want2)
    # scipy input (what the reference's loaders hold) takes the same path
    import scipy.sparse as sp
    a = sp.csr_matrix((host[3].numpy(), host[2].numpy(), host[1].numpy()), shape=(n, n))
    got3 = sgc_b200.precompute_host(a, x_host, 2, standardize=True, out_blocks=blocks, chunk=0, col_slab=col_slab)
    assert np.array_equal(got3.numpy(), want2)
    # row-padded pinned staging buffers (64-byte row pitch) for X and the result: same bits, padding untouched
    from sgc_b200.pipeline import pinned_rows
    xp, op = pinned_rows(n, 77), pinned_rows(n, 77)
    assert xp.stride(0) == 80 and xp.is_pinned()
    xp.copy_(x_host)
    op.fill_(-5.0)
    whole = torch.as_strided(op, (n, 80), (80, 1))
    whole[:, 77:] = -9.0
    got4 = sgc_b200.precompute_host(host, xp, 2, standardize=True, out=op, out_blocks=blocks, chunk=0, col_slab=col_slab)
    assert got4 is op and np.array_equal(op.numpy(), want2) and bool((whole[:, 77:] == -9.0).all())


@pytest.mark.parametrize("col_slab", [None, 32])
def test_precompute_host_uniform_values(dev, monkeypatch, col_slab):
    """A constant-valued adjacency (what the reference's loaders hold: 1.0 for the citation graphs,
    2.0 for Reddit's adj + adj.T) is recognised on the host and its value array filled on the GPU
    instead of uploaded: same bits as the uploading path, nnz x 8 fewer bytes over PCIe; a weighted
    adjacency is uploaded as is and matches the device-resident path."""
    import sgc_b200
    s, graph = powerlaw_case(dev, scale=0.01, seed=11)
    n = s["n"]
    common = float(s["val"][0])
    assert bool((s["val"] == common).all())
    x = s["x"][:, :77].contiguous()
    x_host = x.cpu().pin_memory()
    host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
    want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(x.clone()), graph, 2)
    runs = {}
    for mode in ("force", "0"):
        monkeypatch.setenv("SGC_B200_UNIFORM_VALUES", mode)
        st = {}
        got = sgc_b200.precompute_host(host, x_host, 2, col_slab=col_slab, stats=st)
        assert torch.equal(got, want.cpu())
        runs[mode] = st
    assert runs["force"]["uniform_value"] == common and runs["0"]["uniform_value"] is None
    assert runs["0"]["h2d_bytes"] - runs["force"]["h2d_bytes"] == host[3].numel() * 8
    assert runs["0"]["h2d_bytes"] == sum(t.numel() * t.element_size() for t in host[1:]) + x_host.numel() * 4
    assert runs["0"]["d2h_bytes"] == n * 77 * 4
    # explicit switch wins over the environment
    monkeypatch.setenv("SGC_B200_UNIFORM_VALUES", "force")
    st = {}
    sgc_b200.precompute_host(host, x_host, 1, col_slab=col_slab, stats=st, detect_uniform_values=False)
    assert st["uniform_value"] is None
    # the other constant of the reference's loaders
    ones = torch.ones_like(s["val"])
    g1 = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], ones)
    want1, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(x.clone()), g1, 2)
    st = {}
    got1 = sgc_b200.precompute_host((n, host[1], host[2], ones.cpu().pin_memory()), x_host, 2, col_slab=col_slab,
                                    stats=st)
    assert st["uniform_value"] == 1.0 and torch.equal(got1, want1.cpu())
    # weighted adjacency: one stored value differs (the last one: the scan has to reach it)
    val2 = s["val"].clone()
    val2[-1] = 2.5
    g2 = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], val2)
    want2, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(x.clone()), g2, 2)
    st = {}
    got2 = sgc_b200.precompute_host((n, host[1], host[2], val2.cpu().pin_memory()), x_host, 2, col_slab=col_slab,
                                    stats=st)
    assert st["uniform_value"] is None and torch.equal(got2, want2.cpu())
    assert not torch.equal(got2, want.cpu())


def test_downstream_accuracy_unchanged(dev):
    """Cora / Citeseer: the reference's own training loop (citation.py:35-62, restated here as test
    code, CPU, seed 42, tuned weight decay) on the GPU-propagated features gives the reference's
    validation / test accuracy."""
    import sgc_b200
    import torch.nn.functional as F
    for ds in ("cora", "citeseer"):
        g = load_golden(ds)
        graph = gpu_graph(g, dev)
        y, _ = sgc_b200.sgc_precompute(to_dev(dense_features(g), dev), graph, 2)
        y = y.cpu()
        labels = torch.from_numpy(g["labels"])
        tr, va, te = (torch.from_numpy(g[k]) for k in ("idx_train", "idx_val", "idx_test"))
        torch.manual_seed(42)
        np.random.seed(42)
        model = torch.nn.Linear(y.shape[1], int(labels.max()) + 1)
        # the reference wraps the layer as SGC.W (models.py:7-18); parameter init order is identical
        opt = torch.optim.Adam(model.parameters(), lr=0.2, weight_decay=float(g["weight_decay"]))
        for _ in range(int(g["epochs"])):
            model.train()
            opt.zero_grad()
            F.cross_entropy(model(y[tr]), labels[tr]).backward()
            opt.step()
        model.eval()
        with torch.no_grad():
            acc = lambda idx: (model(y[idx]).max(1)[1] == labels[idx]).double().mean().item()
            assert abs(acc(va) - float(g["acc_val"])) < 1e-12, ds
            assert abs(acc(te) - float(g["acc_test"])) < 1e-12, ds


def test_full_size_reddit_shape_properties(dev):
    """BASELINE configs[2] at full size (232,965 nodes, ~23M nnz, 602 features, K=2): size-
    independent checks -- S.1 equals the row sums of S, linearity of the hop, and the oracle on a
    random sample of output rows."""
    import sgc_b200
    from sgc_b200 import synth
    from oracle import spmm_oracle as so
    s = synth.make_shape("reddit", device=dev)
    graph = sgc_b200.normalize_csr(s["n"], s["row_ptr"], s["col"], s["val"])
    n, f = s["n"], s["f"]
    assert graph.nnz > 20_000_000
    x = s["x"]
    mean, std = sgc_b200.col_stats(x)
    y1 = sgc_b200.hop(graph, x, mean=mean, std=std)
    # (1) S.1 = row sums of S
    ones = torch.ones((n, 8), dtype=torch.float32, device=dev)
    rs = sgc_b200.hop(graph, ones)
    rows = torch.repeat_interleave(torch.arange(n, device=dev), graph.row_ptr[1:] - graph.row_ptr[:-1])
    want_rs = torch.zeros(n, dtype=torch.float64, device=dev).index_add_(0, rows, graph.values32().double())
    assert torch.allclose(rs[:, 0].double(), want_rs, rtol=1e-5, atol=1e-6)
    # (2) linearity: S(2a - b) = 2 S a - S b  (up to fp32 rounding of the three results)
    a, b = x[:, :64].contiguous(), x[:, 64:128].contiguous()
    lhs = sgc_b200.hop(graph, 2 * a - b)
    rhs = 2 * sgc_b200.hop(graph, a) - sgc_b200.hop(graph, b)
    assert torch.allclose(lhs, rhs, rtol=1e-4, atol=1e-4)
    # (3) oracle on 1500 sampled rows of hop 1 (fused standardisation) and hop 2
    rng = np.random.RandomState(0)
    pick = np.sort(rng.choice(n, 1500, replace=False))
    rp = graph.row_ptr.cpu().numpy()
    col, val = graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    sub_ptr = np.zeros(pick.size + 1, dtype=np.int64)
    np.cumsum(rp[pick + 1] - rp[pick], out=sub_ptr[1:])
    take = np.concatenate([np.arange(rp[i], rp[i + 1]) for i in pick])
    want1 = so.spmm_hop(sub_ptr, col[take], val[take], x.cpu().numpy(), mean=mean.cpu().numpy(),
                        std=std.cpu().numpy())
    assert np.allclose(y1.cpu().numpy()[pick], want1, rtol=RTOL, atol=ATOL)
    y2 = sgc_b200.hop(graph, y1)
    want2 = so.spmm_hop(sub_ptr, col[take], val[take], y1.cpu().numpy())
    assert np.allclose(y2.cpu().numpy()[pick], want2, rtol=RTOL, atol=ATOL)
    # (4) the K-hop entry point agrees with hop-by-hop
    feats = sgc_b200.StandardizedFeatures(sgc_b200.to_padded(x.clone()), mean, std)
    yk, secs = sgc_b200.sgc_precompute(feats, graph, 2, std_mode="div")
    assert torch.equal(yk, y2) and secs > 0
    yi, _ = sgc_b200.sgc_precompute(feats, graph, 2)                    # in-place standardisation (default)
    assert torch.equal(yi, y2)


@pytest.mark.parametrize("name", ["reddit", "products"])
def test_full_size_normalisation_vs_oracle(dev, name):
    """K1 + K2 at BASELINE size (Reddit shape: 23.4 M stored entries with 16 k-entry hub rows through the
    element-parallel tile / binary-search scheme of csr_fast.cu; products shape: 126 M): row_ptr, col and the
    fp32 values of S bit for bit against the NumPy restatement of normalization.py:5-12 + utils.py:25, the
    fp64 values and D^-1/2 within the ulp bar of the golden-vector tests."""
    import hashlib
    import sgc_b200
    from sgc_b200 import synth
    from oracle import normalization_oracle as no
    s = synth.make_shape(name, device=dev, with_features=False)
    n = s["n"]
    graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    ref = no.aug_normalized_adjacency_oracle(n, s["row_ptr"].cpu().numpy(), s["col"].cpu().numpy(),
                                             s["val"].cpu().numpy())
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    assert graph.nnz == ref["indices"].size > (20_000_000 if name == "reddit" else 120_000_000)
    assert sha(graph.row_ptr.cpu().numpy()) == sha(ref["indptr"])
    assert sha(graph.col.cpu().numpy()) == sha(ref["indices"].astype(np.int32))
    assert sha(graph.values32().cpu().numpy()) == sha(ref["data"].astype(np.float32))
    d = graph.dinv.cpu().numpy()
    assert np.max(np.abs(d - ref["d"]) / np.spacing(ref["d"])) <= 1.0          # numpy's pow is within 1 ulp itself
    v64 = graph.values64().cpu().numpy()
    assert np.max(np.abs(v64 - ref["data"]) / np.spacing(ref["data"])) <= 4.0


def test_full_size_products_shape_hops(dev):
    """BASELINE configs[3] at full size on one GPU (2.45 M nodes, 126 M stored entries, F = 100, K = 3): the
    path that runs there -- whole-row kernel <32,4,1> with hot-column tags, long rows split -- against the C
    oracle on 1500 sampled rows of hop 1 and of hop 3, within north_star's tolerance; tags on/off and the K-hop
    entry point vs hop by hop bit-identical."""
    import sgc_b200
    from sgc_b200 import synth
    from oracle import spmm_oracle as so
    s = synth.make_shape("products", device=dev)
    n, f, k = s["n"], s["f"], s["k"]
    graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    x = sgc_b200.to_padded(s["x"])
    y1 = sgc_b200.hop(graph, x)
    plan = graph.plan(sgc_b200.propagate.default_chunk())
    assert plan.n_hot > 0 and plan.col_tagged is not None              # the tagged variant is what ran
    assert torch.equal(sgc_b200.hop(graph, x, hints=3), y1)            # tags off: same bits
    y2 = sgc_b200.hop(graph, y1)
    y3 = sgc_b200.hop(graph, y2)
    yk, _ = sgc_b200.sgc_precompute(x, graph, k)
    assert torch.equal(yk, y3)
    rng = np.random.RandomState(1)
    deg = (graph.row_ptr[1:] - graph.row_ptr[:-1]).cpu().numpy()
    pick = np.unique(np.concatenate([rng.choice(n, 1480, replace=False), np.argsort(deg)[-20:]]))   # + the hubs
    rp = graph.row_ptr.cpu().numpy()
    col, val = graph.col.cpu().numpy(), graph.values32().cpu().numpy()
    sub_ptr = np.zeros(pick.size + 1, dtype=np.int64)
    np.cumsum(rp[pick + 1] - rp[pick], out=sub_ptr[1:])
    take = np.concatenate([np.arange(rp[i], rp[i + 1]) for i in pick])
    for src, dst in ((x, y1), (y2, y3)):
        want = so.spmm_hop(sub_ptr, col[take], val[take], np.ascontiguousarray(src.cpu().numpy()))
        assert np.allclose(dst.cpu().numpy()[pick], want, rtol=RTOL, atol=ATOL)
