"""CPU-only checks: the C-ABI library loads and exports every symbol include/sgc_b200.h declares,
the ctypes mirror of its structs matches the C layout, the host-side API mirrors the reference's
names and signatures, and there is no CPU fallback."""
import ctypes
import inspect
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(REPO, "include", "sgc_b200.h")
REFERENCE = "/root/reference"


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sgc_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as entry
    entry.build()
    from sgc_b200 import _ffi
    return _ffi


def test_library_exports_every_declared_symbol(built):
    names = declared_symbols()
    assert len(names) >= 20
    handle = ctypes.CDLL(built.LIB_PATH)
    for name in names:
        assert hasattr(handle, name), "libsgc_b200.so does not export %s" % name
    # and the ctypes table binds exactly the declared surface
    assert sorted(built.SIGNATURES) == names
    assert built.lib().sgc_version().decode().startswith("sgc_b200")


def test_library_is_sm100a_only(built):
    out = subprocess.run(["cuobjdump", "-lelf", built.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_hop_args_layout_matches_c(built, tmp_path):
    src = tmp_path / "layout.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "sgc_b200.h"\n'
        "int main(void){printf(\"%zu\", sizeof(sgc_hop_args_t));\n"
        + "".join('printf(" %%zu", offsetof(sgc_hop_args_t, %s));\n' % f for f in
                  ["aval_d", "dinv_d", "dinv_rows_d", "sval_d", "x_d", "ldx", "y_d", "ldy", "partial_d", "ldp", "mean_d",
                   "std_d", "f", "slab_floats", "flags", "hints", "n_peers", "push_tile_tasks", "peer_y_d", "peer_mask_d"])
        + "return 0;}\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), str(src), "-o", str(exe)])
    nums = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    assert nums[0] == ctypes.sizeof(built.HopArgs)
    got = [getattr(built.HopArgs, name).offset for name, _ in built.HopArgs._fields_]
    assert nums[1:] == got


def test_host_calls_validate_arguments_without_a_gpu(built):
    lib = built.lib()
    assert lib.sgc_scan_workspace_bytes(1 << 20) > 0
    assert lib.sgc_spmm_plan_workspace_bytes(1000, 5000, 0) > 0
    assert lib.sgc_spmm_plan_workspace_bytes(1000, 5000, 64) > lib.sgc_spmm_plan_workspace_bytes(1000, 5000, 0)
    rc = lib.sgc_spmm_hop(None, None, None)
    assert rc == -1 and b"null" in lib.sgc_last_error()
    with pytest.raises(built.SgcError):
        built.check(rc, "sgc_spmm_hop")


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    import sgc_b200
    import scipy.sparse as sp
    from sgc_b200 import _ffi
    with pytest.raises(_ffi.SgcError):
        sgc_b200.aug_normalized_adjacency(sp.eye(4, format="csr"))
    import utils as drop_in
    with pytest.raises(_ffi.SgcError):
        drop_in.load_citation("cora", "AugNormAdj", False)
    with pytest.raises(_ffi.SgcError):
        drop_in.load_reddit_data("AugNormAdj", cuda=False)
    sparse = torch.sparse_coo_tensor(torch.tensor([[0, 1], [1, 0]]), torch.ones(2), (2, 2))
    with pytest.raises(_ffi.SgcError):
        drop_in.sgc_precompute(torch.ones(2, 3), sparse, 2)
    same, secs = drop_in.sgc_precompute(sparse, sparse, 0)      # degree 0: loop does not run
    assert same is sparse and secs == 0.0


def test_product_does_not_import_the_oracle():
    """Nothing under sgc_b200/ (nor the drop-in shims) may import, call or link the oracle."""
    offenders = []
    files = [os.path.join(REPO, "utils.py"), os.path.join(REPO, "normalization.py")]
    for root, _, names in os.walk(os.path.join(REPO, "sgc_b200")):
        files += [os.path.join(root, n) for n in names if n.endswith((".py", ".cu", ".cuh", ".h", "Makefile"))]
    for path in files:
        text = open(path).read()
        if re.search(r"^\s*(from|import)\s+oracle\b|liboracle|oracle/", text, flags=re.M):
            offenders.append(path)
    assert not offenders, offenders


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference tree not present on this machine")
def test_api_mirrors_reference_signatures():
    """Same public names, parameter names, order and defaults as the reference's modules
    (utils.py / normalization.py), checked by loading the reference's source under another name."""
    import importlib.util
    import normalization as mine_norm
    import utils as mine_utils

    def load(name, alias):
        # the reference's utils does `from normalization import ...`: resolve it to the reference's
        sys.path.insert(0, REFERENCE)
        saved = {k: sys.modules.pop(k) for k in ("normalization", "utils") if k in sys.modules}
        try:
            spec = importlib.util.spec_from_file_location(alias, os.path.join(REFERENCE, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
        finally:
            sys.path.remove(REFERENCE)
            for k in ("normalization", "utils"):
                sys.modules.pop(k, None)
            sys.modules.update(saved)

    ref_norm, ref_utils = load("normalization", "_ref_normalization"), load("utils", "_ref_utils")
    for ref, mine, names in (
            (ref_norm, mine_norm, ["aug_normalized_adjacency", "fetch_normalization", "row_normalize"]),
            (ref_utils, mine_utils, ["sgc_precompute", "load_citation", "load_reddit_data",
                                     "sparse_mx_to_torch_sparse_tensor", "preprocess_citation",
                                     "parse_index_file", "set_seed", "loadRedditFromNPZ",
                                     "fetch_normalization", "row_normalize"])):
        for name in names:
            a, b = inspect.signature(getattr(ref, name)), inspect.signature(getattr(mine, name))
            pa = [(p.name, p.default) for p in a.parameters.values()]
            pb = [(p.name, p.default) for p in b.parameters.values() if p.kind != p.VAR_KEYWORD]
            assert pa == pb, (name, pa, pb)
    # unknown normalisation: a zero-argument callable, like the reference
    assert mine_norm.fetch_normalization("NoSuch")() == ref_norm.fetch_normalization("NoSuch")()


@pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "data")), reason="reference data not present")
def test_planetoid_adjacency_matches_reference_loader():
    import pickle
    from conftest import load_golden
    from sgc_b200 import utils as mine
    for ds in ("cora", "citeseer", "pubmed"):
        with open(os.path.join(REFERENCE, "data", "ind.%s.graph" % ds), "rb") as fh:
            graph = pickle.load(fh, encoding="latin1")
        adj = mine._planetoid_adjacency(graph)
        g = load_golden(ds)
        assert np.array_equal(adj.indptr, g["adj_indptr"]) and np.array_equal(adj.indices, g["adj_indices"])
        assert np.array_equal(adj.data.astype(np.float64), g["adj_data"])


def test_row_normalize_has_no_cpu_path():
    """row_normalize (normalization.py:21-28) runs on the GPU (tests/test_gpu_sparse_features.py checks
    it against the golden vectors); without a device it raises instead of falling back to scipy."""
    import scipy.sparse as sp
    import torch
    import normalization as mine
    from sgc_b200 import _ffi
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    with pytest.raises(_ffi.SgcError):
        mine.row_normalize(sp.identity(4, format="csr", dtype=np.float32))


def test_standardized_features_semantics():
    from sgc_b200.features import StandardizedFeatures
    torch.manual_seed(0)
    raw = torch.randn(50, 7) * 3 + 2
    mean, std = raw.mean(0), raw.std(0)
    feats = StandardizedFeatures(raw, mean, std)
    ref = (raw - mean) / std                                  # utils.py:119
    assert feats.size(1) == 7 and feats.shape == ref.shape and len(feats) == 50
    idx = np.array([3, 1, 40])
    sub = feats[idx]                                          # reddit.py:45 indexes with ndarrays
    assert isinstance(sub, StandardizedFeatures) and torch.equal(sub.standardized(), ref[idx])
    lin = torch.nn.Linear(7, 3)
    assert torch.equal(lin(feats), lin(ref))                  # any torch op sees standardised values
    assert torch.equal(feats * 1.0, ref) and type(feats * 1.0) is torch.Tensor
    assert feats.cuda is not None and feats.float() is feats
    rawt, m, s = feats.unwrap()
    assert type(rawt) is torch.Tensor and rawt.data_ptr() == raw.data_ptr() and m is mean and s is std


def test_host_pipeline_schedule_choice(monkeypatch):
    """pipeline.default_col_slab / slab_widths: column slabs only when a slab fits L2 and the block holds
    at least two full slabs (otherwise the contiguous row-block schedule); no CPU path."""
    import torch
    from sgc_b200 import _ffi, pipeline
    monkeypatch.delenv("SGC_B200_COL_SLAB", raising=False)
    n_reddit, n_products = 232965, 2449029
    assert pipeline.default_col_slab(n_reddit, 602) == 128
    assert pipeline.slab_widths(602, 128) == [128, 128, 128, 128, 90]
    assert pipeline.default_col_slab(n_reddit, 301) == 128 and pipeline.slab_widths(301, 128) == [128, 128, 45]
    for f in (80, 138, 160, 255):                  # column blocks of the 4- and 8-GPU feature sharding
        assert pipeline.default_col_slab(n_reddit, f) == 0 and pipeline.slab_widths(f, 0) == [f]
    assert pipeline.default_col_slab(n_products, 100) == 0          # one slab (and none fits L2)
    assert pipeline.default_col_slab(2708, 1433) == 256             # Cora: everything fits, widest slab
    monkeypatch.setenv("SGC_B200_COL_SLAB", "64")
    assert pipeline.default_col_slab(n_reddit, 602) == 64 and pipeline.default_col_slab(n_reddit, 60) == 0
    if not torch.cuda.is_available():
        with pytest.raises(_ffi.SgcError):
            pipeline.precompute_host((2, torch.tensor([0, 1, 2]), torch.tensor([1, 0], dtype=torch.int32),
                                      torch.ones(2, dtype=torch.float64)), torch.ones(2, 3), 1)


def test_host_value_scan(built, monkeypatch):
    """sgc_host_all_equal_f64 (host-side, no GPU work) and pipeline.uniform_value: exact bit compare,
    threaded, position-independent; non-fp64 / strided / weighted arrays are never taken for constant."""
    from sgc_b200 import pipeline
    fn = built.lib().sgc_host_all_equal_f64
    rng = np.random.default_rng(3)
    for n in (0, 1, 5, 65535, 65536, 65537, 1 << 20, (1 << 20) + 3):
        v = np.ones(n, dtype=np.float64)
        for threads in (1, 3, 8, 200):
            assert fn(v.ctypes.data, n, 1.0, threads, 0) == 1
        for pos in ([0, n - 1, n // 2] + list(rng.integers(0, n, 3)) if n else []):
            for bad in (0.0, 2.0, -1.0, np.nan, np.nextafter(1.0, 2.0)):
                w = v.copy()
                w[pos] = bad
                assert fn(w.ctypes.data, n, 1.0, 4, 0) == 0, (n, pos, bad)
    z = np.array([0.0, -0.0], dtype=np.float64)
    assert fn(z.ctypes.data, 2, 0.0, 1, 0) == 0             # bit compare: -0.0 is not 0.0
    assert fn(None, 3, 1.0, 1, 0) == -1 and fn(None, 0, 1.0, 1, 0) == 1
    big = np.ones(1 << 22, dtype=np.float64)
    assert fn(big.ctypes.data, big.size, 1.0, 1, 1) in (1, 2)      # 1 us budget: may give up, never lies
    big[-1] = 7.0
    assert fn(big.ctypes.data, big.size, 1.0, 1, 1) in (0, 2)

    monkeypatch.setenv("SGC_B200_UNIFORM_VALUES", "force")
    twos = torch.full((100000,), 2.0, dtype=torch.float64)
    assert pipeline.uniform_value(twos) == 2.0 and pipeline.uniform_value(twos[:1]) == 2.0
    assert pipeline.uniform_value(torch.ones(70000, dtype=torch.float64)) == 1.0
    assert pipeline.uniform_value(twos[:0]) is None                      # nothing to save
    assert pipeline.uniform_value(twos.float()) is None and pipeline.uniform_value(twos[::2]) is None
    assert pipeline.uniform_value(torch.full((10,), float("nan"), dtype=torch.float64)) is None
    twos[77777] = 3.0
    assert pipeline.uniform_value(twos) is None
    twos[77777] = 2.0
    monkeypatch.setenv("SGC_B200_UNIFORM_VALUES", "0")
    assert pipeline.uniform_value(twos) is None
    # a host that cannot finish inside the budget gives up for THIS call only and says so; nothing is
    # remembered between calls (no process-global state)
    monkeypatch.setenv("SGC_B200_UNIFORM_VALUES", "1")
    assert not hasattr(pipeline, "_scan_gave_up")
    rep = {}
    assert pipeline.uniform_value(twos, hidden_bytes=1 << 30, report=rep) == 2.0 and rep == {"scan_gave_up": False}
    monkeypatch.setattr(pipeline, "PCIE_BYTES_PER_US", 1e12)          # a budget nobody can meet (500 us floor)
    huge = torch.full((1 << 24,), 2.0, dtype=torch.float64)
    monkeypatch.setenv("SGC_B200_HOST_THREADS", "1")
    got = pipeline.uniform_value(huge, report=rep)
    assert (got is None and rep["scan_gave_up"]) or (got == 2.0 and not rep["scan_gave_up"])
    monkeypatch.setattr(pipeline, "PCIE_BYTES_PER_US", 40e3)
    assert pipeline.uniform_value(twos, report=rep) == 2.0 and rep["scan_gave_up"] is False


def test_textsgc_and_sparse_feature_entry_points_have_no_cpu_path():
    import scipy.sparse as sp
    import torch
    from sgc_b200 import _ffi, sparse_features, textsgc
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    a = sp.identity(4, format="csr", dtype=np.float32)
    for fn in (textsgc.preprocess_adj, textsgc.normalize_adj, sparse_features.row_normalize):
        with pytest.raises(_ffi.SgcError):
            fn(a)


def test_bench_reference_arm_contract(tmp_path):
    """`bench.py --impl reference` (the staged, unmodified reference on the host cores): exactly one JSON
    line on stdout with the contract's keys and the same `config` object the own arm prints; the process
    imports nothing of the sgc_b200 package and maps no libsgc_b200.so; every step is a whole K-hop
    utils.sgc_precompute; ranks other than 0 print nothing and exit 0."""
    import json
    env = dict(os.environ, SGC_BENCH_SCALE="0.02")
    env.pop("CUDA_VISIBLE_DEVICES", None)
    probe = tmp_path / "probe.txt"
    wrapper = ("import runpy, sys\n"
               "sys.argv = ['bench.py'] + sys.argv[1:]\n"
               "try:\n"
               "    runpy.run_path(%r, run_name='__main__')\n"
               "finally:\n"
               "    mods = sorted(m for m in sys.modules if m.split('.')[0] in ('sgc_b200', 'oracle'))\n"
               "    maps = [l for l in open('/proc/self/maps') if 'libsgc_b200' in l or 'liboracle' in l]\n"
               "    import torch\n"
               "    open(%r, 'w').write(repr((mods, len(maps), torch.cuda.is_initialized())))\n"
               % (os.path.join(REPO, "bench.py"), str(probe)))
    cmd = [sys.executable, "-c", wrapper, "--impl", "reference", "--steps", "2", "--warmup", "1"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600, cwd=REPO)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["unit"] == "edge*feat/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["gpu_launches"] == 0
    assert d["config"]["workload"].startswith("Reddit-shaped") and "model" not in d["config"]
    staged = os.path.exists(os.path.join(REPO, "baseline", "_ref", "utils.py"))
    assert d["cpu_baseline"]["kind"] == ("reference" if staged else "port")
    assert d["cpu_baseline"]["cores"] >= 1 and "utils.sgc_precompute" in d["cpu_baseline"]["sample"]
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["details"]["steps_requested"] == 2 and len(d["details"]["step_ms"]) == 2
    # `config` is a pure function of the workload: the own arm prints the identical object
    import importlib.util
    spec = importlib.util.spec_from_file_location("_bench_for_test", os.path.join(REPO, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    c = d["config"]
    assert bench.workload_config("reddit", c["nodes"], c["nnz_S"], c["features"], c["k_hops"]) == c
    mods, n_maps, cuda_init = eval(probe.read_text())
    assert n_maps == 0 and not cuda_init, (mods, n_maps, cuda_init)
    assert not [m for m in mods if m.startswith("sgc_b200")], mods
    if staged:
        assert mods == [], mods                       # not even the oracle port
    # a time budget that does not hold the requested steps: fewer WHOLE steps, and the line says so
    r = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--steps", "5",
                        "--warmup", "3"], env=dict(env, SGC_BENCH_REF_BUDGET_S="0.001"), capture_output=True, text=True,
                       timeout=600)
    d = json.loads(r.stdout.strip())
    assert d["steps"] == 1 and d["warmup"] == 1 and d["details"]["steps_requested"] == 5
    assert "whole workload" in d["cpu_baseline"]["sample"] and "1 whole steps timed (5 requested" in d["cpu_baseline"]["sample"]
    r = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       env=dict(env, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_row_blocks_cover_rows_and_balance_entries():
    """pipeline.row_blocks: row-slice views of one CSR (shared col / value arrays, sliced row_ptr),
    128-row aligned cuts, balanced by stored entries rather than rows (power-law graphs)."""
    from sgc_b200 import pipeline
    from sgc_b200.graph import SgcGraph
    rng = np.random.default_rng(5)
    n = 5000
    deg = np.minimum((rng.pareto(1.2, n) * 3).astype(np.int64) + 1, 900)
    deg[:40] = 900                                            # hubs first: equal row counts would be unbalanced
    row_ptr = torch.from_numpy(np.concatenate(([0], np.cumsum(deg))).astype(np.int64))
    nnz = int(row_ptr[-1])
    col = torch.from_numpy(rng.integers(0, n, nnz).astype(np.int32))
    sval = torch.ones(nnz, dtype=torch.float32)
    g = SgcGraph(n, n, row_ptr, col, None, None, sval=sval)
    assert [(a, b) for a, b, _ in pipeline.row_blocks(g, 1)] == [(0, n)] and pipeline.row_blocks(g, 1)[0][2] is g
    for nb in (2, 3, 8, 64):
        blocks = pipeline.row_blocks(g, nb)
        assert 1 <= len(blocks) <= nb
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        sizes = []
        for r0, r1, sub in blocks:
            assert r1 > r0 and (r0 % 128 == 0) and sub.n_rows == r1 - r0 and sub.n_cols == n
            assert sub.col.data_ptr() == col.data_ptr() and sub.sval.data_ptr() == sval.data_ptr()
            assert torch.equal(sub.row_ptr, row_ptr[r0:r1 + 1])          # offsets into the shared arrays
            sizes.append(int(row_ptr[r1] - row_ptr[r0]))
        assert sum(sizes) == nnz
        if nb <= 8:
            assert max(sizes) <= nnz // nb + 128 * 900                    # within one aligned cut of the target


def test_column_block_policy_is_off_unless_asked(monkeypatch):
    """propagate.col_blocks_for: 0 by default (the column-blocked hop measured slower); an explicit
    count is taken as is; 'auto' only names block counts for one-slab-wide operands between 2x and 64x the L2 budget."""
    from sgc_b200 import propagate
    from sgc_b200.graph import L2_SLAB_BUDGET

    class G(object):
        def __init__(self, n_cols):
            self.n_cols = n_cols
    monkeypatch.delenv("SGC_B200_COL_BLOCKS", raising=False)
    assert propagate.col_blocks_for(G(2_449_029), 416) == 0
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "7")
    assert propagate.col_blocks_for(G(10), 416) == 7
    monkeypatch.setenv("SGC_B200_COL_BLOCKS", "auto")
    nb = propagate.col_blocks_for(G(2_449_029), 416)            # products: X = 1 GB
    assert nb >= 2 and 2_449_029 * 416 / nb <= L2_SLAB_BUDGET
    assert propagate.col_blocks_for(G(232_965), 416) == 0        # fits L2 twice over: single pass
    assert propagate.col_blocks_for(G(232_965), 2432) == 0       # wider than one slab: feature slabs instead
    assert propagate.col_blocks_for(G(111_059_956), 512) == 0    # papers100M: y would be re-read too often


def test_grid_cost_model_is_monotone():
    from sgc_b200 import dist as sd
    widths = [sd._block_cost_slabs(w) for w in range(1, 700)]
    assert all(b >= a for a, b in zip(widths, widths[1:]))        # a wider block never costs less
    assert sd._block_cost_slabs(128) == 1.0 and sd._block_cost_slabs(256) == 2.0
    # more row parts: less compute per rank, more pushed bytes per exchanging hop
    n, nnz, f = 232965, 23.4e6, 602
    assert sd.grid_cost_ms(n, nnz, f, 1, 1, 8) < sd.grid_cost_ms(n, nnz, f, 1, 1, 4) < sd.grid_cost_ms(n, nnz, f, 1, 1, 2)
    # feature sharding has no exchange term: K hops cost K times one hop
    assert abs(sd.grid_cost_ms(n, nnz, f, 3, 8, 1) - 3 * sd.grid_cost_ms(n, nnz, f, 1, 8, 1)) < 1e-9
