#!/usr/bin/env python
"""Benchmark of the SGC feature-propagation hot path (BASELINE.json metric: K-hop propagation
edge*feat/s and achieved GB/s vs roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

One "step" = one `sgc_precompute` (K hops of S X on device-resident CSR + standardised X: the span the
reference times at utils.py:93-96) on synthetic inputs of the named BASELINE shape.  N=1 workload:
configs[2], the Reddit-shaped graph (232,965 nodes, ~23M nnz, 602 features, K=2) -- the configuration
north_star's roofline target is quoted on.  The normalisation (K1/K2), the column statistics and the
in-place standardisation (K3) are timed separately (`details`) and are inside `e2e`.  Prints ONE JSON
line (rank 0).  See DESIGN.md "Measurement" for the definitions of every key.

`--impl reference`: the UNMODIFIED reference (baseline/_ref, staged by tools/stage_reference.sh) on the
host cores -- its own aug_normalized_adjacency, sparse_mx_to_torch_sparse_tensor and sgc_precompute; it
imports nothing of this repo's package and never touches a GPU.
"""
import argparse
import contextlib
import importlib.util
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "K-hop propagation edge*feat/s (sgc_precompute)"
UNIT = "edge*feat/s"
REF_DIR = os.path.join(ROOT, "baseline", "_ref")
TOLERANCE = "|a-b| <= 1e-6 + 1e-5*|b| (north_star: 1e-5 relative / 1e-6 absolute per element)"


def env_int(name, default):
    return int(os.environ.get(name, default))


def load_synth():
    """sgc_b200/synth.py loaded by path: the input generator is pure torch, and loading it this way keeps
    the reference arm free of the package (and of libsgc_b200.so)."""
    mod = sys.modules.get("_sgc_synth")
    if mod is None:
        spec = importlib.util.spec_from_file_location("_sgc_synth", os.path.join(ROOT, "sgc_b200", "synth.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["_sgc_synth"] = mod
        spec.loader.exec_module(mod)
    return mod


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload):
    """Per-launch counters of the hop kernel from the committed `ncu --set full` capture of this workload
    (profiles/roofline_traffic.json: dram bytes, L2 hit rate, the capture's file and date), or None."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(path):
        with open(path) as fh:
            v = json.load(fh).get(workload)
        if isinstance(v, (int, float)):
            return {"dram_bytes": int(v)}
        return v
    return None


def kernel_source_sha256():
    """sha256 over the kernel sources the library is built from (sgc_b200/csrc/*.cu, *.cuh, Makefile, in name
    order).  The sources, not the .so: nvcc's output is not byte-reproducible from one build to the next."""
    import glob
    import hashlib
    h = hashlib.sha256()
    base = os.path.join(ROOT, "sgc_b200", "csrc")
    for path in sorted(glob.glob(os.path.join(base, "*.cu")) + glob.glob(os.path.join(base, "*.cuh"))
                       + [os.path.join(base, "Makefile")]):
        h.update(os.path.basename(path).encode() + b"\0")
        with open(path, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def capture_matches_build():
    """True when the kernel sources of the library being timed are the ones the committed ncu captures were
    taken from (profiles/roofline_traffic.json `source_sha256`), False when they differ, None when no hash was
    recorded."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(path) as fh:
            want = json.load(fh).get("source_sha256")
        got = kernel_source_sha256()
    except Exception:
        return None
    return None if not want else (got == want)


def hw_counter_block(cap, secs):
    """Which unit binds the kernel, by the hardware's own accounting: per-unit percent of sustained peak from the
    committed `ncu --set full` capture of this kernel and shape (the counters cannot be read outside a profiler).
    The data-pipe figures are re-scaled to this run's hop time: wavefronts per launch are a property of the
    kernel and its input, only the duration differs between the capture and the timed run."""
    pipes = cap.get("pipes_pct_of_peak_sustained_elapsed")
    if not pipes:
        return None
    out = {"pct_of_peak_sustained_under_ncu": pipes, "busiest_unit": cap.get("busiest_unit"),
           "lts_busiest_slice_pct": cap.get("lts_busiest_slice_pct"), "l1tex_busiest_sm_pct": cap.get("l1tex_busiest_sm_pct"),
           "duration_ms_under_ncu": cap.get("duration_ms_under_ncu"), "capture": cap.get("capture"),
           "capture_matches_timed_build": capture_matches_build()}
    wf, peak_wf, dur = cap.get("lsu_data_pipe_wavefronts"), cap.get("lsu_data_pipe_peak_wavefronts_in_launch"), cap.get("duration_ms_under_ncu")
    if wf and peak_wf and dur:
        peak_rate = peak_wf / (dur * 1e-3) / 1e9                    # Gwavefront/s: one per cycle per SM
        out["lsu_data_pipe"] = {
            "unit": "Gwavefront/s", "achieved": wf / secs / 1e9, "peak": peak_rate, "frac": wf / secs / 1e9 / peak_rate,
            "wavefronts_per_launch": wf, "of_which_shfl_broadcasts": cap.get("shfl_wavefronts"),
            "of_which_global_loads": cap.get("global_load_wavefronts"),
            "what": "L1TEX LSU data-pipe wavefronts per launch (ncu) / hop time of THIS run, against the pipe's "
                    "sustained peak of one wavefront per cycle per SM at the capture's clock"}
    return out


def hop_bytes(n, nnz, f):
    """Algorithmic bytes of one hop, gather model (SURVEY.md 8d / DESIGN.md): per stored entry one
    int32 column, one fp32 weight and one F-float row of X; per row one F-float store and one
    int64 row pointer."""
    return nnz * (4 + 4 + 4 * f) + 4 * f * n + 8 * (n + 1)


def compulsory_bytes(n, nnz, f):
    """Every array once (SURVEY 8d B_min): col + value per entry, row pointers, X read, Y written."""
    return 8 * nnz + 8 * (n + 1) + 8 * f * n


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            "hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
            "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4,
            "hw_power_brake_slowdown": 0x80,
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": float(self.max_mhz) if self.max_mhz else None,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def workload_config(name, n, nnz, f, k):
    """`config` of the JSON line: names the workload and nothing else, so that both arms (and every N)
    print the same object; measurements and schedule descriptions live in `details`."""
    label = {"reddit": "Reddit-shaped synthetic inductive SGC K=2 (BASELINE configs[2])",
             "products": "ogbn-products-shaped synthetic power-law graph K=3 (BASELINE configs[3])",
             "papers100m": "ogbn-papers100M-shaped synthetic K=3 (BASELINE configs[4])"}[name]
    cfg = {"workload": label, "nodes": int(n), "nnz_S": int(nnz), "features": int(f), "k_hops": int(k),
           "l2_hygiene": "inputs larger than L2 (X %.0f MB + CSR %.0f MB vs 126 MB L2); no flush"
                         % (n * f * 4 / 1e6, nnz * 12 / 1e6)}
    if name == "reddit":
        cfg["choice"] = ("largest single-GPU configuration of BASELINE.json and the one SURVEY 8(d) sets the target on; "
                         "configs[1] (Pubmed: X = 39 MB, L2-resident, launch-bound) is other_workloads.pubmed")
    return cfg


# --------------------------------------------------------------------------------------------------
# The reference's CPU implementation of the path
class Reference(object):
    """The reference's own functions for this path.  kind == "reference": the unmodified files staged
    under baseline/_ref (normalization.py, utils.py), imported as they are; kind == "port":
    oracle/reference_port.py, only when nothing is staged (the same calls, restated)."""

    def __init__(self):
        if os.path.exists(os.path.join(REF_DIR, "utils.py")) and os.path.exists(os.path.join(REF_DIR, "normalization.py")):
            saved = {k: sys.modules.pop(k) for k in ("utils", "normalization") if k in sys.modules}
            sys.path.insert(0, REF_DIR)
            try:
                import normalization as ref_norm
                import utils as ref_utils
            finally:
                sys.path.remove(REF_DIR)
                for k in ("utils", "normalization"):
                    sys.modules.pop(k, None)
                sys.modules.update(saved)
            for mod in (ref_norm, ref_utils):
                assert os.path.dirname(os.path.abspath(mod.__file__)) == REF_DIR, mod.__file__
            self.kind = "reference"
            self.where = "baseline/_ref (unmodified normalization.py + utils.py of the reference)"
            self.aug_normalized_adjacency = ref_norm.aug_normalized_adjacency
            self.to_torch_sparse = ref_utils.sparse_mx_to_torch_sparse_tensor
            self.sgc_precompute = ref_utils.sgc_precompute
        else:
            from oracle import reference_port as rp
            self.kind = "port"
            self.where = "oracle/reference_port.py (baseline/_ref is not staged: run tools/stage_reference.sh)"
            self.aug_normalized_adjacency = rp.aug_normalized_adjacency
            self.to_torch_sparse = rp.sparse_mx_to_torch_sparse_tensor
            self.sgc_precompute = rp.sgc_precompute

    def adjacency(self, a_scipy):
        """utils.py:120-122: adj_normalizer(adj) -> sparse_mx_to_torch_sparse_tensor(adj).float()"""
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return self.to_torch_sparse(self.aug_normalized_adjacency(a_scipy)).float()

    @staticmethod
    def standardize(x):
        """utils.py:119, the reference's expression (it sits inside load_reddit_data, not callable alone)."""
        return (x - x.mean(dim=0)) / x.std(dim=0)


def host_graph(s):
    """scipy CSR of A from the synthetic workload dict (host tensors)."""
    import scipy.sparse as sp
    return sp.csr_matrix((s["val"].numpy(), s["col"].numpy(), s["row_ptr"].numpy()), shape=(s["n"], s["n"]))


def run_reference(args):
    """--impl reference: the reference's CPU path on the same config.  Every step is one call of the
    unmodified utils.sgc_precompute(X, adj, K) on the WHOLE workload (all rows, all stored entries, all
    feature columns) -- torch.spmm's cost on the CPU has a large per-call part (it sorts/coalesces the COO
    tensor the reference's loader builds, every hop), so a row or column sample would misstate the rate.
    A whole step takes ~10-20 s on the host cores; when steps+warmup of them exceed the time budget
    (SGC_BENCH_REF_BUDGET_S, default 90 s) fewer whole steps are timed and the line says so: `steps` is the
    number actually timed, `details.steps_requested` what the command asked for."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["CUDA_VISIBLE_DEVICES"] = ""             # this arm never touches a GPU
    ref = Reference()
    synth = load_synth()
    name = args.workload
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    t0 = time.perf_counter()
    s = synth.make_shape(name, device="cpu", scale=float(os.environ.get("SGC_BENCH_SCALE", "1")))
    n, f, k = s["n"], s["f"], s["k"]
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    adj = ref.adjacency(host_graph(s))
    normalise_s = time.perf_counter() - t0
    nnz = int(adj._nnz())
    x = s["x"]
    t0 = time.perf_counter()
    if synth.SHAPES[name]["scaled"]:
        x = ref.standardize(x)
    standardise_s = time.perf_counter() - t0
    del s

    budget = float(os.environ.get("SGC_BENCH_REF_BUDGET_S", "90"))
    warm_times = []
    for _ in range(min(args.warmup, 1)):                 # one whole step; more warm-up is pointless at ~10 s a step
        _, secs = ref.sgc_precompute(x, adj, k)
        warm_times.append(secs)
    steps = args.steps
    if warm_times and warm_times[0] * steps > budget:
        steps = max(1, int(budget / warm_times[0]))
    total, times = 0.0, []
    for _ in range(steps):
        _, secs = ref.sgc_precompute(x, adj, k)
        total += secs
        times.append(secs)
    secs = total / steps
    value = k * nnz * f / secs
    sample = ("every step = utils.sgc_precompute(X, adj, %d) on the whole workload (%d rows, %d stored entries of S, "
              "%d columns), torch.spmm fp32 on %d threads; %d whole steps timed (%d requested, %.0f s budget), "
              "%d warm-up" % (k, n, nnz, f, threads, steps, args.steps, budget, len(warm_times)))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": len(warm_times), "ms_per_step": secs * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, n, nnz, f, k),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": ref.kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "details": {
            "code": ref.where,
            "steps_requested": args.steps, "warmup_requested": args.warmup, "time_budget_s": budget,
            "step_ms": [round(v * 1e3, 1) for v in times], "warmup_ms": [round(v * 1e3, 1) for v in warm_times],
            "outside_the_timed_steps_s": {"synthetic_inputs": gen_s,
                                          "aug_normalized_adjacency+sparse_mx_to_torch_sparse_tensor": normalise_s,
                                          "standardise (utils.py:119)": standardise_s},
            "whole_path_s": normalise_s + standardise_s + secs,
        },
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# Checks of the GPU output against the reference's CPU arithmetic (the checker, never the timed path)
def allclose_report(got, want):
    """north_star tolerance, element-wise; got / want: host fp32 tensors of one shape."""
    diff = (got.double() - want.double()).abs()
    bad = diff > (1e-6 + 1e-5 * want.double().abs())
    return {"elements": int(want.numel()), "violations": int(bad.sum().item()),
            "max_abs_diff": float(diff.max().item()) if diff.numel() else 0.0,
            "bit_identical_elements": int((got == want).sum().item()), "tolerance": TOLERANCE}


def sampled_rows_check(ref, graph, x_in, y_out, rows):
    """Rows `rows` (sorted int64 device tensor) of one hop output y_out = S x_in, recomputed on the CPU by
    the reference's torch.spmm from the GPU's own operand: the sampled rows of S become a small sparse
    matrix over the columns they touch (ascending order kept, so torch.spmm accumulates every element in
    the same sequential order), built by the reference's sparse_mx_to_torch_sparse_tensor."""
    import scipy.sparse as sp
    rp = graph.row_ptr
    beg, end = rp[rows], rp[rows + 1]
    lens = (end - beg)
    tot = int(lens.sum().item())
    seg = torch.repeat_interleave(torch.arange(rows.numel(), device=rows.device), lens)
    starts = torch.cumsum(lens, 0) - lens
    pos = beg[seg] + (torch.arange(tot, device=rows.device) - starts[seg])
    cols = graph.col[pos].long()
    vals = graph.values32()[pos]
    uniq, inv = torch.unique(cols, sorted=True, return_inverse=True)
    indptr = np.concatenate(([0], np.cumsum(lens.cpu().numpy()))).astype(np.int64)
    sub = sp.csr_matrix((vals.cpu().numpy(), inv.cpu().numpy().astype(np.int64), indptr),
                        shape=(rows.numel(), uniq.numel()))
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        adj = ref.to_torch_sparse(sub)
    want, _ = ref.sgc_precompute(x_in[uniq].float().cpu().contiguous(), adj, 1)
    rep = allclose_report(y_out[rows].cpu(), want)
    rep.update(rows=int(rows.numel()), stored_entries=tot)
    return rep


def sample_rows(n, count, device, seed=7):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return torch.sort(torch.randperm(n, generator=g)[:min(count, n)])[0].to(device)


def hopwise_parity(ref, graph, x, k, y_timed, count=1500):
    """Hop 1 and hop K of the timed K-hop output on sampled rows (the intermediate operand is recomputed by
    the same deterministic kernels, so it is bit-identical to what the timed step consumed)."""
    import sgc_b200
    rows = sample_rows(graph.n_rows, count, graph.device)
    h1 = sgc_b200.hop(graph, x)
    out = {"hop1": sampled_rows_check(ref, graph, x, h1, rows)}
    if k > 1:
        prev = h1 if k == 2 else sgc_b200.propagate.propagate(graph, x, k - 1)
        out["hop%d" % k] = sampled_rows_check(ref, graph, prev, y_timed, rows)
    out["ok"] = all(v["violations"] == 0 for v in out.values() if isinstance(v, dict))
    out["checker"] = "the reference's torch.spmm on CPU (%s) on %d sampled rows per hop" % (ref.kind, rows.numel())
    return out


def fp64_gate(adj, x_raw, mean64, std64, k, y_gpu, y_ref, chunk_cols=128):
    """The accuracy gate of the path that is timed and shipped for Reddit (K3 -> hops).  Truth = utils.py:118-119
    and the K hops of utils.py:94-95 in fp64 on the CPU, on the raw X and the reference's own fp32 S values
    promoted to fp64 (so the only differences left are the fp32 roundings each implementation makes).  Every
    element must satisfy   |gpu - truth| <= |reference_fp32 - truth| + 1e-6 + 1e-5*|truth|   -- the GPU is never
    further from the exact result than the unmodified reference is, plus north_star's tolerance.  All rows, all
    columns, in column chunks to bound host memory."""
    t0 = time.perf_counter()
    adj64 = adj.to(torch.float64)
    n, f = x_raw.shape
    out = {"elements": int(n * f), "violations": 0, "max_err_gpu": 0.0, "max_err_reference": 0.0,
           "gpu_at_least_as_close_elements": 0, "sum_err_gpu": 0.0, "sum_err_reference": 0.0}
    for c0 in range(0, f, chunk_cols):
        c1 = min(f, c0 + chunk_cols)
        y64 = (x_raw[:, c0:c1].double() - mean64[c0:c1]) / std64[c0:c1]
        for _ in range(k):
            y64 = torch.spmm(adj64, y64)
        e_gpu = (y_gpu[:, c0:c1].double() - y64).abs()
        e_ref = (y_ref[:, c0:c1].double() - y64).abs()
        out["violations"] += int((e_gpu > e_ref + 1e-6 + 1e-5 * y64.abs()).sum().item())
        out["max_err_gpu"] = max(out["max_err_gpu"], float(e_gpu.max().item()))
        out["max_err_reference"] = max(out["max_err_reference"], float(e_ref.max().item()))
        out["gpu_at_least_as_close_elements"] += int((e_gpu <= e_ref).sum().item())
        out["sum_err_gpu"] += float(e_gpu.sum().item())
        out["sum_err_reference"] += float(e_ref.sum().item())
    out["mean_err_gpu"] = out.pop("sum_err_gpu") / out["elements"]
    out["mean_err_reference"] = out.pop("sum_err_reference") / out["elements"]
    out["ok"] = out["violations"] == 0
    out["seconds"] = time.perf_counter() - t0
    out["what"] = ("truth = (x - mean)/std and K x torch.spmm in fp64 on the CPU (fp32 S values promoted); gate: "
                   "|gpu - truth| <= |reference - truth| + 1e-6 + 1e-5*|truth| on every element; fatal")
    return out


def reference_leg(ref, s, name, k, y_gpu, graph, stats):
    """cpu_baseline + full-size parity in one go (rank 0, N=1): the reference's whole path on the host --
    its aug_normalized_adjacency on the raw A, utils.py:119 on the raw X, ONE whole K-hop
    utils.sgc_precompute (timed).  Its output is compared element by element with
      (1) the GPU's K hops started from the SAME standardised X the reference propagated (north_star's bar:
          same inputs, every element within tolerance) -- S built by this repo's K1/K2 from the raw A;
      (2) the timed GPU output, whose X was standardised on the GPU (K3).  K3 accumulates the column statistics
          in fp64 and rounds once; torch's CPU mean/std reduce in fp32, so the reference's own statistics carry
          a few ulp of rounding noise (reported as `statistics`), which the division by a small std and the
          propagation turn into a per-column offset of the output.  The element-wise difference to the reference
          is reported; what is GATED is accuracy: utils.py:118-119 + the K hops evaluated in fp64 on the CPU
          (same fp32 S values, promoted) give the exact answer, and every element of the timed GPU output must be
          no further from it than the reference's own output is, plus north_star's tolerance (`vs_fp64`)."""
    import sgc_b200
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    host = {kk: s[kk].cpu() for kk in ("row_ptr", "col", "val")}
    host["n"] = s["n"]
    adj = ref.adjacency(host_graph(host))
    x_raw = s["x"].cpu()
    x = x_raw
    synth = load_synth()
    scaled = synth.SHAPES[name]["scaled"]
    if scaled:
        x = ref.standardize(x_raw)
    nnz, f = int(adj._nnz()), int(x.shape[1])
    ref.sgc_precompute(x[:, :8].contiguous(), adj, 1)           # thread pool / page warm-up, 1 % of a step
    want, secs = ref.sgc_precompute(x, adj, k)
    cpu = {"value": k * nnz * f / secs, "unit": UNIT, "cores": threads, "kind": ref.kind,
           "sample": "ONE whole K=%d step (all %d rows, %d stored entries, %d columns) of utils.sgc_precompute, "
                     "torch.spmm fp32 on CPU, %.2f s; code: %s" % (k, s["n"], nnz, f, secs, ref.where)}
    same = sgc_b200.propagate.propagate(graph, sgc_b200.to_padded(x, graph.device), k)
    parity = allclose_report(same.cpu(), want)
    parity["ok"] = parity["violations"] == 0
    parity["what"] = ("GPU: S from the raw A by K1+K2, then K hops from the X the %s standardised (same inputs as its "
                      "torch.spmm path) vs the %s's whole CPU path; all %d x %d elements"
                      % (ref.kind, ref.kind, s["n"], f))
    if scaled:
        timed = allclose_report(y_gpu.cpu(), want)
        timed["what"] = ("the TIMED output (X standardised on the GPU by K3: fp64-accumulated statistics, exact "
                         "(x-mean)/std) vs the reference, whose fp32 mean/std carry rounding noise -- reported only")
        mean_ref, std_ref = x_raw.mean(dim=0), x_raw.std(dim=0)
        m64, s64 = x_raw.double().mean(dim=0), x_raw.double().std(dim=0)
        mean_gpu, std_gpu = (t.cpu() for t in stats)
        ulp = lambda a, b: float(((a.double() - b.double()).abs() / torch.finfo(torch.float32).eps
                                  / b.double().abs().clamp_min(1e-30)).max())
        timed["statistics"] = {
            "gpu_vs_fp64_numpy_like_ulps": {"mean": ulp(mean_gpu, m64), "std": ulp(std_gpu, s64)},
            "reference_vs_fp64_ulps": {"mean": ulp(mean_ref, m64), "std": ulp(std_ref, s64)},
            "max_offset_of_standardised_x": float(((mean_ref - mean_gpu).abs() / std_gpu).max())}
        timed["vs_fp64"] = fp64_gate(adj, x_raw, m64, s64, k, y_gpu.cpu(), want)
        parity["ok"] = parity["ok"] and timed["vs_fp64"]["ok"]
        parity["timed_output_with_gpu_standardisation"] = timed
    else:
        parity["timed_output_bit_identical_to_checked_output"] = bool(torch.equal(same, y_gpu))
    return cpu, parity


# --------------------------------------------------------------------------------------------------
def measure_l2_peak(dev):
    """L2 -> SM read bandwidth of this box, measured now (sgc_l2_read_probe, CUDA events, best of 3):
    L2-resident buffers of 32 and 64 MB read in address order and in hashed order, and the Reddit hop's own
    footprint (512-byte pieces of 232,965 rows at a 2432-byte pitch, hashed order).  `gbs` -- the roofline
    denominator -- is the BEST of them (the most the path was seen to deliver, whatever the pattern)."""
    from sgc_b200 import _ffi
    sink = torch.zeros(1, dtype=torch.float32, device=dev)
    stream = lambda: torch.cuda.current_stream().cuda_stream

    def rate(buf, stride, reps, pattern, moved):
        call = lambda: _ffi.call("sgc_l2_read_probe", buf.data_ptr(), buf.numel() * 4, stride, reps, pattern, 8,
                                 sink.data_ptr(), stream())
        call()
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); call(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return moved * reps / (min(ts) * 1e-3) / 1e9

    points = {}
    for mb in (32, 64):
        buf = torch.ones(mb * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
        for pattern in (0, 1, 2, 3):
            key = "%dMB_%s_%s" % (mb, "hashed" if pattern & 1 else "address", "ldg_hint" if pattern & 2 else "no_allocate")
            points[key] = rate(buf, 0, 40, pattern, buf.numel() * 4)
        del buf
    n, pitch = 232965, 2432
    buf = torch.ones(n * pitch // 4, dtype=torch.float32, device=dev)
    points["reddit_slab_footprint_hashed_no_allocate"] = rate(buf, pitch, 20, 1, n * 512)
    points["reddit_slab_footprint_hashed_ldg_hint"] = rate(buf, pitch, 20, 3, n * 512)
    del buf
    best = max(points, key=points.get)
    return {"gbs": points[best], "best_point": best, "points_gbs": points,
            "how": "sgc_l2_read_probe: every warp keeps eight 512-byte pieces (32 lanes x 16-byte loads; L1::no_allocate, or "
                   "the hop kernel's own L1-allocating load with an L2 evict_last hint) of an L2-resident buffer in "
                   "flight, 8 CTAs x 256 threads per SM, best of 3, CUDA events; measured in this run"}


def run_own(args):
    import torch.distributed as dist
    import sgc_b200
    from sgc_b200 import _ffi
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner to fd 1 from C; keep stdout for the ONE JSON line: everything
        # written to fd 1 from here on goes to stderr, the JSON line is written to the saved descriptor
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        sys.stdout = os.fdopen(saved, "w")
        dist.init_process_group("nccl", device_id=dev)
    if world > 1:
        import bench_dist
        return bench_dist.bench(args, rank, world, dev, globals())

    synth = load_synth()
    name = args.workload
    s = synth.make_shape(name, device=dev)
    n, f, k = s["n"], s["f"], s["k"]
    torch.cuda.synchronize()
    # ---- one-off preparation, timed separately (K1 + K2, K3) ----
    ev = lambda: torch.cuda.Event(enable_timing=True)
    def timed(fn):
        a, b = ev(), ev()
        torch.cuda.synchronize()
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        return out, a.elapsed_time(b)

    _, normalise_cold_ms = timed(lambda: sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"]))
    graph, normalise_ms = timed(lambda: sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"]))
    # features in the layout the loaders produce (utils.load_reddit_data): padded row stride
    x = sgc_b200.to_padded(s["x"])
    if x.data_ptr() == s["x"].data_ptr():
        x = x.clone()                       # s["x"] stays raw for the reference leg
    mean32 = std32 = x_raw = None
    stats_ms = standardise_ms = 0.0
    scaled = synth.SHAPES[name]["scaled"]
    if scaled:
        (mean32, std32), stats_ms = timed(lambda: sgc_b200.col_stats(x))
        x_raw = sgc_b200.features.padded_rows(n, f, dev)   # raw copy kept for the fused-hop variants below
        x_raw.copy_(x)
        _, standardise_ms = timed(lambda: sgc_b200.standardize_(x, mean32, std32))   # utils.py:119, exact, in place
    nnz = graph.nnz
    graph.plan(sgc_b200.propagate.default_chunk())
    sval_ms = 0.0
    if not sgc_b200.propagate.fused_norm():
        _, sval_ms = timed(graph.values32)

    def step():
        # what utils.sgc_precompute times in the reference: K hops on (standardised) features
        return sgc_b200.propagate.propagate(graph, x, k)

    for _ in range(args.warmup):
        y = step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = _ffi.launch_count()
    start, stop = ev(), ev()
    start.record()
    for _ in range(args.steps):
        y = step()
    stop.record()
    torch.cuda.synchronize()
    launches = _ffi.launch_count() - launches0
    total_ms = start.elapsed_time(stop)
    clocks = sampler.stop()
    ms_per_step = total_ms / args.steps
    value = k * nnz * f / (ms_per_step * 1e-3)

    # ---- per-hop kernel time (CUDA events on the launching stream) ----
    def hop_time(fn):
        ts = []
        for _ in range(max(3, min(args.steps, 10))):
            a, b = ev(), ev()
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    ybuf = sgc_b200.features.padded_rows(n, f, dev)
    hop_ms = {"plain": hop_time(lambda: sgc_b200.hop(graph, x, out=ybuf))}
    if scaled and not args.skip_variants:
        for mode in ("rcp", "div"):
            hop_ms["fused_std_" + mode] = hop_time(
                lambda: sgc_b200.hop(graph, x_raw, mean=mean32, std=std32, std_mode=mode, out=ybuf))
    del x_raw, ybuf
    incumbent = None
    if not args.skip_incumbent:
        try:
            incumbent = gpu_incumbent(graph, x, f, hop_time, hop_ms["plain"])
        except Exception as exc:        # reported for context only: never costs the primary line
            incumbent = {"error": repr(exc)[:300]}
        torch.cuda.empty_cache()
    roofline = roofline_block(name, n, nnz, f, ms_per_step / k, dev)
    roofline["hop_ms"] = hop_ms

    # ---- the reference's CPU path on this box's host cores + parity of the timed output ----
    cpu = parity = None
    if not args.no_cpu_baseline:
        ref = Reference()
        cpu, parity = reference_leg(ref, s, name, k, y, graph, (mean32, std32))
        parity["sampled_rows_per_hop"] = hopwise_parity(ref, graph, x, k, y)
        if not (parity["ok"] and parity["sampled_rows_per_hop"]["ok"]):
            print(json.dumps({"error": "GPU output differs from the reference beyond tolerance", "parity": parity}),
                  file=sys.stderr)
            sys.exit(3)

    # ---- end to end through the public API with HOST buffers ----
    e2e = None if args.skip_e2e else run_e2e(args, s, name, k, nnz, dev)

    # ---- the other single-GPU-sized BASELINE config, for the 1/2/4/8-GPU series (not the headline) ----
    others = {}
    extra_names = os.environ.get("SGC_BENCH_EXTRA", "products" if not args.workload_given else "")
    del graph, x, y, s
    torch.cuda.empty_cache()
    for other in [e for e in extra_names.split(",") if e and e != name]:
        try:
            others[other] = single_gpu_extra(other, dev, check=not args.no_cpu_baseline)
        except Exception as exc:        # an optional extra must never cost the primary line
            others[other] = {"error": repr(exc)[:300]}

    if not args.workload_given:
        try:
            others["pubmed"] = pubmed_extra(dev)
        except Exception as exc:
            others["pubmed"] = {"error": repr(exc)[:300]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, n, nnz, f, k),
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "parity": parity,
        "details": {
            "normalise_ms": normalise_ms, "normalise_first_call_ms": normalise_cold_ms,
            "col_stats_ms": stats_ms, "standardise_inplace_ms": standardise_ms,
            "chunk": sgc_b200.propagate.default_chunk(),
            "weights": ("fp32((d_i*a_ij)*d_j) formed in the hop kernel (SGC_B200_FUSED_NORM=1)"
                        if sgc_b200.propagate.fused_norm() else
                        "fp32 S values written once by sgc_norm_values (inside normalise_ms, outside "
                        "the timed hops like the reference's normalisation), streamed by every hop"),
            "timed_region": "K hops on device-resident CSR + standardised X, as utils.py:93-96 times them; "
                            "K1/K2 (normalise_ms), K3 (col_stats_ms + standardise_inplace_ms: one exact in-place "
                            "pass, NOT fused into hop 1 -- measured slower, see roofline.hop_ms.fused_std_*) are "
                            "timed separately here and are inside e2e"},
        "gpu_incumbent": incumbent, "other_workloads": others,
    }
    print(json.dumps(line))


def roofline_block(name, n, nnz, f, hop_ms, dev, per_gpu_fraction=1.0, l2=True):
    """`roofline` of the hop kernel.  The hop is a gather: its algorithmic bytes (SURVEY 8d gather model)
    are served by L2 wherever a row of X is re-read, by DRAM otherwise.  bound = "l2" when a feature slab of
    X is L2-resident (Reddit shape: the kernel is limited by the L2 -> SM path, its DRAM traffic is a
    fraction of the algorithmic bytes), "hbm" when X is far larger than L2 (products / papers shapes)."""
    hbm, hbm_src = peaks()
    b_hop = hop_bytes(n, nnz, f) * per_gpu_fraction
    b_min = compulsory_bytes(n, nnz, f) * per_gpu_fraction
    secs = hop_ms * 1e-3
    gather = b_hop / secs / 1e9
    slab_resident = n * min(f, 32) * 4 <= 128 * 1024 * 1024
    cap = ncu_traffic(name) if per_gpu_fraction == 1.0 else None
    dram = {"peak": hbm, "peak_source": hbm_src, "unit": "GB/s",
            "compulsory_bytes_per_launch": b_min, "compulsory_frac": b_min / secs / 1e9 / hbm}
    traffic = None
    if cap:
        traffic = cap.get("dram_bytes")
        dram.update(achieved=traffic / secs / 1e9, frac=traffic / secs / 1e9 / hbm,
                    traffic_over_compulsory=traffic / b_min, capture=cap)
    out = {"kernel": "spmm_hop_kernel", "unit": "GB/s", "algorithmic_bytes_per_launch": b_hop,
           "gather_model_gbs": gather, "traffic": traffic, "dram": dram}
    if slab_resident and l2:
        peak = measure_l2_peak(dev)
        l2_bytes = cap.get("l2_to_sm_bytes") if cap else None
        l2_rate = (l2_bytes / secs / 1e9) if l2_bytes else gather
        hw = hw_counter_block(cap, secs) if cap else None
        if hw and hw.get("lsu_data_pipe") and hw["busiest_unit"] == "l1tex_lsu_data_pipe":
            # the hardware counters name the SM-side LSU data pipe as the busiest unit (gathers + the SHFL broadcasts
            # of (col, weight)); L2 and HBM have headroom.  The top-level figures are that unit's, in its own unit.
            pipe = hw["lsu_data_pipe"]
            out.update(bound="l1tex_lsu_data_pipe", unit=pipe["unit"], achieved=pipe["achieved"], peak=pipe["peak"],
                       frac=pipe["frac"], peak_source="hardware: 1 wavefront/cycle/SM (ncu pct_of_peak_sustained of the "
                                                     "committed capture), re-scaled to this run's hop time",
                       bound_is="the unit with the highest percent of sustained peak in hw_counters")
        else:
            out.update(bound="l2", achieved=gather, peak=peak["gbs"], frac=gather / peak["gbs"],
                       peak_source="L2 -> SM read bandwidth measured in this run (sgc_l2_read_probe, best point)")
        out["hw_counters"] = hw
        out["hbm_model"] = {"unit": "GB/s", "achieved": gather, "peak": hbm, "frac": gather / hbm, "peak_source": hbm_src,
                            "what": "SURVEY 8(d): algorithmic (gather-model) bytes per launch / hop time against the "
                                    "measured HBM copy peak; above 1 = re-reads of the L2-resident slab served by L2, "
                                    "not skipped work (all outputs are checked)"}
        out.update(l2=dict(
                       achieved=l2_rate, peak=peak["gbs"], frac=l2_rate / peak["gbs"], probe=peak,
                       achieved_is=("ncu l1tex__m_xbar2l1tex_read_bytes of the committed capture / hop time of this run"
                                    if l2_bytes else "gather-model bytes / hop time"),
                       note="secondary figure: the probe reads L2 through the same LSU path as the kernel, so it is an "
                            "SM-side ceiling for ONE access pattern, not L2's peak -- the kernel can exceed it.  The "
                            "gather-model bytes exceed the bytes that cross L2 -> SM by the L1 hits on hub rows (8 % in "
                            "the capture)"))
        out["frac_of_hbm_peak_gather_model"] = gather / hbm
    else:
        # X does not stay in L2: the kernel is DRAM-bound; the fraction is real DRAM traffic over the copy peak
        # when a capture exists, else the gather model (an upper bound on the traffic the kernel may cause)
        ach = dram.get("achieved", gather)
        out.update(bound="hbm", achieved=ach, peak=hbm, frac=ach / hbm, peak_source=hbm_src,
                   achieved_is="ncu dram bytes / hop time" if "achieved" in dram else "gather-model bytes / hop time")
        out["hw_counters"] = hw_counter_block(cap, secs) if cap else None
    return out


def gpu_incumbent(graph, x, f, hop_time, ours_ms):
    """the on-box GPU incumbent for one hop on the same S and X, as context for the
    kernel time (library calls, outside every timed region of this bench): the reference's own CUDA path
    -- torch.spmm on the uncoalesced fp32 COO tensor its loaders build (utils.py:23-30, 95) -- and
    cuSPARSE on a CSR tensor of the same matrix."""
    coo = graph.to_torch_sparse()
    xd = x[:, :f].contiguous()
    torch.spmm(coo, xd)
    coo_ms = hop_time(lambda: torch.spmm(coo, xd))
    csr = coo.coalesce().to_sparse_csr()
    torch.mm(csr, xd)
    csr_ms = hop_time(lambda: torch.mm(csr, xd))
    return {"unit": "ms per hop", "torch_spmm_coo": coo_ms, "torch_mm_csr_cusparse": csr_ms,
            "spmm_hop_kernel": ours_ms,
            "note": "torch.spmm(COO) is what utils.py:95 runs on a GPU in the reference; library calls, reported only"}


def single_gpu_extra(name, dev, check=True):
    """K hops of another named workload on this one GPU (device-timed, 3 warm-ups + 5 steps), with the
    DRAM-side roofline and a sampled-row check of the timed output against the reference's torch.spmm."""
    import sgc_b200
    synth = load_synth()
    s = synth.make_shape(name, device=dev)
    n, f, k = s["n"], s["f"], s["k"]
    graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
    x = sgc_b200.to_padded(s["x"])
    del s
    for _ in range(3):
        sgc_b200.propagate.propagate(graph, x, k)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        y = sgc_b200.propagate.propagate(graph, x, k)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    nnz = graph.nnz
    out = {"workload": workload_config(name, n, nnz, f, k)["workload"], "n_gpus": 1, "value": k * nnz * f / (ms * 1e-3),
           "unit": UNIT, "ms_per_step": ms, "nnz_S": nnz, "roofline": roofline_block(name, n, nnz, f, ms / k, dev)}
    if check:
        out["parity"] = hopwise_parity(Reference(), graph, x, k, y)
    del graph, x, y
    torch.cuda.empty_cache()
    return out


def pubmed_extra(dev):
    """BASELINE configs[1]: Pubmed SGC K=2 on one B200 -- the real graph (tests/golden/pubmed.npz:
    19,717 nodes, 108,365 stored entries of S) with the synthetic 500-dim TF-IDF-like features the
    golden vectors were generated with (ind.pubmed.allx is missing from the reference snapshot).
    X (39 MB) fits L2: the hops are launch/latency-bound, reported for completeness."""
    import sgc_b200
    synth = load_synth()
    g = np.load(os.path.join(ROOT, "tests", "golden", "pubmed.npz"), allow_pickle=False)
    n = g["adj_indptr"].size - 1
    graph = sgc_b200.normalize_csr(n, *(torch.from_numpy(np.ascontiguousarray(g["adj_" + kk])).to(dev)
                                        for kk in ("indptr", "indices", "data")))
    x = sgc_b200.to_padded(synth.tfidf_features(n, 500, density=0.1, seed=1).to(dev))
    for _ in range(5):
        sgc_b200.propagate.propagate(graph, x, 2)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50):
        y = sgc_b200.propagate.propagate(graph, x, 2)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 50
    import hashlib
    ok = hashlib.sha256(y.cpu().numpy().tobytes()).hexdigest() == str(g["Y_sha"])
    return {"workload": "Pubmed SGC K=2 (BASELINE configs[1]; real graph, synthetic features)", "n_gpus": 1,
            "value": 2 * graph.nnz * 500 / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "nnz_S": graph.nnz,
            "matches_reference_sha256": bool(ok), "l2_hygiene": "X (39 MB) is L2-resident: warm-cache, launch-bound"}


@contextlib.contextmanager
def env_override(**kw):
    old = {k: os.environ.get(k) for k in kw}
    os.environ.update({k: str(v) for k, v in kw.items()})
    try:
        yield
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def run_e2e(args, s, name, k, nnz_s, dev):
    """Same metric through the reference-facing API with host buffers: every step uploads the raw
    adjacency (CSR of A) and the raw features from pinned host memory, normalises on the GPU
    (K1+K2), computes the column statistics (K3), runs the K hops and reads the propagated
    features back to pinned host memory.  Measured twice: with the exact host scan that recognises a
    constant-valued adjacency (the reference's loaders only produce those) and fills its value array on the
    GPU, and with the scan off (`weighted_adjacency`: what an adjacency with real weights costs)."""
    import sgc_b200
    synth = load_synth()
    n, f = s["n"], s["f"]
    host = {kk: s[kk].cpu().pin_memory() for kk in ("row_ptr", "col", "val", "x")}
    out_host = torch.empty((n, f), dtype=torch.float32).pin_memory()
    stats = {}

    def step():
        # sgc_b200.precompute_host: CSR(A) up first, K1/K2 overlap the upload of X, the last hop runs
        # by row blocks whose D2H overlaps the next block (sgc_b200/pipeline.py)
        sgc_b200.precompute_host((n, host["row_ptr"], host["col"], host["val"]), host["x"], k,
                                 standardize=bool(synth.SHAPES[name]["scaled"]), out=out_host, stats=stats)

    widths = sgc_b200.pipeline.slab_widths(f, sgc_b200.pipeline.default_col_slab(n, f))
    schedule = ("column slabs of %s floats, strided 2-D DMA copies" % "+".join(str(w) for w in widths)
                if len(widths) > 1 else
                "one contiguous copy; last hop in %d row blocks" % sgc_b200.pipeline.default_out_blocks())
    steps = max(2, min(args.steps, 5))

    def measure():
        step()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        return (time.perf_counter() - t0) / steps

    secs = measure()
    st = dict(stats)
    with env_override(SGC_B200_UNIFORM_VALUES="0"):
        secs_w = measure()
    st_w = dict(stats)
    # same call, X and the result in pinned staging buffers whose row pitch is a multiple of 64 bytes
    # (sgc_b200.pipeline.pinned_rows: 2432 instead of 2408 bytes): the strided slab copies then start and end on
    # cache-line boundaries
    aligned = None
    if len(widths) > 1:
        xa, oa = sgc_b200.pipeline.pinned_rows(n, f), sgc_b200.pipeline.pinned_rows(n, f)
        xa.copy_(host["x"])
        plain_x, plain_out = host["x"], out_host

        def step_aligned():
            sgc_b200.precompute_host((n, host["row_ptr"], host["col"], host["val"]), xa, k,
                                     standardize=bool(synth.SHAPES[name]["scaled"]), out=oa, stats=stats)

        step_aligned()
        t0 = time.perf_counter()
        for _ in range(steps):
            step_aligned()
        secs_a = (time.perf_counter() - t0) / steps
        aligned = {"value": k * nnz_s * f / secs_a, "ms_per_step": secs_a * 1e3,
                   "same_bits_as_contiguous_layout": bool(torch.equal(oa, plain_out)),
                   "what": "same call with X and the result in sgc_b200.pipeline.pinned_rows staging buffers (host row "
                           "pitch %d bytes instead of %d: slab pieces start and end on 64-byte lines)"
                           % (4 * xa.stride(0), 4 * f)}
        del xa, oa
    host_bytes = sum(t.numel() * t.element_size() for t in host.values())
    values = ("the %d stored values of A are all %g (adj + adj.T of a 0/1 matrix, utils.py:116; found by a threaded "
              "exact scan on the host, inside the timed region, every step): filled on the GPU instead of uploaded "
              "-- %d of the %d input bytes cross PCIe"
              % (host["val"].numel(), st["uniform_value"], st["h2d_bytes"], host_bytes)) \
        if st["uniform_value"] is not None else \
        "all %d input bytes cross PCIe" % host_bytes
    return {"value": k * nnz_s * f / secs, "unit": UNIT, "h2d_bytes_per_step": int(st["h2d_bytes"]),
            "d2h_bytes_per_step": int(st["d2h_bytes"]), "ms_per_step": secs * 1e3, "steps": steps,
            "host_input_bytes": int(host_bytes),
            "weighted_adjacency": {"value": k * nnz_s * f / secs_w, "ms_per_step": secs_w * 1e3,
                                   "h2d_bytes_per_step": int(st_w["h2d_bytes"]),
                                   "what": "same call with SGC_B200_UNIFORM_VALUES=0: the value array of A is "
                                           "uploaded like any weighted adjacency's"},
            "aligned_host_rows": aligned,
            "scan_gave_up": bool(st.get("scan_gave_up", False)),
            "path": "sgc_b200.precompute_host: pinned host CSR(A)+X -> H2D (row_ptr + col first; X in %s; %s) -> "
                    "aug_normalized_adjacency (K1,K2, overlapped with the upload of X) -> per slab: col_stats + "
                    "standardise (K3), K hops, strided D2H into the pinned host result while the next slab is "
                    "uploaded and computed" % (schedule, values)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default=None, choices=[None, "reddit", "products", "papers100m"])
    ap.add_argument("--no-cpu-baseline", action="store_true",
                    help="profiling runs only: skips the reference's CPU leg and the parity check it feeds")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only: leaves e2e null")
    ap.add_argument("--skip-incumbent", action="store_true", help="do not time torch.spmm / cuSPARSE beside the hop")
    ap.add_argument("--skip-variants", action="store_true", help="do not time the fused-standardisation hop variants")
    args = ap.parse_args()
    args.workload_given = args.workload is not None
    if args.workload is None:
        args.workload = "reddit"
    args.warmup = max(args.warmup, 3) if args.impl == "own" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
