"""Diagnostic for the Reddit F1 parity test: run-to-run spread of the reference's stock CUDA path and of the
drop-in on the synthetic Reddit, and the element-wise difference of the features both produce.
    python tools/reddit_f1_diag.py [epochs...]"""
import os, re, shutil, subprocess, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_reference_scripts as T
from pathlib import Path
from sgc_b200 import synth

tmp = Path(tempfile.mkdtemp())
stock, dropin = T.run_dir(tmp, "stock", True), T.run_dir(tmp, "dropin", False)
data = tmp / "reddit_data"
print(synth.write_reddit_npz(str(data), n=24000, f=602, n_classes=41, seed=0))
for d in (stock, dropin):
    os.unlink(d / "data"); os.symlink(str(data), d / "data")
DUMP = ("import sys; sys.path.insert(0, ''); import torch, utils\n"
        "adj, tadj, x, y, it, iv, ite = utils.load_reddit_data('AugNormAdj')\n"
        "pf, _ = utils.sgc_precompute(x, adj, 2)\n"
        "tf, _ = utils.sgc_precompute(x[it], tadj, 2)\n"
        "torch.save({'pf': pf.cpu(), 'tf': tf.cpu(), 'x': (x * 1.0).cpu()}, sys.argv[1])\n")
T.run_script(stock, ["-c", DUMP, str(tmp / "stock.pt")], drop_in=False)
T.run_script(dropin, ["-c", DUMP, str(tmp / "own.pt")], drop_in=True)
a, b = torch.load(tmp / "stock.pt"), torch.load(tmp / "own.pt")
for k in ("x", "pf", "tf"):
    d = (a[k].double() - b[k].double()).abs()
    viol = (d > 1e-6 + 1e-5 * a[k].double().abs()).sum().item()
    print("%s: shape %s max|diff| %.3e mean|diff| %.3e violations %d of %d" % (k, tuple(a[k].shape), d.max(), d.mean(), viol, d.numel()))
T.run_script(stock, ["-c", DUMP, str(tmp / "stock2.pt")], drop_in=False)
a2 = torch.load(tmp / "stock2.pt")
for k in ("x", "pf", "tf"):
    print("stock run-to-run %s: max|diff| %.3e" % (k, (a[k] - a2[k]).abs().max()))
epochs_list = sys.argv[1:] or ["2"]
for ep in epochs_list:
    for flags in (["--inductive", "--test"], ["--test"], []):
        fl = flags + ["--epochs", ep]
        ref = [T.reddit_f1(T.run_script(stock, ["reddit.py"] + fl, drop_in=False)) for _ in range(3)]
        own = [T.reddit_f1(T.run_script(dropin, ["reddit.py"] + fl, drop_in=True)) for _ in range(3)]
        print("epochs %s %-22s stock %s   drop-in %s" % (ep, " ".join(flags), ref, own), flush=True)
shutil.rmtree(tmp, ignore_errors=True)
