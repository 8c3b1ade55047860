import sys, time, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sgc_b200
from sgc_b200 import synth
dev = torch.device("cuda", 0)
s = synth.make_shape("products", device=dev)
n, f, k = s["n"], s["f"], s["k"]
host = (n, s["row_ptr"].cpu().pin_memory(), s["col"].cpu().pin_memory(), s["val"].cpu().pin_memory())
xh = s["x"].cpu().pin_memory()
graph = sgc_b200.normalize_csr(n, s["row_ptr"], s["col"], s["val"])
want, _ = sgc_b200.sgc_precompute(sgc_b200.to_padded(s["x"]), graph, k)
want = want.cpu()
del s, graph
torch.cuda.empty_cache()
out = torch.empty((n, f), dtype=torch.float32).pin_memory()
for blocks in (1, 8):
    sgc_b200.precompute_host(host, xh, k, out=out, out_blocks=blocks)
    t0 = time.perf_counter()
    for _ in range(3):
        sgc_b200.precompute_host(host, xh, k, out=out, out_blocks=blocks)
    print("products precompute_host out_blocks=%d: %.1f ms  identical=%s" % (blocks, (time.perf_counter() - t0) / 3 * 1e3, torch.equal(out, want)), flush=True)
