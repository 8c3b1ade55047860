#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/summary.txt
tail -2 gpurun_out/smoke.log
timeout 300 python tools/sweep_persistent.py > gpurun_out/sweep_persistent.log 2>&1; echo "sweep rc=$?" | tee -a gpurun_out/summary.txt
grep -v "Warn\|warn" gpurun_out/sweep_persistent.log | tail -16
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/summary.txt
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print("ms_per_step", d["ms_per_step"], "value", d["value"], "roofline frac", d["roofline"]["frac"], d["roofline"]["peak"], d["roofline"]["hop_ms"])
print("e2e", d["e2e"]["ms_per_step"], d["e2e"]["weighted_adjacency"]["ms_per_step"], "parity", d["parity"]["ok"], d["parity"]["violations"])
print("products", d["other_workloads"]["products"]["ms_per_step"], d["other_workloads"]["products"]["parity"]["ok"], "pubmed", d["other_workloads"]["pubmed"]["ms_per_step"], d["other_workloads"]["pubmed"]["matches_reference_sha256"])
PY
