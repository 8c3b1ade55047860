#!/bin/bash
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 tools/nvlink_probe2.py products > gpurun_out/nvlink_probe2_n2.log 2>&1; echo "probe rc=$?" | tee gpurun_out/summary_n2b.txt
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/nvlink_probe2_n2.log | tail -30
SECONDS=0
SGC_BENCH_LAMBDAS=0 timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2b.json 2> gpurun_out/bench_n2b.err; echo "bench n2 rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary_n2b.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_n2b.json'))
print(d["ms_per_step"], d["parity"]["reddit"], d["strong_scaling_efficiency"])
p=d["other_workloads"]["products"]; print(p["ms_per_step"], p["strong_scaling_efficiency"], p["parity"]["ok"], p["per_hop_breakdown"])
PY
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n2b.err | tail -8
