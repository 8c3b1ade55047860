#!/bin/bash
# 2-GPU run: multi-device parity tests; Reddit-shaped leg on the 1x2 (rows) and 2x1 (features) grids with the
# peer-tile size swept; host-to-host trace of the feature-sharded path.
set -u
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q > gpurun_out/pytest_n2.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/summary_n2.txt
tail -6 gpurun_out/pytest_n2.log
port=29530
for cfg in "2x1 0" "1x2 0" "1x2 1" "1x2 2" "1x2 4"; do
  set -- $cfg
  port=$((port+1))
  SGC_B200_GRID=$1 SGC_B200_PUSH_ROWS=$2 timeout 600 $TR --master-port $port bench.py --gpus 2 --steps 10 --warmup 3 --workload reddit --skip-e2e \
     > gpurun_out/bench_n2_grid$1_rows$2.json 2> gpurun_out/bench_n2_grid$1_rows$2.err
  echo "grid $1 push_rows $2 rc=$?" | tee -a gpurun_out/summary_n2.txt
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/bench_n2_grid$1_rows$2.json").read().strip().splitlines()[-1])
    print("  grid $1 push_rows $2: %.3f ms/step  eff %.3f  parity %s  breakdown %s" % (d["ms_per_step"], d["strong_scaling_efficiency"], d["parity"].get("reddit"), d["details"].get("per_hop_breakdown")))
except Exception as e:
    print("  no line:", e)
P
  grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n2_grid$1_rows$2.err | tail -4
done
for rows in 0 4 2; do
  port=$((port+1))
  SGC_BENCH_LAMBDAS=0 SGC_B200_PUSH_ROWS=$rows timeout 600 $TR --master-port $port bench.py --gpus 2 --steps 5 --warmup 3 --workload products --skip-e2e \
     > gpurun_out/bench_n2_products_rows$rows.json 2> gpurun_out/bench_n2_products_rows$rows.err
  echo "products push_rows $rows rc=$?" | tee -a gpurun_out/summary_n2.txt
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/bench_n2_products_rows$rows.json").read().strip().splitlines()[-1])
    print("  products push_rows $rows: %.3f ms/step  eff %.3f  parity %s  breakdown %s" % (d["ms_per_step"], d["strong_scaling_efficiency"], d["parity"].get("products"), d["details"].get("per_hop_breakdown")))
except Exception as e:
    print("  no line:", e)
P
done
timeout 300 $TR --master-port 29540 tools/e2e_dist_trace.py > gpurun_out/e2e_trace_n2.log 2>&1; echo "trace rc=$?" | tee -a gpurun_out/summary_n2.txt
grep "^rank" gpurun_out/e2e_trace_n2.log
