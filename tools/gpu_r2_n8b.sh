#!/bin/bash
# N-GPU run (N = $1, default 8): Reddit-shaped leg on the candidate grids (column blocks x row parts), device-timed
# only, then the default line of that workload with the host-to-host figure.
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n$N.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29550
: > gpurun_out/summary_n$N.txt
for grid in ${GRIDS:-"2x4 1x8 4x2"}; do
  port=$((port+1))
  SGC_B200_GRID=$grid timeout 400 $TR --master-port $port bench.py --gpus $N --steps 10 --warmup 3 --workload reddit --skip-e2e \
     > gpurun_out/bench_n${N}_grid$grid.json 2> gpurun_out/bench_n${N}_grid$grid.err
  echo "grid $grid rc=$?" | tee -a gpurun_out/summary_n$N.txt
  python - <<P
import json
try:
    d = json.loads(open("gpurun_out/bench_n${N}_grid$grid.json").read().strip().splitlines()[-1])
    print("  grid $grid: %.3f ms/step  eff %.3f  parity %s  breakdown %s" % (d["ms_per_step"], d["strong_scaling_efficiency"], d["parity"].get("reddit"), d["details"].get("per_hop_breakdown")))
except Exception as e:
    print("  no line:", e)
P
  grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n${N}_grid$grid.err | tail -4
done
port=$((port+1))
SGC_B200_GRID_AUTO=1 timeout 600 $TR --master-port $port bench.py --gpus $N --steps 10 --warmup 3 ${FINAL_ARGS:-"--workload reddit"} > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "default line rc=$?" | tee -a gpurun_out/summary_n$N.txt
head -c 2500 gpurun_out/bench_n$N.json; echo
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n$N.err | tail -6
