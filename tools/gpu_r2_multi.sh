#!/bin/bash
# One 8-GPU box, several world sizes: the Reddit-shaped leg on candidate grids (column blocks x row parts) at
# N = 8, 4 and 2 (torchrun on the first N GPUs), device-timed and verified, then the default N = 8 line with the
# host-to-host figure.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_multi.txt
: > gpurun_out/summary_multi.txt
port=29560
run() {   # N grid push_rows
  local N=$1 grid=$2 rows=$3
  port=$((port+1))
  local tag=n${N}_grid${grid}_rows${rows}
  SGC_B200_GRID=$grid SGC_B200_PUSH_ROWS=$rows timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
     --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 10 --warmup 3 --workload reddit --skip-e2e \
     > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
  echo "N=$N grid $grid push_rows $rows rc=$?" | tee -a gpurun_out/summary_multi.txt
  python - <<P | tee -a gpurun_out/summary_multi.txt
import json
try:
    d = json.loads(open("gpurun_out/bench_$tag.json").read().strip().splitlines()[-1])
    print("  N=$N grid $grid push_rows $rows: %.3f ms/step  eff %.3f  parity %s  breakdown %s" % (d["ms_per_step"], d["strong_scaling_efficiency"], d["parity"].get("reddit"), d["details"].get("per_hop_breakdown")))
except Exception as e:
    print("  no line:", e)
P
  grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_$tag.err | tail -3
}
CONFIGS=${CONFIGS:-"8:2x4:0 8:1x8:0 8:4x2:0 4:1x4:0 4:2x2:0 2:1x2:0"}
for cfg in $CONFIGS; do
  N=${cfg%%:*}; rest=${cfg#*:}; grid=${rest%%:*}; rows=${rest#*:}
  run $N $grid $rows
done
port=$((port+1))
SGC_B200_GRID_AUTO=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
   bench.py --gpus 8 --steps 10 --warmup 3 --workload reddit > gpurun_out/bench_n8_auto.json 2> gpurun_out/bench_n8_auto.err
echo "N=8 auto grid with e2e rc=$?" | tee -a gpurun_out/summary_multi.txt
head -c 2000 gpurun_out/bench_n8_auto.json; echo
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n8_auto.err | tail -5
