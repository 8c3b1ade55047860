#!/bin/bash
# 8-GPU run: exchange probe (products), then the bench line at N=8 (Reddit feature-sharded, products and
# papers100M row-sharded at lambda 0 / 0.8; every leg verifies its output).
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n8.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 tools/nvlink_probe2.py products > gpurun_out/nvlink_probe2_n8.log 2>&1; echo "probe rc=$?" | tee gpurun_out/summary_n8.txt
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/nvlink_probe2_n8.log | tail -30
SECONDS=0
timeout 1100 $TR --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary_n8.txt
head -c 3000 gpurun_out/bench_n8.json; echo
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n8.err | tail -15
