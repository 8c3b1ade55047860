#!/bin/bash
# 8-GPU run: the default bench line at N=8 as the driver launches it (Reddit feature-sharded; products and
# papers100M row-sharded at lambda 0 / 0.8; every leg verifies its output).
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n8.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
SECONDS=0
timeout 800 $TR --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$? in ${SECONDS}s" | tee gpurun_out/summary_n8.txt
head -c 3000 gpurun_out/bench_n8.json; echo
grep -v "Warning\|warn\|\*\*\*\|OMP_NUM" gpurun_out/bench_n8.err | tail -15
