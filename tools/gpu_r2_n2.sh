#!/bin/bash
# 2-GPU check: the multi-device parity tests, then the bench line at N=2 (its legs verify their outputs).
set -u
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus.txt
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q > gpurun_out/pytest_n2.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/summary_n2.txt
tail -6 gpurun_out/pytest_n2.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?" | tee -a gpurun_out/summary_n2.txt
head -c 14000 gpurun_out/bench_n2.json; echo
grep -v "Warning\|warn" gpurun_out/bench_n2.err | tail -15
