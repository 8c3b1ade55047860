#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/summary.txt
tail -12 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
head -c 12000 gpurun_out/bench.json; echo
tail -5 gpurun_out/bench.err
timeout 120 python tools/l2_peak.py > gpurun_out/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a gpurun_out/summary.txt
tail -8 gpurun_out/l2_peak.log
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,serial,uuid --format=csv
