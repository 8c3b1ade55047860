#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q > gpurun_out/pytest_dist.log 2>&1; echo "pytest dist rc=$?" | tee gpurun_out/summary.txt
tail -4 gpurun_out/pytest_dist.log
bash tools/gpu_sanitize.sh memcheck
