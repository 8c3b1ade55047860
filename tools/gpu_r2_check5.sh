#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "full_size" > gpurun_out/pytest_full.log 2>&1; echo "pytest full_size rc=$?" | tee gpurun_out/summary.txt
tail -5 gpurun_out/pytest_full.log
timeout 300 python tools/sweep_persistent.py > gpurun_out/sweep_persistent.log 2>&1; echo "sweep rc=$?" | tee -a gpurun_out/summary.txt
grep -v "Warn\|warn" gpurun_out/sweep_persistent.log | tail -14
