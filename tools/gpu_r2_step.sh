#!/bin/bash
# Short 1-GPU step: the tests of the kernels touched last, then the row-part sweep.
set -u
mkdir -p gpurun_out
SECONDS=0
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dist.py -m gpu -q -x -k "not full_size" > gpurun_out/pytest_step.log 2>&1; echo "pytest rc=$? in ${SECONDS}s" | tee gpurun_out/summary_step.txt
tail -5 gpurun_out/pytest_step.log
timeout 400 python tools/sweep_rowpart.py > gpurun_out/sweep_rowpart.log 2>&1; echo "rowpart rc=$?" | tee -a gpurun_out/summary_step.txt
grep -v "Warn\|warn" gpurun_out/sweep_rowpart.log | tail -14
