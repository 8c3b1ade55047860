#!/bin/bash
# Last GPU call of the round (bounded: ~5.5 min): parity suite, A/B of the host value scan, bench line.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/unit_mode.txt
timeout 150 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/summary.txt
tail -4 gpurun_out/pytest_gpu.log
timeout 60 python tools/e2e_unit_ab.py > gpurun_out/unit_ab.log 2>&1; echo "ab rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/unit_ab.log | tail -45
MODE=$(cat gpurun_out/unit_mode.txt 2>/dev/null || echo 0)
echo "bench mode SGC_B200_UNIFORM_VALUES=$MODE" | tee -a gpurun_out/summary.txt
SGC_B200_UNIFORM_VALUES=$MODE timeout 160 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench.json
tail -3 gpurun_out/bench.err
