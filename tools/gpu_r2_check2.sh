#!/bin/bash
# Round-2 GPU check, second pass: bench line, L2 curve with the fixed hashed order, Reddit F1 diagnostics.
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee gpurun_out/summary.txt
head -c 9000 gpurun_out/bench.json; echo
tail -5 gpurun_out/bench.err
timeout 120 python tools/l2_peak.py > gpurun_out/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a gpurun_out/summary.txt
tail -48 gpurun_out/l2_peak.log
timeout 900 python tools/reddit_f1_diag.py 2 10 > gpurun_out/reddit_f1_diag.log 2>&1; echo "diag rc=$?" | tee -a gpurun_out/summary.txt
grep -v Warning gpurun_out/reddit_f1_diag.log | tail -30
