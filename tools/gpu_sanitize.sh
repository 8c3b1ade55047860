#!/bin/bash
# compute-sanitizer on the small cases of tools/sanitize_case.py: ONE tool per gpurun call (B200_PROFILING.md).
#   bash tools/gpu_sanitize.sh memcheck|racecheck
set -u
TOOL="${1:-memcheck}"
mkdir -p gpurun_out
timeout 300 python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1; rc=$?; echo "plain rc=$rc" | tee gpurun_out/summary_sanitize.txt
tail -8 gpurun_out/sanitize_plain.log
if [ $rc -ne 0 ]; then echo "not sanitizing: the plain run failed"; exit 1; fi
timeout 1500 compute-sanitizer --tool "$TOOL" --error-exitcode 7 --print-limit 50 python tools/sanitize_case.py > gpurun_out/sanitize_$TOOL.log 2>&1; echo "$TOOL rc=$?" | tee -a gpurun_out/summary_sanitize.txt
grep -v "Warning\|warn" gpurun_out/sanitize_$TOOL.log | tail -25
