#!/bin/bash
# Round-2 GPU check (1 GPU, no profiler): smoke, the whole GPU parity suite, the bench line (own arm and
# reference arm), the L2 bandwidth curve.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.csv 2>&1
nproc > gpurun_out/host.txt; free -g | head -2 >> gpurun_out/host.txt; lscpu | grep -E "Model name|NUMA|Socket" >> gpurun_out/host.txt
nvidia-smi topo -m >> gpurun_out/host.txt 2>&1
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/summary.txt
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/summary.txt
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
head -c 6000 gpurun_out/bench.json; echo
tail -5 gpurun_out/bench.err
timeout 120 python tools/l2_peak.py > gpurun_out/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/l2_peak.log | tail -40
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" | tee -a gpurun_out/summary.txt
head -c 3000 gpurun_out/bench_ref.json; echo
