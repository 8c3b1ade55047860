#!/bin/bash
# Round-2 re-entry check (1 GPU, no profiler): smoke, the whole GPU suite (timed), the bench line of both arms,
# the PCIe duplex probe.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.csv 2>&1
nproc > gpurun_out/host.txt; free -g | head -2 >> gpurun_out/host.txt; lscpu | grep -E "Model name|NUMA|Socket|Thread" >> gpurun_out/host.txt
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/summary.txt
SECONDS=0
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary.txt
tail -30 gpurun_out/pytest_gpu.log
SECONDS=0
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary.txt
head -c 1500 gpurun_out/bench.json; echo
tail -5 gpurun_out/bench.err
timeout 300 python tools/pcie_duplex_probe.py > gpurun_out/pcie_duplex.log 2>&1; echo "pcie rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/pcie_duplex.log | tail -40
SECONDS=0
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary.txt
head -c 1500 gpurun_out/bench_ref.json; echo
cat gpurun_out/summary.txt
timeout 400 python tools/sweep_tma_gather.py > gpurun_out/tma_gather.log 2>&1; echo "tma rc=$?" | tee -a gpurun_out/summary.txt
grep -v "Warn\|warn" gpurun_out/tma_gather.log | tail -45
timeout 500 python tools/sweep_col_blocks.py products > gpurun_out/col_blocks.log 2>&1; echo "col blocks rc=$?" | tee -a gpurun_out/summary.txt
grep -v "Warn\|warn" gpurun_out/col_blocks.log | tail -16
