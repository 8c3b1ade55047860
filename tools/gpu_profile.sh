#!/bin/bash
# Runs on the GPU box (via gpurun): bench line, then ncu launch list of the same command and ncu --set
# full captures of the hop kernel and of the normalisation kernels.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e --skip-incumbent --workload reddit"
$CMD > gpurun_out/bench_short.json 2>&1; echo "short rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/profile_norm.py > gpurun_out/profile_norm.log 2>&1; echo "norm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"aug_|degree_warp|norm_values" -c 12 -o gpurun_out/norm_full -f python tools/profile_norm.py > gpurun_out/ncu_norm.log 2>&1; echo "ncu norm rc=$?"
ncu -i gpurun_out/norm_full.ncu-rep --page raw --csv > gpurun_out/norm_full_raw.csv 2>/dev/null
python tools/profile_hop.py > gpurun_out/profile_hop.log 2>&1; echo "hop rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_hop -s 3 -c 1 -o gpurun_out/hop_full -f python tools/profile_hop.py > gpurun_out/ncu_hop.log 2>&1; echo "ncu hop rc=$?"
ncu -i gpurun_out/hop_full.ncu-rep --page raw --csv > gpurun_out/hop_full_raw.csv 2>/dev/null
ls -la gpurun_out
