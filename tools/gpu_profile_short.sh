# Reduced refresh of profiles/: launch list of the bench + full capture of the normalisation kernels.
# ncu's -k filter matches kernel base names (no namespace), hence the explicit list.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e --skip-incumbent --workload reddit"
$CMD > gpurun_out/bench_short.json 2>&1; echo "short rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_|aug_|degree_|norm_|scan_|plan_|col_stats|standardize|store_total|row_ptr_ends" -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"aug_|degree_warp|norm_values" -c 6 -o gpurun_out/norm_full -f python tools/profile_norm.py > gpurun_out/ncu_norm.log 2>&1; echo "ncu norm rc=$?"
ncu -i gpurun_out/norm_full.ncu-rep --page raw --csv > gpurun_out/norm_full_raw.csv 2>/dev/null
wc -l gpurun_out/launches.csv gpurun_out/norm_full_raw.csv
