#!/bin/bash
# Round-2 profile refresh (1 GPU): TMA-gather experiment (plain), then ncu: launch list of the short bench
# command and `--set full` captures of the hop kernel on the Reddit and products shapes.
set -u
mkdir -p gpurun_out
timeout 600 python tools/sweep_tma_gather.py > gpurun_out/tma_gather.log 2>&1; echo "tma rc=$?" | tee gpurun_out/summary.txt
grep -v "Warn\|warn" gpurun_out/tma_gather.log | tail -30
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e --skip-incumbent --skip-variants --workload reddit"
$CMD > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_|aug_|degree_|norm_|scan_|plan_|col_stats|standardize|store_total|row_ptr_ends|l2_read" -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?" | tee -a gpurun_out/summary.txt
python tools/profile_plain.py reddit > gpurun_out/profile_reddit.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_hop -s 1 -c 1 -o gpurun_out/hop_reddit -f python tools/profile_plain.py reddit > gpurun_out/ncu_hop_reddit.log 2>&1; echo "ncu reddit rc=$?" | tee -a gpurun_out/summary.txt
ncu -i gpurun_out/hop_reddit.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_select.py > gpurun_out/hop_reddit_sel.csv
cat gpurun_out/hop_reddit_sel.csv
python tools/profile_plain.py products > gpurun_out/profile_products.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_hop -s 1 -c 1 -o gpurun_out/hop_products -f python tools/profile_plain.py products > gpurun_out/ncu_hop_products.log 2>&1; echo "ncu products rc=$?" | tee -a gpurun_out/summary.txt
ncu -i gpurun_out/hop_products.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_select.py > gpurun_out/hop_products_sel.csv
cat gpurun_out/hop_products_sel.csv
tail -3 gpurun_out/ncu_hop_reddit.log
