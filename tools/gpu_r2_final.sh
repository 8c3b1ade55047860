#!/bin/bash
# Final 1-GPU pass of round 2: smoke, the whole GPU suite, both bench arms, the host-pipeline sweep, then the ncu
# evidence of the same build (launch list of the short bench command, --set full of one hop launch per shape).
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/summary_final.txt
SECONDS=0
timeout 1500 python -m pytest tests -m gpu -q -x --durations=12 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? in ${SECONDS}s" | tee -a gpurun_out/summary_final.txt
tail -22 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary_final.txt
head -c 1200 gpurun_out/bench.json; echo; tail -3 gpurun_out/bench.err
[ "${SKIP_EXTRAS:-0}" = 1 ] || timeout 300 python tools/e2e_sweep.py > gpurun_out/e2e_sweep.log 2>&1; echo "e2e sweep rc=$?" | tee -a gpurun_out/summary_final.txt
grep -v "Warn\|warn" gpurun_out/e2e_sweep.log | tail -36
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e --skip-incumbent --skip-variants --workload reddit"
$CMD > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_|aug_|degree_|norm_|scan_|plan_|col_stats|standardize|store_total|row_ptr_ends|l2_read" -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?" | tee -a gpurun_out/summary_final.txt
for shape in reddit products; do
  python tools/profile_plain.py $shape > gpurun_out/profile_$shape.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_hop -s 1 -c 1 -o gpurun_out/hop_$shape -f python tools/profile_plain.py $shape > gpurun_out/ncu_hop_$shape.log 2>&1; echo "ncu $shape rc=$?" | tee -a gpurun_out/summary_final.txt
  ncu -i gpurun_out/hop_$shape.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_select.py > gpurun_out/hop_${shape}_sel.csv
  cat gpurun_out/hop_${shape}_sel.csv
done
[ "${SKIP_EXTRAS:-0}" = 1 ] || timeout 400 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" | tee -a gpurun_out/summary_final.txt
head -c 600 gpurun_out/bench_ref.json; echo
cat gpurun_out/summary_final.txt
