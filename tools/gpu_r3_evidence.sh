#!/bin/bash
# Round-3 1-GPU evidence pass: the papers100M weak-scaling yardstick (1/8-size generator on one GPU), the launch
# list of the short bench command, and `ncu --set full` of one hop launch per shape with the pipe-utilisation
# counters selected (tools/ncu_select.py).  Each ncu pass runs only after its command exited 0 without ncu.
set -u
mkdir -p gpurun_out
sha256sum sgc_b200/libsgc_b200.so | tee gpurun_out/r03_build_sha.txt
timeout 600 python - > gpurun_out/r03_weak_ref_papers.log 2>&1 <<'PY'
import json, torch, bench, bench_dist
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
for lam in (0.0, 0.8):
    r = bench_dist.weak_scaling_reference("papers100m", 8, lam, dev, vars(bench))
    print(json.dumps(r), flush=True)
print("peak memory GB", torch.cuda.max_memory_allocated() / 1e9)
PY
echo "weak ref rc=$?"; tail -4 gpurun_out/r03_weak_ref_papers.log
timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -m gpu -k "fp64 or two_streams or row_slice" > gpurun_out/r03_pytest_new.log 2>&1; echo "new tests rc=$?"; tail -3 gpurun_out/r03_pytest_new.log
timeout 900 python bench.py > gpurun_out/r03_bench_gate.json 2> gpurun_out/r03_bench_gate.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r03_bench_gate.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r03_bench_gate.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["ms_per_step"], json.dumps(d["parity"]["timed_output_with_gpu_standardisation"]["vs_fp64"]))
PY
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e --skip-incumbent --skip-variants --workload reddit"
$CMD > gpurun_out/r03_bench_short.json 2> gpurun_out/r03_bench_short.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_|aug_|degree_|norm_|scan_|plan_|col_stats|standardize|store_total|row_ptr_ends|l2_read" -c 200 --csv --log-file gpurun_out/r03_launches.csv $CMD > gpurun_out/r03_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
for shape in reddit products; do
  python tools/profile_plain.py $shape > gpurun_out/r03_profile_$shape.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_hop -s 1 -c 1 -o gpurun_out/r03_hop_$shape -f python tools/profile_plain.py $shape > gpurun_out/r03_ncu_hop_$shape.log 2>&1; echo "ncu $shape rc=$?"
  ncu -i gpurun_out/r03_hop_$shape.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_select.py > gpurun_out/r03_ncu_full_spmm_hop_${shape}.csv
  wc -c gpurun_out/r03_ncu_full_spmm_hop_${shape}.csv
done
