mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r03_smoke.log 2>&1; echo "smoke rc=$?"
python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_gpu_final.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r03_pytest_gpu_final.log
python bench.py > gpurun_out/r03_bench_final.json 2> gpurun_out/r03_bench_final.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r03_bench_final.json").read().strip().splitlines()[-1])
r = d["roofline"]
print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["parity"]["ok"], d["gpu_launches"])
print({k: r[k] for k in ("bound", "unit", "achieved", "peak", "frac", "traffic")}, r["hbm_model"]["frac"], r["hw_counters"]["capture_matches_timed_build"], r["l2"]["frac"])
print(d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"])
PY
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r03_bench_ref.json 2> gpurun_out/r03_bench_ref.err; echo "ref rc=$?"; head -c 400 gpurun_out/r03_bench_ref.json
