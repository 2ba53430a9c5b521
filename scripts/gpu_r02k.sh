#!/bin/bash
# HEAD after PASS_BATCH 8 and the no-alloc entry point: every GPU test, smoke, the default bench line (C3) and C2,
# then the DRAM traffic of four pass launches (the traffic figure of the roofline object)
set -u
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?"
timeout 120 python bench.py --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"
for W in c3 c2; do python - <<PY
import json
try:
    j = json.load(open("gpurun_out/bench_$W.json")); print("$W", j["value"], j["ms_per_step"], j["run"]["ms_per_step_spread"], j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], "e2e", j["e2e"]["value"], j["e2e"]["ms_per_step"], j["parity_ok"], j.get("cpu_baseline", {}).get("value"))
except Exception as e: print("$W no result", e)
PY
done
CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
timeout 120 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:k_depth_pass --launch-skip 4 --launch-count 4 --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
echo "ncu traffic rc=$?"
