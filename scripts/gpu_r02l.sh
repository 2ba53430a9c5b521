#!/bin/bash
# HEAD with the bulk-copy depth stores: the parity tests that cover the pass, smoke, bench lines for C3 (with the oracle
# check of cpu_baseline), C2 and C4
set -u
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py tests/test_gpu_margin.py tests/test_zz_no_alloc.py -m gpu -q -x --timeout 90 -p no:cacheprovider > gpurun_out/pytest_gpu_l.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_gpu_l.log
timeout 60 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 120 python bench.py --steps 20 --warmup 5 --cpu-budget 6 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?"
timeout 60 python bench.py --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"
timeout 60 python bench.py --workload c4 --no-cpu-baseline --repeats 5 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"
for W in c3 c2 c4; do python - <<PY
import json
try:
    j = json.load(open("gpurun_out/bench_$W.json")); print("$W", j["value"], j["ms_per_step"], j["run"]["ms_per_step_spread"], j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], "e2e", j["e2e"]["value"], j["e2e"]["ms_per_step"], j["parity_ok"], j["parity"].get("oracle"))
except Exception as e: print("$W no result", e)
PY
done
