#!/bin/bash
# sweep of pass CTAs per SM x pass overlap on a workload; prints one line per setting
set -u
mkdir -p gpurun_out
W=${W:-c3}
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
for ov in 0 1; do for n in 2 3; do
  CG_PASS_OVERLAP=$ov CG_PASS_CTAS=$n timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 ${EXTRA:-} > gpurun_out/sweep_${W}_ov${ov}_ctas$n.json 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err
  python - <<PY
import json
try:
    j = json.load(open("gpurun_out/sweep_${W}_ov${ov}_ctas$n.json")); print("$W overlap=$ov ctas=$n step", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], j["run"]["ms_per_step_spread"], j["parity_ok"])
except Exception as e: print("no result", e)
PY
done; done
