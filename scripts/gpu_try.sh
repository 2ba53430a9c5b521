#!/bin/bash
# one kernel change, checked and measured: the parity tests that cover it, then the step on C3, one rank's share of
# the 8-GPU job, and C2.  usage: gpu_try.sh <tag> [workloads...]
set -u
TAG=${1:-try}; shift || true
WLS=${*:-c3 c3@0/8 c2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py tests/test_gpu_margin.py -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
for W in $WLS; do
  NAME=${W%@*}; EM=""; [ "$W" != "$NAME" ] && EM="--emulate-rank ${W#*@}"
  OUT=gpurun_out/${TAG}_$(echo $W | tr '@/' '__').json
  timeout 300 python bench.py --workload $NAME $EM --no-cpu-baseline --repeats 5 > $OUT 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err
  python - $OUT "$W" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print(sys.argv[2], "step", j["ms_per_step"], j["run"]["ms_per_step_spread"], "e2e", j["e2e"]["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"].get("kernel_ms_per_step"), "launches", j["gpu_launches"], j["parity_ok"])
except Exception as e: print(sys.argv[2], "no result", e)
PY
done
