#!/bin/bash
# samples per submission (resident leg) on one rank's share of an N-GPU job
set -u
mkdir -p gpurun_out
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print(sys.argv[2], "step", j["ms_per_step"], j["run"]["ms_per_step_spread"], "ksub", j["run"]["samples_per_submission"], "frac", j["roofline"]["frac"], j["roofline"].get("kernel_ms_per_step"), "launches/step", j["gpu_launches_per_step"], j["parity_ok"])
except Exception as e: print(sys.argv[2], "no result", e)
PY
}
for W in 0/8 0/4 0/2; do for K in 1 0; do
  timeout 300 python bench.py --workload c3 --emulate-rank $W --samples-per-submission $K --no-cpu-baseline --repeats 5 > gpurun_out/ksub.json 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err
  show gpurun_out/ksub.json "c3 $W K=$K"
done; done
timeout 300 python bench.py --workload c2 --emulate-rank 0/8 --no-cpu-baseline --repeats 5 > gpurun_out/ksub.json 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err
show gpurun_out/ksub.json "c2 0/8 auto"
