#!/bin/bash
# on an 8-GPU box: the default workload at N = 8 and 4, then C2 and C5 at N = 8 (own arm)
mkdir -p gpurun_out
run() { N=$1; TAG=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --no-cpu-baseline --repeats 5 "$@" > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err
  echo "bench $TAG rc=$?"; python - gpurun_out/bench_${TAG}.json <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print("  value", j["value"], "step", j["ms_per_step"], j["run"]["ms_per_step_spread"], "ksub", j["run"]["samples_per_submission"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], "parity", j["parity_ok"], j["parity"])
except Exception as e: print("  no result", e)
PY
}
run 8 c3_n8
run 4 c3_n4
run 8 c2_n8 --workload c2
run 8 c5_n8 --workload c5
run 8 c3_n8_k1 --samples-per-submission 1
