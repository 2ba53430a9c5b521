#!/bin/bash
# record groups of a chunk asked for one chunk ahead (PASS_PRE): variants built into conga_b200/_variants
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print(sys.argv[2], "step", j["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"].get("kernel_ms_per_step"), j["parity_ok"])
except Exception as e: print(sys.argv[2], "no result", e)
PY
}
mkdir -p gpurun_out
for W in c4 c3 c2; do
for V in p3 p4; do
  CG_LIB=$PWD/conga_b200/_variants/libconga_gpu_$V.so timeout 200 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/t.json 2>gpurun_out/sweep.err || tail -n 3 gpurun_out/sweep.err
  show gpurun_out/t.json "$V [$W]"
done; done
