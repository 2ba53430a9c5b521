#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
for W in c3 c2; do for G in 8 1 4 16 32; do
  LIB=conga_b200/_variants/libconga_gpu_g$G.so; [ $G = 8 ] && LIB=conga_b200/libconga_gpu.so
  CG_LIB=$PWD/$LIB CG_PASS_CTAS=3 timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/geno_${W}_g$G.json 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err
  python - <<PY
import json
try:
    j = json.load(open("gpurun_out/geno_${W}_g$G.json")); print("$W G=$G step", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], j["parity_ok"])
except Exception as e: print("no result", e)
PY
done; done
CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
CG_PASS_CTAS=3 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
