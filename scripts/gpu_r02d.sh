#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[2])); print(sys.argv[1], "step", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], j["run"]["ms_per_step_spread"], j["parity_ok"])
except Exception as e: print(sys.argv[1], "no result", e)
PY
}
for W in c3 c2; do
  timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/d_${W}.json 2>gpurun_out/sweep.err || tail -n 5 gpurun_out/sweep.err; show "$W default" gpurun_out/d_${W}.json
  for G in 1 4; do CG_GENO_G=$G timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/d_${W}_g$G.json 2>/dev/null; show "$W G=$G" gpurun_out/d_${W}_g$G.json; done
  CG_PASS_CTAS=2 timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/d_${W}_ctas2.json 2>/dev/null; show "$W ctas=2" gpurun_out/d_${W}_ctas2.json
  CG_PASS_CTAS=4 timeout 300 python bench.py --workload $W --no-cpu-baseline --repeats 5 > gpurun_out/d_${W}_ctas4.json 2>/dev/null; show "$W ctas=4" gpurun_out/d_${W}_ctas4.json
done
CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
for K in k_genotype k_mark; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 -o gpurun_out/prof_$K -f $CMD > gpurun_out/ncu_full_$K.log 2>&1
  echo "ncu $K rc=$?"
done
