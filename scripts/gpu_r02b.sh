#!/bin/bash
# One gpurun call (round 2): parity tests of the submission path, bench c3 / c2, CTA sweep, ncu launch list and
# --set full captures of the pass, the record sweep and the genotype kernel.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
timeout 500 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?"; tail -n 3 gpurun_out/bench_c3.err; cat gpurun_out/bench_c3.json
timeout 400 python bench.py --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"; cat gpurun_out/bench_c2.json
for n in 1 3; do
  CG_PASS_CTAS=$n timeout 300 python bench.py --no-cpu-baseline --repeats 4 > gpurun_out/bench_c3_ctas$n.json 2>/dev/null; echo "ctas=$n rc=$?"
  python - <<PY
import json
try:
    j = json.load(open("gpurun_out/bench_c3_ctas$n.json")); print("ctas $n", j["ms_per_step"], j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"])
except Exception as e: print("no result", e)
PY
done
if [ "${NCU:-1}" = "1" ]; then
  CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  for K in k_depth_pass k_mark k_genotype; do
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 \
        -o gpurun_out/prof_$K -f $CMD > gpurun_out/ncu_full_$K.log 2>&1
    echo "ncu $K rc=$?"
  done
fi
