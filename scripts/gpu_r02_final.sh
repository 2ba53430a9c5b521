#!/bin/bash
# One gpurun call at HEAD: every GPU test, smoke, both bench arms on the default workload, the other workloads,
# then (each only after the plain run exited 0) the ncu launch list, a DRAM-traffic capture and one --set full capture
# of k_depth_pass.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 400 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_c3_ref.json 2> gpurun_out/bench_c3_ref.err; echo "ref rc=$?"
for W in c2 c4; do timeout 600 python bench.py --workload $W > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err; echo "$W rc=$?"; done
timeout 900 python bench.py --workload c5 --steps 64 --warmup 5 --repeats 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
for W in c3 c2 c4 c5; do python - <<PY
import json
try:
    j = json.load(open("gpurun_out/bench_$W.json")); print("$W", j["value"], j["ms_per_step"], j["roofline"]["frac"], j["roofline"]["kernel_ms_per_step"], "e2e", j["e2e"]["value"], j["e2e"]["ms_per_step"], j["e2e"]["bytes_per_read"], j["parity_ok"])
except Exception as e: print("$W no result", e)
PY
done
python -c "
import json; j=json.load(open('gpurun_out/bench_c3_ref.json')); print('ref', j['value'], j['ms_per_step'], j['run'])"
if [ "${NCU:-1}" = "1" ]; then
  CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
      -k regex:k_depth_pass --launch-skip 4 --launch-count 4 --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
  echo "ncu traffic rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_depth_pass -s 4 -c 1 -o gpurun_out/prof_depth_pass -f $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
