#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench line, then the ncu launch list and one
# --set full capture of the fused depth pass.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
if [ "${NCU:-1}" = "1" ]; then
  CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
      --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_depth_pass -c 2 \
      -o gpurun_out/prof_depth_pass -f $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_genotype -c 2 \
      -o gpurun_out/prof_genotype -f $CMD > gpurun_out/ncu_full_geno.log 2>&1
  echo "ncu geno rc=$?"
fi
tail -n 5 gpurun_out/pytest_gpu.log; tail -n 5 gpurun_out/smoke.log
