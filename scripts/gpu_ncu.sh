#!/bin/bash
# ncu launch list + one --set full capture of kernels matching $1 (default k_depth_pass)
K=${1:-k_depth_pass}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -c 2 \
    -o gpurun_out/prof_$K -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
