#!/bin/bash
# DRAM bytes of every k_depth_pass launch of ONE sample (22 launches, after 3 warm-up samples):
# the denominator check for roofline.traffic.  Output: gpurun_out/traffic.csv
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_traffic.log 2>&1 &&
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:k_depth_pass --launch-skip 66 --launch-count 22 --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
echo "ncu traffic rc=$?"
