#!/bin/bash
# DRAM traffic of four launches of the bulk-store pass, then the small whole-program parity cases
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --repeats 2 --no-cpu-baseline"
timeout 50 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:k_depth_pass --launch-skip 4 --launch-count 4 --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
echo "ncu traffic rc=$?"
timeout 45 python -m pytest tests/test_cli_parity.py -m gpu -q -x -k "not full_size" --timeout 40 -p no:cacheprovider > gpurun_out/pytest_cli.log 2>&1; echo "cli rc=$?"; tail -n 2 gpurun_out/pytest_cli.log
