#!/bin/bash
# k_depth_pass with one bulk copy shared -> global per chunk (cp.async.bulk, -DPASS_BULK_STORE) instead of eight
# 16-byte streaming stores per lane: parity of the variant, then pass and step per workload
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print(sys.argv[2], "step", j["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"].get("kernel_ms_per_step"), j["parity_ok"])
except Exception as e: print(sys.argv[2], "no result", e)
PY
}
mkdir -p gpurun_out
V=$PWD/conga_b200/_variants
CG_LIB=$V/libconga_gpu_bulk1.so timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -q -x --timeout 120 -p no:cacheprovider > gpurun_out/pytest_bulk.log 2>&1; echo "pytest(bulk1) rc=$?"; tail -n 2 gpurun_out/pytest_bulk.log
for RUN in "bulk1 c3" "bulk1 c2" "bulk1 c4" "bulk1h c3" "bulk1h c2" "bulk2 c3" "bulk2 c2"; do
  set -- $RUN
  CG_LIB=$V/libconga_gpu_$1.so timeout 100 python bench.py --workload $2 --no-cpu-baseline --repeats 5 > gpurun_out/t.json 2>gpurun_out/sweep.err || tail -n 3 gpurun_out/sweep.err
  show gpurun_out/t.json "$1 [$2]"
done
