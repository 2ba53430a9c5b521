show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); print(sys.argv[2], "step", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"], "frac", j["roofline"]["frac"], j["roofline"].get("kernel_ms_per_step"), j["parity_ok"])
except Exception as e: print(sys.argv[2], "no result", e)
PY
}
mkdir -p gpurun_out
for V in base b3 b6 ov; do
  LIB=$PWD/conga_b200/libconga_gpu.so; OV=0
  [ $V = b3 ] && LIB=$PWD/conga_b200/_variants/libconga_gpu_b3.so
  [ $V = b6 ] && LIB=$PWD/conga_b200/_variants/libconga_gpu_b6.so
  [ $V = ov ] && OV=1
  for W in "--emulate-rank 0/8" "--emulate-rank 0/4" ""; do
    CG_LIB=$LIB CG_PASS_OVERLAP=$OV timeout 300 python bench.py --workload c3 $W --no-cpu-baseline --repeats 5 > gpurun_out/t.json 2>gpurun_out/sweep.err || tail -n 3 gpurun_out/sweep.err
    show gpurun_out/t.json "$V [$W]"
  done
done
