#!/bin/bash
# one ncu --set full capture of k_genotype on a C3 sample (source counters on), for profiles/
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_genotype --launch-skip 6 -c 1 -f -o gpurun_out/prof_k_genotype \
  python bench.py --workload c3 --steps 2 --warmup 1 --repeats 1 --no-cpu-baseline > gpurun_out/prof_geno.log 2>&1
echo "ncu rc=$?"; tail -n 3 gpurun_out/prof_geno.log; ls -la gpurun_out/prof_k_genotype.ncu-rep
