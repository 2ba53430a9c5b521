#!/bin/bash
# conga_b200/_variants/libconga_gpu_checked.so: the library with -DCG_CHECKED (every index asserted against its array,
# cg_internal.h).  Run the parity suite against it with CG_LIB=$PWD/conga_b200/_variants/libconga_gpu_checked.so.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/conga_b200/csrc; OUT=$ROOT/conga_b200/_variants; mkdir -p $OUT/checked
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=false -DCG_CHECKED"
for f in conga_gpu depth_pass genotype flatten split_read gc_profile no_alloc; do nvcc $FL -c $SRC/$f.cu -o $OUT/checked/$f.o & done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libconga_gpu_checked.so $OUT/checked/*.o -cudart shared
echo built $OUT/libconga_gpu_checked.so
