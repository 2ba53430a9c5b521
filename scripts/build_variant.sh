#!/bin/bash
# build_variant.sh NAME DEPTH_PASS_CU [EXTRA NVCC FLAGS...] -> build/variants/libconga_gpu_NAME.so
set -e
NAME=$1; DP=$2; shift 2
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/build/variants; mkdir -p $OUT/obj_$NAME
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=false -I$ROOT/conga_b200/csrc $*"
cp $DP $OUT/obj_$NAME/depth_pass.cu
for f in conga_gpu genotype flatten; do nvcc $FL -c $ROOT/conga_b200/csrc/$f.cu -o $OUT/obj_$NAME/$f.o & done
nvcc $FL -c $OUT/obj_$NAME/depth_pass.cu -o $OUT/obj_$NAME/depth_pass.o
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libconga_gpu_$NAME.so $OUT/obj_$NAME/*.o -cudart shared
echo built $OUT/libconga_gpu_$NAME.so
