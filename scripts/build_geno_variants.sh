#!/bin/bash
# libconga_gpu variants that differ in CNVs per warp of the genotype kernel -> conga_b200/_variants/libconga_gpu_gN.so
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/conga_b200/csrc; OUT=$ROOT/conga_b200/_variants; mkdir -p $OUT
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=false"
make -C $SRC -j4 >/dev/null
for G in ${LIST:-1 4 16 32}; do
  nvcc $FL -DGENO_G=$G -c $SRC/genotype.cu -o $OUT/genotype_g$G.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libconga_gpu_g$G.so $SRC/conga_gpu.o $SRC/depth_pass.o $SRC/flatten.o $SRC/split_read.o $SRC/gc_profile.o $OUT/genotype_g$G.o -cudart shared
  echo built $OUT/libconga_gpu_g$G.so
done
