#!/bin/bash
# build_variant_u32.sh NAME "-D..." : build/variants/NAME/libthrs_b200.so with only inst_u32.cu recompiled under
# the given defines; every other object comes from the product build (tinyhipradixsort_b200/csrc/build).  Dev tuning only:
# LD_PRELOAD=build/variants/NAME/libthrs_b200.so scripts/devbench ...
set -e
cd "$(dirname "$0")/../tinyhipradixsort_b200/csrc"
out=../../build/variants/$1; mkdir -p $out
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v $2 -c inst_u32.cu -o $out/inst_u32.o 2> $out/ptxas.log
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libthrs_b200.so build/capi.o build/capi_mg.o build/capi_xchg.o $out/inst_u32.o build/inst_u64.o build/inst_f32.o build/inst_f64.o -Xlinker -soname,libthrs_b200.so
rm -f $out/inst_u32.o
