#!/bin/bash
# build_variant.sh NAME "-DTHRS_TUNE_THREADS=.. ..." : builds build/variants/NAME/libthrs_b200.so (dev tuning only)
set -e
cd "$(dirname "$0")/../tinyhipradixsort_b200/csrc"
out=../../build/variants/$1; mkdir -p $out/obj
for f in capi capi_mg capi_xchg inst_u32 inst_u64 inst_f32 inst_f64; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden $2 -c $f.cu -o $out/obj/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libthrs_b200.so $out/obj/*.o -Xlinker -soname,libthrs_b200.so
rm -rf $out/obj
