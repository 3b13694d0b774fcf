#!/bin/bash
# build_dev.sh : builds the developer tools (scripts/devbench, scripts/microbench/mb_rank) for sm_100a
set -e
cd "$(dirname "$0")/.."
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a"
$NV -o scripts/microbench/mb_rank scripts/microbench/mb_rank.cu &
$NV -o scripts/devbench scripts/devbench.cu -Ltinyhipradixsort_b200 -lthrs_b200 -ldl -Xlinker -rpath,'$ORIGIN/../tinyhipradixsort_b200' &
$NV -o scripts/microbench/mb_order scripts/microbench/mb_order.cu &
wait
