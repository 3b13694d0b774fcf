#!/bin/bash
# round-2 GPU job A: parity suite, ranking/LSU variant sweep, PDL cost, sanitizer, bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
{
for v in base r1d8 r1d4 r1d16 r1d8k56 dbl_atoms dbl_stage dbl_lds dbl_stg; do
  pre=""; [ $v != base ] && pre=$PWD/build/variants/$v/libthrs_b200.so
  cub=0; case $v in r1*) cub=1;; esac
  for pairs in 1 0; do
    echo "variant=$v pairs=$pairs"
    LD_PRELOAD=$pre timeout 120 scripts/devbench 28 $pairs uniform 5 $cub
  done
done
for v in base r1d8; do
  pre=""; [ $v != base ] && pre=$PWD/build/variants/$v/libthrs_b200.so
  for dist in equal range16 unique8 window; do
    for pairs in 1 0; do echo "variant=$v C4 $dist pairs=$pairs"; LD_PRELOAD=$pre timeout 120 scripts/devbench 27 $pairs $dist 5 1; done
  done
done
} > gpurun_out/r2a_variants.txt 2>&1
{
for n in 13 16 18 20 22; do for pairs in 0 1; do for pdl in 0 1; do echo "pdl=$pdl"; THRS_PDL=$pdl THRS_NOPROFILE=1 timeout 60 scripts/devbench $n $pairs uniform 20 1; done; done; done
} > gpurun_out/r2a_pdl.txt 2>&1
timeout 900 bash scripts/sanitize.sh > gpurun_out/r2a_sanitize.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?" >> gpurun_out/r2a_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err
tail -3 gpurun_out/r2a_pytest.log; grep -E "thrs [0-9.]+ ms" gpurun_out/r2a_variants.txt | head -40
