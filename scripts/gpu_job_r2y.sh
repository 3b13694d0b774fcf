#!/bin/bash
# round-2 single-GPU record (bounded to ~20 min of box time): parity suite incl. the no-alloc entry point, bench line
# (ours + reference arm), ncu launch list + counters + one full capture, devbench C2/C4/PDL sweep, compute-sanitizer.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2y_smi.txt
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2y_pytest.log
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2y_pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2y_pytest_noalloc_alone.log
timeout 500 python bench.py --steps 10 --warmup 3 > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err; echo "bench rc=$?" >> gpurun_out/r2y_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2y_bench_ref.json 2> gpurun_out/r2y_bench_ref.err
# after the plain runs exited 0: the launch list of the same bench command, then counters and one full capture of a C2 pass
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2y_launches.csv python bench.py --steps 2 --warmup 3 --no-extra --no-reference-gpu --cpu-log2n 20 --e2e-steps 1 > gpurun_out/r2y_ncu_launch.log 2>&1
M=$(cat scripts/ncu_metrics.txt)
timeout 200 ncu --metrics $M --clock-control none -k regex:onesweep -s 9 -c 1 --csv --log-file gpurun_out/r2y_ncu_counters_pairs.csv scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2y_ncu_counters.log 2>&1
timeout 200 ncu --metrics $M --clock-control none -k regex:onesweep -s 9 -c 1 --csv --log-file gpurun_out/r2y_ncu_counters_keys.csv scripts/devbench 28 0 uniform 1 0 >> gpurun_out/r2y_ncu_counters.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:onesweep -s 9 -c 1 -o gpurun_out/r2y_full -f scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2y_ncu_full.log 2>&1
for d in uniform; do for p in 1 0; do timeout 60 scripts/devbench 28 $p $d 5 1; done; done > gpurun_out/r2y_devbench_c2.txt 2>&1
for d in equal range16 unique8 window; do for p in 1 0; do timeout 60 scripts/devbench 27 $p $d 5 1; done; done > gpurun_out/r2y_devbench_c4.txt 2>&1
{ for n in 13 16 18 20 22; do for pairs in 0 1; do for pdl in 0 1; do echo "pdl=$pdl"; THRS_PDL=$pdl THRS_NOPROFILE=1 timeout 60 scripts/devbench $n $pairs uniform 20 1; done; done; done; } > gpurun_out/r2y_devbench_pdl.txt 2>&1
# compute-sanitizer: every kernel path once (scripts/dev/paths), then the C++ harness under memcheck
CS=/usr/local/cuda/bin/compute-sanitizer
for job in "racecheck scripts/dev/paths 300000" "memcheck scripts/dev/paths 300000" "synccheck scripts/dev/paths 300000" "memcheck harness/unittest 1"; do
  set -- $job; tool=$1; shift; name=$(basename $1)
  out=gpurun_out/sanitize_${tool}_${name}.txt
  echo "== compute-sanitizer --tool $tool $*" > $out
  timeout 130 $CS --tool $tool --print-limit 20 "$@" > $out.full 2>&1
  echo "exit code $?" >> $out
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Hazard|Error|error|OK  |FAILED|failed" $out.full | sort | uniq -c | sort -rn | head -40 >> $out
  tail -3 $out.full >> $out
  head -c 200000 $out.full > $out.full.head; rm -f $out.full
done
tail -3 gpurun_out/r2y_pytest.log; tail -3 gpurun_out/r2y_pytest_noalloc_alone.log; tail -2 gpurun_out/r2y_bench.err; head -c 700 gpurun_out/r2y_bench.json; echo; head -30 gpurun_out/sanitize_racecheck_paths.txt
