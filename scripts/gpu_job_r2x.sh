#!/bin/bash
# round-2 single-GPU record: parity suite, bench line (ours + reference arm), sanitizer, ncu launch list + counters + full capture
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2x_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2x_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?" >> gpurun_out/r2x_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2x_bench_ref.json 2> gpurun_out/r2x_bench_ref.err
timeout 900 bash scripts/sanitize.sh > gpurun_out/r2x_sanitize.log 2>&1
# after the plain runs exited 0: the launch list of the same bench command, then counters and one full capture of a C2 pass
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2x_launches.csv python bench.py --steps 2 --warmup 3 --no-extra --no-reference-gpu --cpu-log2n 20 --e2e-steps 1 > gpurun_out/r2x_ncu_launch.log 2>&1
M=$(cat scripts/ncu_metrics.txt)
timeout 300 ncu --metrics $M --clock-control none -k regex:onesweep -s 9 -c 1 --csv --log-file gpurun_out/r2x_ncu_counters_pairs.csv scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2x_ncu_counters.log 2>&1
timeout 300 ncu --metrics $M --clock-control none -k regex:onesweep -s 9 -c 1 --csv --log-file gpurun_out/r2x_ncu_counters_keys.csv scripts/devbench 28 0 uniform 1 0 >> gpurun_out/r2x_ncu_counters.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:onesweep -s 9 -c 1 -o gpurun_out/r2x_full -f scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2x_ncu_full.log 2>&1
for d in uniform; do for p in 1 0; do scripts/devbench 28 $p $d 5 1; done; done > gpurun_out/r2x_devbench_c2.txt 2>&1
for d in equal range16 unique8 window; do for p in 1 0; do scripts/devbench 27 $p $d 5 1; done; done > gpurun_out/r2x_devbench_c4.txt 2>&1
{ for n in 13 16 18 20 22; do for pairs in 0 1; do for pdl in 0 1; do echo "pdl=$pdl"; THRS_PDL=$pdl THRS_NOPROFILE=1 timeout 60 scripts/devbench $n $pairs uniform 20 1; done; done; done; } > gpurun_out/r2x_devbench_pdl.txt 2>&1
tail -3 gpurun_out/r2x_pytest.log; tail -2 gpurun_out/r2x_bench.err; head -c 600 gpurun_out/r2x_bench.json; echo; ls -la gpurun_out/sanitize_* | head; cat gpurun_out/sanitize_racecheck_paths.txt | head -20
