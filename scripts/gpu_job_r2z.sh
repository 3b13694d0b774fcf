#!/bin/bash
# ncu captures of one C2 pass (pairs and keys-only).  thrs_create's production ranking check launches 32 onesweep
# kernels before the first user sort, so the launch to capture is found by skipping those (and two warm-up sorts),
# and three launches are captured: the summaries keep the ones with the full persistent grid (296 blocks).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
M=$(cat scripts/ncu_metrics.txt)
timeout 200 ncu --metrics $M --clock-control none -k regex:onesweep -s 40 -c 3 --csv --log-file gpurun_out/r2z_ncu_counters_pairs.csv scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2z_ncu_counters.log 2>&1
timeout 200 ncu --metrics $M --clock-control none -k regex:onesweep -s 40 -c 3 --csv --log-file gpurun_out/r2z_ncu_counters_keys.csv scripts/devbench 28 0 uniform 1 0 >> gpurun_out/r2z_ncu_counters.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:onesweep -s 41 -c 1 -o gpurun_out/r2z_full -f scripts/devbench 28 1 uniform 1 0 > gpurun_out/r2z_ncu_full.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:onesweep -s 41 -c 1 -o gpurun_out/r2z_full_keys -f scripts/devbench 28 0 uniform 1 0 > gpurun_out/r2z_ncu_full_keys.log 2>&1
grep -c "296" gpurun_out/r2z_ncu_counters_pairs.csv; tail -3 gpurun_out/r2z_ncu_full.log
