#!/bin/bash
# protocol-check build (-DTHRS_PROTOCOL_CHECK, build/variants/check) under production contention: devbench compares every
# result with CUB bit for bit and prints the violation counters of the look-back / ticket protocol (kernels.cuh)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
V=$PWD/build/variants/check/libthrs_b200.so
{
for spec in "28 1 uniform" "28 0 uniform" "27 1 unique8" "27 0 equal" "22 1 uniform" "20 0 uniform"; do
  echo "== devbench $spec 2 1"
  LD_PRELOAD=$V timeout 25 scripts/devbench $spec 2 1 2>&1 | grep -E "protocol check|mismatches|error|Error"
done
} > gpurun_out/r2p_protocol_check.txt 2>&1
cat gpurun_out/r2p_protocol_check.txt
