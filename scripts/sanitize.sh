#!/bin/bash
# sanitize.sh — compute-sanitizer passes of SURVEY.md 5 ("race detection") on the GPU box: memcheck, racecheck
# and synccheck over harness/unittest (2 iterations per case) and over scripts/dev/paths (one sort per kernel
# path: one-block, cluster 2/4/8, mid, full, both rankings, scatter).  Summaries land in gpurun_out/sanitize_*.txt.
#   gpurun --timeout 1500 -- 'bash scripts/sanitize.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck synccheck; do
  for prog in "scripts/dev/paths 300000" "harness/unittest 2"; do
    name=$(echo $prog | awk '{print $1}' | xargs basename)
    out=gpurun_out/sanitize_${tool}_${name}.txt
    echo "== $tool: $prog" | tee $out
    timeout 200 $CS --tool $tool --print-limit 20 $prog > $out.full 2>&1
    echo "exit code $?" >> $out
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Hazard|Error|error|OK  |FAILED|failed" $out.full | sort | uniq -c | sort -rn | head -40 >> $out
    tail -3 $out.full >> $out
    rm -f $out.full.tmp
  done
done
ls -la gpurun_out/sanitize_*
