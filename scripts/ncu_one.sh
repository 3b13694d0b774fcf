#!/bin/bash
# ncu_one.sh TAG [LD_PRELOAD lib] : metric pass over the 3rd onesweep launch of devbench (dev tool)
tag=$1; lib=$2
M=$(cat scripts/ncu_metrics.txt)
LD_PRELOAD=$lib ncu --metrics $M --clock-control none -k regex:onesweep -s 9 -c 1 --csv --log-file gpurun_out/ncu_$tag.csv ./scripts/devbench 28 1 uniform 1 0 > gpurun_out/ncu_$tag.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/ncu_$tag.csv")) if len(r)>10]
h=rows[0]; 
for r in rows[1:]:
    d=dict(zip(h,r)); print("$tag", d["Metric Name"], d["Metric Value"], d["Metric Unit"])
PY
