#!/bin/bash
# split tuning: gpu_job_mg4.sh N TAG "splits" "splits" ...
N=$1; TAG=$2; shift; shift
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for sp in "$@"; do
  i=$((i+1))
  THRS_XCHG_SPLITS=$sp timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_s$i.json 2> gpurun_out/${TAG}_s$i.err
  python - gpurun_out/${TAG}_s$i.json "$sp" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "->", "%.2f Gkeys/s"%(d["value"]/1e9), "ms %.3f"%d["ms_per_step"], d["verified"], {k[:10]:round(v,3) for k,v in d["config"]["stage_ms_rank0"].items()})
except Exception as e:
    print(sys.argv[2], "no line:", e)
PY
done
