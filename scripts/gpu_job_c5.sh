#!/bin/bash
# C5 literal: 2^32 pairs over N GPUs (strong scaling) + the default weak-scaling line: gpu_job_c5.sh N TAG
N=$1; TAG=$2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for cfg in "c5:--log2n-total 32 --steps 3 --warmup 3" "weak:--steps 5 --warmup 3"; do
  name=${cfg%%:*}; extra=${cfg#*:}
  timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N $extra > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  echo "$name rc=$?" >> gpurun_out/${TAG}_${name}.err
  python - gpurun_out/${TAG}_${name}.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["n_gpus"], d["scaling"], d["config"]["mode"], "n_total", d["config"]["n_total"], "%.2f Gkeys/s"%(d["value"]/1e9), "ms %.3f"%d["ms_per_step"], d["verified"], {k[:10]:round(v,3) for k,v in d["config"]["stage_ms_rank0"].items()}, "e2e", d["e2e"]["value"])
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
  tail -2 gpurun_out/${TAG}_${name}.err
done
