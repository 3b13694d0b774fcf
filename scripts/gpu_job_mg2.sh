#!/bin/bash
# quick multi-GPU A/B: gpu_job_mg2.sh N TAG  -> exchange with 4 / 8 / 16 groups, one run with the host trace
N=$1; TAG=$2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 100 scripts/dev/paths 300000 | tail -2
run() { name=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  echo "$name rc=$?" >> gpurun_out/${TAG}_${name}.err
  python - gpurun_out/${TAG}_${name}.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    x=d["config"].get("exchange",{})
    print(sys.argv[1], d["n_gpus"], "%.2f Gkeys/s"%(d["value"]/1e9), "ms %.3f"%d["ms_per_step"], d["verified"], {k[:12]:round(v,3) for k,v in d["config"]["stage_ms_rank0"].items()}, "seg_alone", x.get("segmented_sort_alone_ms"), "e2e", d["e2e"]["value"])
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
run g8 THRS_MG_GROUPS=8
run g4 THRS_MG_GROUPS=4
run g16 THRS_MG_GROUPS=16
run g8_trace THRS_MG_GROUPS=8 THRS_XCHG_TRACE=1

grep "xchg r0" gpurun_out/${TAG}_g8_trace.err | tail -20

