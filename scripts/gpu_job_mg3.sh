#!/bin/bash
# device timeline of the exchange: gpu_job_mg3.sh N TAG [groups]
N=$1; TAG=$2; G=${3:-4}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
THRS_MG_GROUPS=$G THRS_XCHG_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${TAG}_trace.json 2> gpurun_out/${TAG}_trace.err
grep "timeline" gpurun_out/${TAG}_trace.err | grep "r0\]" | tail -10
grep "timeline" gpurun_out/${TAG}_trace.err | grep "r1\]" | sed -n '9,16p'
