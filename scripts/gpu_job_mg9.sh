#!/bin/bash
# exchange-configuration sweep on N GPUs in ONE torchrun launch (scripts/dev_xchg_sweep.py): gpu_job_mg9.sh N CONFIG...
N=${1:-8}; shift
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 \
  scripts/dev_xchg_sweep.py --steps 5 --warmup 3 "$@" > gpurun_out/mg9_n$N.txt 2> gpurun_out/mg9_n$N.err
echo "rc=$?"; cat gpurun_out/mg9_n$N.txt; tail -5 gpurun_out/mg9_n$N.err
