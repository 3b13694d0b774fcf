#!/bin/bash
# multi-GPU confirmation with the defaults: the tests that need 2+ real GPUs, then the judged bench command: gpu_job_mg10.sh N [pytest]
N=${1:-4}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ "$2" = "pytest" ]; then
  timeout 300 python -m pytest tests/test_mg_capi_gpu.py tests/test_exchange_gpu.py tests/test_scatter_gpu.py -m gpu -q -x > gpurun_out/mg10_n${N}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/mg10_n${N}_pytest.log
  tail -4 gpurun_out/mg10_n${N}_pytest.log
fi
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 5 --warmup 3 --e2e-steps 1 > gpurun_out/mg10_n$N.json 2> gpurun_out/mg10_n$N.err
echo "bench rc=$?"; python - gpurun_out/mg10_n$N.json <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith("{"):
        d=json.loads(l); print(d["n_gpus"], "GPUs", round(d["value"]/1e9,1), "Gkeys/s", round(d["ms_per_step"],3), "ms", d["verified"], d["config"]["mode"], d["config"]["stage_ms_rank0"], "e2e", d["e2e"]["value"])
PY
tail -3 gpurun_out/mg10_n$N.err
