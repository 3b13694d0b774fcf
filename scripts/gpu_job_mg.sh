#!/bin/bash
# multi-GPU job: gpu_job_mg.sh N TAG [extra bench args]  (run under gpurun --gpus N)
N=$1; TAG=$2; shift; shift
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_exchange_gpu.py tests/test_mg_capi_gpu.py -x -q --timeout 400 > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 5 --warmup 3 $EXTRA > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  echo "$name rc=$?" >> gpurun_out/${TAG}_${name}.err
}
EXTRA="$*"
run exchange THRS_MG_MODE=exchange
run exchange_g8 THRS_MG_MODE=exchange THRS_MG_GROUPS=8
run exchange_g2 THRS_MG_MODE=exchange THRS_MG_GROUPS=2
run scatter THRS_MG_MODE=scatter
tail -3 gpurun_out/${TAG}_pytest.log
for f in gpurun_out/${TAG}_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["n_gpus"], "value %.2f Gkeys/s"%(d["value"]/1e9), "ms %.3f"%d["ms_per_step"], d["verified"], json.dumps(d["config"]["stage_ms_rank0"]), json.dumps(d["config"].get("exchange"))[:600])
except Exception as e:
    print("no line:", e)
PY
done
tail -5 gpurun_out/${TAG}_exchange.err
