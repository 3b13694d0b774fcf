#!/bin/bash
# N-GPU confirmation: gpu_job_mg8.sh N TAG : exchange test at world N, weak bench (3 and 4 groups), C5 (2^32 total)
N=$1; TAG=$2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_exchange_gpu.py -x -q --timeout 300 -k "[$N]" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log; tail -3 gpurun_out/${TAG}_pytest.log
run() { name=$1; extra=$2; shift; shift
  env "$@" timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 5 --warmup 3 $extra > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  echo "$name rc=$?" >> gpurun_out/${TAG}_${name}.err
  python - gpurun_out/${TAG}_${name}.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    x=d["config"].get("exchange",{})
    print(sys.argv[1], d["n_gpus"], d["scaling"], "%.2f Gkeys/s"%(d["value"]/1e9), "ms %.3f"%d["ms_per_step"], d["verified"], {k[:10]:round(v,3) for k,v in d["config"]["stage_ms_rank0"].items()}, "seg_alone", x.get("segmented_sort_alone_ms"), "e2e", d["e2e"]["value"])
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
run weak_g3 "" THRS_MG_GROUPS=3
run weak_g4 "" THRS_MG_GROUPS=4
run c5_total32 "--log2n-total 32" THRS_MG_GROUPS=3
THRS_MG_GROUPS=3 THRS_XCHG_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/${TAG}_trace.json 2> gpurun_out/${TAG}_trace.err
grep "r0\]" gpurun_out/${TAG}_trace.err | grep timeline | tail -4
tail -2 gpurun_out/${TAG}_c5_total32.err
