"""The one-process-per-GPU multi-GPU sort (DistributedRadixSort over thrs_xchg_*, CUDA IPC windows,
copy-engine bucket exchange, segmented sort per received group) on REAL devices: needs >= 2 GPUs
(skipped on a single-GPU box, where tests/test_mg_capi_gpu.py runs the same exchange layer with several
ranks on device 0).  Launched with torch.distributed.run like bench.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
def test_exchange_sort_on_real_gpus(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29610 + world), os.path.join(root, "tests", "mp_exchange_worker.py")]
    r = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "EXCHANGE-OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
