"""Developer timing of thrs_scatter_pairs with all destinations on the SAME device (SM-side cost of
the multi-GPU partition pass without NVLink), for 2/4/8 destination bins."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import make_sorter  # noqa: E402

n = 1 << 28
g = torch.Generator(device="cuda"); g.manual_seed(1)
keys = torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int32, device="cuda", generator=g)
vals = torch.arange(n, dtype=torch.int32, device="cuda")
s = make_sorter("u32", 4)
tmp = torch.empty(s.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
hist = torch.zeros(256, dtype=torch.int32, device="cuda")
s.digitHistogram(keys, n, 24, hist, tmp)
h = hist.cpu().numpy().astype(np.int64)
for bins in (1, 2, 4, 8):
    per = 256 // bins
    bin_first = np.arange(bins) * per
    counts = np.array([h[b * per:(b + 1) * per].sum() for b in range(bins)])
    first = np.concatenate(([0], np.cumsum(counts)[:-1]))
    ok = torch.empty(n, dtype=torch.int32, device="cuda"); ov = torch.empty(n, dtype=torch.int32, device="cuda")
    kd = np.full(bins, ok.data_ptr(), dtype=np.uint64); vd = np.full(bins, ov.data_ptr(), dtype=np.uint64)
    ts = []
    for r in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); s.scatterPairs(keys, vals, n, tmp, 24, bin_first, kd, vd, first); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    print(f"{bins} bins, local destinations, n=2^28 pairs: {min(ts):.3f} ms (memset + table upload + kernel)", flush=True)
