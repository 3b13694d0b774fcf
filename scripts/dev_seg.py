"""Developer timing of the segmented sort on the GPU box: plain 3-pass sort on bits [0,24) vs
thrs_sort_segments_pairs over the same 2^28 pairs cut into 1 / 32 / 128 / 256 equal segments."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tinyhipradixsort_b200 as thrs  # noqa: E402

n = 1 << int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 28
g = torch.Generator(device="cuda"); g.manual_seed(1)
k0 = torch.randint(-(1 << 31), (1 << 31) - 1, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
v0 = torch.arange(n, dtype=torch.int32, device="cuda")
s = thrs.RadixSort([], thrs.RadixSort.Config().configureWithKeyPair("u32", 4))
k, v = k0.clone(), v0.clone()
st = torch.cuda.current_stream()


def timed(fn, reps=4):
    best = None
    for _ in range(reps):
        k.copy_(k0); v.copy_(v0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st); fn(); b.record(st); b.synchronize()
        ms = a.elapsed_time(b)
        best = ms if best is None else min(best, ms)
    return best


tmp = torch.empty(s.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
print(f"plain sortPairs bits[0,24) (3 passes + copy-back): {timed(lambda: s.sortPairs(k, v, n, tmp, 0, 24, st)):.3f} ms", flush=True)
print(f"plain sortPairs bits[0,32) (4 passes):              {timed(lambda: s.sortPairs(k, v, n, tmp, 0, 32, st)):.3f} ms", flush=True)
del tmp
for nseg in (1, 32, 128, 256) if n >= (1 << 20) else (1,):
    if n // nseg >= (1 << 30):
        continue
    edges = [i * (n // nseg) for i in range(nseg)] + [n]
    stmp = torch.empty(s.getSegmentsTemporaryBufferBytes(n, nseg, True), dtype=torch.uint8, device="cuda")
    print(f"segmented {nseg:4d} segments bits[0,24) (3 passes + copy-back): {timed(lambda: s.sortSegments(k, v, edges, stmp, 0, 24, st)):.3f} ms", flush=True)
    del stmp
