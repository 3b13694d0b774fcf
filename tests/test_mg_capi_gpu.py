"""thrs_mg_* (single-process multi-GPU entry points of the C ABI): the concatenated outputs must be
the oracle's stable sort of the concatenated inputs.  With one visible GPU the device id is
repeated, which runs the real kernels and peer-pointer tables with several ranks on one device;
with more GPUs the first min(count, 4) real devices are used as well."""
import numpy as np
import pytest
import torch

import oracle
from tinyhipradixsort_b200.multigpu import SingleProcessMultiGpuSort

pytestmark = pytest.mark.gpu
_NP = {"u32": np.uint32, "f32": np.uint32, "u64": np.uint64, "f64": np.uint64}


def _run(ids, key, vbytes, desc, sizes, case):
    rng = oracle.SplitMix64(3 + len(ids))
    slices = [rng.randomize(key, n, mode=1) for n in sizes]
    if case == "equal":
        for s in slices:
            s[:] = 0xDEADBEEF
    elif case == "skew":
        for s in slices:
            s[: len(s) // 2] &= 0xFFFF
    vdt = np.uint32 if vbytes == 4 else np.uint64
    base = np.cumsum([0] + list(sizes))
    vals = [np.arange(n, dtype=vdt) + vdt(base[i]) for i, n in enumerate(sizes)]
    total = int(base[-1])
    kd = torch.int32 if slices[0].dtype == np.uint32 else torch.int64
    vd = torch.int32 if vbytes == 4 else torch.int64
    dk, dv, ok, ov = [], [], [], []
    for g, dev in enumerate(ids):
        with torch.cuda.device(dev):
            dk.append(torch.from_numpy(slices[g].view(np.int32 if kd == torch.int32 else np.int64)).cuda())
            dv.append(torch.from_numpy(vals[g].view(np.int32 if vd == torch.int32 else np.int64)).cuda())
            ok.append(torch.empty(max(total, 1), dtype=kd, device="cuda"))  # room for the worst split
            ov.append(torch.empty(max(total, 1), dtype=vd, device="cuda"))
    mg = SingleProcessMultiGpuSort(ids, key, vbytes, desc)
    counts = mg.sort_pairs(dk, dv, ok, ov)
    assert sum(counts) == total
    got_k = np.concatenate([ok[g][:counts[g]].cpu().numpy().view(_NP[key]) for g in range(len(ids))])
    got_v = np.concatenate([ov[g][:counts[g]].cpu().numpy().view(vdt) for g in range(len(ids))])
    exp_k, exp_v = oracle.sort(key, np.concatenate(slices), np.concatenate(vals), 0, None, desc)
    assert np.array_equal(got_k, exp_k), "keys differ from the stable sort of the whole input"
    assert np.array_equal(got_v, exp_v), "global stability broken"
    for g in range(len(ids)):  # the inputs are only read
        assert np.array_equal(dk[g].cpu().numpy().view(_NP[key]), slices[g])
    mg.close()
    return counts


@pytest.mark.parametrize("nranks,key,vbytes,desc,case", [(1, "u32", 4, False, "uniform"), (2, "u32", 4, False, "uniform"),
                                                         (3, "f32", 4, True, "uniform"), (4, "u64", 8, False, "skew"),
                                                         (2, "u32", 4, False, "equal"), (8, "u32", 4, False, "uniform")])
def test_mg_sort_pairs_emulated_ranks_on_device_0(nranks, key, vbytes, desc, case):
    sizes = [200_003, 3000, 0, 77_777, 150_000, 1, 90_000, 40_960][:nranks]
    counts = _run([0] * nranks, key, vbytes, desc, sizes, case)
    if case == "uniform" and nranks in (2, 8):
        assert max(counts) < 1.3 * sum(counts) / nranks + 4096, "uniform keys must split evenly"


def test_mg_sort_pairs_real_devices():
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("one visible GPU")
    _run(list(range(n)), "u32", 4, False, [300_000 + 1000 * g for g in range(n)], "uniform")


def test_mg_capacity_error_reports_needed_sizes():
    mg = SingleProcessMultiGpuSort([0, 0], "u32", 4)
    k = [torch.zeros(1000, dtype=torch.int32, device="cuda"), torch.ones(500, dtype=torch.int32, device="cuda")]
    v = [torch.zeros(1000, dtype=torch.int32, device="cuda"), torch.zeros(500, dtype=torch.int32, device="cuda")]
    small = [torch.empty(10, dtype=torch.int32, device="cuda") for _ in range(2)]
    with pytest.raises(RuntimeError):
        mg.sort_pairs(k, v, small, [torch.empty(10, dtype=torch.int32, device="cuda") for _ in range(2)])
    assert sum(mg.last_needed) == 1500
    mg.close()
