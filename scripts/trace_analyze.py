"""Dev tool: timing of look-back events per tile from gpurun_out/trace.bin (scripts/devbench with a
-DTHRS_PHASE_TIMING build of the library writes it).  python scripts/trace_analyze.py [file] [tiles]"""
import sys
import numpy as np

path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/trace.bin"
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 26214
tr = np.fromfile(path, dtype=np.uint64).reshape(6, 32768).astype(np.int64)
t0 = tr[0, 1:nt][tr[0, 1:nt] > 0].min()
ev = tr[:, :nt] - t0
names = ["draw", "scan", "far", "near", "res0", "res1"]
lo, hi = 2000, nt - 2000
sl = slice(lo, hi)
print("kernel span us", ev[5, :nt].max() / 1e3)
for a, b in [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (0, 5)]:
    d = (ev[b, sl] - ev[a, sl]) / 1e3
    print(f"{names[a]}->{names[b]}: mean {d.mean():.2f} us, p10 {np.percentile(d, 10):.2f}, p50 {np.percentile(d, 50):.2f}"
          f" p90 {np.percentile(d, 90):.2f} p99 {np.percentile(d, 99):.2f}")
scan = ev[1]
for w in (7, 16, 64, 256):
    T = np.arange(lo, hi, 7)
    m = np.array([scan[t - w:t].max() for t in T])
    lf, ln, lr = (m - ev[2, T]) / 1e3, (m - ev[3, T]) / 1e3, (m - ev[4, T]) / 1e3
    print(f"window {w}: latest predecessor scan - my far issue: mean {lf.mean():.2f} us, late {(lf > 0).mean():.2f} |"
          f" - near issue: mean {ln.mean():.2f}, late {(ln > 0).mean():.2f} | - resolve start: mean {lr.mean():.2f}, late {(lr > 0).mean():.2f}")
x = np.arange(lo, hi)
p = np.polyfit(x, scan[sl], 1)
res = (scan[sl] - np.polyval(p, x)) / 1e3
print(f"scan time vs tile index: {p[0]:.1f} ns per tile, residual std {res.std():.2f} us, min {res.min():.1f}, max {res.max():.1f}")
print("evolution (1000-tile windows): residual std of scan time [us], mean resolve duration [us], mean draw->scan [us]")
for a in range(0, nt - 1000, 2000):
    w = slice(a + 1, a + 1001)
    xx = np.arange(a + 1, a + 1001)
    pp = np.polyfit(xx, scan[w], 1)
    rr = (scan[w] - np.polyval(pp, xx)) / 1e3
    print(f"  tiles {a:6d}+: slope {pp[0]:6.1f} ns/tile, jitter {rr.std():5.2f}, resolve {((ev[5, w] - ev[4, w]) / 1e3).mean():5.2f}, draw->scan {((ev[1, w] - ev[0, w]) / 1e3).mean():6.2f}")
