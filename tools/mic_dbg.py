"""Reads mic_dbg.bin (MF_MIC_DBG=<sweep number>, csrc/mic.cuh) and prints where a MIC(0) tile-wavefront sweep spends its time."""
import sys

import numpy as np

raw = open(sys.argv[1] if len(sys.argv) > 1 else "mic_dbg.bin", "rb").read()
tx, ty, tz, mode = np.frombuffer(raw[:16], np.int32)
d = np.frombuffer(raw[16:], np.uint64).reshape(-1, 8).astype(np.int64)
t0 = d[:, 0].min()
st = d[:, :6] - t0
names = ["static", "wait", "faces", "sweep", "publish"]
has_pred = st[:, 2] > 0
print("tiles %d x %d x %d, mode %d, span %.1f us, SMs used %d" % (tx, ty, tz, mode, (st[:, 5].max()) / 1e3, len(set(d[:, 6]))))
for q, n in enumerate(names):
    a, b = st[:, q], st[:, q + 1]
    if n == "wait":
        b = np.where(has_pred, b, a)
    if n == "faces":
        a = np.where(has_pred, a, st[:, 1])
    dt = (b - a)
    print("  %-8s mean %7.2f us   median %7.2f   p90 %7.2f   max %7.2f" % (n, dt.mean() / 1e3, np.median(dt) / 1e3, np.percentile(dt, 90) / 1e3, dt.max() / 1e3))
# wavefront progress: completion time of the last tile of each tile wavefront
idx = np.arange(d.shape[0])
bi, bj, bk = idx % tx, (idx // tx) % ty, idx // (tx * ty)
if mode == 2:
    bi, bj, bk = tx - 1 - bi, ty - 1 - bj, tz - 1 - bk
w = bi + bj + bk
print("  wavefront: tiles, first ticket, last published (us)")
for ww in range(0, w.max() + 1, max(1, (w.max() + 1) // 16)):
    m = w == ww
    print("   %3d %5d %8.1f %8.1f   sweep-start min %8.1f" % (ww, m.sum(), st[m, 0].min() / 1e3, st[m, 5].max() / 1e3, st[m, 3].min() / 1e3))
# critical chain through the main diagonal tiles (b, b, b)
for b in range(min(tx, ty, tz)):
    t = b + tx * (b + ty * b) if mode != 2 else (tx - 1 - b) + tx * ((ty - 1 - b) + ty * (tz - 1 - b))
    print("   diag %2d: ticket %8.1f static %8.1f flags %8.1f faces %8.1f swept %8.1f published %8.1f  (sm %d, q %d)" % ((b,) + tuple(st[t, :6] / 1e3) + (d[t, 6], d[t, 7])))
