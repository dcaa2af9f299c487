"""One sequence through the C++ orchestrator on the GPU: host seconds per phase (whole run), per frame after the bootstrap"""
import sys
import numpy as np
sys.path.insert(0, "/root/repo")
import bench
from ldso_b200 import capi, system
import torch
n = 150
data = bench.render_sequences([1001], n)
gpu = capi.load()
pin = [torch.from_numpy(f).pin_memory() for f in data[1001][0]]
s = system.System(gpu, bench.CALIB, npoints_counter_mode=1)
ctx = s.ctx()
import time
t_first = None
for k in range(n):
    st = s.add_frame_ptr(pin[k].data_ptr())
    if st[0] and t_first is None:
        t_first, k_first, ph0 = time.perf_counter(), k, s.phase_seconds()
ctx.sync()
t1 = time.perf_counter()
ph = s.phase_seconds()
nf = n - 1 - k_first
print("frames after init:", nf, "keyframes", s.counters()["keyframes"], "wall ms/frame %.3f" % (1e3 * (t1 - t_first) / nf))
for k_, v in ph.items():
    print("  %-45s %.1f us/frame" % (k_, 1e6 * (v - ph0[k_]) / nf))
# per-frame time of the trackNewestCoarse calls (second system, same frames)
s2 = system.System(gpu, bench.CALIB, npoints_counter_mode=1)
prev, rows = 0.0, []
for k in range(n):
    st = s2.add_frame_ptr(pin[k].data_ptr())
    cur = s2.phase_seconds()["[trackNewestCoarse calls]"]
    rows.append((k, int(st[8]), int(st[1]), round(1e6 * (cur - prev))))
    prev = cur
print("frame, tries, kf, us in trackNewestCoarse:", [r for r in rows if r[3] > 400][:40])
print("median us", np.median([r[3] for r in rows if r[1] == 1]))
