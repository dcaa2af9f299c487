"""Steady-state cost of a keyframe: host wall time per phase of FullSystem::makeKeyFrame (one sequence on the GPU, no L2
flush), averaged over the keyframes after bench.PRIME; and the cost of a plain tracked frame."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ldso_b200 import capi, system  # noqa: E402

n = bench.PRIME + 60
data = bench.render_sequences([1001], n)
import torch  # noqa: E402
gpu = capi.load()
pin = [torch.from_numpy(f).pin_memory() for f in data[1001][0]]
s = system.System(gpu, bench.CALIB, npoints_counter_mode=bench.DENSITY_MODE)
ctx = s.ctx()
for k in range(bench.PRIME):
    s.add_frame_ptr(pin[k].data_ptr())
ctx.sync()
kf_ph, fr_t, kf_t = {}, [], []
for k in range(bench.PRIME, n):
    p0 = s.phase_seconds()
    t0 = time.perf_counter()
    st = s.add_frame_ptr(pin[k].data_ptr())
    ctx.sync()
    dt = time.perf_counter() - t0
    if st[1]:
        p1 = s.phase_seconds()
        kf_t.append(dt)
        for key in p1:
            kf_ph[key] = kf_ph.get(key, 0.0) + p1[key] - p0[key]
    else:
        fr_t.append(dt)
print("plain frames %d: %.1f us each; keyframes %d: %.1f us each" % (len(fr_t), 1e6 * sum(fr_t) / len(fr_t), len(kf_t), 1e6 * sum(kf_t) / len(kf_t)))
for key, v in kf_ph.items():
    if v > 0:
        print("  %-45s %.1f us per keyframe" % (key, 1e6 * v / len(kf_t)))
print(s.keyframe_log())
