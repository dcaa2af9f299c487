"""Profiling target: ONE configs[1] sequence through ldso_gpu_system_add_active_frame; the steady-state frames (after
bench.PRIME + 5) run between cudaProfilerStart / Stop, so `ncu --profile-from-start off` sees exactly the timed workload
of bench.py's single_sequence leg (same frames, same keyframes)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ldso_b200 import capi, system  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
data = bench.render_sequences([1001], bench.PRIME + 5 + steps + 1)
import torch  # noqa: E402
gpu = capi.load()
pin = [torch.from_numpy(bench.to_raw8(f)).pin_memory() for f in data[1001][0]]   # the 8-bit sensor image, as bench.py's e2e
s = system.System(gpu, bench.CALIB, npoints_counter_mode=bench.DENSITY_MODE)
ctx = s.ctx()
ctx.undistort_set(bench.CALIB["w"], bench.CALIB["h"], photometric_mode=0)
for k in range(bench.PRIME + 5):
    s.add_frame_ptr(pin[k].data_ptr(), bench.RAW8)
ctx.sync()
torch.cuda.profiler.start()
nkf = 0
kfl = []
for k in range(bench.PRIME + 5, bench.PRIME + 5 + steps):
    kf = int(s.add_frame_ptr(pin[k].data_ptr(), bench.RAW8)[1])
    nkf += kf
    if kf:
        kfl.append(s.keyframe_log())
ctx.sync()
torch.cuda.profiler.stop()
print("profiled frames", steps, "keyframes", nkf, "launches total", ctx.launch_count())
print("keyframe logs (activated, currentMinActDist, ...):", kfl)
