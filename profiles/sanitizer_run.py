"""Address / undefined-behaviour sanitizer pass over the host code that both arms share (include/ldso_system.hpp: orchestrator,
batch runner, fibers) and over the oracle.  CPU only:
  g++ -std=c++17 -O1 -g -march=x86-64-v3 -DENABLE_SSE -fPIC -shared -pthread -fsanitize=address,undefined \
      -fno-omit-frame-pointer -ffp-contract=off -w -o build/liboracle_asan.so oracle/*.cpp
  LD_PRELOAD=$(g++ -print-file-name=libasan.so):$(g++ -print-file-name=libubsan.so) ASAN_OPTIONS=detect_leaks=0 \
      UBSAN_OPTIONS=print_stacktrace=1 python profiles/sanitizer_run.py
Covers: bootstrap, keyframes with marginalisation and FEJ (configs[1], 100 frames), the 10-keyframe / 8000-point shape of
configs[3], a lost system, batch_run in two calls with three fibers on one worker.  Last run (head of round 2): no report
(ASan's note that it does not fully support swapcontext is printed once)."""
import sys
sys.path.insert(0, "/root/repo")
import numpy as np
from ldso_b200 import capi, synth, system
from tests import util
api = capi.Api("/root/repo/build/liboracle_asan.so", "orc_")
c = util.CALIB["small"]
kw = dict(desired_point_density=600, desired_immature_density=500, npoints_counter_mode=1)
frames, depths, poses = synth.make_sequence(1001, 70, start=0, stride=2, **c)
s, sts, kfs = system.run(api, c, frames, **kw)
print("single: kfs", len(kfs), "lost", sts[-1]["lost"], s.counters())
# lost path
flat = np.zeros_like(frames[0])
for k in range(2): print("lost?", s.add_frame(flat)[0]["lost"])
s.close()
# batch with fibers, two calls
seqs = [synth.make_sequence(1001 + i, 30, start=0, **c)[0] for i in range(3)]
imgs = [[np.ascontiguousarray(f, np.float32) for f in fr] for fr in seqs]
ss = [system.System(api, c, **kw) for _ in seqs]
b = system.Batch(ss, 1)
for lo in (0, 15):
    st = b.run_ptrs([[im.ctypes.data for im in row[lo:lo + 15]] for row in imgs], sync_after=True)
print("batch ok", st[:, -1, 0])
b.close()
for x in ss: x.close()

c = util.CALIB["c1"]
frames, depths, poses = synth.make_sequence(1001, 100, start=0, stride=3, **c)
s, sts, kfs = system.run(api, c, frames, npoints_counter_mode=1)
print("c1: kfs", len(kfs), "lost", sts[-1]["lost"], s.counters())
HM, bM = s.prior(); print(HM.shape, s.window(), len(s.trajectory()[0]))
s.close()
# config 4 shape: 10-keyframe window, 8000 points
c4 = util.CALIB["c4"]
frames, depths, poses = synth.make_sequence(1001, 50, start=0, stride=3, **c4)
s, sts, kfs = system.run(api, c4, frames, npoints_counter_mode=1, desired_point_density=8000, desired_immature_density=6000, max_frames=10, min_frames=7)
print("c4: kfs", len(kfs), "lost", sts[-1]["lost"], s.counters())
s.close()

print("done")
