"""The orchestrator's own host work per frame: wall clock minus the seconds spent inside the C ABI.  CPU only: the oracle
library built with the measurement switch of include/ldso_system.hpp,
  g++ -std=c++17 -O3 -march=x86-64-v3 -DENABLE_SSE -DLDSO_API_TIMING -fPIC -shared -pthread -fopenmp -w -o build/liboracle_timing.so oracle/*.cpp
LDSO_SYS_PROF=1 adds the per-phase table (wall, and host outside the C ABI) when the system is destroyed."""
import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np
import ctypes as C
from ldso_b200 import capi, synth, system
from tests import util
api = capi.Api("/root/repo/build/liboracle_timing.so", "orc_")
c = util.CALIB["c1"]
N = 150
frames, depths, poses = synth.make_sequence(1001, N, start=0, stride=3, **c)
s = system.System(api, c, npoints_counter_mode=1)
rows = []
for i, f in enumerate(frames):
    p = f.ctypes.data_as(C.c_void_p)
    a0 = s.phase_seconds()["[unused]"]; t0 = time.perf_counter()
    st = s.add_frame_ptr(p, False, float(i))
    w = time.perf_counter() - t0
    a = s.phase_seconds()["[unused]"] - a0
    rows.append((i, int(st[1]), w, a))
r = np.array(rows[110:])
kf = r[:, 1] == 1
print("non-KF frames %d: wall %.1f us api %.1f us orchestrator %.1f us" % ((~kf).sum(), 1e6 * r[~kf, 2].mean(), 1e6 * r[~kf, 3].mean(), 1e6 * (r[~kf, 2] - r[~kf, 3]).mean()))
print("KF frames %d: wall %.1f us api %.1f us orchestrator %.1f us" % (kf.sum(), 1e6 * r[kf, 2].mean(), 1e6 * r[kf, 3].mean(), 1e6 * (r[kf, 2] - r[kf, 3]).mean()))
