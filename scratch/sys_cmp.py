"""GPU vs oracle through the C++ FullSystem: keyframe logs side by side, trajectory difference, timing"""
import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
from ldso_b200 import build, capi, system, synth
from tests import util
calib = sys.argv[1] if len(sys.argv) > 1 else "c1"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 150
stride = int(sys.argv[3]) if len(sys.argv) > 3 else 3
mode = int(sys.argv[4]) if len(sys.argv) > 4 else 1
c = util.CALIB[calib]
frames, depths, poses = synth.make_sequence(1001, n, start=0, stride=stride, **c)
res = {}
for name, api in (("gpu", capi.load()), ("orc", capi.Api(build.ORACLE_LIB, "orc_"))):
    t0 = time.time()
    s, sts, kfs = system.run(api, c, frames, npoints_counter_mode=mode)
    dt = time.time() - t0
    ids, cen, kf = s.centres()
    gt = np.array([poses[i][1] for i in ids])
    ate, scale = system.ate_rmse(cen, gt)
    res[name] = dict(sts=sts, kfs=kfs, ids=ids, cen=cen, ate=ate, scale=scale, dt=dt, cnt=s.counters(), ph=s.phase_seconds())
    print(name, "time %.2fs ATE %.3e scale %.4f" % (dt, ate, scale), s.counters())
    print(name, {k: round(v, 3) for k, v in s.phase_seconds().items()})
g, o = res["gpu"], res["orc"]
for a, b in zip(g["kfs"], o["kfs"]):
    print("KF gpu f%d pts %d res %d marg %d lin %d act %d rmse %.4f win %s" % (a["frame"], a["points"], a["residuals"], a["marg_points"], a["lin_residuals"], a["activated"], a["rmse"], a["window"]))
    print("   orc f%d pts %d res %d marg %d lin %d act %d rmse %.4f win %s" % (b["frame"], b["points"], b["residuals"], b["marg_points"], b["lin_residuals"], b["activated"], b["rmse"], b["window"]))
if len(g["ids"]) == len(o["ids"]):
    print("max |centre gpu - orc| %.3e" % np.abs(g["cen"] - o["cen"]).max())
print("tries gpu", [s["tries"] for s in g["sts"]])
print("tries orc", [s["tries"] for s in o["sts"]])
