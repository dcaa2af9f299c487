import sys, time, numpy as np
sys.path.insert(0,'/root/repo')
from ldso_b200 import capi, synth, odometry
from tests import util
api = capi.Api('/root/repo/oracle/liboracle_fast.so','orc_')
calib = util.CALIB[sys.argv[1] if len(sys.argv)>1 else "small"]
n = int(sys.argv[2]) if len(sys.argv)>2 else 30
frames, depths, poses = synth.make_sequence(1001, n, start=40, **calib)
t=time.time()
od, est, gt, ids = odometry.run(api, "x", frames, depths, poses, calib, desired_points=int(sys.argv[3]) if len(sys.argv)>3 else 600, desired_immature=int(sys.argv[4]) if len(sys.argv)>4 else 500)
print("time", time.time()-t)
for l in od.log: print(l)
print("ATE", odometry.ate_rmse(est, gt), "path", np.linalg.norm(np.diff(gt,axis=0),axis=1).sum())
print(np.abs(est-gt).max(0))
