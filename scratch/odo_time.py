import sys, time, os, numpy as np
sys.path.insert(0,'/root/repo')
os.environ["OMP_WAIT_POLICY"]="passive"; os.environ["ORC_THREADS"]="6"
from ldso_b200 import capi, synth, odometry, build
from tests import util
calib = util.CALIB["c1"]
n = 200
frames, depths, poses = synth.make_sequence(1001, n, start=40, **calib)
for name, api in (("gpu", capi.load()), ("oracle6", capi.Api(build.ORACLE_FAST_LIB, "orc_"))):
    od = odometry.Odometry(api, calib)
    R0,t0 = poses[0]
    od.add_frame(frames[0], gt_depth=depths[0], gt_w2c=(R0.T, -R0.T@t0))
    for i in range(1, 21): od.add_frame(frames[i])
    t=time.perf_counter(); tk = 0; nk = 0
    for i in range(21, n):
        t1=time.perf_counter(); r = od.add_frame(frames[i]); 
        if r["kf"]: tk += time.perf_counter()-t1; nk += 1
    dt=time.perf_counter()-t
    ids = sorted(od.poses); est = np.array([-od.poses[i][0].T @ od.poses[i][1] for i in ids]); gt = np.array([poses[i][1] for i in ids])
    print(name, "frames/s", (n-21)/dt, "ms/frame", 1e3*dt/(n-21), "kf", nk, "ms/kf", 1e3*tk/max(nk,1), "non-kf ms", 1e3*(dt-tk)/(n-21-nk), "ATE", odometry.ate_rmse(est, gt)[0])
