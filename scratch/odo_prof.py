import sys, time, os, numpy as np, cProfile, pstats
sys.path.insert(0,'/root/repo')
from ldso_b200 import capi, synth, odometry, build
from tests import util
calib = util.CALIB["c1"]
n = 90
frames, depths, poses = synth.make_sequence(1001, n, start=40, **calib)
api = capi.Api(build.ORACLE_FAST_LIB, "orc_")
od = odometry.Odometry(api, calib)
R0,t0 = poses[0]
od.add_frame(frames[0], gt_depth=depths[0], gt_w2c=(R0.T, -R0.T@t0))
from scipy import ndimage
pr = cProfile.Profile(); pr.enable()
for i in range(1, n): od.add_frame(frames[i])
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
