import sys, numpy as np, ctypes as C
sys.path.insert(0,'/root/repo')
from ldso_b200 import capi
from tests import util
gpu = capi.load()
IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
for dbg in (0, 1, 2, 3):
    cfg_kw = dict(tracker_lm_mode=0)
    c = util.make_ctx(gpu, "c1", **cfg_kw)
    c.cfg.reserved[0] = dbg
    # cfg is copied at create: recreate with the field set
    c.close()
    cc = util.CALIB["c1"]
    cfg = gpu.default_config(w=cc["w"], h=cc["h"], fx=cc["fx"], fy=cc["fy"], cx=cc["cx"], cy=cc["cy"], affine_opt_mode_a=0.0, affine_opt_mode_b=0.0, tracker_lm_mode=0)
    cfg.reserved[0] = dbg
    c = gpu.create(cfg)
    util.setup_tracker(c, "c1", 2000)
    for rep in range(3):
        c.profile_enable(True)
        r = c.tracker_track(0, 1, IDENT, [0, 0])
        ms, n, units = c.profile_get(0)
        print("dbg", dbg, "rep", rep, "launches", n, "total_us", round(ms*1e3,1), "iters", r["iters"], "us/eval", round(ms*1e3/max(r["iters"][0] if dbg==0 else 12*n,1),2))
    c.close()
