"""Phase clocks of the device-resident tracker (LDSO_LM_DBG=3 semantics, cfg.reserved[0] = 3): SM clocks of CTA 0 / thread 0
spent in evaluation + push, barrier, fold, decide, solve, candidate.  Modes: 0 = thread-block cluster, 2 = cooperative grid."""
import sys
import numpy as np
sys.path.insert(0, "/root/repo")
from ldso_b200 import capi
from tests import util

gpu = capi.load()
IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
cc = util.CALIB["c1"]
for mode in (2, 0):
    for dbg in (0, 3):
        cfg = gpu.default_config(w=cc["w"], h=cc["h"], fx=cc["fx"], fy=cc["fy"], cx=cc["cx"], cy=cc["cy"], affine_opt_mode_a=0.0,
                                 affine_opt_mode_b=0.0, tracker_lm_mode=mode)
        cfg.reserved[0] = dbg
        c = gpu.create(cfg)
        util.setup_tracker(c, "c1", 2000)
        for rep in range(4):
            c.profile_enable(True)
            r = c.tracker_track(0, 1, IDENT, [0, 0])
            ms, n, units = c.profile_get(0)
            print("mode", mode, "dbg", dbg, "rep", rep, "total_us", round(ms * 1e3, 1), "iters", r["iters"].tolist(), "us/eval",
                  round(ms * 1e3 / max(r["iters"][0], 1), 2), "pose_t", r["pose"][9:].round(6).tolist(), flush=True)
        c.close()
