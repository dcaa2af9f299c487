"""Host wall time of the C-ABI calls of one non-keyframe frame (pinned host image): makeImages, trackNewestCoarse,
traceNewCoarse (queued), drain.  Shows where the per-frame time goes beyond the kernels."""
import sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import torch
from ldso_b200 import capi
from tests import util

gpu = capi.load()
c = util.make_ctx(gpu, "c1")
gt = util.setup_tracker(c, "c1", 2000)
S = util.immature_scene("c1", pts_per_host=700)
(img0, _, _, _), (img1, _, _, _) = util.frames("c1", 1001, (40, 42))
pin = torch.from_numpy(np.ascontiguousarray(img1, np.float32)).pin_memory()
ptr = pin.data_ptr()
import ctypes as C
fn = gpu.fn("frame_make_images")
IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
acc = {"make_images": 0.0, "track": 0.0, "sync": 0.0}
N = 200
for it in range(N + 20):
    if it == 20:
        acc = {k: 0.0 for k in acc}
    t0 = time.perf_counter()
    fn(c.h, C.c_int(1), C.c_void_p(ptr))
    t1 = time.perf_counter()
    r = c.tracker_track(0, 1, gt, [0, 0])
    t2 = time.perf_counter()
    c.sync()
    t3 = time.perf_counter()
    acc["make_images"] += t1 - t0; acc["track"] += t2 - t1; acc["sync"] += t3 - t2
print({k: round(1e6 * v / N, 1) for k, v in acc.items()}, "us per call; iters", r["iters"].tolist())
c.profile_enable(True)
for it in range(50):
    fn(c.h, C.c_int(1), C.c_void_p(ptr)); c.tracker_track(0, 1, gt, [0, 0])
print("kernel us: k34 %.1f k1 %.1f" % (1e3 * c.profile_get(0)[0] / 50, 1e3 * c.profile_get(5)[0] / 50))
