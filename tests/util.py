"""Shared scenario builders for the parity tests (seeded, deterministic)."""
import functools

import numpy as np

from ldso_b200 import synth

CALIB = {
    "c1": dict(w=640, h=480, fx=384.0, fy=384.0, cx=319.5, cy=239.5),           # BASELINE configs 1/2/5
    "c3": dict(w=1232, h=368, fx=718.856, fy=718.856, cx=602.69, cy=181.22),    # KITTI crop (Kitti00-02.txt)
    "c4": dict(w=752, h=480, fx=460.0, fy=460.0, cx=375.5, cy=239.5),           # EuRoC-shaped
    "small": dict(w=160, h=128, fx=96.0, fy=96.0, cx=79.5, cy=63.5),            # 3 levels, fast
}


@functools.lru_cache(maxsize=16)
def frames(calib="c1", seed=1001, idx=(40, 42)):
    c = CALIB[calib]
    sc = synth.Scene(seed)
    out = []
    for i in idx:
        R, t = synth.trajectory(i)
        img, dep = synth.render(sc, c["w"], c["h"], c["fx"], c["fy"], c["cx"], c["cy"], R, t)
        out.append((img, dep, R, t))
    return out


pick_points = synth.pick_points


def make_ctx(api, calib="c1", **kw):
    c = CALIB[calib]
    cfg = api.default_config(w=c["w"], h=c["h"], fx=c["fx"], fy=c["fy"], cx=c["cx"], cy=c["cy"],
                             affine_opt_mode_a=0.0, affine_opt_mode_b=0.0, **kw)   # mode=1 (run_dso_tum_mono.cc:283-288)
    return api.create(cfg)


def setup_tracker(ctx, calib="c1", n_points=2000, idx=(40, 42), seed=1001, idepth_noise=0.0):
    """slot 0 = reference frame, slot 1 = new frame; tracker 0 referenced on slot 0."""
    (img0, dep0, R0, t0), (img1, dep1, R1, t1) = frames(calib, seed, idx)
    ctx.make_images(0, img0)
    ctx.make_images(1, img1)
    u, v, idepth, hdi = pick_points(img0, dep0, n_points)
    if idepth_noise > 0:
        rng = np.random.default_rng(3)
        idepth = (idepth * (1 + idepth_noise * rng.standard_normal(idepth.size))).astype(np.float32)
    ctx.tracker_set_ref(0, 0, u, v, idepth, hdi)
    gt = synth.rel_pose(R0, t0, R1, t1)
    return gt


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
