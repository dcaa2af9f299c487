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


def immature_scene(calib="small", host_ids=(40, 44), new_ids=(46, 48, 50), pts_per_host=300, seed=1001):
    """Keyframes hosting fresh immature points + later frames to trace them on (FullSystem::traceNewCoarse)."""
    c = CALIB[calib]
    fr = frames(calib, seed, tuple(host_ids) + tuple(new_ids))
    K = np.array([[c["fx"], 0, c["cx"]], [0, c["fy"], c["cy"]], [0, 0, 1]], dtype=np.float32)
    Ki = np.linalg.inv(K.astype(np.float64)).astype(np.float32)
    nh = len(host_ids)
    U, V, H, GT = [], [], [], []
    for hi in range(nh):
        img, dep, _, _ = fr[hi]
        u, v, idp, _ = pick_points(img, dep, pts_per_host, seed=hi + 3, border=8)
        U.append(np.round(u)); V.append(np.round(v)); H.append(np.full(len(u), hi, np.int32)); GT.append(idp)
    tables = []
    for ni in range(len(new_ids)):
        _, _, Rn, tn = fr[nh + ni]
        KRKi, Kt, aff = [], [], []
        for hi in range(nh):
            _, _, Rh, th = fr[hi]
            R = (Rn.T @ Rh).astype(np.float32); t = (Rn.T @ (th - tn)).astype(np.float32)
            KRKi.append((K @ R @ Ki).astype(np.float32)); Kt.append((K @ t).astype(np.float32)); aff.append([1.0, 0.0])
        tables.append((np.array(KRKi, np.float32), np.array(Kt, np.float32), np.array(aff, np.float32)))
    w2c = [(R.T, -R.T @ t) for (_, _, R, t) in fr]
    return dict(calib=c, frames=fr, n_hosts=nh, u=np.concatenate(U).astype(np.float32), v=np.concatenate(V).astype(np.float32),
                host=np.concatenate(H), gt_idepth=np.concatenate(GT).astype(np.float32), tables=tables, w2c=w2c)


def run_immature(ctx, S, activate=True, async_trace=False):
    """Drive create -> set -> trace on every new frame -> activate through one C ABI (oracle or GPU)."""
    from ldso_b200 import window_host as wh
    nh, fr = S["n_hosts"], S["frames"]
    for i, f in enumerate(fr):
        ctx.make_images(i, f[0])
    cols, wgts, ghs, eths = [], [], [], []
    for hi in range(nh):
        m = S["host"] == hi
        col, wgt, gh, eth = ctx.immature_create(hi, S["u"][m], S["v"][m])
        cols.append(col); wgts.append(wgt); ghs.append(gh); eths.append(eth)
    col, wgt, gh, eth = [np.concatenate(x) for x in (cols, wgts, ghs, eths)]
    n = len(S["u"])
    ctx.immature_set(S["host"], S["u"], S["v"], col, wgt, gh, eth, np.zeros(n, np.float32), np.full(n, np.nan, np.float32),
                     np.full(n, 10000.0, np.float32), np.full(n, 5, np.int32))
    out = dict(color=col, weights=wgt, grad_h=gh, energy_th=eth, traces=[], counts=[])
    for ni, (KRKi, Kt, aff) in enumerate(S["tables"]):
        if async_trace:                      # queued without waiting; immature_get completes it, the counts come afterwards
            ctx.immature_trace_async(nh + ni, KRKi, Kt, aff)
            out["traces"].append(ctx.immature_get())
            out["counts"].append(ctx.immature_trace_counts())
        else:
            out["counts"].append(ctx.immature_trace(nh + ni, KRKi, Kt, aff))
            out["traces"].append(ctx.immature_get())
    if activate:
        c = S["calib"]
        nf = len(fr)
        w2c = S["w2c"]
        pc = wh.precalc_table(w2c, w2c, (c["fx"], c["fy"], c["cx"], c["cy"]), [(0.0, 0.0)] * nf, [(0.0, 0.0)] * nf, [1.0] * nf)
        W = dict(n=nf, precalc=pc, frame_energy_th=np.full(nf, 512.0, np.float32),
                 calib=np.array([c["fx"], c["fy"], c["cx"], c["cy"]], np.float32))
        ctx.window_set_frames(W)
        last = out["traces"][-1]
        cand = np.nonzero((last["status"] == 0) | (last["status"] == 3) | (last["status"] == 4))[0].astype(np.int32)
        out["cand"] = cand
        out["act"] = ctx.immature_activate(cand, min_obs=1)
    return out
