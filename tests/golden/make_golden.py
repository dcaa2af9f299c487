#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the CPU oracle (strict build) on seeded synthetic inputs.

The reference ships no golden vectors for this path and cannot be built in this image, so
these fixtures pin the ORACLE, not the reference: they catch an accidental change of the oracle or of the GPU path
relative to the state that was argued correct.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from ldso_b200 import build, capi, synth  # noqa: E402
from tests import util  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def frontend(api):
    """K1 + K2 + K3/K4 + tracker on the 160x128 'small' calibration"""
    ctx = util.make_ctx(api, "small")
    gt = util.setup_tracker(ctx, "small", n_points=600)
    out = {}
    for lvl in range(ctx.levels):
        out[f"dI_l{lvl}"] = ctx.download(1, lvl, 0).astype(np.float32)
        out[f"abs_l{lvl}"] = ctx.download(1, lvl, 1).astype(np.float32)
        u, v, idp, col = ctx.tracker_get_pc(0, lvl)
        out[f"pc_l{lvl}"] = np.stack([u, v, idp, col], 1)
    ident = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
    for lvl in range(ctx.levels):
        out[f"res_l{lvl}"] = ctx.tracker_calc_res(0, 1, lvl, ident, 0.0, 0.0, 20.0)
        H, b = ctx.tracker_calc_gs(0, lvl, ident, 0.0, 0.0)
        out[f"H_l{lvl}"], out[f"b_l{lvl}"] = H, b
    r = ctx.tracker_track(0, 1, ident, [0.0, 0.0])
    out["track_pose"], out["track_aff"], out["track_res"], out["track_ok"] = r["pose"], r["aff"], r["last_residuals"], np.array([r["ok"]])
    out["gt_pose"] = gt
    return out


def selector(api):
    ctx = util.make_ctx(api, "small")
    img = util.frames("small")[0][0]
    ctx.make_images(0, img)
    out = {}
    pot = 3
    for k, density in enumerate((300.0, 300.0, 1200.0, 40.0)):
        m, n, pot = ctx.select_make_maps(0, density, potential=pot)
        ys, xs = np.nonzero(m)
        out[f"sel{k}_xyv"] = np.stack([xs, ys, m[ys, xs]], 1).astype(np.int32)
        out[f"sel{k}_n_pot"] = np.array([n, pot], np.int32)
    return out


def immature(api):
    S = util.immature_scene("small")
    ctx = util.make_ctx(api, "small", n_frame_slots=8)
    o = util.run_immature(ctx, S)
    out = dict(color=o["color"], weights=o["weights"], grad_h=o["grad_h"], energy_th=o["energy_th"], cand=o["cand"],
               act_idepth=o["act"][0], act_ok=o["act"][1], act_state=o["act"][2])
    for k, (c, t) in enumerate(zip(o["counts"], o["traces"])):
        out[f"counts{k}"] = c
        for key, val in t.items():
            out[f"trace{k}_{key}"] = val
    return out


def backend(api):
    """K7-K10 + H1 on a 4-keyframe window of the 'small' calibration"""
    c = util.CALIB["small"]
    W = synth.make_window(frame_ids=(40, 44, 48, 52), pts_per_frame=120, **c)
    ctx = util.make_ctx(api, "small", n_frame_slots=8)
    for i, img in enumerate(W["imgs"]):
        ctx.make_images(i, img)
    ctx.window_load(W)
    out = {}
    st, th = ctx.window_linearize(False)
    out["lin_stats"], out["lin_th"] = st, np.array([th], np.float32)
    out["J"] = ctx.window_residual_floats(0)
    out["new_state"] = ctx.window_residual_ints(1)
    ctx.window_apply_res()
    acc = ctx.window_accumulate()
    for k in ("HA", "bA", "HL", "bL", "Hsc", "bsc", "counts"):
        out[k] = acc[k]
    x = ctx.window_solve(0, 1e-5, acc, W["HM"] * 1e6, W["bM"] * 0, W["delta"] * 0)
    out["x"] = x
    out["step"] = ctx.window_resubstitute(x, 1e-5)
    out["HdiF"] = ctx.window_point_floats(6)
    return out


def closed_loop(api):
    """FullSystem::addActiveFrame over 60 frames of the 'small' sequence through the C++ orchestrator (include/ldso_system.hpp):
    bootstrap, keyframes, activation, window optimisation.  Deterministic on the CPU oracle only (the GPU's atomic sums move
    the last bits), so this part pins the oracle + orchestrator pair and is not in PARTS"""
    from ldso_b200 import system
    c = util.CALIB["small"]
    frames, depths, poses = synth.make_sequence(1001, 60, start=0, stride=1, **c)
    s, sts, kfs = system.run(api, c, frames, desired_point_density=600, desired_immature_density=500)
    ids, tr, kf = s.trajectory()
    out = dict(ids=ids.astype(np.int32), poses=tr, is_kf=kf.astype(np.uint8),
               kf_frames=np.array([k["frame"] for k in kfs], np.int32),
               kf_points=np.array([k["points"] for k in kfs], np.int32),
               kf_residuals=np.array([k["residuals"] for k in kfs], np.int32),
               kf_activated=np.array([k["activated"] for k in kfs], np.int32),
               kf_rmse=np.array([k["rmse"] for k in kfs], np.float64),
               window=np.array(s.window(), np.int32), cpu=np.array(cpu_model()))
    s.close()
    return out


def cpu_model():
    try:
        return [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except (OSError, IndexError):
        return "unknown"


PARTS = dict(frontend=frontend, selector=selector, immature=immature, backend=backend)
CPU_PARTS = dict(closed_loop=closed_loop)

if __name__ == "__main__":
    build.build_oracle()
    api = capi.Api(build.ORACLE_LIB, "orc_")
    for name, fn in list(PARTS.items()) + list(CPU_PARTS.items()):
        d = fn(api)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, sum(v.nbytes for v in d.values()) // 1024, "KiB raw")
