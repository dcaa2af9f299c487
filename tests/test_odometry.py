"""Closed loop: the whole hot path in the reference's order (track -> trace -> keyframe: activate, window BA, drop, new
tracking reference, new immature points) through ldso_b200/odometry.py, which restates the control flow of FullSystem.
Gate of SURVEY.md section 8d: Sim(3)-aligned ATE below 1 % of the path length, and the GPU run within 10 % of the
oracle's ATE (or both far below a millimetre)."""
import functools

import numpy as np
import pytest

from ldso_b200 import odometry, synth
from tests import util


def _seq(calib, n):
    c = util.CALIB[calib]
    return (c,) + synth.make_sequence(1001, n, start=40, **c)


def test_oracle_closed_loop_tracks_ground_truth(oracle):
    c, frames, depths, poses = _seq("small", 60)
    od, est, gt, ids = odometry.run(oracle, "small", frames, depths, poses, c, desired_points=600, desired_immature=500)
    assert len(ids) == 60 and not any(l.get("lost") for l in od.log)
    kfs = [l["frame"] for l in od.log if l.get("kf")]
    assert len(kfs) >= 2
    ate, scale = odometry.ate_rmse(est, gt)
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    assert ate < 0.03 * path, (ate, path)               # 160x128 images: coarse, but it is tracking the trajectory
    assert abs(scale - 1) < 0.05


@pytest.mark.gpu
def test_gpu_closed_loop_matches_oracle(gpu, oracle):
    c, frames, depths, poses = _seq("c1", 130)
    og, eg, gt, ids_g = odometry.run(gpu, "c1", frames, depths, poses, c)
    oo, eo, _, ids_o = odometry.run(oracle, "c1", frames, depths, poses, c)
    assert ids_g == ids_o and len(ids_g) == 130
    kf_g = [l["frame"] for l in og.log if l.get("kf")]
    kf_o = [l["frame"] for l in oo.log if l.get("kf")]
    # the keyframe test is a threshold on a continuous score (FullSystem.cc:160-176): a 1e-4 difference in the flow
    # indicators may move a keyframe by a frame or two, after which the two runs are different (equally valid) systems
    assert kf_g[0] == kf_o[0] and abs(len(kf_g) - len(kf_o)) <= 1 and len(kf_g) >= 4, (kf_g, kf_o)
    assert all(abs(a - b) <= 6 for a, b in zip(kf_g, kf_o)), (kf_g, kf_o)
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    ate_g, sg = odometry.ate_rmse(eg, gt)
    ate_o, so = odometry.ate_rmse(eo, gt)
    print("ATE gpu %.3e oracle %.3e path %.3f max|gpu-oracle| %.3e keyframes %s %s" % (ate_g, ate_o, path, np.abs(eg - eo).max(), kf_g, kf_o))
    assert ate_g < 0.01 * path and ate_o < 0.01 * path, (ate_g, ate_o, path)
    assert abs(ate_g - ate_o) < max(0.1 * ate_o, 2e-4), (ate_g, ate_o)
    assert np.abs(eg - eo).max() < 2e-3                  # the two trajectories themselves stay together
    # up to the first keyframe where the decisions differ the two runs are the same system: same point counts
    for lg, lo in zip([l for l in og.log if l.get("kf")], [l for l in oo.log if l.get("kf")]):
        if lg["frame"] != lo["frame"]:
            break
        assert abs(lg["points"] - lo["points"]) <= 0.02 * lo["points"] + 5


def test_oracle_closed_loop_from_the_initializer(oracle):
    """no ground-truth depth anywhere: CoarseInitializer (N1) bootstraps the map, so the trajectory is up to scale"""
    c = util.CALIB["c1"]
    frames, depths, poses = synth.make_sequence(1001, 45, start=0, **c)
    od, est, gt, ids = odometry.run(oracle, "c1", frames, depths, poses, c, bootstrap="init")
    done = [l for l in od.log if l.get("init_done")]
    assert len(done) == 1 and done[0]["frame"] <= 12 and done[0]["init_points"] > 1500
    assert not any(l.get("lost") for l in od.log) and len(ids) >= 45 - done[0]["frame"]
    ate, scale = odometry.ate_rmse(est, gt)
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    assert ate < 0.01 * path, (ate, path)


@pytest.mark.gpu
def test_gpu_closed_loop_from_the_initializer_matches_oracle(gpu, oracle):
    c = util.CALIB["c1"]
    frames, depths, poses = synth.make_sequence(1001, 70, start=0, **c)
    og, eg, gt, ids_g = odometry.run(gpu, "c1", frames, depths, poses, c, bootstrap="init")
    oo, eo, _, ids_o = odometry.run(oracle, "c1", frames, depths, poses, c, bootstrap="init")
    assert ids_g == ids_o
    dg, do = [l for l in og.log if l.get("init_done")][0], [l for l in oo.log if l.get("init_done")][0]
    assert dg["frame"] == do["frame"] and dg["init_points"] == do["init_points"] and abs(dg["rescale"] / do["rescale"] - 1) < 1e-3
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    ate_g, sg = odometry.ate_rmse(eg, gt)
    ate_o, so = odometry.ate_rmse(eo, gt)
    print("ATE gpu %.3e oracle %.3e path %.3f scale %.3f %.3f" % (ate_g, ate_o, path, sg, so))
    assert ate_g < 0.01 * path and ate_o < 0.01 * path
    assert abs(sg / so - 1) < 2e-2 and abs(ate_g - ate_o) < max(0.2 * ate_o, 3e-4)


@functools.lru_cache(maxsize=2)
def _c3_sequence(n):
    c = util.CALIB["c3"]
    return (c,) + tuple(synth.make_sequence(1001, n, start=0, **c))


def test_oracle_closed_loop_config3_shape_with_corner_detector(oracle):
    """BASELINE configs[2] shape: 1232x368 (KITTI crop, 5 pyramid levels), setting_pointSelection = 1 (new immature points
    from FeatureDetector::DetectCorners, N3), CoarseInitializer bootstrap.  Loop closing (DBoW3 / g2o) is host code outside
    the path and not available in this image."""
    c, frames, depths, poses = _c3_sequence(48)
    od, est, gt, ids = odometry.run(oracle, "c3", frames, depths, poses, c, bootstrap="init", point_selection=1)
    assert not any(l.get("lost") for l in od.log)
    kfs = [l for l in od.log if l.get("kf")]
    assert len(kfs) >= 2 and all(l["new_immature"] > 800 for l in kfs)
    ate, _ = odometry.ate_rmse(est, gt)
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    assert ate < 0.02 * path, (ate, path)


@pytest.mark.gpu
def test_gpu_closed_loop_config3_shape_matches_oracle(gpu, oracle):
    c, frames, depths, poses = _c3_sequence(48)
    og, eg, gt, ids_g = odometry.run(gpu, "c3", frames, depths, poses, c, bootstrap="init", point_selection=1)
    oo, eo, _, ids_o = odometry.run(oracle, "c3", frames, depths, poses, c, bootstrap="init", point_selection=1)
    assert ids_g == ids_o
    kg, ko = [l for l in og.log if l.get("kf")], [l for l in oo.log if l.get("kf")]
    assert kg[0]["frame"] == ko[0]["frame"] and kg[0]["new_immature"] == ko[0]["new_immature"]
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    ate_g, sg = odometry.ate_rmse(eg, gt)
    ate_o, so = odometry.ate_rmse(eo, gt)
    print("c3 ATE gpu %.3e oracle %.3e path %.3f scale %.3f %.3f" % (ate_g, ate_o, path, sg, so))
    assert ate_g < 0.02 * path and ate_o < 0.02 * path
    assert abs(sg / so - 1) < 3e-2


@pytest.mark.gpu
def test_gpu_closed_loop_config4_dense_matches_oracle(gpu, oracle):
    """BASELINE configs[3] shape: 752x480, setting_desiredPointDensity = 8000, desiredImmatureDensity = 6000, window of 10"""
    c = util.CALIB["c4"]
    frames, depths, poses = synth.make_sequence(1001, 70, start=0, **c)
    kw = dict(bootstrap="init", desired_points=8000, desired_immature=6000, max_frames=10, max_immature=80000)
    og, eg, gt, ids_g = odometry.run(gpu, "c4", frames, depths, poses, c, **kw)
    oo, eo, _, ids_o = odometry.run(oracle, "c4", frames, depths, poses, c, **kw)
    assert ids_g == ids_o
    kg, ko = [l for l in og.log if l.get("kf")], [l for l in oo.log if l.get("kf")]
    assert kg[0]["frame"] == ko[0]["frame"] and kg[0]["init_points"] == ko[0]["init_points"] > 7000
    assert abs(kg[0]["points"] - ko[0]["points"]) <= 0.01 * ko[0]["points"]
    path = np.linalg.norm(np.diff(gt, axis=0), axis=1).sum()
    ate_g, sg = odometry.ate_rmse(eg, gt)
    ate_o, so = odometry.ate_rmse(eo, gt)
    print("c4 ATE gpu %.3e oracle %.3e path %.3f scale %.3f %.3f" % (ate_g, ate_o, path, sg, so))
    assert ate_g < 0.01 * path and ate_o < 0.01 * path
    assert abs(sg / so - 1) < 2e-2
