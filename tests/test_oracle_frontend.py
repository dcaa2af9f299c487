"""CPU tests of the oracle's front-end restatement (no GPU): structural invariants of the reference
(the reference ships no tests, so these pin the restatement against the source's own
statements and against ground truth of the synthetic scene)."""
import numpy as np

from tests import util


def test_pyramid_level_rule(oracle):
    # src/internal/GlobalCalib.cc:20-28 — SURVEY.md trap T1/T2
    for calib, lv, coarsest in (("c1", 4, (80, 60)), ("c3", 5, (77, 23)), ("c4", 5, (47, 30)), ("small", 3, (40, 32))):
        ctx = util.make_ctx(oracle, calib)
        assert ctx.levels == lv
        assert ctx.sizes[-1] == coarsest
        ctx.close()
    cfg = oracle.default_config(w=1241, h=376)
    ctx = oracle.create(cfg)
    assert ctx.levels == 1   # odd width: no halving at all
    ctx.close()


def test_make_images_matches_numpy_restatement(oracle):
    # independent numpy statement of FrameHessian.cc:59-115 (including the flat-index wrap of columns 0 / w-1)
    ctx = util.make_ctx(oracle, "small")
    rng = np.random.default_rng(0)
    w, h = ctx.sizes[0]
    img = np.floor(rng.uniform(0, 256, (h, w))).astype(np.float32)
    ctx.make_images(0, img)
    I = img.copy()
    for lvl in range(ctx.levels):
        wl, hl = ctx.sizes[lvl]
        if lvl > 0:
            I = np.float32(0.25) * (((I[0::2, 0::2] + I[0::2, 1::2]) + I[1::2, 0::2]) + I[1::2, 1::2])
        flat = I.reshape(-1)
        dx = np.zeros(wl * hl, np.float32); dy = np.zeros(wl * hl, np.float32)
        idx = np.arange(wl, wl * (hl - 1))
        dx[idx] = np.float32(0.5) * (flat[idx + 1] - flat[idx - 1])
        dy[idx] = np.float32(0.5) * (flat[idx + wl] - flat[idx - wl])
        got = ctx.download(0, lvl, 0).reshape(-1, 3)
        ag = ctx.download(0, lvl, 1).reshape(-1)
        assert np.array_equal(got[:, 0], flat)
        assert np.array_equal(got[:, 1], dx)
        assert np.array_equal(got[:, 2], dy)
        assert np.array_equal(ag, dx * dx + dy * dy)   # identity gamma: gw = 1
    ctx.close()


def test_gamma_weight(oracle):
    ctx = util.make_ctx(oracle, "small")
    B = (np.arange(256, dtype=np.float32) ** 1.1).astype(np.float32)
    ctx.set_gamma(B)
    rng = np.random.default_rng(1)
    w, h = ctx.sizes[0]
    img = np.floor(rng.uniform(0, 256, (h, w))).astype(np.float32)
    ctx.make_images(0, img)
    d = ctx.download(0, 0, 0).reshape(-1, 3)
    ag = ctx.download(0, 0, 1).reshape(-1)
    c = np.clip((d[:, 0] + np.float32(0.5)).astype(np.int32), 5, 250)
    gw = B[c + 1] - B[c]
    assert np.array_equal(ag, (d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]) * (gw * gw))
    ctx.close()


def test_coarse_depth_invariants(oracle):
    ctx = util.make_ctx(oracle, "c1")
    util.setup_tracker(ctx, "c1", n_points=2000)
    for lvl in range(ctx.levels):
        u, v, idp, col = ctx.tracker_get_pc(0, lvl)
        w, h = ctx.sizes[lvl]
        assert len(u) > 0
        assert u.min() >= 2 and u.max() <= w - 3 and v.min() >= 2 and v.max() <= h - 3   # CoarseTracker.cc:513-514
        key = v.astype(np.int64) * w + u.astype(np.int64)
        assert np.all(np.diff(key) > 0)        # row-major compaction order
        assert np.all(idp > 0) and np.all(np.isfinite(col))
    n0 = len(ctx.tracker_get_pc(0, 0)[0])
    assert 2000 <= n0 <= 5 * 2000 + 2000       # each point + up to 4 diagonal neighbours
    ctx.close()


def test_tracker_recovers_ground_truth_pose(oracle):
    # end-to-end sanity of K2+K3+K4+LM: from the identity guess the oracle must find the true relative pose
    ctx = util.make_ctx(oracle, "c1")
    gt = util.setup_tracker(ctx, "c1", n_points=2000, idx=(40, 42))
    ident = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
    r = ctx.tracker_track(0, 1, ident, [0, 0])
    assert r["ok"]
    assert np.abs(r["pose"][9:] - gt[9:]).max() < 2e-3, (r["pose"][9:], gt[9:])
    assert np.abs(r["pose"][:9] - gt[:9]).max() < 1e-3
    assert r["last_residuals"][0] < 6.0
    # a and b trade off against each other around the mean brightness; their net effect must be ~identity
    assert abs(np.exp(r["aff"][0]) * 128 + r["aff"][1] - 128) < 2.0
    ctx.close()


def test_accumulator_count_padding(oracle):
    # trap T7: H,b are divided by the padded warped count.  H[7][7] = sum(w * 1 * 1) / n_pad * SCALE_B^2
    ctx = util.make_ctx(oracle, "c1")
    gt = util.setup_tracker(ctx, "c1", n_points=1500)
    out6 = ctx.tracker_calc_res(0, 1, 0, gt, 0.0, 0.0, 20.0)
    H, b = ctx.tracker_calc_gs(0, 0, gt, 0.0, 0.0)
    assert np.allclose(H, H.T)
    assert 0 < H[7, 7] / 1e6 <= 1.0 + 1e-6      # mean huber weight <= 1
    assert out6[1] > 0 and 0 <= out6[5] < 0.6
    ctx.close()
