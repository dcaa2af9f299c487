"""GPU parity tests (K1, K2, K3, K4, trackNewestCoarse): CUDA path through the C ABI vs the CPU oracle
on identical seeded inputs.  Bars: bit-exact pyramids / point-cloud indexing; 1e-4
relative on residual statistics, H, b and pose increments."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])


@pytest.mark.parametrize("calib", ["c1", "c3", "c4", "small"])
def test_k1_make_images_bit_exact(gpu, oracle, calib):
    g = util.make_ctx(gpu, calib)
    o = util.make_ctx(oracle, calib)
    assert g.levels == o.levels and g.sizes == o.sizes
    img = util.frames(calib, 1001, (40, 42))[0][0]
    g.make_images(2, img)
    o.make_images(2, img)
    for lvl in range(g.levels):
        a, b = g.download(2, lvl, 0), o.download(2, lvl, 0)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"dIp lvl {lvl}"
        a, b = g.download(2, lvl, 1), o.download(2, lvl, 1)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"absSquaredGrad lvl {lvl}"
    g.close(); o.close()


def test_k1_random_noise_and_gamma_bit_exact(gpu, oracle):
    # worst case for the border/wrap logic and the NaN / |d|>255 clamps: white noise incl. out-of-range values
    g = util.make_ctx(gpu, "c1")
    o = util.make_ctx(oracle, "c1")
    rng = np.random.default_rng(5)
    img = rng.uniform(-100, 700, (480, 640)).astype(np.float32)
    img[100, 100] = np.nan
    B = (np.arange(256, dtype=np.float32) ** 1.05).astype(np.float32)
    g.set_gamma(B); o.set_gamma(B)
    g.make_images(0, img); o.make_images(0, img)
    for lvl in range(g.levels):
        for what in (0, 1):
            a, b = g.download(0, lvl, what), o.download(0, lvl, what)
            nan = np.isnan(b)   # NaN payload bits are implementation-defined; positions must agree
            assert np.array_equal(np.isnan(a), nan), (lvl, what)
            assert np.array_equal(a.view(np.uint32)[~nan], b.view(np.uint32)[~nan]), (lvl, what)
    g.close(); o.close()


@pytest.mark.parametrize("calib,npts", [("c1", 2000), ("c4", 8000), ("c3", 2000), ("small", 300)])
def test_k2_coarse_depth_indexing_bit_exact(gpu, oracle, calib, npts):
    g = util.make_ctx(gpu, calib)
    o = util.make_ctx(oracle, calib)
    util.setup_tracker(g, calib, npts)
    util.setup_tracker(o, calib, npts)
    for lvl in range(g.levels):
        gu, gv, gi, gc = g.tracker_get_pc(0, lvl)
        ou, ov, oi, oc = o.tracker_get_pc(0, lvl)
        assert len(gu) == len(ou), f"pc_n[{lvl}]"
        assert np.array_equal(gu, ou) and np.array_equal(gv, ov)
        assert np.array_equal(gc, oc)
        assert np.allclose(gi, oi, rtol=1e-6, atol=0)   # atomicAdd order at >=3-way pixel collisions only
    g.close(); o.close()


def test_k2_empty_point_list(gpu, oracle):
    g = util.make_ctx(gpu, "small"); o = util.make_ctx(oracle, "small")
    img = util.frames("small")[0][0]
    for c in (g, o):
        c.make_images(0, img)
        c.tracker_set_ref(0, 0, [], [], [], [])
        assert all(len(c.tracker_get_pc(0, l)[0]) == 0 for l in range(c.levels))
    g.close(); o.close()


def _poses(gt):
    rng = np.random.default_rng(11)
    out = [IDENT.copy(), gt.copy()]
    p = gt.copy(); p[9:] += rng.normal(0, 0.01, 3); out.append(p)
    return out


@pytest.mark.parametrize("calib,npts", [("c1", 2000), ("c4", 8000)])
def test_k3_k4_residual_and_hessian_parity(gpu, oracle, calib, npts):
    g = util.make_ctx(gpu, calib); o = util.make_ctx(oracle, calib)
    gt = util.setup_tracker(g, calib, npts)
    util.setup_tracker(o, calib, npts)
    for pose in _poses(gt):
        for lvl in range(g.levels):
            for aff in ((0.0, 0.0), (0.02, -3.0)):
                for cutoff in (20.0, 40.0):
                    r_g = g.tracker_calc_res(0, 1, lvl, pose, aff[0], aff[1], cutoff)
                    r_o = o.tracker_calc_res(0, 1, lvl, pose, aff[0], aff[1], cutoff)
                    n = max(r_o[1], 1.0)
                    assert abs(r_g[1] - r_o[1]) <= max(1, 0.005 * n), (lvl, r_g, r_o)       # counts: <=0.5 %
                    assert abs(r_g[0] - r_o[0]) <= 1e-4 * abs(r_o[0]) + 0.005 * n * 400 * (r_g[1] != r_o[1]), (lvl, r_g, r_o)
                    assert np.allclose(r_g[2:5], r_o[2:5], rtol=1e-4, atol=1e-6)
                    assert abs(r_g[5] - r_o[5]) < 5e-3
                    H_g, b_g = g.tracker_calc_gs(0, lvl, pose, aff[0], aff[1])
                    H_o, b_o = o.tracker_calc_gs(0, lvl, pose, aff[0], aff[1])
                    if r_g[1] == r_o[1]:
                        assert util.rel_err(H_g, H_o) < 1e-4, (lvl, util.rel_err(H_g, H_o))
                        assert util.rel_err(b_g, b_o) < 1e-4 + 1e-6 * np.linalg.norm(H_o) / max(np.linalg.norm(b_o), 1e-9), (lvl,)
    g.close(); o.close()


@pytest.mark.parametrize("calib,npts,idx", [("c1", 2000, (40, 42)), ("c1", 2000, (60, 65)), ("c4", 8000, (40, 42)),
                                            ("c3", 2000, (40, 41))])
def test_track_newest_coarse_parity(gpu, oracle, calib, npts, idx):
    g = util.make_ctx(gpu, calib); o = util.make_ctx(oracle, calib)
    gt = util.setup_tracker(g, calib, npts, idx)
    util.setup_tracker(o, calib, npts, idx)
    rg = g.tracker_track(0, 1, IDENT, [0, 0])
    ro = o.tracker_track(0, 1, IDENT, [0, 0])
    assert rg["ok"] == ro["ok"] == True
    # pose result: within 1e-4 of the oracle's (relative to the motion magnitude), and both at ground truth
    scale = max(np.linalg.norm(ro["pose"][9:]), 1e-3)
    assert np.linalg.norm(rg["pose"][9:] - ro["pose"][9:]) / scale < 1e-4, (rg["pose"][9:], ro["pose"][9:])   # 1e-4
    assert np.abs(rg["pose"][:9] - ro["pose"][:9]).max() < 1e-4
    assert np.allclose(rg["last_residuals"], ro["last_residuals"], rtol=1e-4, equal_nan=True)
    assert np.abs(rg["pose"][9:] - gt[9:]).max() < 1e-2   # sanity vs ground truth (both implementations)
    g.close(); o.close()


def test_track_single_iteration_increment_parity(gpu, oracle):
    """Per-iteration pose increment: one LM step from the same state, compared at 1e-4."""
    import numpy.linalg as la
    g = util.make_ctx(gpu, "c1"); o = util.make_ctx(oracle, "c1")
    util.setup_tracker(g, "c1", 2000); util.setup_tracker(o, "c1", 2000)
    for lvl in range(g.levels):
        incs = []
        for c in (g, o):
            c.tracker_calc_res(0, 1, lvl, IDENT, 0.0, 0.0, 20.0)
            H, b = c.tracker_calc_gs(0, lvl, IDENT, 0.0, 0.0)
            Hl = H.copy(); Hl[np.diag_indices(8)] *= 1.01
            incs.append(la.solve(Hl, -b))
        assert util.rel_err(incs[0], incs[1]) < 1e-4, (lvl, incs)
    g.close(); o.close()


def test_track_failure_signals(gpu, oracle):
    """minResForAbort and lost-tracking behave as in the reference (return false, NaN residuals above the abort level)."""
    g = util.make_ctx(gpu, "c1"); o = util.make_ctx(oracle, "c1")
    util.setup_tracker(g, "c1", 2000); util.setup_tracker(o, "c1", 2000)
    for c in (g, o):
        r = c.tracker_track(0, 1, IDENT, [0, 0], min_res=[0.01] * 5)
        assert not r["ok"]
        assert np.isfinite(r["last_residuals"][c.levels - 1]) and np.isnan(r["last_residuals"][0])
    g.close(); o.close()


def test_two_contexts_are_independent(gpu, oracle):
    # BASELINE config 5: many sequences in one process, no shared device state
    a = util.make_ctx(gpu, "small"); b = util.make_ctx(gpu, "small"); o = util.make_ctx(oracle, "small")
    f = util.frames("small", 1001, (40, 42)); f2 = util.frames("small", 1002, (40, 42))
    a.make_images(0, f[0][0]); b.make_images(0, f2[0][0]); a.make_images(1, f[1][0])
    o.make_images(0, f[0][0])
    assert np.array_equal(a.download(0, 1, 0), o.download(0, 1, 0))
    o.make_images(0, f2[0][0])
    assert np.array_equal(b.download(0, 1, 0), o.download(0, 1, 0))
    a.close(); b.close(); o.close()


@pytest.mark.parametrize("calib,npts,idx", [("c1", 2000, (40, 42)), ("c1", 2000, (60, 65)), ("c4", 8000, (40, 42)), ("small", 600, (40, 42))])
def test_device_resident_lm_matches_host_driven_loop(gpu, calib, npts, idx):
    """tracker_lm_mode 0 (one cooperative launch per level, LM state machine on the device) against mode 1 (host loop,
    one launch per evaluation): same accept/reject path, same pose up to the device's fp64 sin/cos/exp rounding."""
    res = []
    for mode in (0, 1):
        c = util.make_ctx(gpu, calib, tracker_lm_mode=mode)
        util.setup_tracker(c, calib, npts, idx)
        l0 = c.launch_count()
        r = c.tracker_track(0, 1, IDENT, [0, 0])
        res.append((r, c.launch_count() - l0))
        c.close()
    (rd, ld), (rh, lh) = res
    assert rd["ok"] == rh["ok"] == True
    assert np.array_equal(rd["iters"], rh["iters"]), (rd["iters"], rh["iters"])      # evaluations, accepted steps
    # device expf / fp64 sin, cos differ from glibc by <= 1 ulp; through the affine brightness term that moves the pose by ~1e-7
    assert np.abs(rd["pose"] - rh["pose"]).max() < 2e-6
    assert np.allclose(rd["last_residuals"], rh["last_residuals"], rtol=1e-4, equal_nan=True)
    assert np.allclose(rd["flow"], rh["flow"], rtol=1e-4)
    assert lh == rh["iters"][0] and ld <= 6 and ld < lh                                # one launch per level (+ a repeat)
