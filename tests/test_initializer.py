"""N1 — CoarseInitializer (reference src/frontend/CoarseInitializer.cc): oracle behaviour against ground truth, the exact
kNN graph against scipy, and the CUDA path against the oracle."""
import numpy as np
import pytest

from ldso_b200 import synth
from tests import util


def sequence(calib, n, start=0):
    c = util.CALIB[calib]
    return synth.make_sequence(1001, n, start=start, **c)


def run_init(ctx, frames, max_frames):
    ctx.make_images(0, frames[0])
    n = ctx.init_set_first(0)
    hist = []
    for i in range(1, max_frames):
        ctx.make_images(1, frames[i])
        r = ctx.init_track_frame(1)
        hist.append(r)
        if r["ready"]:
            break
    return n, hist


def test_knn_graph_matches_scipy(oracle):
    from scipy.spatial import cKDTree
    frames, _, _ = sequence("small", 1)
    ctx = util.make_ctx(oracle, "small", n_frame_slots=3)
    ctx.make_images(0, frames[0])
    n = ctx.init_set_first(0)
    assert n[0] > 100
    for lvl in range(ctx.levels):
        P = ctx.init_get_points(lvl)
        nb, par = ctx.init_get_graph(lvl)
        xy = np.stack([P["u"], P["v"]], 1).astype(np.float64)
        k = min(10, len(xy))
        d, _ = cKDTree(xy).query(xy, k=k)
        mine = np.sqrt(((xy[nb[:, :k]] - xy[:, None, :]) ** 2).sum(-1))
        assert np.allclose(np.sort(mine, 1), d, atol=1e-5)        # same distance multiset; ties may pick another index
        assert np.array_equal(nb[:, 0], np.arange(len(xy)))       # a point is its own nearest neighbour
        assert np.all(np.diff(mine, axis=1) >= -1e-6)             # ascending
        if lvl + 1 < ctx.levels:
            Pu = ctx.init_get_points(lvl + 1)
            xyu = np.stack([Pu["u"], Pu["v"]], 1).astype(np.float64)
            q = xy * 0.5 - 0.25
            dpar, _ = cKDTree(xyu).query(q, k=1)
            assert np.allclose(np.sqrt(((xyu[par] - q) ** 2).sum(-1)), dpar, atol=1e-5)
        else:
            assert np.all(par == -1)
    ctx.close()


def test_oracle_initializer_snaps_and_recovers_depth(oracle):
    """sideways motion over a textured room: the initializer snaps, reports ready 5 frames later, its translation
    points along the true baseline and its inverse depths are the true ones up to ONE global scale"""
    frames, depths, poses = sequence("small", 16)
    ctx = util.make_ctx(oracle, "small", n_frame_slots=3)
    n, hist = run_init(ctx, frames, 16)
    assert hist[-1]["ready"] and hist[-1]["snapped"]
    first_snapped = next(i for i, r in enumerate(hist) if r["snapped"])
    assert len(hist) - 1 == first_snapped + 6 or len(hist) - 1 == first_snapped + 5
    i = len(hist)
    gt = synth.rel_pose(poses[0][0], poses[0][1], poses[i][0], poses[i][1])
    t_est, t_gt = hist[-1]["pose"][9:], gt[9:]
    assert np.dot(t_est, t_gt) / (np.linalg.norm(t_est) * np.linalg.norm(t_gt)) > 0.99
    P = ctx.init_get_points(0)
    g = P["isGood"]
    assert g.mean() > 0.9
    d = depths[0][P["v"].astype(int), P["u"].astype(int)]
    prod = P["iR"][g] * d[g]
    scale = np.median(prod)
    assert np.median(np.abs(prod / scale - 1)) < 0.05
    assert abs(scale * np.linalg.norm(t_est) / np.linalg.norm(t_gt) - 1) < 0.3      # the same scale explains the baseline (160x128: coarse)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib", ["small", "c1"])
def test_gpu_set_first_bit_exact(gpu, oracle, calib):
    frames, _, _ = sequence(calib, 1)
    g, o = util.make_ctx(gpu, calib, n_frame_slots=3), util.make_ctx(oracle, calib, n_frame_slots=3)
    out = []
    for ctx in (g, o):
        ctx.make_images(0, frames[0])
        n = ctx.init_set_first(0)
        lv = []
        for lvl in range(ctx.levels):
            P = ctx.init_get_points(lvl)
            nb, par = ctx.init_get_graph(lvl)
            lv.append((P["u"], P["v"], P["type"], P["idepth"], P["iR"], P["isGood"], nb, par))
        out.append((n, lv))
    assert np.array_equal(out[0][0], out[1][0]) and out[0][0][0] > 100
    for a, b in zip(out[0][1], out[1][1]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    g.close(); o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib", ["small", "c1"])
def test_gpu_calc_res_and_gs_matches_oracle(gpu, oracle, calib):
    """one calcResAndGS per level on the fresh point set: H, b, Schur blocks and energies within 1e-4 (fp32 sums in a
    different order); the point set is bit-identical, so the per-point terms are"""
    frames, _, poses = sequence(calib, 6)
    g, o = util.make_ctx(gpu, calib, n_frame_slots=3), util.make_ctx(oracle, calib, n_frame_slots=3)
    rel = synth.rel_pose(poses[0][0], poses[0][1], poses[5][0], poses[5][1])
    rel[9:] *= 0.2                                             # the initializer's scale: unit inverse depth
    res = []
    for ctx in (g, o):
        ctx.make_images(0, frames[0]); ctx.make_images(1, frames[5])
        ctx.init_set_first(0)
        res.append([ctx.init_calc_res_and_gs(1, lvl, rel, (0.01, 0.5)) for lvl in range(ctx.levels)])
    for a, b in zip(*res):
        for x, y in zip(a[:4], b[:4]):
            assert util.rel_err(x, y) < 1e-4
        assert np.allclose(a[4], b[4], rtol=1e-4)
        assert b[4][0] > 0 and np.abs(b[2]).max() > 0      # points took part: a photometric energy and a Schur complement exist
    g.close(); o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib", ["small", "c1"])
def test_gpu_track_frame_matches_oracle(gpu, oracle, calib):
    """the whole initialisation: same snap frame, same ready frame, poses and regularised inverse depths within
    tolerance (LM accept decisions compare fp32 energies summed in a different order, so trajectories may differ in the
    last digits but not in outcome)"""
    frames, depths, poses = sequence(calib, 16)
    g, o = util.make_ctx(gpu, calib, n_frame_slots=3), util.make_ctx(oracle, calib, n_frame_slots=3)
    (ng, hg), (no, ho) = run_init(g, frames, 16), run_init(o, frames, 16)
    assert np.array_equal(ng, no)
    assert len(hg) == len(ho) and hg[-1]["ready"] and ho[-1]["ready"]
    assert [r["snapped"] for r in hg] == [r["snapped"] for r in ho]
    for a, b in zip(hg, ho):
        assert util.rel_err(a["pose"], b["pose"]) < 2e-3
        assert abs(a["res"][0] / b["res"][0] - 1) < 2e-2
    Pg, Po = g.init_get_points(0), o.init_get_points(0)
    assert (Pg["isGood"] == Po["isGood"]).mean() > 0.995
    both = Pg["isGood"] & Po["isGood"]
    assert np.median(np.abs(Pg["iR"][both] / Po["iR"][both] - 1)) < 1e-3
    assert g.launch_count() > 50
    g.close(); o.close()
