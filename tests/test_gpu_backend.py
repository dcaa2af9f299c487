"""GPU parity tests of the windowed BA path (K7 linearize, K8 top Hessian, K9 Schur complement, K10 resubstitute,
H1 stitch + solve): CUDA through the C ABI vs the CPU oracle on the same frozen synthetic window."""
import numpy as np
import pytest

from ldso_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu


def load(api, W, calib="c1"):
    ctx = util.make_ctx(api, calib, max_points=max(4096, len(W["u"]) + 16), max_residuals=max(32768, len(W["target"]) + 16),
                        n_frame_slots=max(10, W["n"]))
    for i, img in enumerate(W["imgs"]):
        ctx.make_images(i, img)
    ctx.window_load(W)
    return ctx


@pytest.fixture(scope="module")
def window():
    return synth.make_window()            # 7 keyframes, ~300 points each, ~12 k residuals (BASELINE config 2 shape)


@pytest.fixture(scope="module")
def pair(gpu, oracle, window):
    g, o = load(gpu, window), load(oracle, window)
    yield g, o
    g.close(); o.close()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_k7_linearize_bit_exact(pair, window):
    g, o = pair
    sg, thg = g.window_linearize(False)
    so, tho = o.window_linearize(False)
    ns_g, ns_o = g.window_residual_ints(1), o.window_residual_ints(1)
    assert np.array_equal(ns_g, ns_o)                                   # IN / OOB / OUTLIER tallies identical
    assert sg[2] == so[2] and abs(sg[0] - so[0]) <= 1e-9 * abs(so[0])
    assert thg == tho                                                   # setNewFrameEnergyTH
    act = ((window["flags"] & 1) == 0) & (ns_o != 1)
    Jg, Jo = g.window_residual_floats(0), o.window_residual_floats(0)
    assert np.array_equal(bits(Jg[act]), bits(Jo[act]))                 # RawResidualJacobian, all 72 floats
    for what in (3, 4, 5):                                              # centerProjectedTo, energies
        a, b = g.window_residual_floats(what), o.window_residual_floats(what)
        assert np.array_equal(bits(a[act]), bits(b[act])), what
    assert np.bincount(ns_o[(window["flags"] & 1) == 0], minlength=3)[0] > 0.6 * act.sum()


def test_apply_res_and_take_data(pair, window):
    g, o = pair
    g.window_apply_res(); o.window_apply_res()
    assert np.array_equal(g.window_residual_ints(0), o.window_residual_ints(0))
    fg, fo = g.window_residual_ints(2), o.window_residual_ints(2)
    assert np.array_equal(fg, fo)
    act = (fo & 2) != 0
    assert np.array_equal(bits(g.window_residual_floats(2)[act]), bits(o.window_residual_floats(2)[act]))   # JpJdF


def test_k8_k9_accumulate_parity(pair, window):
    g, o = pair
    ag, ao = g.window_accumulate(), o.window_accumulate()
    assert np.array_equal(ag["counts"], ao["counts"])
    for k in ("HA", "bA", "HL", "bL", "Hsc", "bsc"):
        assert util.rel_err(ag[k], ao[k]) < 1e-4, (k, util.rel_err(ag[k], ao[k]))
    # block-wise (8x8 pose blocks), as SURVEY.md §8d asks: Frobenius per block
    n = window["n"]
    for k in ("HA", "HL", "Hsc"):
        for h in range(n):
            for t in range(n):
                bg = ag[k][4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t]; bo = ao[k][4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t]
                assert np.linalg.norm(bg - bo) <= 1e-4 * np.linalg.norm(bo) + 1e-7 * np.linalg.norm(ao[k]), (k, h, t)
    # per-point scalars are summed in residual order on both sides: bit-exact
    for what in (0, 1, 2, 3, 4, 5, 6, 7, 8):
        assert np.array_equal(bits(g.window_point_floats(what)), bits(o.window_point_floats(what))), what
    pair[0].acc, pair[1].acc = ag, ao


def test_h1_solve_and_k10_resubstitute(pair, window):
    g, o = pair
    # the marginalisation prior fixes the 7 gauge directions; without a firm prior the system's condition number
    # (~1e5 after diagonal scaling) amplifies the 1e-6 accumulation-order differences of H beyond any fixed bound
    HMs = window["HM"] * 1e4
    xg = g.window_solve(0, 1e-5, g.acc, HMs, window["bM"], window["delta"])
    xo = o.window_solve(0, 1e-5, o.acc, HMs, window["bM"], window["delta"])
    assert util.rel_err(xg, xo) < 1e-4, util.rel_err(xg, xo)
    xo = o.window_solve(0, 1e-5, o.acc, window["HM"], window["bM"], window["delta"])
    xs = g.window_solve(0, 1e-5, o.acc, window["HM"], window["bM"], window["delta"])      # same inputs -> same host algebra
    assert util.rel_err(xs, xo) < 1e-10
    N = np.random.default_rng(0).normal(size=(8 * window["n"] + 4, 7))
    assert util.rel_err(g.window_solve(3, 1e-5, o.acc, window["HM"], window["bM"], window["delta"], N),
                        o.window_solve(3, 1e-5, o.acc, window["HM"], window["bM"], window["delta"], N)) < 1e-9
    sg, so = g.window_resubstitute(xo, 1e-5), o.window_resubstitute(xo, 1e-5)
    assert np.array_equal(bits(sg), bits(so))        # same x, same per-point inputs, same order: bit-exact
    assert np.abs(so).max() > 0


def test_k7_fix_linearization_and_new_residual_bookkeeping(gpu, oracle, window):
    g, o = load(gpu, window), load(oracle, window)
    for c in (g, o):
        c.window_linearize(True)
    assert np.array_equal(g.window_residual_ints(2), o.window_residual_ints(2))
    assert np.array_equal(g.window_residual_ints(0), o.window_residual_ints(0))
    assert np.array_equal(bits(g.window_point_floats(9)), bits(o.window_point_floats(9)))      # maxRelBaseline
    assert np.array_equal(g.window_point_floats(10), o.window_point_floats(10))                # numGoodResiduals
    assert o.window_point_floats(10).sum() > 0
    fl = o.window_residual_ints(2)
    mask = ((fl & 2) != 0) & ((fl & 1) == 0) & (np.arange(len(fl)) % 3 == 0)
    g.window_fix_linearization(mask.astype(np.uint8)); o.window_fix_linearization(mask.astype(np.uint8))
    assert np.array_equal(bits(g.window_residual_floats(1)[mask]), bits(o.window_residual_floats(1)[mask]))   # res_toZeroF
    assert np.array_equal(g.window_residual_ints(2), o.window_residual_ints(2))
    ag, ao = g.window_accumulate(), o.window_accumulate()
    assert np.array_equal(ag["counts"], ao["counts"]) and ao["counts"][1] > (window["flags"] & 1).sum() * 0.5
    for k in ("HA", "bA", "HL", "bL", "Hsc", "bsc"):
        assert util.rel_err(ag[k], ao[k]) < 1e-4, k
    g.close(); o.close()


def test_gauss_newton_iterations_track_the_oracle(gpu, oracle, window):
    """three damped GN steps on point depths (frames frozen): energies and depth updates follow the oracle"""
    g, o = load(gpu, window), load(oracle, window)
    idp = {id(g): window["idepth"].copy(), id(o): window["idepth"].copy()}
    hist = {id(g): [], id(o): []}
    for it in range(3):
        for c in (g, o):
            st, _ = c.window_linearize(False)
            c.window_apply_res()
            acc = c.window_accumulate()
            x = c.window_solve(it, 1e-5, acc, window["HM"] * 1e6, window["bM"] * 0, window["delta"] * 0)
            step = c.window_resubstitute(x, 1e-5)
            idp[id(c)] = (idp[id(c)] + step).astype(np.float32)      # doStepFromBackup for the point depths
            c.window_set_idepth(idp[id(c)], window["idepth_zero"])
            hist[id(c)].append((st[0], st[2]))
    hg, ho = np.array(hist[id(g)]), np.array(hist[id(o)])
    assert np.allclose(hg[:, 0], ho[:, 0], rtol=1e-4)
    assert np.all(np.abs(hg[:, 1] - ho[:, 1]) <= 0.005 * ho[:, 1])
    assert util.rel_err(idp[id(g)] - window["idepth"], idp[id(o)] - window["idepth"]) < 1e-3
    assert ho[-1, 0] < ho[0, 0]                                         # the optimisation does reduce the energy
    g.close(); o.close()


def k7_bit_exact(g, o, W):
    """every float of RawResidualJacobian, centerProjectedTo and the energies of the residuals linearize() touched"""
    ns_g, ns_o = g.window_residual_ints(1), o.window_residual_ints(1)
    assert np.array_equal(ns_g, ns_o)
    act = ((W["flags"] & 1) == 0) & (ns_o != 1)
    assert act.sum() > 0.3 * len(act)
    for what in (0, 3, 4, 5):
        a, b = g.window_residual_floats(what), o.window_residual_floats(what)
        assert np.array_equal(bits(a[act]), bits(b[act])), what


@pytest.mark.parametrize("calib,frame_ids,pts,kw", [
    ("c4", tuple(range(40, 100, 6)), 800, dict(max_points=9000, max_residuals=80000, n_frame_slots=10)),   # configs[3]: 10 keyframes, 8000 points
    ("c3", tuple(range(40, 75, 5)), 300, dict(max_points=4096, max_residuals=32768, n_frame_slots=10)),    # configs[2]: 1232x368, 7 keyframes
])
def test_large_window_shapes(gpu, oracle, calib, frame_ids, pts, kw):
    """BASELINE configs[3] (752x480, 10 keyframes, 8000 points) and configs[2] (KITTI crop) windows: K7 bit-exact, states equal,
    stitched K8 / K9 systems at 1e-4"""
    W = synth.make_window(frame_ids=frame_ids, pts_per_frame=pts, **util.CALIB[calib])
    g = util.make_ctx(gpu, calib, **kw); o = util.make_ctx(oracle, calib, **kw)
    for c in (g, o):
        for i, img in enumerate(W["imgs"]):
            c.make_images(i, img)
        c.window_load(W)
        c.window_linearize(False)
    k7_bit_exact(g, o, W)
    for c in (g, o):
        c.window_apply_res()
    assert np.array_equal(g.window_residual_ints(0), o.window_residual_ints(0))
    ag, ao = g.window_accumulate(), o.window_accumulate()
    assert np.array_equal(ag["counts"], ao["counts"]) and ao["counts"][0] > 1000
    for k in ("HA", "bA", "HL", "bL", "Hsc", "bsc"):
        assert util.rel_err(ag[k], ao[k]) < 1e-4, (k, util.rel_err(ag[k], ao[k]))
    g.close(); o.close()


def linearised(api, W):
    ctx = load(api, W)
    ctx.window_linearize(True)
    fl = ctx.window_residual_ints(2)
    ctx.window_fix_linearization(((fl & 2) != 0).astype(np.uint8))
    return ctx


def test_marginalize_points_parity(gpu, oracle, window):
    """EnergyFunctional::marginalizePointsF: addPoint<2> + SC addPoint(p, false) over a subset of the points"""
    g, o = linearised(gpu, window), linearised(oracle, window)
    np_ = len(window["u"])
    for rnd, sel in enumerate((np.arange(np_) % 3 == 0, np.arange(np_) % 3 == 1)):      # twice: the second call sees removed points
        m = sel.astype(np.uint8)
        Hg, bg, ng = g.window_marginalize_points(m)
        Ho, bo, no = o.window_marginalize_points(m)
        assert ng == no and no > 500
        assert util.rel_err(Hg, Ho) < 1e-4 and util.rel_err(bg, bo) < 1e-4, (util.rel_err(Hg, Ho), util.rel_err(bg, bo))
        n = window["n"]
        for h in range(n):
            for t in range(n):
                a, b = Hg[4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t], Ho[4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t]
                assert np.linalg.norm(a - b) <= 1e-4 * np.linalg.norm(b) + 1e-7 * np.linalg.norm(Ho), (rnd, h, t)
        for what in (0, 1, 2, 3, 4, 5, 6, 7, 8):                                         # per-point fields of the marginalised points
            a, b = g.window_point_floats(what), o.window_point_floats(what)
            assert np.array_equal(bits(a[sel]), bits(b[sel])), (rnd, what)
        assert np.array_equal(g.window_residual_ints(2), o.window_residual_ints(2))       # removePoint
    # the remaining third still accumulates like the oracle's
    ag, ao = g.window_accumulate(), o.window_accumulate()
    assert np.array_equal(ag["counts"], ao["counts"]) and ao["counts"][1] > 500
    for k in ("HL", "bL", "Hsc", "bsc"):
        assert util.rel_err(ag[k], ao[k]) < 1e-4, k
    g.close(); o.close()


@pytest.mark.parametrize("n,idx", [(7, 0), (7, 4), (7, 6), (10, 2), (1, 0)])
def test_marginalize_frame_parity(gpu, oracle, n, idx):
    rng = np.random.default_rng(n * 10 + idx)
    D = 8 * n + 4
    A = rng.normal(size=(D, 2 * D))
    HM = A @ A.T * rng.uniform(1, 1e4, size=(D, 1)) ** 0.5 * rng.uniform(1, 1e4, size=(1, D)) ** 0.5
    HM = 0.5 * (HM + HM.T)
    bM = rng.normal(size=D) * 100
    prior, dprior = rng.uniform(0, 50, 8), rng.normal(size=8)
    g, o = util.make_ctx(gpu, "c1"), util.make_ctx(oracle, "c1")
    Hg, bg = g.window_marginalize_frame(n, idx, HM, bM, prior, dprior)
    Ho, bo = o.window_marginalize_frame(n, idx, HM, bM, prior, dprior)
    assert util.rel_err(Hg, Ho) < 1e-10 and util.rel_err(bg, bo) < 1e-10
    g.close(); o.close()


def test_calc_energy_parity(gpu, oracle, window):
    g, o = linearised(gpu, window), linearised(oracle, window)
    eg = g.window_calc_energy(window["HM"], window["bM"], window["delta"])
    eo = o.window_calc_energy(window["HM"], window["bM"], window["delta"])
    assert abs(eg[0] - eo[0]) <= 1e-4 * abs(eo[0]) and eo[0] != 0, (eg, eo)
    assert abs(eg[1] - eo[1]) <= 1e-12 * max(abs(eo[1]), 1.0)
    g.close(); o.close()


def test_batched_window_parity(gpu, oracle, window):
    """10 x the window in one launch per kernel (20 k points, 120 k residuals): the work list is long enough for
    k8_top_stream_kernel (chunks of 384+ ids, cp.async tiles) in all three modes and for several staging tiles per
    K9 CTA.  Same checks as the single window."""
    W = synth.replicate_window(window, 10)
    g, o = load(gpu, W), load(oracle, W)
    for c in (g, o):
        c.window_linearize(True)
    fl = o.window_residual_ints(2)
    assert np.array_equal(g.window_residual_ints(2), fl)
    mask = (((fl & 2) != 0) & ((fl & 1) == 0) & (np.arange(len(fl)) % 3 == 0)).astype(np.uint8)
    g.window_fix_linearization(mask); o.window_fix_linearization(mask)
    ag, ao = g.window_accumulate(), o.window_accumulate()                      # modes 0 and 1
    assert np.array_equal(ag["counts"], ao["counts"]) and ao["counts"][0] > 40000 and ao["counts"][1] > 20000
    for k in ("HA", "bA", "HL", "bL", "Hsc", "bsc"):
        assert util.rel_err(ag[k], ao[k]) < 1e-4, (k, util.rel_err(ag[k], ao[k]))
    n = W["n"]
    for k in ("HA", "HL", "Hsc"):
        for h in range(n):
            for t in range(n):
                bg = ag[k][4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t]; bo = ao[k][4 + 8 * h:12 + 8 * h, 4 + 8 * t:12 + 8 * t]
                assert np.linalg.norm(bg - bo) <= 1e-4 * np.linalg.norm(bo) + 1e-7 * np.linalg.norm(ao[k]), (k, h, t)
    for what in range(9):                                                      # per-point sums in residual order: bit-exact
        assert np.array_equal(bits(g.window_point_floats(what)), bits(o.window_point_floats(what))), what
    x = o.window_solve(0, 1e-5, ao, W["HM"] * 1e4, W["bM"], W["delta"])
    assert np.array_equal(bits(g.window_resubstitute(x, 1e-5)), bits(o.window_resubstitute(x, 1e-5)))
    g.close(); o.close()
    # mode 2: marginalise half of the points of a fully linearised window
    g, o = linearised(gpu, W), linearised(oracle, W)
    sel = np.arange(len(W["u"])) % 2 == 0
    Hg, bg, ng = g.window_marginalize_points(sel.astype(np.uint8))
    Ho, bo, no = o.window_marginalize_points(sel.astype(np.uint8))
    assert ng == no and no > 5000
    assert util.rel_err(Hg, Ho) < 1e-4 and util.rel_err(bg, bo) < 1e-4, (util.rel_err(Hg, Ho), util.rel_err(bg, bo))
    for what in range(9):
        a, b = g.window_point_floats(what), o.window_point_floats(what)
        assert np.array_equal(bits(a[sel]), bits(b[sel])), what
    g.close(); o.close()


def test_contexts_with_different_window_sizes_share_the_kernels(gpu, window):
    """several sequences live in one process (configs[4]) and their windows have different sizes at the same time: a context with
    a small window must not lower the dynamic shared-memory opt-in another context's K9 launch needs (regression: the batch
    failed with 'invalid argument' while the windows of its sequences were still growing)"""
    small = synth.make_window(frame_ids=(40, 46, 52), pts_per_frame=200)
    a = load(gpu, window)
    a.window_linearize(False); a.window_apply_res()
    ref = a.window_accumulate()
    b = load(gpu, small)                     # created after a: its first K9 launch asks for less shared memory
    b.window_linearize(False); b.window_apply_res()
    accb = b.window_accumulate()
    again = a.window_accumulate()            # the 7-frame window again
    assert util.rel_err(again["Hsc"], ref["Hsc"]) < 1e-6 and np.abs(accb["Hsc"]).max() > 0
    a.close(); b.close()
