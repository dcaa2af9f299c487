"""CPU tests of the oracle's BA restatement (K7-K10, H1) on a synthetic frozen window."""
import numpy as np
import pytest

from ldso_b200 import synth
from tests import util


@pytest.fixture(scope="module")
def window():
    return synth.make_window(pts_per_frame=120, frame_ids=(40, 46, 52, 58, 64))


def load(api, W):
    ctx = util.make_ctx(api, "c1")
    for i, img in enumerate(W["imgs"]):
        ctx.make_images(i, img)
    ctx.window_load(W)
    return ctx


def test_linearize_states_and_jacobian_identities(oracle, window):
    ctx = load(oracle, window)
    stats, th = ctx.window_linearize(False)
    J = ctx.window_residual_floats(0)
    new_state = ctx.window_residual_ints(1)
    lin = (window["flags"] & 1) != 0
    assert stats[2] == np.sum((new_state == 0) & ~lin) > 0.5 * np.sum(~lin)
    ok = (new_state == 0) & ~lin
    # shorthand blocks are the inner products of the per-pixel columns (RawResidualJacobian.h:31-38)
    JI0, JI1, Ja0, Ja1 = J[ok, 8:16], J[ok, 16:24], J[ok, 24:32], J[ok, 32:40]
    assert np.allclose(J[ok, 62], (JI0 * JI0).sum(1), rtol=1e-5) and np.allclose(J[ok, 63], (JI0 * JI1).sum(1), rtol=1e-4, atol=1e-3)
    assert np.allclose(J[ok, 65], (Ja0 * JI0).sum(1), rtol=1e-4, atol=1e-2) and np.allclose(J[ok, 71], (Ja1 * Ja1).sum(1), rtol=1e-5)
    assert 0 < th < 1e5
    ctx.close()


def test_jacobian_matches_finite_differences_in_idepth(oracle, window):
    """d resF / d idepth = JIdx . Jpdd: check against a central difference of the residual itself (FEJ point == state)."""
    W = dict(window)
    W["idepth_zero"] = W["idepth"].copy()
    # make the FEJ pose equal to the current pose so that Jpdd (evaluated at the FEJ point) matches the residual's slope
    W["precalc"] = W["precalc"].copy()
    from ldso_b200 import window_host as wh
    aff = [(W["precalc"][h + W["n"] * h, 24] * 0, 0.0) for h in range(W["n"])]
    W["precalc"] = wh.precalc_table(W["w2c"], W["w2c"], W["calib"], aff, aff, [1.0] * W["n"])
    ctx = load(oracle, W)
    ctx.window_linearize(False)
    J0 = ctx.window_residual_floats(0)
    s0 = ctx.window_residual_ints(1)
    eps = 1e-3 * W["idepth"]
    ctx.window_set_idepth(W["idepth"] + eps, W["idepth_zero"]); ctx.window_linearize(False)
    Jp = ctx.window_residual_floats(0); sp = ctx.window_residual_ints(1)
    ctx.window_set_idepth(W["idepth"] - eps, W["idepth_zero"]); ctx.window_linearize(False)
    Jm = ctx.window_residual_floats(0); sm = ctx.window_residual_ints(1)
    rp = np.repeat(np.arange(len(W["u"])), np.diff(W["res_begin"]))
    ok = (s0 == 0) & (sp == 0) & (sm == 0) & ((W["flags"] & 1) == 0)
    # compare on un-saturated pixels (huber weight changes the scaling): use the sum over the pattern
    num = (Jp[ok, 0:8] - Jm[ok, 0:8]).sum(1) / (2 * eps[rp][ok])
    ana = (J0[ok, 8:16] * J0[ok, 60:61] + J0[ok, 16:24] * J0[ok, 61:62]).sum(1)
    good = np.abs(ana) > 50
    rel = np.abs(num[good] - ana[good]) / np.abs(ana[good])
    assert np.median(rel) < 0.15, np.median(rel)
    ctx.close()


def test_accumulate_solve_resubstitute(oracle, window):
    ctx = load(oracle, window)
    ctx.window_linearize(False)
    ctx.window_apply_res()
    acc = ctx.window_accumulate()
    D = 8 * window["n"] + 4
    for k in ("HA", "HL", "Hsc"):
        assert np.allclose(acc[k], acc[k].T, rtol=1e-9, atol=1e-6 * np.abs(acc[k]).max()), k
    assert acc["counts"][0] > 0 and acc["counts"][1] > 0
    assert np.all(np.linalg.eigvalsh(acc["HA"]) > -1e-6 * np.abs(acc["HA"]).max())
    # Schur complement of a PSD system is PSD: HA + HL - Hsc >= 0 (up to float accumulation noise)
    S = acc["HA"] + acc["HL"] - acc["Hsc"]
    assert np.linalg.eigvalsh(S).min() > -1e-4 * np.abs(S).max()
    x = ctx.window_solve(0, 1e-5, acc, window["HM"], window["bM"], window["delta"])
    # numpy cross-check of the scaled LDLT solve
    lam = 1e-5
    Hf = acc["HL"] + window["HM"] + acc["HA"]
    Hf[np.diag_indices(D)] *= (1 + lam)
    Hf = Hf - acc["Hsc"] * (1.0 / (1 + lam))
    Hf = np.triu(Hf) + np.triu(Hf, 1).T
    bf = acc["bL"] + (window["bM"] + window["HM"] @ window["delta"]) + acc["bA"] - acc["bsc"] / (1 + lam)
    s = 1 / np.sqrt(np.diag(Hf) + 10)
    xr = s * np.linalg.solve(s[:, None] * Hf * s[None, :], s * bf)
    assert util.rel_err(x, xr) < 1e-8
    step = ctx.window_resubstitute(x, 0.0)
    assert np.all(np.isfinite(step)) and np.abs(step).max() > 0
    # orthogonalisation removes the components along N
    N = np.random.default_rng(0).normal(size=(D, 7))
    xo = ctx.window_solve(3, 1e-5, acc, window["HM"], window["bM"], window["delta"], N)
    Q, _ = np.linalg.qr(N)
    assert np.abs(Q.T @ xo).max() < 1e-9 * max(1.0, np.abs(x).max())
    assert util.rel_err(xo, x - Q @ (Q.T @ x)) < 1e-8
    ctx.close()


def test_fix_linearization_roundtrip(oracle, window):
    """res_toZeroF = resF - J*delta, and mode-1 accumulation adds J*delta back: with unchanged deltas the
    linearised residual reproduces resF (Residuals.cc:219-245 vs AccumulatedTopHessian.cc:47-64)."""
    ctx = load(oracle, window)
    ctx.window_linearize(False)
    ctx.window_apply_res()
    J = ctx.window_residual_floats(0)
    mask = ((window["flags"] & 1) == 0) & ((ctx.window_residual_ints(2) & 2) != 0)
    ctx.window_fix_linearization(mask.astype(np.uint8))
    rtz = ctx.window_residual_floats(1)
    fl = ctx.window_residual_ints(2)
    assert np.all((fl[mask] & 1) == 1)
    assert not np.allclose(rtz[mask], J[mask, 0:8])
    ctx.close()


def np_marginalize_frame(HM, bM, n, idx, prior, dprior):
    """numpy restatement of EnergyFunctional::marginalizeFrame (EnergyFunctional.cc:109-191)"""
    D = 8 * n + 4
    io = 4 + 8 * idx
    perm = [i for i in range(D) if not io <= i < io + 8] + list(range(io, io + 8))
    H, b = HM[np.ix_(perm, perm)].copy(), bM[perm].copy()
    H[-8:, -8:] += np.diag(prior)
    b[-8:] += prior * dprior
    sv = np.sqrt(np.abs(np.diag(H)) + 10)
    Hs, bs = H / sv[:, None] / sv[None, :], b / sv
    hp = Hs[-8:, -8:]
    s2 = 1 / np.sqrt(np.abs(np.diag(hp)) + 10)
    U, S, Vt = np.linalg.svd(s2[:, None] * np.triu(hp) * s2[None, :] + (s2[:, None] * np.triu(hp, 1) * s2[None, :]).T)
    inv = s2[:, None] * (Vt.T @ np.diag(np.where(S > 0, 1 / S, 0)) @ U.T) * s2[None, :]
    inv = np.triu(inv) + np.triu(inv, 1).T
    bli = Hs[-8:, :-8].T @ inv
    Hn = Hs[:-8, :-8] - bli @ Hs[-8:, :-8]
    bn = bs[:-8] - bli @ bs[-8:]
    Hn, bn = Hn * sv[:-8, None] * sv[None, :-8], bn * sv[:-8]
    return 0.5 * (Hn + Hn.T), bn


@pytest.mark.parametrize("n,idx", [(7, 0), (7, 3), (7, 6), (2, 0), (1, 0)])
def test_marginalize_frame_matches_numpy_and_the_schur_complement(oracle, n, idx):
    rng = np.random.default_rng(n * 10 + idx)
    D = 8 * n + 4
    A = rng.normal(size=(D, 3 * D))
    HM = A @ A.T * rng.uniform(1, 1e4, size=(D, 1)) ** 0.5 * rng.uniform(1, 1e4, size=(1, D)) ** 0.5
    HM = 0.5 * (HM + HM.T)
    bM = rng.normal(size=D) * 100
    prior, dprior = rng.uniform(0, 50, 8), rng.normal(size=8)
    ctx = util.make_ctx(oracle, "c1")
    Ho, bo = ctx.window_marginalize_frame(n, idx, HM, bM, prior, dprior)
    Hr, br = np_marginalize_frame(HM, bM, n, idx, prior, dprior)
    assert util.rel_err(Ho, Hr) < 1e-9 and util.rel_err(bo, br) < 1e-9
    # it is the Schur complement onto the remaining variables
    io = 4 + 8 * idx
    keep = [i for i in range(D) if not io <= i < io + 8]
    H22 = HM[io:io + 8, io:io + 8] + np.diag(prior)
    S = HM[np.ix_(keep, keep)] - HM[keep, io:io + 8] @ np.linalg.solve(H22, HM[io:io + 8, keep])
    bs = bM[keep] - HM[keep, io:io + 8] @ np.linalg.solve(H22, bM[io:io + 8] + prior * dprior)
    assert util.rel_err(Ho, S) < 1e-6 and util.rel_err(bo, bs) < 1e-6
    assert np.array_equal(Ho, Ho.T)
    ctx.close()


def linearised_window(api, W):
    ctx = load(api, W)
    ctx.window_linearize(True)          # applyRes + bookkeeping
    fl = ctx.window_residual_ints(2)
    mask = (fl & 2) != 0                # every active residual gets fixed: all of them become "linearised"
    ctx.window_fix_linearization(mask.astype(np.uint8))
    return ctx


def test_marginalize_points_equals_the_linearised_system_of_those_points(oracle, window):
    """with zero deltas, addPoint<2> over ALL points (prior factor 1) is addPoint<1> without the priors, and the SC part
    is the same call: H = HL - priors - Hsc, b likewise (EnergyFunctional.cc:233-246 vs :670-749)"""
    W = dict(window)
    W["idepth_zero"] = W["idepth"].copy()
    W["adHTdelta"] = np.zeros_like(W["adHTdelta"]); W["c_delta"] = np.zeros(4, np.float32)
    W["frame_delta_prior"] = np.zeros_like(W["frame_delta_prior"])
    ctx = linearised_window(oracle, W)
    acc = ctx.window_accumulate()
    assert acc["counts"][0] == 0 and acc["counts"][1] > 1000
    np_ = len(W["u"])
    H, b, nres = ctx.window_marginalize_points(np.ones(np_, np.uint8), fac=1.0)
    assert nres == acc["counts"][1]
    prior = np.concatenate([W["c_prior"], np.asarray(W["frame_prior"]).reshape(-1)])
    Hexp = acc["HL"] - np.diag(prior) - acc["Hsc"]
    assert util.rel_err(H, Hexp) < 1e-6, util.rel_err(H, Hexp)
    assert util.rel_err(b, acc["bL"] - acc["bsc"]) < 1e-5
    # removePoint: nothing is active any more
    assert np.all((ctx.window_residual_ints(2) & 2) == 0)
    acc2 = ctx.window_accumulate()
    assert acc2["counts"].sum() == 0 and np.abs(acc2["Hsc"]).max() == 0
    ctx.close()
    # a subset: H is PSD-ish (a Schur complement of a PSD system) and the prior factor only shrinks the SC part
    ctx = linearised_window(oracle, W)
    m = (np.arange(np_) % 3 == 0).astype(np.uint8)
    H1, b1, n1 = ctx.window_marginalize_points(m, fac=600.0 * 600.0)
    ctx.close()
    assert 0 < n1 < nres
    assert np.linalg.eigvalsh(0.5 * (H1 + H1.T)).min() > -1e-5 * np.abs(H1).max()


def test_calc_energy_matches_numpy(oracle, window):
    ctx = linearised_window(oracle, window)
    J, rtz, fl = ctx.window_residual_floats(0).astype(np.float64), ctx.window_residual_floats(1).astype(np.float64), ctx.window_residual_ints(2)
    W = window
    n = W["n"]
    pt = np.repeat(np.arange(len(W["u"])), np.diff(W["res_begin"]))
    dp = W["adHTdelta"].astype(np.float64)[W["host"][pt] + n * W["target"]]
    dd = (W["idepth"].astype(np.float64) - W["idepth_zero"])[pt]
    dc = W["c_delta"].astype(np.float64)
    Jx = (J[:, 40:46] * dp[:, :6]).sum(1) + J[:, 52:56] @ dc + J[:, 60] * dd
    Jy = (J[:, 46:52] * dp[:, :6]).sum(1) + J[:, 56:60] @ dc + J[:, 61] * dd
    Jd = J[:, 8:16] * Jx[:, None] + J[:, 16:24] * Jy[:, None] + J[:, 24:32] * dp[:, 6:7] + J[:, 32:40] * dp[:, 7:8]
    use = ((fl & 1) != 0) & ((fl & 2) != 0)
    E = ((2 * rtz + Jd) * Jd)[use].sum()
    ddp = W["idepth"].astype(np.float64) - W["idepth_zero"]
    E += (ddp * ddp * W["prior"]).sum()
    fp, fd = np.asarray(W["frame_prior"], float), np.asarray(W["frame_delta_prior"], float)
    E += (fd * fp * fd).sum() + (dc * np.asarray(W["c_prior"], float) * dc).sum()
    eL, eM = ctx.window_calc_energy(W["HM"], W["bM"], W["delta"])
    assert use.sum() > 1000
    assert abs(eL - E) < 1e-4 * max(abs(E), 1.0), (eL, E)
    assert abs(eM - W["delta"] @ (2 * W["bM"] + W["HM"] @ W["delta"])) <= 1e-9 * max(abs(eM), 1.0)
    ctx.close()
