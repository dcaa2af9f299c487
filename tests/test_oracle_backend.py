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
