"""Second, independent witness for K8 / K9 / K10: instead of the reference's accumulator classes (AccumulatedTopHessianSSE,
AccumulatedSCHessianSSE, src/internal/OptimizationBackend/) the whole Gauss-Newton system of the window is built DENSELY in
fp64 numpy — one Jacobian row per pattern pixel over [calibration (4) | frames (8 n) | points (np)], assembled from the
RawResidualJacobian records and the adjoints — and the quantities the accumulators produce are read off it:
    HA, bA     = the calibration / frame block of J^T J, J^T r (the priors ride on the linearised system HL, bL),
    Hsc, bsc   = H_xp H_pp^-1 H_px, H_xp H_pp^-1 b_p      (the Schur complement of the point block, EnergyFunctional.cc:683-749),
    step       = -H_pp^-1 (b_p - H_px x)                   (resubstituteFPt, EnergyFunctional.cc:638-667).
Only active, non-linearised residuals (addPoint<0>); shares no code with oracle/."""
import numpy as np

from ldso_b200 import synth
from tests import util
from tests.test_oracle_backend import load

CPARS = 4


def dense_system(W, J, active, deltaF):
    n, npnt = W["n"], len(W["u"])
    D = CPARS + 8 * n
    rows, rhs = [], []
    adH = np.asarray(W["adH"], np.float64).reshape(n * n, 8, 8)
    adT = np.asarray(W["adT"], np.float64).reshape(n * n, 8, 8)
    point_of = np.repeat(np.arange(npnt), np.diff(W["res_begin"]))
    for r in np.nonzero(active)[0]:
        p, t = point_of[r], int(W["target"][r])
        h = int(W["host"][p])
        rec = J[r].astype(np.float64)
        res, gx, gy, ja, jb = rec[0:8], rec[8:16], rec[16:24], rec[24:32], rec[32:40]
        jxi, jc, jd = rec[40:52].reshape(2, 6), rec[52:60].reshape(2, 4), rec[60:62]
        a = h + n * t
        for k in range(8):
            row = np.zeros(D + npnt)
            rel = np.concatenate([gx[k] * jxi[0] + gy[k] * jxi[1], [ja[k], jb[k]]])
            row[0:CPARS] = gx[k] * jc[0] + gy[k] * jc[1]
            row[CPARS + 8 * h:CPARS + 8 * h + 8] += adH[a] @ rel        # d(relative) / d(host) = adHost^T
            row[CPARS + 8 * t:CPARS + 8 * t + 8] += adT[a] @ rel
            row[D + p] = gx[k] * jd[0] + gy[k] * jd[1]
            rows.append(row); rhs.append(res[k])
    Jd = np.array(rows); rv = np.array(rhs)
    return Jd.T @ Jd, Jd.T @ rv, D


def test_oracle_accumulators_match_the_dense_normal_equations(oracle):
    W = synth.make_window(frame_ids=(40, 46, 52, 58), pts_per_frame=60, lin_frac=0.0)     # small: the dense J has ~5 k rows
    ctx = load(oracle, W)
    ctx.window_linearize(False)
    ctx.window_apply_res()
    acc = ctx.window_accumulate()
    J = ctx.window_residual_floats(0)
    fl = ctx.window_residual_ints(2)
    active = (fl & 2) != 0
    assert active.sum() > 300 and acc["counts"][0] == active.sum() and acc["counts"][1] == 0
    H, b, D = dense_system(W, J, active, None)
    n, npnt = W["n"], len(W["u"])
    # ---- K8 + stitch: the frame block of the active residuals (accumulateAF_MT stitches without the priors); the priors ride on the
    # linearised system (accumulateLF_MT, AccumulatedTopHessian.cc:245-254), which holds nothing else here
    prior = np.concatenate([W["c_prior"], np.asarray(W["frame_prior"], np.float64).reshape(-1)])
    dprior = np.concatenate([W["c_delta"].astype(np.float64), np.asarray(W["frame_delta_prior"], np.float64).reshape(-1)])
    assert util.rel_err(acc["HA"], H[:D, :D]) < 1e-4, util.rel_err(acc["HA"], H[:D, :D])
    assert util.rel_err(acc["bA"], b[:D]) < 1e-4, util.rel_err(acc["bA"], b[:D])
    assert util.rel_err(acc["HL"], np.diag(prior)) < 1e-12 and util.rel_err(acc["bL"], prior * dprior) < 1e-6
    # ---- K9: Schur complement of the point block (points without an active residual do not take part)
    Hpp = np.diag(H)[D:] + W["prior"].astype(np.float64)
    deltaF = (W["idepth"] - W["idepth_zero"]).astype(np.float64)
    bp = b[D:] + W["prior"].astype(np.float64) * deltaF                                # shiftPriorToZero (AccumulatedSCHessian.cc:28)
    point_of = np.repeat(np.arange(npnt), np.diff(W["res_begin"]))
    has = np.zeros(npnt, bool); has[point_of[active]] = True
    Hpp = np.where(Hpp < 1e-10, 1e-10, Hpp)
    inv = np.where(has, 1.0 / Hpp, 0.0)
    Hxp = H[:D, D:]
    Hsc = (Hxp * inv) @ Hxp.T
    bsc = Hxp @ (inv * bp)
    assert util.rel_err(acc["Hsc"], Hsc) < 1e-4, util.rel_err(acc["Hsc"], Hsc)
    assert util.rel_err(acc["bsc"], bsc) < 1e-4, util.rel_err(acc["bsc"], bsc)
    # ---- K10: back-substitution of an arbitrary frame step (lambda = 0)
    x = np.random.default_rng(1).normal(0, 1e-3, D)
    step = ctx.window_resubstitute(x, 0.0)
    want = -inv * (bp - Hxp.T @ x)
    assert util.rel_err(step[has], want[has]) < 1e-4, util.rel_err(step[has], want[has])
    ctx.close()
