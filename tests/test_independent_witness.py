"""Second, independent witnesses (numpy, written from the formulas, not from the oracle's code) for the rows that had
only the line-by-line restatement: K7's pose / calibration / idepth Jacobians and photometric columns
(PointFrameResidual::linearize, src/internal/Residuals.cc:15-217) and the tracker's K3 / K4 sums
(CoarseTracker::calcRes :540-673, calcGSSSE :675-741).  Geometry columns are checked against central finite differences
of the projection itself; photometric values against a vectorised numpy evaluation of the same image model."""
import numpy as np
import pytest

from ldso_b200 import synth
from tests import util

PATTERN = synth.PATTERN


def se3_exp(xi):
    u, w = np.asarray(xi[:3], float), np.asarray(xi[3:], float)
    th = np.linalg.norm(w)
    W = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th < 1e-12:
        return np.eye(3) + W, u
    A, B, Cc = np.sin(th) / th, (1 - np.cos(th)) / th ** 2, (th - np.sin(th)) / th ** 3
    return np.eye(3) + A * W + B * W @ W, (np.eye(3) + B * W + Cc * W @ W) @ u


def project(R, t, K, u, v, idepth):
    """pixel of host pixel (u, v) at inverse depth idepth in the target: K (R K^-1 [u v 1]^T + t idepth), dehomogenised"""
    p = K @ (R @ (np.linalg.inv(K) @ np.array([u, v, 1.0])) + t * idepth)
    return p[:2] / p[2]


def bilinear(img, x, y):
    ix, iy = np.floor(x).astype(int), np.floor(y).astype(int)
    dx, dy = x - ix, y - iy
    return (img[iy, ix] * (1 - dx) * (1 - dy) + img[iy, ix + 1] * dx * (1 - dy) + img[iy + 1, ix] * (1 - dx) * dy +
            img[iy + 1, ix + 1] * dx * dy)


@pytest.fixture(scope="module")
def lin_window(oracle):
    W = synth.make_window(pts_per_frame=80, frame_ids=(40, 46, 52, 58))
    ctx = util.make_ctx(oracle, "c1")
    for i, img in enumerate(W["imgs"]):
        ctx.make_images(i, img)
    ctx.window_load(W)
    ctx.window_linearize(False)
    J = ctx.window_residual_floats(0)
    st = ctx.window_residual_ints(1)
    dI = [ctx.download(i, 0, 0) for i in range(W["n"])]       # [h, w, 3] = I, dx, dy of makeImages (K1 has its own numpy twin)
    ctx.close()
    return W, J, st, dI


def test_k7_geometry_jacobians_match_finite_differences(lin_window):
    """Jpdxi (2x6), Jpdc (2x4), Jpdd (2) of RawResidualJacobian are d(Ku, Kv) / d(xi, c, idepth) of the centre pixel at the
    FEJ point (Residuals.cc:60-113): xi = left increment of the host->target pose, c = (fx, fy, cx, cy) / (SCALE_F, SCALE_F,
    SCALE_C, SCALE_C) = 50, idepth with SCALE_IDEPTH = 1"""
    W, J, st, _ = lin_window
    n = W["n"]
    fx, fy, cx, cy = [float(c) for c in W["calib"]]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    rp = np.repeat(np.arange(len(W["u"])), np.diff(W["res_begin"]))
    rng = np.random.default_rng(0)
    picks = rng.choice(np.nonzero(st == 0)[0], 60, replace=False)
    for r in picks:
        p, t_idx = rp[r], W["target"][r]
        pre = W["precalc"][W["host"][p] + n * t_idx].astype(np.float64)
        R0, t0 = pre[0:9].reshape(3, 3), pre[9:12]                  # PRE_RTll_0, PRE_tTll_0: the FEJ evaluation point
        u, v, idz = float(W["u"][p]), float(W["v"][p]), float(W["idepth_zero"][p])
        ana_xi = np.stack([J[r, 40:46], J[r, 46:52]])
        ana_c = np.stack([J[r, 52:56], J[r, 56:60]])
        ana_d = J[r, 60:62]
        num_xi = np.zeros((2, 6))
        for k in range(6):
            e = np.zeros(6); e[k] = 1e-6
            Rp, tp = se3_exp(e); Rm, tm = se3_exp(-e)
            num_xi[:, k] = (project(Rp @ R0, Rp @ t0 + tp, K, u, v, idz) - project(Rm @ R0, Rm @ t0 + tm, K, u, v, idz)) / 2e-6
        num_c = np.zeros((2, 4))
        for k, scale in enumerate((50.0, 50.0, 50.0, 50.0)):
            dpar = np.zeros(4); dpar[k] = 1e-7 * scale                # a step of 1e-7 in the unscaled parameter
            def Kof(s):
                f = np.array([fx, fy, cx, cy]) + s * dpar
                return np.array([[f[0], 0, f[2]], [0, f[1], f[3]], [0, 0, 1.0]])
            num_c[:, k] = (project(R0, t0, Kof(+1), u, v, idz) - project(R0, t0, Kof(-1), u, v, idz)) / 2e-7
        h = 1e-6 * max(idz, 1e-3)
        num_d = (project(R0, t0, K, u, v, idz + h) - project(R0, t0, K, u, v, idz - h)) / (2 * h)
        for ana, num, name in ((ana_xi, num_xi, "Jpdxi"), (ana_c, num_c, "Jpdc"), (ana_d, num_d, "Jpdd")):
            assert np.abs(ana - num).max() <= 2e-4 * max(1.0, np.abs(num).max()), (name, r, ana, num)


def test_k7_photometric_columns_match_numpy(lin_window):
    """resF, JIdx, JabF and the energy of Residuals.cc:128-190 evaluated with numpy on the same pyramid"""
    W, J, st, dI = lin_window
    n = W["n"]
    rp = np.repeat(np.arange(len(W["u"])), np.diff(W["res_begin"]))
    ok = np.nonzero(st == 0)[0]
    worst = 0.0
    for r in ok[::7]:
        p, t_idx = rp[r], W["target"][r]
        pre = W["precalc"][W["host"][p] + n * t_idx].astype(np.float64)
        KRKi, Kt, aff, b0 = pre[12:21].reshape(3, 3), pre[21:24], pre[24:26], pre[26]
        px = W["u"][p] + PATTERN[:, 0]; py = W["v"][p] + PATTERN[:, 1]
        ptp = (KRKi @ np.stack([px, py, np.ones(8)])) + Kt[:, None] * float(W["idepth"][p])
        Ku, Kv = ptp[0] / ptp[2], ptp[1] / ptp[2]
        hit = np.stack([bilinear(dI[t_idx][:, :, c].astype(np.float64), Ku, Kv) for c in range(3)])
        color, wgt = W["color"][p].astype(np.float64), W["weights"][p].astype(np.float64)
        residual = hit[0] - (aff[0] * color + aff[1])
        w = 0.5 * (np.sqrt(2500.0 / (2500.0 + hit[1] ** 2 + hit[2] ** 2)) + wgt)
        hw = np.where(np.abs(residual) < 9, 1.0, 9 / np.abs(residual))
        hwJ = np.where(hw < 1, np.sqrt(hw), hw) * w
        exp_res = residual * hwJ
        exp = np.concatenate([exp_res, hit[1] * hwJ, hit[2] * hwJ, (color - b0) * hwJ, hwJ])
        err = np.abs(J[r, 0:40] - exp).max() / max(1.0, np.abs(exp).max())
        worst = max(worst, err)
    assert worst < 2e-4, worst                                        # float32 evaluation against float64


def test_k3_k4_sums_match_numpy(oracle):
    """calcRes' Vec6 and calcGSSSE's H, b (CoarseTracker.cc:540-741) from a vectorised numpy evaluation over the same
    point cloud: E, numTermsInE, saturated ratio, and the 8x8 system (scaled, divided by the padded count, trap T7)"""
    ctx = util.make_ctx(oracle, "c1")
    gt = util.setup_tracker(ctx, "c1", 1500)
    c = util.CALIB["c1"]
    for lvl, pose, aff, cutoff in ((0, gt, (0.0, 0.0), 20.0), (1, gt, (0.03, -2.0), 20.0), (2, gt, (-0.02, 3.0), 40.0)):      # (the identity pose would put border points exactly ON the Ku > 2 thresholds)
        out6 = ctx.tracker_calc_res(0, 1, lvl, pose, aff[0], aff[1], cutoff)
        H, b = ctx.tracker_calc_gs(0, lvl, pose, aff[0], aff[1])
        u, v, idp, col = [a.astype(np.float64) for a in ctx.tracker_get_pc(0, lvl)]
        img = ctx.download(1, lvl, 0).astype(np.float64)
        s = 2.0 ** lvl
        fx, fy, cx, cy = c["fx"] / s, c["fy"] / s, (c["cx"] + 0.5) / s - 0.5, (c["cy"] + 0.5) / s - 0.5
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
        R, t = pose[:9].reshape(3, 3), pose[9:]
        pt = (R @ np.linalg.inv(K)) @ np.stack([u, v, np.ones_like(u)]) + t[:, None] * idp
        uu, vv = pt[0] / pt[2], pt[1] / pt[2]
        Ku, Kv, nid = fx * uu + cx, fy * vv + cy, idp / pt[2]
        h, w = img.shape[:2]
        inb = (Ku > 2) & (Kv > 2) & (Ku < w - 3) & (Kv < h - 3) & (nid > 0)
        Ku, Kv, uu, vv, nid, colr = Ku[inb], Kv[inb], uu[inb], vv[inb], nid[inb], col[inb]
        hit = np.stack([bilinear(img[:, :, k], Ku, Kv) for k in range(3)])
        a = np.exp(aff[0] - 0.0)                                      # reference frame affine (0, 0), exposures 1
        res = hit[0] - (a * colr + aff[1])
        hw = np.where(np.abs(res) < 9, 1.0, 9 / np.abs(res))
        sat = np.abs(res) > cutoff
        E = np.where(sat, 2 * 9 * cutoff - 81, hw * res * res * (2 - hw)).sum()
        assert out6[1] == len(res)
        assert abs(out6[0] - E) <= 2e-4 * E
        assert abs(out6[5] - sat.mean()) < 1e-6
        g = ~sat
        dx, dy = hit[1][g] * fx, hit[2][g] * fy
        Jm = np.stack([nid[g] * dx, nid[g] * dy, -nid[g] * (uu[g] * dx + vv[g] * dy), -(uu[g] * vv[g] * dx + dy * (1 + vv[g] ** 2)),
                       uu[g] * vv[g] * dy + dx * (1 + uu[g] ** 2), uu[g] * dy - vv[g] * dx, a * (0.0 - colr[g]), -np.ones(g.sum())])
        npad = (g.sum() + 3) // 4 * 4
        sc = np.array([0.5, 0.5, 0.5, 1, 1, 1, 10, 1000.0])
        Hn = (Jm * hw[g]) @ Jm.T / npad * np.outer(sc, sc)
        bn = (Jm * hw[g]) @ res[g] / npad * sc
        assert np.linalg.norm(H - Hn) <= 3e-4 * np.linalg.norm(Hn), (lvl, np.linalg.norm(H - Hn) / np.linalg.norm(Hn))
        # b sums signed terms (cancellation): the oracle's float32 SSE accumulators lose ~1e-3 of |b| against float64
        assert np.linalg.norm(b - bn) <= 2e-3 * np.linalg.norm(bn) + 1e-6
    ctx.close()
