"""Second, independent witness for N1's per-point part: CoarseInitializer::calcResAndGS (src/frontend/CoarseInitializer.cc:186-342
with calculateAllResiduals :391-518, calculateAlphaEnergy :376-389, calculateSC :349-374) in vectorised fp64 numpy, on the
fresh point set of setFirst (inverse depth 1 everywhere), for a small translation (alpha term active: alphaOpt = alphaW) and a
larger one (alphaOpt = 0: coupling to iR): the 8x8 system, its Schur complement over the inverse depths and the energies within
1e-4 of the oracle (fp32 sums there, fp64 here).  The SE(3) logarithm comes from scipy's matrix logarithm.  Shares no code with
oracle/."""
import numpy as np
import pytest
import scipy.linalg

from ldso_b200 import synth
from tests import util
from tests.test_initializer import sequence

PATTERN = synth.PATTERN
HUBER, ALPHA_K, ALPHA_W, COUPLING, OUTLIER_TH = 9.0, 2.5 * 2.5, 150.0 * 150.0, 1.0, 8 * 12.0 * 12.0


def lerp(img, x, y):
    """getInterpolatedElement33 on [h, w, 3] for float arrays x, y"""
    ix, iy = np.floor(x).astype(int), np.floor(y).astype(int)
    dx, dy = (x - ix)[..., None], (y - iy)[..., None]
    return (dx * dy * img[iy + 1, ix + 1] + (dy - dx * dy) * img[iy + 1, ix] + (dx - dx * dy) * img[iy, ix + 1]
            + (1 - dx - dy + dx * dy) * img[iy, ix])


def calc_res_and_gs(P, ref, new, K, R, t, aff, wl, hl):
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    RKi = R @ np.linalg.inv(K)
    a0, a1 = np.exp(aff[0]), aff[1]
    n = len(P["u"])
    u0 = P["u"][:, None] + PATTERN[None, :, 0]; v0 = P["v"][:, None] + PATTERN[None, :, 1]       # [n, 8]
    idp = P["idepth"][:, None]
    pt = np.einsum("ij,nkj->nki", RKi, np.stack([u0, v0, np.ones_like(u0)], -1)) + t[None, None, :] * idp[..., None]
    u, v = pt[..., 0] / pt[..., 2], pt[..., 1] / pt[..., 2]
    Ku, Kv = fx * u + cx, fy * v + cy
    nid = idp / pt[..., 2]
    inb = (Ku > 1) & (Kv > 1) & (Ku < wl - 2) & (Kv < hl - 2) & (nid > 0)
    Kuc, Kvc = np.where(inb, Ku, 2.0), np.where(inb, Kv, 2.0)
    hit = lerp(new, Kuc, Kvc)
    rlR = lerp(ref, u0.astype(float), v0.astype(float))[..., 0]
    fin = np.isfinite(rlR) & np.isfinite(hit[..., 0])
    good = P["isGood"] & (inb & fin).all(1)
    res = hit[..., 0] - a0 * rlR - a1
    hw = np.where(np.abs(res) < HUBER, 1.0, HUBER / np.maximum(np.abs(res), 1e-30))
    energy = (hw * res * res * (2 - hw)).sum(1)
    good &= ~(energy > OUTLIER_TH * 20)
    dxdd, dydd = (t[0] - t[2] * u) / pt[..., 2], (t[1] - t[2] * v) / pt[..., 2]
    hs = np.where(hw < 1, np.sqrt(hw), hw)
    dxi, dyi = hs * hit[..., 1] * fx, hs * hit[..., 2] * fy
    dp = np.stack([nid * dxi, nid * dyi, -nid * (u * dxi + v * dyi), -u * v * dxi - (1 + v * v) * dyi,
                   (1 + u * u) * dxi + u * v * dyi, -v * dxi + u * dyi, -hs * a0 * rlR, -hs * np.ones_like(u)], -1)   # [n, 8 px, 8]
    dd = dxi * dxdd + dyi * dydd
    r = hs * res
    g = good
    H = np.einsum("nki,nkj->ij", dp[g], dp[g]); b = np.einsum("nki,nk->i", dp[g], r[g])
    accE = energy[g].sum() + P["energy"][~g].sum()
    # alpha energy: (idepth - 1)^2 of the good points, the stored energy[1] (0 on a fresh set) of the others
    alphaE = (t @ t) * n + ((P["idepth"][g] - 1) ** 2).sum()
    alphaE *= ALPHA_W
    if alphaE > ALPHA_K * n:
        alphaOpt, alphaE = 0.0, ALPHA_K * n
    else:
        alphaOpt = ALPHA_W
    Jb = np.einsum("nki,nk->ni", dp[g], dd[g])                   # sum over the pattern of dp * dd
    Jb8 = (r[g] * dd[g]).sum(1) + alphaOpt * (P["idepth"][g] - 1)
    Jb9 = (dd[g] * dd[g]).sum(1) + alphaOpt
    if alphaOpt == 0:
        Jb8 = Jb8 + COUPLING * (P["idepth"][g] - P["iR"][g]); Jb9 = Jb9 + COUPLING
    wgt = 1.0 / (1.0 + Jb9)
    Hsc = np.einsum("ni,nj,n->ij", Jb, Jb, wgt); bsc = np.einsum("ni,n,n->i", Jb, Jb8, wgt)
    T = np.eye(4); T[:3, :3] = R; T[:3, 3] = t
    tlog = np.real(scipy.linalg.logm(T))[:3, 3]
    H[np.arange(3), np.arange(3)] += alphaOpt * n
    b[:3] += tlog * alphaOpt * n
    return H, b, Hsc, bsc, np.array([accE, alphaE, n]), alphaOpt


@pytest.mark.parametrize("tscale", [0.2, 1.5])
def test_oracle_calc_res_and_gs_matches_the_numpy_witness(oracle, tscale):
    calib = "small"
    c = util.CALIB[calib]
    frames, _, poses = sequence(calib, 6)
    ctx = util.make_ctx(oracle, calib, n_frame_slots=3)
    rel = synth.rel_pose(poses[0][0], poses[0][1], poses[5][0], poses[5][1])
    rel[9:] *= tscale
    ctx.make_images(0, frames[0]); ctx.make_images(1, frames[5])
    ctx.init_set_first(0)
    aff = (0.01, 0.5)
    seen = set()
    for lvl in range(ctx.levels):
        H, b, Hsc, bsc, res = ctx.init_calc_res_and_gs(1, lvl, rel, aff)
        P = ctx.init_get_points(lvl)
        P = {k: np.asarray(v, np.float64) if k != "isGood" else v for k, v in P.items()}
        wl, hl = c["w"] >> lvl, c["h"] >> lvl
        s = 0.5 ** lvl
        K = np.array([[c["fx"] * s, 0, (c["cx"] + 0.5) * s - 0.5], [0, c["fy"] * s, (c["cy"] + 0.5) * s - 0.5], [0, 0, 1]])   # makeK :858-885
        ref = ctx.download(0, lvl, 0).astype(np.float64); new = ctx.download(1, lvl, 0).astype(np.float64)
        R, t = rel[:9].reshape(3, 3), rel[9:]
        Hw, bw, Hscw, bscw, resw, alphaOpt = calc_res_and_gs(P, ref, new, K, R, t, aff, wl, hl)
        seen.add(alphaOpt)
        assert util.rel_err(H, Hw) < 1e-4 and util.rel_err(b, bw) < 1e-4, (lvl, util.rel_err(H, Hw), util.rel_err(b, bw))
        assert util.rel_err(Hsc, Hscw) < 1e-4 and util.rel_err(bsc, bscw) < 1e-4, (lvl, util.rel_err(Hsc, Hscw), util.rel_err(bsc, bscw))
        assert np.allclose(res, resw, rtol=1e-4), (lvl, res, resw)
    assert seen == ({ALPHA_W} if tscale < 1 else {0.0}), seen
    ctx.close()
