"""Second, independent witness for K5 / K6: the ImmaturePoint constructor (src/internal/ImmaturePoint.cc:14-39) and
ImmaturePoint::traceOn (:47-317) written out in numpy straight from the reference's text — scalar float32 arithmetic for the
geometry that feeds discrete decisions (the sub-pixel start of the search is `uMin * 1000 - floor(uMin * 1000)` in float),
small numpy vectors over the 8 pattern pixels for the photometric part — and compared with the oracle on two consecutive
traces of the synthetic scene: first from an uninitialised interval (idepth_max = NaN: the search runs along the epipolar
direction for maxPixSearch pixels), then from the interval the first trace produced.

This file shares no code with oracle/ (it goes through the C ABI only to obtain the oracle's results)."""
import numpy as np

from ldso_b200 import synth
from tests import util

F = np.float32
PATTERN = synth.PATTERN                      # src/Setting.cc:275 (staticPattern[8])
GOOD, OOB, OUTLIER, SKIPPED, BADCONDITION, UNINITIALIZED = range(6)   # include/internal/ImmaturePoint.h:31-38
# src/Setting.cc
MAX_PIX_SEARCH, OUTLIER_TH, OUTLIER_TH_SUM, MIN_TEST_RADIUS, HUBER_TH = F(0.027), F(12 * 12), F(50 * 50), 2, F(9)
STEPSIZE, GN_ITS, GN_TH, EXTRA_SLACK, SLACK_INTERVAL, MIN_IMPROVEMENT = F(1.0), 3, F(0.1), F(1.2), F(1.5), F(2)


def lerp3(dI, w, x, y):
    """getInterpolatedElement33 (include/internal/GlobalFuncs.h:90-103) for vectors x, y: [n, 3]"""
    ix, iy = x.astype(np.int32), y.astype(np.int32)
    dx, dy = x - ix.astype(F), y - iy.astype(F)
    dxdy = dx * dy
    bp = ix + iy * w
    return (dxdy[:, None] * dI[bp + 1 + w] + (dy - dxdy)[:, None] * dI[bp + w] + (dx - dxdy)[:, None] * dI[bp + 1]
            + (F(1) - dx - dy + dxdy)[:, None] * dI[bp])


def ctor(dI, w, u, v):
    """ImmaturePoint::ImmaturePoint: colour, weights, gradH, energyTH of one point (getInterpolatedElement33BiLin :186-207)"""
    color, weights, gradH = np.zeros(8, F), np.zeros(8, F), np.zeros((2, 2), F)
    for idx, (px, py) in enumerate(PATTERN):
        x, y = F(u + px), F(v + py)
        ix, iy = int(x), int(y)
        bp = ix + iy * w
        tl, tr, bl, br = dI[bp, 0], dI[bp + 1, 0], dI[bp + w, 0], dI[bp + w + 1, 0]
        dx, dy = F(x - ix), F(y - iy)
        top, bot = dx * tr + (F(1) - dx) * tl, dx * br + (F(1) - dx) * bl
        left, right = dy * bl + (F(1) - dy) * tl, dy * br + (F(1) - dy) * tr
        c, gx, gy = dx * right + (F(1) - dx) * left, right - left, bot - top
        color[idx] = c
        if not np.isfinite(c):
            return color, weights, gradH, F(np.nan)
        g = np.array([gx, gy], F)
        gradH += np.outer(g, g)
        weights[idx] = np.sqrt(OUTLIER_TH_SUM / (OUTLIER_TH_SUM + (gx * gx + gy * gy)))
    return color, weights, gradH, F(8) * OUTLIER_TH


def trace_on(P, dI, w, h, KRKi, Kt, aff):
    """ImmaturePoint::traceOn on the point state P (a dict, updated in place); returns the status"""
    def out(status, uv=(-1.0, -1.0), interval=0.0):
        P["uv"], P["interval"], P["status"] = (F(uv[0]), F(uv[1])), F(interval), status
        return status
    if P["status"] == OOB:
        return OOB
    maxPix = F(w + h) * MAX_PIX_SEARCH
    u, v = P["u"], P["v"]
    pr = np.array([KRKi[r, 0] * u + KRKi[r, 1] * v + KRKi[r, 2] for r in range(3)], F)
    ptpMin = pr + Kt * P["idepth_min"]
    uMin, vMin = ptpMin[0] / ptpMin[2], ptpMin[1] / ptpMin[2]
    inside = lambda a, b: a > 4 and b > 4 and a < w - 5 and b < h - 5
    if not inside(uMin, vMin):
        return out(OOB)
    finite_max = bool(np.isfinite(P["idepth_max"]))
    if finite_max:
        ptpMax = pr + Kt * P["idepth_max"]
        uMax, vMax = ptpMax[0] / ptpMax[2], ptpMax[1] / ptpMax[2]
        if not inside(uMax, vMax):
            return out(OOB)
        dist = np.sqrt((uMin - uMax) * (uMin - uMax) + (vMin - vMax) * (vMin - vMax))
        if dist < SLACK_INTERVAL:
            return out(SKIPPED, ((uMax + uMin) * F(0.5), (vMax + vMin) * F(0.5)), dist)
    else:
        dist = maxPix
        ptpMax = pr + Kt * F(0.01)
        uMax, vMax = ptpMax[0] / ptpMax[2], ptpMax[1] / ptpMax[2]
        ddx, ddy = uMax - uMin, vMax - vMin
        d = F(1) / np.sqrt(ddx * ddx + ddy * ddy)
        uMax, vMax = uMin + dist * ddx * d, vMin + dist * ddy * d
        if not inside(uMax, vMax):
            return out(OOB)
    if not (P["idepth_min"] < 0 or (ptpMin[2] > 0.75 and ptpMin[2] < 1.5)):
        return out(OOB)
    dx, dy = STEPSIZE * (uMax - uMin), STEPSIZE * (vMax - vMin)
    gH = P["gradH"]
    a = dx * (gH[0, 0] * dx + gH[0, 1] * dy) + dy * (gH[1, 0] * dx + gH[1, 1] * dy)
    b = dy * (gH[0, 0] * dy - gH[0, 1] * dx) - dx * (gH[1, 0] * dy - gH[1, 1] * dx)
    errorInPixel = F(0.2) + F(0.2) * (a + b) / a
    if errorInPixel * MIN_IMPROVEMENT > dist and finite_max:
        return out(BADCONDITION, ((uMax + uMin) * F(0.5), (vMax + vMin) * F(0.5)), dist)
    if errorInPixel > 10:
        errorInPixel = F(10)
    dx, dy = dx / dist, dy / dist
    if dist > maxPix:
        dist = maxPix
    numSteps = int(F(1.9999) + dist / STEPSIZE)
    Rplane = KRKi[:2, :2]
    randShift = uMin * F(1000) - np.floor(uMin * F(1000))
    ptx, pty = uMin - randShift * dx, vMin - randShift * dy
    rot = np.array([[Rplane[0, 0] * px + Rplane[0, 1] * py, Rplane[1, 0] * px + Rplane[1, 1] * py] for px, py in PATTERN.astype(F)], F)
    if not (np.isfinite(dx) and np.isfinite(dy)):
        return out(OOB)
    if numSteps >= 100:
        numSteps = 99
    color, weights = P["color"], P["weights"]
    errors = np.zeros(numSteps, F)
    bestU = bestV = F(0)
    bestEnergy, bestIdx = F(1e10), -1

    def huber_energy(res):
        hw = np.where(np.abs(res) < HUBER_TH, F(1), HUBER_TH / np.abs(res)).astype(F)
        return hw, hw * res * res * (F(2) - hw)
    for i in range(numSteps):
        hit = lerp3(dI, w, ptx + rot[:, 0], pty + rot[:, 1])[:, 0]
        res = hit - (aff[0] * color + aff[1])
        _, e = huber_energy(res)
        e = np.where(np.isfinite(hit), e, F(1e5))
        energy = F(0)
        for q in range(8):
            energy = F(energy + e[q])
        errors[i] = energy
        if energy < bestEnergy:
            bestU, bestV, bestEnergy, bestIdx = ptx, pty, energy, i
        ptx, pty = F(ptx + dx), F(pty + dy)
    idx = np.arange(numSteps)
    far = (idx < bestIdx - MIN_TEST_RADIUS) | (idx > bestIdx + MIN_TEST_RADIUS)
    secondBest = errors[far].min() if far.any() else F(1e10)
    if secondBest > F(1e10):
        secondBest = F(1e10)
    newQuality = secondBest / bestEnergy
    if newQuality < P["quality"] or numSteps > 10:
        P["quality"] = newQuality
    uBak, vBak, stepBack = bestU, bestV, F(0)
    bestEnergy = F(1e5)
    for it in range(GN_ITS):
        hit = lerp3(dI, w, bestU + rot[:, 0], bestV + rot[:, 1])
        ok = np.isfinite(hit[:, 0])
        res = hit[:, 0] - (aff[0] * color + aff[1])
        dRes = dx * hit[:, 1] + dy * hit[:, 2]
        hw, e = huber_energy(res)
        H, bb, energy = F(1), F(0), F(0)
        for q in range(8):
            if not ok[q]:
                energy = F(energy + F(1e5))
                continue
            H = F(H + hw[q] * dRes[q] * dRes[q])
            bb = F(bb + hw[q] * res[q] * dRes[q])
            energy = F(energy + weights[q] * weights[q] * e[q])
        if energy > bestEnergy:
            stepBack = F(stepBack * F(0.5))
            bestU, bestV = F(uBak + stepBack * dx), F(vBak + stepBack * dy)
        else:
            step = -bb / H
            step = F(min(max(step, F(-0.5)), F(0.5))) if np.isfinite(step) else F(0)
            uBak, vBak, stepBack = bestU, bestV, step
            bestU, bestV, bestEnergy = F(bestU + step * dx), F(bestV + step * dy), energy
        if abs(stepBack) < GN_TH:
            break
    if not (bestEnergy < P["energyTH"] * EXTRA_SLACK):
        return out(OOB if P["status"] == OUTLIER else OUTLIER)
    if dx * dx > dy * dy:
        lo, hi = bestU - errorInPixel * dx, bestU + errorInPixel * dx
        imin, imax = (pr[2] * lo - pr[0]) / (Kt[0] - Kt[2] * lo), (pr[2] * hi - pr[0]) / (Kt[0] - Kt[2] * hi)
    else:
        lo, hi = bestV - errorInPixel * dy, bestV + errorInPixel * dy
        imin, imax = (pr[2] * lo - pr[1]) / (Kt[1] - Kt[2] * lo), (pr[2] * hi - pr[1]) / (Kt[1] - Kt[2] * hi)
    if imin > imax:
        imin, imax = imax, imin
    P["idepth_min"], P["idepth_max"] = F(imin), F(imax)
    if not np.isfinite(imin) or not np.isfinite(imax) or imax < 0:
        return out(OUTLIER)
    return out(GOOD, (bestU, bestV), F(2) * errorInPixel)


def test_oracle_constructor_and_trace_match_the_numpy_witness(oracle):
    S = util.immature_scene("small", pts_per_host=90)
    c = S["calib"]
    w, h = c["w"], c["h"]
    ctx = util.make_ctx(oracle, "small", n_frame_slots=8)
    res = util.run_immature(ctx, S, activate=False)
    dIs = [ctx.download(i, 0, 0).reshape(-1, 3).astype(F) for i in range(len(S["frames"]))]
    n, nh = len(S["u"]), S["n_hosts"]
    # ---- K6: the constructor
    pts = []
    for i in range(n):
        col, wgt, gH, eth = ctor(dIs[S["host"][i]], w, S["u"][i], S["v"][i])
        assert np.allclose(col, res["color"][i], rtol=1e-6, atol=1e-5) and np.allclose(wgt, res["weights"][i], rtol=1e-6)
        assert np.allclose([gH[0, 0], gH[0, 1], gH[1, 1]], res["grad_h"][i], rtol=1e-5, atol=1e-4) and eth == res["energy_th"][i]
        pts.append(dict(u=F(S["u"][i]), v=F(S["v"][i]), color=col, weights=wgt, gradH=gH, energyTH=eth, idepth_min=F(0), idepth_max=F(np.nan),
                        quality=F(10000), status=UNINITIALIZED, uv=(F(-1), F(-1)), interval=F(0)))
    # ---- K5: two traces in a row (uninitialised interval, then the interval the first trace left)
    seen = set()
    for ti in range(2):
        KRKi, Kt, aff = S["tables"][ti]
        t = res["traces"][ti]
        st = np.array([trace_on(P, dIs[nh + ti], w, h, KRKi[S["host"][i]], Kt[S["host"][i]], aff[S["host"][i]]) for i, P in enumerate(pts)])
        same = st == t["status"]
        assert same.mean() > 0.97, (ti, same.mean(), np.bincount(st, minlength=6), np.bincount(t["status"], minlength=6))
        good = same & (st == GOOD)
        assert good.sum() > 0.3 * n
        mine = lambda k: np.array([P[k] for P in pts], F)
        for k in ("idepth_min", "idepth_max", "quality", "interval"):
            assert np.allclose(mine(k)[good], t[k][good], rtol=2e-3, atol=1e-5), (ti, k, np.abs(mine(k)[good] - t[k][good]).max())
        uv = np.array([P["uv"] for P in pts], F)
        assert np.abs(uv[good] - t["uv"][good]).max() < 2e-2
        seen |= set(np.unique(t["status"]).tolist())
        # the next trace starts from the oracle's state wherever the two disagreed, so one flip does not cascade
        for i in np.nonzero(~same)[0]:
            pts[i].update(status=int(t["status"][i]), idepth_min=t["idepth_min"][i], idepth_max=t["idepth_max"][i], quality=t["quality"][i])
    assert GOOD in seen and len(seen) >= 3      # several exits of traceOn were exercised (OOB / OUTLIER / SKIPPED / BADCONDITION)
    ctx.close()


# ---- K6, second half: FullSystem::optimizeImmaturePoint (src/frontend/FullSystem.cc:1035-1155) over
# ImmaturePoint::linearizeResidual (src/internal/ImmaturePoint.cc:319-389), projectPoint / derive_idepth
# (include/internal/ResidualProjections.h:12-92), scalar float32 -------------------------------------------------------------
IN, ST_OOB, ST_OUTLIER = 0, 1, 2                      # ResState (include/internal/Residuals.h)
MIN_IDEPTH_H_ACT, GN_ITS_ACT = F(100), 3              # src/Setting.cc:25, :47


def linearize_residual(P, T, slack, tmp, Hdd, bd, idepth):
    """-> (energy, Hdd, bd); tmp = [state, energy, new_state, new_energy] of this target, updated in place"""
    if tmp[0] == ST_OOB:
        tmp[2] = ST_OOB
        return tmp[1], Hdd, bd
    R, t, aff, dI, fx, fy, cx, cy, fxi, fyi, w, h = T
    energy = F(0)
    for idx in range(8):
        dx, dy = int(PATTERN[idx][0]), int(PATTERN[idx][1])
        k0, k1 = F(F(P["u"] + dx - cx) * fxi), F(F(P["v"] + dy - cy) * fyi)
        ptp = [F(F(F(R[r, 0] * k0) + F(R[r, 1] * k1)) + F(R[r, 2] * F(1))) + F(t[r] * idepth) for r in range(3)]
        ptp = [F(x) for x in ptp]
        drescale = F(F(1) / ptp[2])
        if not (drescale > 0):
            tmp[2] = ST_OOB
            return tmp[1], Hdd, bd
        u, v = F(ptp[0] * drescale), F(ptp[1] * drescale)
        Ku, Kv = F(F(u * fx) + cx), F(F(v * fy) + cy)
        if not (Ku > F(1.1) and Kv > F(1.1) and Ku < w - 3 and Kv < h - 3):
            tmp[2] = ST_OOB
            return tmp[1], Hdd, bd
        hit = lerp3(dI, w, np.array([Ku], F), np.array([Kv], F))[0]
        if not np.isfinite(hit[0]):
            tmp[2] = ST_OOB
            return tmp[1], Hdd, bd
        residual = F(hit[0] - F(F(aff[0] * P["color"][idx]) + aff[1]))
        hw = F(1) if abs(residual) < HUBER_TH else F(HUBER_TH / abs(residual))
        wgt = P["weights"][idx]
        energy = F(energy + F(F(F(F(F(wgt * wgt) * hw) * residual) * residual) * F(F(2) - hw)))
        dxi, dyi = F(hit[1] * fx), F(hit[2] * fy)
        d_idepth = F(F(F(dxi * drescale) * F(t[0] - F(t[2] * u))) + F(F(dyi * drescale) * F(t[1] - F(t[2] * v))))   # * SCALE_IDEPTH = 1
        hw = F(hw * F(wgt * wgt))
        Hdd = F(Hdd + F(F(hw * d_idepth) * d_idepth))
        bd = F(bd + F(F(hw * residual) * d_idepth))
    lim = F(P["energyTH"] * F(slack))
    if energy > lim:
        energy = lim
        tmp[2] = ST_OUTLIER
    else:
        tmp[2] = IN
    tmp[3] = energy
    return energy, Hdd, bd


def optimize_immature_point(P, host, targets, min_obs):
    """-> (idepth or None, final residual states per frame (-1 for the host))"""
    n = len(targets)
    tmp = {f: [IN, F(0), ST_OUTLIER, F(0)] for f in range(n) if f != host}
    lastE = lastH = lastb = F(0)
    cur = F(F(P["idepth_max"] + P["idepth_min"]) * F(0.5))
    for f in tmp:
        e, lastH, lastb = linearize_residual(P, targets[f], 1000, tmp[f], lastH, lastb, cur)
        lastE = F(lastE + e)
        tmp[f][0], tmp[f][1] = tmp[f][2], tmp[f][3]
    if not np.isfinite(lastE) or lastH < MIN_IDEPTH_H_ACT:
        return None, None
    lam = F(0.1)
    for it in range(GN_ITS_ACT):
        H = F(lastH * F(F(1) + lam))
        step = F(np.float64(1.0) / np.float64(H) * np.float64(lastb))
        new_id = F(cur - step)
        newH = newb = newE = F(0)
        for f in tmp:
            e, newH, newb = linearize_residual(P, targets[f], 1, tmp[f], newH, newb, new_id)
            newE = F(newE + e)
        if not np.isfinite(lastE) or newH < MIN_IDEPTH_H_ACT:
            return None, None
        if newE < lastE:
            cur, lastH, lastb, lastE = new_id, newH, newb, newE
            for f in tmp:
                tmp[f][0], tmp[f][1] = tmp[f][2], tmp[f][3]
            lam = F(lam * F(0.5))
        else:
            lam = F(lam * F(5))
        if abs(step) < np.float64(0.0001) * np.float64(cur):
            break
    if not np.isfinite(cur):
        return None, None
    states = [-1 if f == host else tmp[f][0] for f in range(n)]
    if sum(1 for s in states if s == IN) < min_obs:
        return None, states
    return cur, states


def test_oracle_activation_matches_the_numpy_witness(oracle):
    from ldso_b200 import window_host as wh
    S = util.immature_scene("small", pts_per_host=90)
    c = S["calib"]
    w, h = c["w"], c["h"]
    ctx = util.make_ctx(oracle, "small", n_frame_slots=8)
    res = util.run_immature(ctx, S, activate=True)
    nf = len(S["frames"])
    dIs = [ctx.download(i, 0, 0).reshape(-1, 3).astype(F) for i in range(nf)]
    pc = wh.precalc_table(S["w2c"], S["w2c"], (c["fx"], c["fy"], c["cx"], c["cy"]), [(0.0, 0.0)] * nf, [(0.0, 0.0)] * nf, [1.0] * nf)
    fx, fy, cx, cy = F(c["fx"]), F(c["fy"]), F(c["cx"]), F(c["cy"])
    fxi, fyi = F(F(1) / fx), F(F(1) / fy)
    last = res["traces"][-1]
    idepth_o, ok_o, st_o = res["act"]
    cand = res["cand"]
    assert len(cand) > 60
    n_ok = n_equal = 0
    for q, pt in enumerate(cand):
        host = int(S["host"][pt])
        P = dict(u=F(S["u"][pt]), v=F(S["v"][pt]), color=res["color"][pt], weights=res["weights"][pt], energyTH=res["energy_th"][pt],
                 idepth_min=last["idepth_min"][pt], idepth_max=last["idepth_max"][pt])
        targets = []
        for f in range(nf):
            e = pc[host + nf * f]
            targets.append((e[27:36].reshape(3, 3), e[36:39], e[24:26], dIs[f], fx, fy, cx, cy, fxi, fyi, w, h))
        idp, states = optimize_immature_point(P, host, targets, 1)
        if ok_o[q]:
            n_ok += 1
            assert idp is not None and states == st_o[q].tolist(), (q, states, st_o[q])
            assert idp == idepth_o[q] or abs(idp - idepth_o[q]) <= 2e-6 * abs(idepth_o[q]), (q, idp, idepth_o[q])
            n_equal += idp == idepth_o[q]
        else:
            assert idp is None, (q, idp)
    assert n_ok > 40 and n_equal > 0.95 * n_ok, (n_ok, n_equal)
    ctx.close()
