"""Second, independent witness for A2: PixelSelector's constructor, makeHists, select and makeMaps
(src/frontend/PixelSelector2.cc:9-316) written out in scalar float32 Python straight from the reference's text — glibc's own
rand() for randomPattern (through ctypes), the 32x32-block gradient histograms and their smoothed thresholds, the four nested
potential grids with their first-maximum rule and level fallbacks, the potential recursion and the random thinning — and
compared bit for bit with the oracle's selection map on the small calibration (pure Python loops: small images only).
Shares no code with oracle/."""
import ctypes

import numpy as np

from tests import util

F = np.float32
DIRS = np.array([(0, 1.0), (0.3827, 0.9239), (0.1951, 0.9808), (0.9239, 0.3827), (0.7071, 0.7071), (0.3827, -0.9239),
                 (0.8315, 0.5556), (0.8315, -0.5556), (0.5556, -0.8315), (0.9808, 0.1951), (0.9239, -0.3827), (0.7071, -0.7071),
                 (0.5556, 0.8315), (0.9808, -0.1951), (1.0, 0.0), (0.1951, -0.9808)], F)
MIN_GRAD_HIST_CUT, MIN_GRAD_HIST_ADD, DOWNWEIGHT = F(0.5), F(7), F(0.75)     # src/Setting.cc:83-86


def random_pattern(n):
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(3141592)                                   # PixelSelector::PixelSelector (:11-13)
    return np.array([libc.rand() & 0xFF for _ in range(n)], np.uint8)


def make_hists(abs0, w, h):
    """thsSmoothed [h32][w32] (:36-109); the histogram bins the reference never initialises (50..90) count as empty"""
    w32, h32 = w // 32, h // 32
    ths = np.zeros((h32, w32), F)
    for y in range(h32):
        for x in range(w32):
            hist = np.zeros(92, np.int64)
            for j in range(32):
                for i in range(32):
                    it, jt = i + 32 * x, j + 32 * y
                    if it > w - 2 or jt > h - 2 or it < 1 or jt < 1:
                        continue
                    g = min(int(np.sqrt(F(abs0[jt, it]))), 48)
                    hist[g + 1] += 1
                    hist[0] += 1
            th = int(F(hist[0]) * MIN_GRAD_HIST_CUT + F(0.5))
            q = 90
            for i in range(90):
                th -= hist[i + 1]
                if th < 0:
                    q = i
                    break
            ths[y, x] = F(q) + MIN_GRAD_HIST_ADD
    sm = np.zeros((h32, w32), F)
    for y in range(h32):
        for x in range(w32):
            s, num = F(0), F(0)
            for dx, dy in ((-1, -1), (-1, 1), (-1, 0), (1, -1), (1, 1), (1, 0), (0, -1), (0, 1), (0, 0)):   # the reference's order
                xx, yy = x + dx, y + dy
                if 0 <= xx < w32 and 0 <= yy < h32:
                    num = F(num + 1); s = F(s + ths[yy, xx])
            sm[y, x] = F((s / num) * (s / num))
    return sm


def select(dI, abs0, abs1, abs2, ths, rp, w, h, pot, th_factor):
    """PixelSelector::select (:171-316) -> (map, n2, n3, n4)"""
    w1, w2 = w // 2, w // 4
    out = np.zeros(w * h, F)
    a0, a1, a2 = abs0.reshape(-1), abs1.reshape(-1), abs2.reshape(-1)
    gx, gy = dI[..., 1].reshape(-1), dI[..., 2].reshape(-1)
    dw1 = DOWNWEIGHT; dw2 = F(dw1 * dw1)
    thf = F(th_factor)
    n2 = n3 = n4 = 0
    for y4 in range(0, h, 4 * pot):
        for x4 in range(0, w, 4 * pot):
            my3, mx3 = min(4 * pot, h - y4), min(4 * pot, w - x4)
            best4, val4 = -1, F(0)
            d4 = DIRS[rp[n2] & 0xF]
            for y3 in range(0, my3, 2 * pot):
                for x3 in range(0, mx3, 2 * pot):
                    x34, y34 = x3 + x4, y3 + y4
                    my2, mx2 = min(2 * pot, h - y34), min(2 * pot, w - x34)
                    best3, val3 = -1, F(0)
                    d3 = DIRS[rp[n2] & 0xF]
                    for y2 in range(0, my2, pot):
                        for x2 in range(0, mx2, pot):
                            x234, y234 = x2 + x34, y2 + y34
                            my1, mx1 = min(pot, h - y234), min(pot, w - x234)
                            best2, val2 = -1, F(0)
                            d2 = DIRS[rp[n2] & 0xF]
                            for y1 in range(my1):
                                for x1 in range(mx1):
                                    xf, yf = x1 + x234, y1 + y234
                                    idx = xf + w * yf
                                    if xf < 4 or xf >= w - 5 or yf < 4 or yf > h - 4:
                                        continue
                                    t0 = ths[yf >> 5, xf >> 5]
                                    t1 = F(t0 * dw1); t2 = F(t1 * dw2)
                                    if a0[idx] > F(t0 * thf):
                                        dn = abs(F(F(gx[idx] * d2[0]) + F(gy[idx] * d2[1])))
                                        if dn > val2:
                                            val2, best2, best3, best4 = dn, idx, -2, -2
                                    if best3 == -2:
                                        continue
                                    if a1[int(F(xf * F(0.5)) + F(0.25)) + int(F(yf * F(0.5)) + F(0.25)) * w1] > F(t1 * thf):
                                        dn = abs(F(F(gx[idx] * d3[0]) + F(gy[idx] * d3[1])))
                                        if dn > val3:
                                            val3, best3, best4 = dn, idx, -2
                                    if best4 == -2:
                                        continue
                                    if a2[int(xf * 0.25 + 0.125) + int(yf * 0.25 + 0.125) * w2] > F(t2 * thf):
                                        dn = abs(F(F(gx[idx] * d4[0]) + F(gy[idx] * d4[1])))
                                        if dn > val4:
                                            val4, best4 = dn, idx
                            if best2 > 0:
                                out[best2] = 1; val3 = F(1e10); n2 += 1
                    if best3 > 0:
                        out[best3] = 2; val4 = F(1e10); n3 += 1
            if best4 > 0:
                out[best4] = 4; n4 += 1
    return out, n2, n3, n4


def make_maps(dI, absg, ths, rp, w, h, density, potential, recursions, th_factor=1.0):
    """PixelSelector::makeMaps (:111-169) -> (map, numHaveSub, new currentPotential)"""
    while True:
        m, n2, n3, n4 = select(dI, absg[0], absg[1], absg[2], ths, rp, w, h, potential, th_factor)
        num_have = F(n2 + n3 + n4)
        quotia = F(F(density) / num_have)
        K = F(F(num_have * F(potential + 1)) * F(potential + 1))
        ideal = max(int(F(np.sqrt(F(K / F(density)))) - F(1)), 1)
        if recursions > 0 and quotia > 1.25 and potential > 1:
            potential = potential - 1 if ideal >= potential else ideal
            recursions -= 1
            continue
        if recursions > 0 and quotia < 0.25:
            potential = potential + 1 if ideal <= potential else ideal
            recursions -= 1
            continue
        break
    have = int(num_have)
    if quotia < 0.95:
        char_th = int(F(255) * quotia) & 0xFF
        rn = 0
        for i in np.nonzero(m)[0]:
            if rp[rn] > char_th:
                m[i] = 0; have -= 1
            rn += 1
    return m.reshape(h, w), have, ideal


def _images():
    img = util.frames("small")[0][0]
    # a low-texture image: flat areas and wide soft ramps that only the coarser pyramid levels see (selection types 2 and 4)
    rng = np.random.default_rng(3)
    soft = np.full(img.shape, 100.0, F)
    soft[:, 60:] += 40.0
    soft[80:, :] += 30.0
    soft[20:50, 90:150] += np.floor(rng.uniform(0, 60, (30, 60))).astype(F)
    for _ in range(6):
        soft[1:-1, 1:-1] = 0.2 * (soft[1:-1, 1:-1] + soft[:-2, 1:-1] + soft[2:, 1:-1] + soft[1:-1, :-2] + soft[1:-1, 2:])
    return [("textured", img), ("soft", np.floor(soft).astype(F))]


def test_oracle_selection_maps_match_the_python_witness(oracle):
    c = util.CALIB["small"]
    w, h = c["w"], c["h"]
    rp = random_pattern(w * h)
    seen_types = set()
    for name, img in _images():
        ctx = util.make_ctx(oracle, "small")
        ctx.make_images(0, img)
        dI = ctx.download(0, 0, 0)
        absg = [ctx.download(0, l, 1) for l in range(3)]
        ths = make_hists(absg[0], w, h)
        pot_o = pot_w = 3
        for density, rec in ((300.0, 1), (300.0, 1), (1500.0, 1), (60.0, 1), (400.0, 0)):     # the potential carries over between calls
            mo, no, pot_o = ctx.select_make_maps(0, density, potential=pot_o, recursions=rec)
            mw, nw, pot_w = make_maps(dI, absg, ths, rp, w, h, density, pot_w, rec)
            assert (no, pot_o) == (nw, pot_w), (name, density, no, nw, pot_o, pot_w)
            assert np.array_equal(mo, mw), (name, density, int((mo != mw).sum()))
            seen_types |= set(np.unique(mo).tolist())
        ctx.close()
    assert {0.0, 1.0, 2.0}.issubset(seen_types) and (4.0 in seen_types or 2.0 in seen_types)
