"""Second, independent witness for K2: CoarseTracker::makeCoarseDepthL0 (src/frontend/CoarseTracker.cc:357-538) written out in
vectorised float32 numpy from the reference's text — weighted scatter of the active points' projections, 2x2 pooling down the
pyramid, one dilation per level (diagonal neighbours on levels 0 and 1, direct neighbours above), normalisation and the row-major
point-cloud lists — and compared with the oracle through the C ABI.  Shares no code with oracle/."""
import numpy as np
import pytest

from tests import util

F = np.float32


def coarse_depth_l0(w, h, levels, u, v, new_idepth, hdi, color_of_level):
    """-> per level (pc_u, pc_v, pc_idepth, pc_color) as the reference fills them"""
    ws_, hs_ = [w >> l for l in range(levels)], [h >> l for l in range(levels)]
    idepth = [np.zeros(ws_[l] * hs_[l], F) for l in range(levels)]
    wsum = [np.zeros(ws_[l] * hs_[l], F) for l in range(levels)]
    # :364-383  weighted scatter, in list order
    weight = np.sqrt(F(1e-3) / (hdi.astype(F) + F(1e-12))).astype(F)
    cell = (u + F(0.5)).astype(np.int32) + ws_[0] * (v + F(0.5)).astype(np.int32)
    np.add.at(idepth[0], cell, (new_idepth.astype(F) * weight).astype(F))
    np.add.at(wsum[0], cell, weight)
    # :385-412  2x2 sums, ((a + b) + c) + d
    for l in range(1, levels):
        for dst, src in ((idepth, idepth), (wsum, wsum)):
            m = src[l - 1].reshape(hs_[l - 1], ws_[l - 1])
            a, b = m[0:2 * hs_[l]:2, 0:2 * ws_[l]:2], m[0:2 * hs_[l]:2, 1:2 * ws_[l]:2]
            c_, d = m[1:2 * hs_[l]:2, 0:2 * ws_[l]:2], m[1:2 * hs_[l]:2, 1:2 * ws_[l]:2]
            dst[l] = (((a + b).astype(F) + c_).astype(F) + d).astype(F).reshape(-1)
    # :414-499  dilation: cells without weight take the mean of their (diagonal / direct) neighbours that have one
    for l in range(levels):
        wl, n = ws_[l], ws_[l] * hs_[l]
        offs = (1 + wl, -1 - wl, wl - 1, -wl + 1) if l < 2 else (1, -1, wl, -wl)
        bak = wsum[l].copy()
        src_id = idepth[l].copy()          # only cells with bak > 0 are read, only cells with bak <= 0 written: a copy is equivalent
        i = np.arange(wl, n - wl)
        s = np.zeros(len(i), F); num = np.zeros(len(i), F); numn = np.zeros(len(i), F)
        for o in offs:
            j = i + o
            inside = (j >= 0) & (j < n)    # the reference reads one element past its planes for two cells that never reach the cloud
            jj = np.clip(j, 0, n - 1)
            has = inside & (bak[jj] > 0)
            s = np.where(has, (s + src_id[jj]).astype(F), s)
            num = np.where(has, (num + bak[jj]).astype(F), num)
            numn = np.where(has, numn + F(1), numn)
        fill = (bak[i] <= 0) & (numn > 0)
        with np.errstate(divide="ignore", invalid="ignore"):
            idepth[l][i] = np.where(fill, (s / numn).astype(F), idepth[l][i])
            wsum[l][i] = np.where(fill, (num / numn).astype(F), wsum[l][i])
    # :501-537  normalise, list the interior cells row by row
    out = []
    for l in range(levels):
        wl, hl = ws_[l], hs_[l]
        ys, xs = np.mgrid[2:hl - 2, 2:wl - 2]
        i = (xs + ys * wl).reshape(-1)
        with np.errstate(divide="ignore", invalid="ignore"):
            idn = (idepth[l][i] / wsum[l][i]).astype(F)
        col = color_of_level[l].reshape(-1)[i]
        ok = (wsum[l][i] > 0) & np.isfinite(col) & (idn > 0)
        out.append((xs.reshape(-1)[ok].astype(F), ys.reshape(-1)[ok].astype(F), idn[ok], col[ok]))
    return out


@pytest.mark.parametrize("calib,npts,idx", [("small", 400, (40, 42)), ("c1", 2000, (40, 42)), ("c3", 2000, (40, 41))])
def test_oracle_make_coarse_depth_matches_the_numpy_witness(oracle, calib, npts, idx):
    c = util.CALIB[calib]
    ctx = util.make_ctx(oracle, calib)
    (img0, dep0, _, _), _ = util.frames(calib, 1001, idx)
    u, v, idepth, hdi = util.pick_points(img0, dep0, npts)
    # the same point several times and points next to each other: weighted sums in one cell, dense neighbourhoods for the dilation
    u = np.concatenate([u, u[:50], u[:50] + 1]).astype(F); v = np.concatenate([v, v[:50], v[:50]]).astype(F)
    idepth = np.concatenate([idepth, idepth[:50] * F(1.1), idepth[:50]]).astype(F); hdi = np.concatenate([hdi, hdi[:50] * F(2), hdi[:50]]).astype(F)
    util.setup_tracker(ctx, calib, npts, idx)
    ctx.tracker_set_ref(0, 0, u, v, idepth, hdi)
    colors = [ctx.download(0, l, 0)[..., 0] for l in range(ctx.levels)]
    wit = coarse_depth_l0(c["w"], c["h"], ctx.levels, u, v, idepth, hdi, colors)
    for l in range(ctx.levels):
        pu, pv, pid, pcol = ctx.tracker_get_pc(0, l)
        wu, wv, wid, wcol = wit[l]
        assert len(pu) == len(wu) > 50, (l, len(pu), len(wu))
        assert np.array_equal(pu, wu) and np.array_equal(pv, wv) and np.array_equal(pcol, wcol)
        assert np.allclose(pid, wid, rtol=2e-6, atol=0), (l, np.abs(pid - wid).max())
    ctx.close()
