"""N2: CoarseDistanceMap (CoarseTracker.cc:795-930) and the candidate walk of activatePointsMT (FullSystem.cc:1226-1298).
CPU tests pin the oracle's list-based BFS against an independent numpy formulation (ring dilation) and a pure-Python
walk; the GPU tests compare the CUDA kernels with the oracle bit for bit."""
import numpy as np
import pytest

from tests import util


def level1_tables(S, newest):
    """KRKi = K[1] R Ki[0], Kt = K[1] t of host -> frame `newest` (FullSystem.cc:1236-1238)"""
    c = S["calib"]
    K1 = np.array([[c["fx"] * 0.5, 0, (c["cx"] + 0.5) / 2 - 0.5], [0, c["fy"] * 0.5, (c["cy"] + 0.5) / 2 - 0.5], [0, 0, 1]], np.float32)
    K0 = np.array([[c["fx"], 0, c["cx"]], [0, c["fy"], c["cy"]], [0, 0, 1]], np.float64)
    Ki0 = np.linalg.inv(K0).astype(np.float32)
    fr = S["frames"]
    _, _, Rn, tn = fr[newest]
    KRKi, Kt = [], []
    for hi in range(S["n_hosts"]):
        _, _, Rh, th = fr[hi]
        R = (Rn.T @ Rh).astype(np.float32); t = (Rn.T @ (th - tn)).astype(np.float32)
        KRKi.append(K1 @ R @ Ki0); Kt.append(K1 @ t)
    return np.array(KRKi, np.float32), np.array(Kt, np.float32)


def np_distmap(w1, h1, cells):
    """ring k: every cell still above k with a non-border neighbour at k - 1 (8-connected for odd k, 4 for even k)"""
    D = np.full((h1, w1), 1000.0, np.float32)
    for x, y in cells:
        D[y, x] = 0
    inner = np.zeros((h1, w1), bool); inner[1:-1, 1:-1] = True
    for k in range(1, 40):
        src = (D == k - 1) & inner
        grow = np.zeros_like(src)
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if (dx == 0 and dy == 0) or (k % 2 == 0 and dx != 0 and dy != 0):
                    continue
                sh = np.zeros_like(src)
                ys, yd = (slice(0, h1 - 1), slice(1, h1)) if dy == 1 else (slice(1, h1), slice(0, h1 - 1)) if dy == -1 else (slice(None), slice(None))
                xs, xd = (slice(0, w1 - 1), slice(1, w1)) if dx == 1 else (slice(1, w1), slice(0, w1 - 1)) if dx == -1 else (slice(None), slice(None))
                sh[yd, xd] = src[ys, xs]
                grow |= sh
        D[grow & (D > k)] = k
    return D


def project_cells(KRKi, Kt, host, u, v, idp, w1, h1):
    out = []
    for i in range(len(u)):
        K, t = KRKi[host[i]], Kt[host[i]]
        p = [np.float32(np.float32(np.float32(K[r, 0] * u[i]) + np.float32(K[r, 1] * v[i])) + K[r, 2]) + np.float32(t[r] * idp[i]) for r in range(3)]
        p = [np.float32(x) for x in p]
        x, y = int(np.float32(p[0] / p[2]) + np.float32(0.5)), int(np.float32(p[1] / p[2]) + np.float32(0.5))
        if 0 < x < w1 and 0 < y < h1:
            out.append((x, y))
    return out


def scene_with_traces(ctx, S):
    out = util.run_immature(ctx, S, activate=False)
    return out


def active_subset(S, every=5):
    m = np.arange(len(S["u"])) % every == 0
    return S["host"][m], S["u"][m], S["v"][m], S["gt_idepth"][m]


def test_single_source_is_the_octagonal_metric(oracle):
    ctx = util.make_ctx(oracle, "small")
    w1, h1 = ctx.sizes[1]
    K = np.eye(3, dtype=np.float32)[None]
    cx, cy = w1 // 2, h1 // 2
    ctx.distmap_make(K, np.zeros((1, 3), np.float32), [0], [cx], [cy], [1.0])
    D = ctx.distmap_get()
    yy, xx = np.mgrid[0:h1, 0:w1]
    a, b = np.maximum(abs(xx - cx), abs(yy - cy)), np.minimum(abs(xx - cx), abs(yy - cy))
    exp = np.full((h1, w1), 1000.0, np.float32)
    for k in range(39, -1, -1):                      # smallest k with a <= k and a + b <= k + ceil(k / 2)
        exp[(a <= k) & (a + b <= k + (k + 1) // 2)] = k
    assert np.array_equal(D, exp)
    ctx.close()


def test_oracle_distance_map_matches_ring_dilation(oracle):
    S = util.immature_scene("small")
    ctx = util.make_ctx(oracle, "small")
    w1, h1 = ctx.sizes[1]
    KRKi, Kt = level1_tables(S, S["n_hosts"] + 2)
    h, u, v, idp = active_subset(S, 7)
    ctx.distmap_make(KRKi, Kt, h, u, v, idp)
    D = ctx.distmap_get()
    cells = project_cells(KRKi, Kt, h, u, v, idp, w1, h1)
    assert len(cells) > 20
    assert np.array_equal(D, np_distmap(w1, h1, cells))
    # border sources do not propagate, and points of the newest frame (host -1) are ignored
    ctx.distmap_make(np.eye(3, dtype=np.float32)[None], np.zeros((1, 3), np.float32), [0, 0, -1], [0.6, w1 - 1, 20], [5, 7, 20], [1, 1, 1])
    D = ctx.distmap_get()
    assert D[5, 1] == 0 and D[7, w1 - 1] == 0 and D[7, w1 - 2] == 1000 and D[20, 20] != 0
    ctx.close()


def py_walk(D, P, KRKi, Kt, skip, flagged, min_dist, min_q=3.0):
    """pure-Python restatement of FullSystem.cc:1226-1298 with set-based BFS rings"""
    h1, w1 = D.shape
    D = D.copy()
    act = np.zeros(len(P["host"]), np.uint8)
    f32 = np.float32
    for i in range(len(P["host"])):
        h = P["host"][i]
        if skip[h]:
            continue
        st, imax, imin = P["status"][i], P["idepth_max"][i], P["idepth_min"][i]
        if not np.isfinite(imax) or st == 2:
            act[i] = 1; continue
        can = st in (0, 3, 4, 1) and P["interval"][i] < 8 and P["quality"][i] > min_q and f32(imax + imin) > 0
        if not can:
            if flagged[h] or st == 1:
                act[i] = 1
            continue
        K, t = KRKi[h], Kt[h]
        idp = f32(f32(0.5) * f32(imax + imin))
        p = [f32(f32(f32(f32(K[r, 0] * P["u"][i]) + f32(K[r, 1] * P["v"][i])) + K[r, 2]) + f32(t[r] * idp)) for r in range(3)]
        x, y = int(f32(p[0] / p[2]) + f32(0.5)), int(f32(p[1] / p[2]) + f32(0.5))
        if not (0 < x < w1 and 0 < y < h1):
            act[i] = 1; continue
        dist = f32(D[y, x] + f32(p[0] - np.floor(p[0])))
        if dist >= f32(f32(min_dist) * P["type"][i]):
            act[i] = 2
            D[y, x] = 0
            front = [(x, y)]
            for k in range(1, 40):
                nxt = []
                for (a, b) in front:
                    if a == 0 or b == 0 or a == w1 - 1 or b == h1 - 1:
                        continue
                    nb = [(a + 1, b), (a - 1, b), (a, b + 1), (a, b - 1)]
                    if k % 2:
                        nb += [(a + 1, b + 1), (a - 1, b + 1), (a - 1, b - 1), (a + 1, b - 1)]
                    for (c, d) in nb:
                        if D[d, c] > k:
                            D[d, c] = k; nxt.append((c, d))
                front = nxt
                if not front:
                    break
    return act, D


def walk_inputs(ctx, S, newest):
    t = ctx.immature_get()
    n = len(S["u"])
    types = np.array([1, 2, 4, 1, 1, 2], np.float32)[np.arange(n) % 6]
    ctx.immature_set_types(0, types)
    P = dict(host=S["host"], u=S["u"], v=S["v"], status=t["status"], idepth_min=t["idepth_min"], idepth_max=t["idepth_max"],
             quality=t["quality"], interval=t["interval"], type=types)
    KRKi, Kt = level1_tables(S, newest)
    return P, KRKi, Kt


@pytest.mark.parametrize("min_dist,flag", [(2.0, 0), (0.5, 1), (4.0, 0)])
def test_oracle_walk_matches_python(oracle, min_dist, flag):
    S = util.immature_scene("small", pts_per_host=250)
    ctx = util.make_ctx(oracle, "small")
    scene_with_traces(ctx, S)
    newest = S["n_hosts"] + 2
    P, KRKi, Kt = walk_inputs(ctx, S, newest)
    h, u, v, idp = active_subset(S, 9)
    ctx.distmap_make(KRKi, Kt, h, u, v, idp)
    D0 = ctx.distmap_get()
    skip, flagged = np.zeros(S["n_hosts"], np.uint8), np.zeros(S["n_hosts"], np.uint8)
    flagged[0] = flag
    act, cnt = ctx.immature_select_activation(KRKi, Kt, skip, flagged, min_dist)
    ref, Dref = py_walk(D0, P, KRKi, Kt, skip, flagged, min_dist)
    assert np.array_equal(act, ref)
    assert cnt == (int((ref == 2).sum()), int((ref == 1).sum())) and cnt[0] > 10
    assert np.array_equal(ctx.distmap_get(), Dref)
    # the newest frame's own points are not visited
    skip[1] = 1
    ctx.distmap_make(KRKi, Kt, h, u, v, idp)
    act2, _ = ctx.immature_select_activation(KRKi, Kt, skip, flagged, min_dist)
    assert np.all(act2[S["host"] == 1] == 0)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib,pts", [("small", 250), ("c1", 1500)])
def test_gpu_distance_map_and_walk_bit_exact(gpu, oracle, calib, pts):
    S = util.immature_scene(calib, pts_per_host=pts)
    g, o = util.make_ctx(gpu, calib), util.make_ctx(oracle, calib)
    newest = S["n_hosts"] + 2
    res = []
    for ctx in (g, o):
        scene_with_traces(ctx, S)
        P, KRKi, Kt = walk_inputs(ctx, S, newest)
        h, u, v, idp = active_subset(S, 9)
        ctx.distmap_make(KRKi, Kt, h, u, v, idp)
        D0 = ctx.distmap_get()
        out = [D0]
        for md in (2.0, 0.7):
            skip, flagged = np.zeros(S["n_hosts"], np.uint8), np.array([1] + [0] * (S["n_hosts"] - 1), np.uint8)
            act, cnt = ctx.immature_select_activation(KRKi, Kt, skip, flagged, md)
            out += [act, np.array(cnt), ctx.distmap_get()]
        res.append(out)
    for a, b in zip(*res):
        assert np.array_equal(a, b)
    assert res[1][2][0] > 20 and (res[1][0] < 1000).mean() > 0.5
    # pool maintenance keeps my_type with its point
    for ctx in (g, o):
        ctx.immature_compact(drop_host=-1, drop_oob=False, remove=np.nonzero(res[1][1] != 0)[0])
    h, u, v, idp = active_subset(S, 9)
    outs = []
    for ctx in (g, o):
        ctx.distmap_make(KRKi, Kt, h, u, v, idp)
        outs.append(ctx.immature_select_activation(KRKi, Kt, np.zeros(S["n_hosts"], np.uint8), np.zeros(S["n_hosts"], np.uint8), 0.3))
    assert np.array_equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]
    g.close(); o.close()
