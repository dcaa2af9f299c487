"""K5 (ImmaturePoint::traceOn) and K6 (ImmaturePoint ctor, optimizeImmaturePoint): oracle sanity on CPU, GPU parity.

Reference: src/internal/ImmaturePoint.cc:14-389, src/frontend/FullSystem.cc:1035-1195."""
import numpy as np
import pytest

from tests import util

GOOD, OOB, OUTLIER, SKIPPED, BADCONDITION, UNINIT = range(6)


def _run(api, calib="small", **kw):
    S = util.immature_scene(calib, **kw)
    ctx = util.make_ctx(api, calib, n_frame_slots=8)
    return S, util.run_immature(ctx, S)


def test_oracle_ctor_matches_numpy(oracle):
    S, o = _run(oracle)
    img = S["frames"][0][0].astype(np.float32)
    m = S["host"] == 0
    u, v = S["u"][m].astype(int), S["v"][m].astype(int)
    # integer positions: the bilinear sample is the pixel itself (K1 stores the image unchanged at level 0)
    for k, (dx, dy) in enumerate([(0, -2), (-1, -1), (1, -1), (-2, 0), (0, 0), (2, 0), (-1, 1), (0, 2)]):
        np.testing.assert_array_equal(o["color"][m][:, k], img[v + dy, u + dx])
    assert np.all(o["energy_th"] == 8 * 12 * 12)          # patternNum * setting_outlierTH * overallEnergyTHWeight^2
    assert np.all((o["weights"] > 0) & (o["weights"] <= 1))
    assert np.all(o["grad_h"][:, 0] >= 0) and np.all(o["grad_h"][:, 2] >= 0)


def test_oracle_trace_brackets_ground_truth(oracle):
    S, o = _run(oracle)
    gt = S["gt_idepth"]
    widths = []
    for cnt, t in zip(o["counts"], o["traces"]):
        assert cnt.sum() == len(gt)
        assert np.array_equal(np.bincount(t["status"], minlength=6), cnt)
        g = t["status"] == GOOD
        assert g.sum() > 0.5 * len(gt)
        inside = (t["idepth_min"][g] <= gt[g]) & (gt[g] <= t["idepth_max"][g])
        assert inside.mean() > 0.85
        widths.append(np.median(t["idepth_max"][g] - t["idepth_min"][g]))
    assert widths[-1] < widths[0]                         # the depth interval shrinks as the baseline grows
    # OOB is absorbing (ImmaturePoint.cc:53)
    for a, b in zip(o["traces"][:-1], o["traces"][1:]):
        assert np.all(b["status"][a["status"] == OOB] == OOB)


def test_oracle_activation_recovers_depth(oracle):
    S, o = _run(oracle)
    idp, ok, st = o["act"]
    cand = o["cand"]
    assert ok.sum() > 0.9 * len(cand)
    gt = S["gt_idepth"][cand]
    rel = np.abs(idp[ok == 1] - gt[ok == 1]) / gt[ok == 1]
    assert np.median(rel) < 0.15
    # the host frame's own column is -1, every other entry of an accepted point is a ResState
    for q in np.nonzero(ok == 1)[0][:50]:
        h = S["host"][cand[q]]
        assert st[q, h] == -1 and np.all(np.delete(st[q], h) >= 0)
    assert np.all(idp[ok == 0] == 0)


def test_oracle_trace_edge_cases(oracle):
    """empty pool; points whose patch leaves the image; a point already OOB; an idepth interval too narrow (SKIPPED)."""
    S = util.immature_scene("small", pts_per_host=40)
    ctx = util.make_ctx(oracle, "small", n_frame_slots=8)
    for i, f in enumerate(S["frames"]):
        ctx.make_images(i, f[0])
    KRKi, Kt, aff = S["tables"][0]
    z = np.zeros(0, np.float32)
    ctx.immature_set(np.zeros(0, np.int32), z, z, np.zeros((0, 8)), np.zeros((0, 8)), np.zeros((0, 3)), z, z, z, z, np.zeros(0, np.int32))
    assert ctx.immature_trace(2, KRKi, Kt, aff).sum() == 0
    m = S["host"] == 0
    u, v, gt = S["u"][m][:6].copy(), S["v"][m][:6].copy(), S["gt_idepth"][m][:6]
    col, wgt, gh, eth = ctx.immature_create(0, u, v)
    n = len(u)
    idmin = np.zeros(n, np.float32); idmax = np.full(n, np.nan, np.float32); status = np.full(n, UNINIT, np.int32)
    status[0] = OOB                                       # stays OOB, untouched
    idmin[1] = gt[1] * 0.999; idmax[1] = gt[1] * 1.001    # < setting_trace_slackInterval px -> SKIPPED
    idmin[2] = 50.0; idmax[2] = 60.0                      # projects far outside -> OOB
    ctx.immature_set(np.zeros(n, np.int32), u, v, col, wgt, gh, eth, idmin, idmax, np.full(n, 1e4, np.float32), status)
    cnt = ctx.immature_trace(2, KRKi, Kt, aff)
    t = ctx.immature_get()
    assert t["status"][0] == OOB and t["status"][1] == SKIPPED and t["status"][2] == OOB
    assert t["idepth_min"][1] == idmin[1] and t["idepth_max"][1] == idmax[1]
    assert np.all(t["uv"][2] == -1) and t["interval"][2] == 0
    assert cnt.sum() == n


@pytest.mark.gpu
@pytest.mark.parametrize("calib,kw", [("small", {}), ("c1", dict(pts_per_host=1500, host_ids=(40, 44, 48), new_ids=(50, 52, 54)))])
def test_gpu_immature_parity(oracle, gpu, calib, kw):
    S, a = _run(oracle, calib, **kw)
    _, b = _run(gpu, calib, **kw)
    for key in ("color", "weights", "grad_h", "energy_th"):
        assert np.array_equal(a[key], b[key], equal_nan=True), key
    for ca, cb, ta, tb in zip(a["counts"], b["counts"], a["traces"], b["traces"]):
        assert np.array_equal(ca, cb)
        assert np.array_equal(ta["status"], tb["status"])
        for key in ("idepth_min", "idepth_max", "quality", "uv", "interval"):
            assert np.array_equal(ta[key], tb[key], equal_nan=True), key   # strict float policy: bit-exact per point
    assert np.array_equal(a["cand"], b["cand"])
    ia, oka, sa = a["act"]
    ib, okb, sb = b["act"]
    assert np.array_equal(oka, okb) and np.array_equal(sa, sb)
    assert np.array_equal(ia, ib)


def _run_async(api, calib="small", **kw):
    S = util.immature_scene(calib, **kw)
    ctx = util.make_ctx(api, calib, n_frame_slots=8)
    return S, util.run_immature(ctx, S, async_trace=True)


def test_oracle_async_trace_pair(oracle):
    """immature_trace_async + immature_trace_counts give what immature_trace gives"""
    _, a = _run(oracle)
    _, b = _run_async(oracle)
    for ca, cb, ta, tb in zip(a["counts"], b["counts"], a["traces"], b["traces"]):
        assert np.array_equal(ca, cb) and ca.sum() > 0
        for key in ("status", "idepth_min", "idepth_max", "quality", "uv", "interval"):
            assert np.array_equal(ta[key], tb[key], equal_nan=True), key


@pytest.mark.gpu
def test_gpu_async_trace_matches_oracle(oracle, gpu):
    """the trace queued without waiting (ldso_gpu_immature_trace_async): pool state and counts equal the oracle's"""
    _, a = _run(oracle)
    _, b = _run_async(gpu)
    for ca, cb, ta, tb in zip(a["counts"], b["counts"], a["traces"], b["traces"]):
        assert np.array_equal(ca, cb)
        for key in ("status", "idepth_min", "idepth_max", "quality", "uv", "interval"):
            assert np.array_equal(ta[key], tb[key], equal_nan=True), key
    assert np.array_equal(a["cand"], b["cand"]) and np.array_equal(a["act"][0], b["act"][0])


@pytest.mark.gpu
def test_gpu_immature_edge_cases(oracle, gpu):
    """the same hand-made pool (OOB, SKIPPED, far OOB, NaN patch) through both libraries"""
    S = util.immature_scene("small", pts_per_host=40)
    res = []
    for api in (oracle, gpu):
        ctx = util.make_ctx(api, "small", n_frame_slots=8)
        img0 = S["frames"][0][0].copy()
        for i, f in enumerate(S["frames"]):
            ctx.make_images(i, f[0])
        KRKi, Kt, aff = S["tables"][0]
        m = S["host"] == 0
        u, v, gt = S["u"][m][:8].copy(), S["v"][m][:8].copy(), S["gt_idepth"][m][:8]
        col, wgt, gh, eth = ctx.immature_create(0, u, v)
        n = len(u)
        idmin = np.zeros(n, np.float32); idmax = np.full(n, np.nan, np.float32); status = np.full(n, UNINIT, np.int32)
        status[0] = OOB
        idmin[1] = gt[1] * 0.999; idmax[1] = gt[1] * 1.001
        idmin[2] = 50.0; idmax[2] = 60.0
        status[3] = OUTLIER; col[3] += 100.0              # second outlier verdict -> OOB (:273-274)
        ctx.immature_set(np.zeros(n, np.int32), u, v, col, wgt, gh, eth, idmin, idmax, np.full(n, 1e4, np.float32), status)
        cnt = ctx.immature_trace(2, KRKi, Kt, aff)
        res.append((cnt, ctx.immature_get(), col, wgt, gh, eth))
    (ca, ta, *ra), (cb, tb, *rb) = res
    assert ta["status"][3] == OOB
    assert np.array_equal(ca, cb)
    for key in ta:
        assert np.array_equal(ta[key], tb[key], equal_nan=True), key


def _pool_scenario(ctx, S):
    """create -> append (device-resident) -> trace -> compact -> trace again; returns everything the host can observe"""
    nh = S["n_hosts"]
    for i, f in enumerate(S["frames"]):
        ctx.make_images(i, f[0])
    z = np.zeros(0, np.float32)
    ctx.immature_set(np.zeros(0, np.int32), z, z, np.zeros((0, 8)), np.zeros((0, 8)), np.zeros((0, 3)), z, z, z, z, np.zeros(0, np.int32))
    sizes = []
    for hi in range(nh):
        m = S["host"] == hi
        sizes.append(ctx.immature_append(hi, hi, S["u"][m], S["v"][m]))
    out = dict(sizes=sizes)
    KRKi, Kt, aff = S["tables"][0]
    out["c0"] = ctx.immature_trace(nh, KRKi, Kt, aff)
    out["t0"] = ctx.immature_get()
    st = out["t0"]["status"]
    remove = np.nonzero(st == GOOD)[0][::3].astype(np.int32)          # "activated" points
    out["n1"] = ctx.immature_compact(drop_host=0, drop_oob=True, remove=remove)
    out["t1"] = ctx.immature_get()
    # the window slid by one: old host 1 is now host 0
    KRKi2, Kt2, aff2 = S["tables"][1]
    out["c2"] = ctx.immature_trace(nh + 1, KRKi2[1:], Kt2[1:], aff2[1:])
    out["t2"] = ctx.immature_get()
    keep = (S["host"] != 0) & (st != OOB)
    keep[remove] = False
    out["expect_keep"] = keep
    return out


def test_oracle_pool_compact_and_append(oracle):
    S = util.immature_scene("small")
    ctx = util.make_ctx(oracle, "small", n_frame_slots=8)
    o = _pool_scenario(ctx, S)
    assert o["sizes"][-1] == len(S["u"])
    keep = o["expect_keep"]
    assert o["n1"] == keep.sum() == len(o["t1"]["status"])
    for key in ("idepth_min", "idepth_max", "quality", "status", "interval"):
        assert np.array_equal(o["t1"][key], o["t0"][key][keep], equal_nan=True), key    # order preserved, fields carried over
    assert o["c2"].sum() == o["n1"]


@pytest.mark.gpu
def test_gpu_pool_compact_and_append_parity(oracle, gpu):
    S = util.immature_scene("small")
    a = _pool_scenario(util.make_ctx(oracle, "small", n_frame_slots=8), S)
    b = _pool_scenario(util.make_ctx(gpu, "small", n_frame_slots=8), S)
    assert a["sizes"] == b["sizes"] and a["n1"] == b["n1"]
    for k in ("c0", "c2"):
        assert np.array_equal(a[k], b[k])
    for k in ("t0", "t1", "t2"):
        for key in a[k]:
            assert np.array_equal(a[k][key], b[k][key], equal_nan=True), (k, key)


@pytest.mark.gpu
@pytest.mark.parametrize("calib", ["small", "c1", "c4"])
def test_gpu_select_points_matches_map(oracle, gpu, calib):
    """the compacted list equals np.nonzero of the map in row-major order, on both libraries"""
    res = []
    for api in (oracle, gpu):
        c1 = util.make_ctx(api, calib); c2 = util.make_ctx(api, calib)
        img = util.frames(calib)[0][0]
        c1.make_images(0, img); c2.make_images(0, img)
        m, n, pot_m = c1.select_make_maps(0, 1500.0, potential=3)
        u, v, t, pot_l = c2.select_points(0, 1500.0, potential=3, border=5)
        mm = m.copy(); mm[:5] = 0; mm[-5:] = 0; mm[:, :5] = 0; mm[:, -5:] = 0
        ys, xs = np.nonzero(mm)
        assert pot_m == pot_l
        assert np.array_equal(u, xs.astype(np.float32)) and np.array_equal(v, ys.astype(np.float32))
        assert np.array_equal(t, mm[ys, xs])
        res.append((u, v, t))
    for x, y in zip(res[0], res[1]):
        assert np.array_equal(x, y)
