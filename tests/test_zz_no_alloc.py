"""GPU: the allocation-free entry point (ldso_gpu_tracker_calc_res_noalloc, no counterpart in the reference; operators'
request) against the existing ldso_gpu_tracker_calc_res through the C ABI.

Same kernel, same fold order: the six outputs must be bit-identical.  Two calls in a row (the shared file-scope scratch
keeps the previous call's contents), then the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])


def _free_bytes():
    import torch
    return torch.cuda.mem_get_info(0)[0]


def test_noalloc_calc_res_matches_the_allocating_entry_point_twice_in_a_row(gpu):
    g = util.make_ctx(gpu, "c1")
    gt = util.setup_tracker(g, "c1", 2000)
    # (level, pose, affine, cutoff) of two consecutive calls: different level sizes, so the second call sees stale partials
    calls = [(0, gt, (0.0, 0.0), 20.0), (2, IDENT, (0.02, -1.5), 40.0), (0, IDENT, (0.0, 0.0), 20.0)]
    want = [g.tracker_calc_res(0, 1, lvl, pose, a[0], a[1], cut).copy() for lvl, pose, a, cut in calls]
    H_want = g.tracker_calc_gs(0, 0, IDENT, 0.0, 0.0)          # state left by the last allocating call
    g.sync()
    g.tracker_calc_res_noalloc(0, 1, 0, gt, 0.0, 0.0, 20.0)     # first use resolves the symbol address (no allocation either)
    free0 = _free_bytes()
    for (lvl, pose, a, cut), w in zip(calls, want):
        got, sums = g.tracker_calc_res_noalloc(0, 1, lvl, pose, a[0], a[1], cut)
        assert np.array_equal(got, w), (lvl, got, w)             # tolerance: none (integer counts and identical float sums)
        assert got[1] > 100 and sums[48] == got[1] and np.isfinite(sums).all()
    assert _free_bytes() == free0, "the allocation-free entry point changed the device's free memory"
    # the new entry point leaves the tracker's calc_gs state alone
    H_after = g.tracker_calc_gs(0, 0, IDENT, 0.0, 0.0)
    assert np.array_equal(H_want[0], H_after[0]) and np.array_equal(H_want[1], H_after[1])
    g.close()


def test_noalloc_rejects_sizes_beyond_its_maximum(gpu):
    limit = 131072                                               # LDSO_NOALLOC_MAX_POINTS, include/ldso_gpu.h
    g = util.make_ctx(gpu, "c1")
    (img0, dep0, _, _), (img1, _, _, _) = util.frames("c1", 1001, (40, 42))
    g.make_images(0, img0)
    g.make_images(1, img1)
    # a dense reference: every 2nd pixel of the interior (both directions) carries a point, the diagonal dilation of makeCoarseDepthL0 fills level 0
    ys, xs = np.mgrid[4:476:2, 4:636:2]
    u, v = xs.reshape(-1).astype(np.float32), ys.reshape(-1).astype(np.float32)
    g.tracker_set_ref(0, 0, u, v, (1.0 / dep0[ys, xs]).reshape(-1).astype(np.float32), np.full(u.size, 1e-3, np.float32))
    n0 = len(g.tracker_get_pc(0, 0)[0])
    assert n0 > limit, n0
    out = np.empty(6, np.float32)
    rc = gpu.fn("tracker_calc_res_noalloc")(g.h, C.c_int(0), C.c_int(1), C.c_float(1.0), C.c_int(0),
                                            IDENT.ctypes.data_as(C.POINTER(C.c_double)), C.c_float(0), C.c_float(0),
                                            C.c_float(20.0), out.ctypes.data_as(C.POINTER(C.c_float)), None)
    assert rc == -2
    assert b"LDSO_NOALLOC_MAX_POINTS" in gpu.fn("last_error")(g.h)
    # a coarser level of the same reference is within the limit and still agrees
    n2 = len(g.tracker_get_pc(0, 2)[0])
    assert 0 < n2 <= limit
    want = g.tracker_calc_res(0, 1, 2, IDENT, 0.0, 0.0, 20.0)
    got, _ = g.tracker_calc_res_noalloc(0, 1, 2, IDENT, 0.0, 0.0, 20.0)
    assert np.array_equal(got, want)
    g.close()
