"""N3 — FeatureDetector::DetectCorners up to the corner flags (reference src/frontend/FeatureDetector.cc:34-124):
the oracle against a numpy restatement, the CUDA path against the oracle (bit-exact)."""
import numpy as np
import pytest

from tests import util


def np_detect(dIp, grad, n_features):
    """dIp [h][w][3], grad [h][w]; plain loops over cells, numpy inside"""
    f32 = np.float32
    h, w = grad.shape
    g = int(np.sqrt(f32(w * h // n_features)) + 0.5)
    gridX, gridY = w // g + 1, h // g + 1
    nf = f32(n_features) / f32(w * h) * f32(g * g)
    skip = 30 // g + 1
    dx, dy = dIp[..., 1].astype(f32), dIp[..., 2].astype(f32)
    feats, max_score = [], f32(0)

    def score(u, v):
        if u - 4 < 1 or u + 4 >= w - 1 or v - 4 < 1 or v + 4 >= h - 1:
            return f32(0)
        a, b = dx[v - 4:v + 4, u - 4:u + 4].reshape(-1), dy[v - 4:v + 4, u - 4:u + 4].reshape(-1)
        sxx = syy = sxy = f32(0)
        for k in range(64):                                   # the reference's float running sums, row-major
            sxx = f32(sxx + f32(a[k] * a[k])); syy = f32(syy + f32(b[k] * b[k])); sxy = f32(sxy + f32(a[k] * b[k]))
        sxx, syy, sxy = f32(np.float64(sxx) / 128.0), f32(np.float64(syy) / 128.0), f32(np.float64(sxy) / 128.0)
        s = f32(sxx + syy)
        q = f32(f32(sxx * syy) - f32(sxy * sxy))
        return f32(0.5) * f32(s - np.sqrt(f32(f32(s * s) - f32(f32(4) * q))))

    for gx in range(skip, gridX - skip):
        for gy in range(skip, gridY - skip):
            cell = grad[gy * g:(gy + 1) * g, gx * g:(gx + 1) * g]
            mx = max(f32(0), cell.max())
            th = f32(0.5) * mx if f32(0.5) * mx > 5 else f32(5)
            cand = []
            for x in range(g):
                for y in range(g):
                    if cell[y, x] > th:
                        s = score(gx * g + x, gy * g + y)
                        cand.append((y * g + x, s))
                        if s > max_score:
                            max_score = s
            cand.sort(key=lambda p: -p[1])                    # Python's sort is stable
            picked = 0
            for idx, s in cand:
                feats.append((gx * g + idx % g, gy * g + idx // g, s))
                picked += 1
                if picked > nf:
                    break
    x = np.array([f[0] for f in feats], f32); y = np.array([f[1] for f in feats], f32); s = np.array([f[2] for f in feats], f32)
    th = f32(0.01 * np.float64(max_score))
    corner = s > th
    ids = np.nonzero(corner)[0]
    for a in range(len(ids)):
        for b in range(a + 1, len(ids)):
            i, j = ids[a], ids[b]
            if np.sqrt(f32(f32((x[i] - x[j]) ** 2) + f32((y[i] - y[j]) ** 2))) < 5:
                if s[i] > s[j]:
                    corner[j] = False
                else:
                    corner[i] = False
    return x, y, s, corner


@pytest.mark.parametrize("calib,nfeat", [("small", 300), ("small", 120)])
def test_oracle_detect_corners_matches_numpy(oracle, calib, nfeat):
    (img, _, _, _), = util.frames(calib, 1001, (40,))
    ctx = util.make_ctx(oracle, calib, n_frame_slots=3)
    ctx.make_images(0, img)
    w, h = ctx.sizes[0]
    dIp = ctx.download(0, 0, 0).reshape(h, w, 3)
    grad = ctx.download(0, 0, 1).reshape(h, w)
    x, y, s, c, nc = ctx.detect_corners(0, nfeat)
    ex, ey, es, ec = np_detect(dIp, grad, nfeat)
    assert len(x) == len(ex) > 20
    assert np.array_equal(x, ex) and np.array_equal(y, ey) and np.array_equal(s, es) and np.array_equal(c, ec)
    assert nc == int(ec.sum()) and 0 < nc <= len(x)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib,nfeat", [("c1", 1500), ("c1", 600), ("c3", 2000), ("c4", 6000), ("small", 300)])
def test_gpu_detect_corners_bit_exact(gpu, oracle, calib, nfeat):
    (img, _, _, _), = util.frames(calib, 1001, (40,))
    res = []
    for api in (gpu, oracle):
        ctx = util.make_ctx(api, calib, n_frame_slots=3)
        ctx.make_images(0, img)
        res.append(ctx.detect_corners(0, nfeat))
        ctx.close()
    for a, b in zip(*res):
        assert np.array_equal(a, b)
    assert len(res[0][0]) > 50 and res[0][4] > 10
