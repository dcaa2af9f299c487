"""N4 — Undistort::undistort + PhotometricUndistorter::processFrame (reference src/frontend/Undistort.cc:190-230, :366-458)
in front of makeImages: the oracle against a numpy restatement, the CUDA path against the oracle (bit-exact)."""
import numpy as np
import pytest

from tests import util


def scene(calib, seed=3, bits=8):
    c = util.CALIB[calib]
    w, h = c["w"], c["h"]
    w_org, h_org = w + 24, h + 16                       # raw sensor larger than the rectified output
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h_org, 0:w_org]
    tex = 120 + 60 * np.sin(xx * 0.11) * np.cos(yy * 0.07) + 30 * np.sin((xx + yy) * 0.31) + rng.normal(0, 4, (h_org, w_org))
    top = 255 if bits == 8 else 65535
    raw = np.clip(tex * (top / 255.0), 0, top).astype(np.uint8 if bits == 8 else np.uint16)
    G = (255.0 * (np.arange(top + 1) / float(top)) ** 0.8).astype(np.float32)          # inverse response
    r2 = ((xx - w_org / 2) ** 2 + (yy - h_org / 2) ** 2) / float(w_org * w_org)
    vig_inv = (1.0 / (1.0 - 0.6 * r2)).astype(np.float32)
    # a smooth radial lookup; a band on the left falls outside the sensor (remapX = -1 as Undistort.cc:352-357 marks it)
    oy, ox = np.mgrid[0:h, 0:w].astype(np.float64)
    dx, dy = ox - w / 2, oy - h / 2
    k = 1 + 0.08 * (dx * dx + dy * dy) / float(w * w)
    rx = (w_org / 2 + dx * k).astype(np.float32)
    ry = (h_org / 2 + dy * k).astype(np.float32)
    bad = (rx < 0.5) | (ry < 0.5) | (rx > w_org - 2.5) | (ry > h_org - 2.5) | (ox < 3)
    rx[bad] = -1; ry[bad] = -1
    return dict(w=w, h=h, w_org=w_org, h_org=h_org, raw=raw, G=G, vig=vig_inv, rx=rx, ry=ry)


def np_undistort(S, mode, exposure, factor):
    f32 = np.float32
    raw = S["raw"].astype(np.int64)
    if mode == 0 or exposure <= 0:
        data = (f32(factor) * raw.astype(f32)).astype(f32)
    else:
        data = S["G"][raw]
        if mode == 2:
            data = (data * S["vig"]).astype(f32)
    rx, ry = S["rx"].copy(), S["ry"].copy()
    out = np.zeros((S["h"], S["w"]), f32)
    ok = rx >= 0
    xi, yi = rx[ok].astype(np.int64), ry[ok].astype(np.int64)
    xf, yf = (rx[ok] - xi.astype(f32)).astype(f32), (ry[ok] - yi.astype(f32)).astype(f32)
    xy = (xf * yf).astype(f32)
    s00, s01, s10, s11 = data[yi, xi], data[yi, xi + 1], data[yi + 1, xi], data[yi + 1, xi + 1]
    v = (xy * s11).astype(f32)
    v = (v + ((yf - xy).astype(f32) * s10).astype(f32)).astype(f32)
    v = (v + ((xf - xy).astype(f32) * s01).astype(f32)).astype(f32)
    v = (v + ((((f32(1) - xf).astype(f32) - yf).astype(f32) + xy).astype(f32) * s00).astype(f32)).astype(f32)
    out[ok] = v
    return out


@pytest.mark.parametrize("mode,exposure,bits", [(2, 0.02, 8), (1, 0.02, 8), (0, 0.02, 8), (2, 0.0, 8), (2, 0.02, 16)])
def test_oracle_undistort_matches_numpy(oracle, mode, exposure, bits):
    S = scene("small", bits=bits)
    ctx = util.make_ctx(oracle, "small", n_frame_slots=3)
    ctx.undistort_set(S["w_org"], S["h_org"], S["G"], S["vig"], S["rx"], S["ry"], photometric_mode=mode)
    ctx.make_images_raw(0, S["raw"], exposure, 0.5)
    img = ctx.undistort_get_image()
    assert np.array_equal(img, np_undistort(S, mode, exposure, 0.5))
    assert (img[:, :3] == 0).all() and img[:, 3:].std() > 5
    # the pyramid built behind it is makeImages of that image
    ref = util.make_ctx(oracle, "small", n_frame_slots=3)
    ref.make_images(0, img)
    for lvl in range(ctx.levels):
        assert np.array_equal(ctx.download(0, lvl, 0), ref.download(0, lvl, 0))
    ctx.close(); ref.close()


def test_oracle_passthrough(oracle):
    c = util.CALIB["small"]
    raw = (np.arange(c["w"] * c["h"]) % 251).astype(np.uint8).reshape(c["h"], c["w"])
    ctx = util.make_ctx(oracle, "small", n_frame_slots=3)
    ctx.undistort_set(c["w"], c["h"])                                         # no calibration at all: data = factor * raw
    ctx.make_images_raw(0, raw, 1.0, 1.0)
    assert np.array_equal(ctx.undistort_get_image(), raw.astype(np.float32))
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib,mode,exposure,bits", [("c1", 2, 0.02, 8), ("c4", 1, 0.02, 8), ("small", 0, 0.02, 8), ("c1", 2, 0.02, 16),
                                                      ("small", 2, 0.0, 8)])
def test_gpu_undistort_bit_exact(gpu, oracle, calib, mode, exposure, bits):
    S = scene(calib, bits=bits)
    outs = []
    for api in (gpu, oracle):
        ctx = util.make_ctx(api, calib, n_frame_slots=3)
        ctx.undistort_set(S["w_org"], S["h_org"], S["G"], S["vig"], S["rx"], S["ry"], photometric_mode=mode)
        ctx.make_images_raw(0, S["raw"], exposure, 0.5)
        outs.append([ctx.undistort_get_image()] + [ctx.download(0, lvl, k) for lvl in range(ctx.levels) for k in (0, 1)])
        ctx.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)
    assert outs[0][0].std() > 5


@pytest.mark.gpu
def test_gpu_passthrough_and_pinned_raw(gpu, oracle):
    import torch
    c = util.CALIB["c1"]
    raw = (np.random.default_rng(0).integers(0, 256, (c["h"], c["w"]))).astype(np.uint8)
    pinned = torch.from_numpy(raw).pin_memory()
    res = []
    for api, buf in ((gpu, pinned.numpy()), (gpu, raw), (oracle, raw)):
        ctx = util.make_ctx(api, "c1", n_frame_slots=3)
        ctx.undistort_set(c["w"], c["h"])
        ctx.make_images_raw(0, buf, 1.0, 1.0)
        res.append((ctx.undistort_get_image(), ctx.download(0, 1, 0)))
        ctx.close()
    for a in res[:2]:
        assert np.array_equal(a[0], res[2][0]) and np.array_equal(a[1], res[2][1])
