"""A2 PixelSelector: oracle sanity on CPU, and bit-exact GPU parity of the selection maps (SURVEY.md §8d gate)."""
import numpy as np
import pytest

from tests import util


def test_oracle_selector_properties(oracle):
    ctx = util.make_ctx(oracle, "c1")
    img = util.frames("c1")[0][0]
    ctx.make_images(0, img)
    m, n, pot = ctx.select_make_maps(0, 2000 * 0.8)          # makeNewTraces asks for desiredImmatureDensity-scaled counts
    assert set(np.unique(m)).issubset({0.0, 1.0, 2.0, 4.0})
    assert n == int((m != 0).sum())
    assert 0.5 * 1600 < n < 1.6 * 1600
    ys, xs = np.nonzero(m)
    assert xs.min() >= 4 and xs.max() < 640 - 5 and ys.min() >= 4 and ys.max() <= 480 - 4   # PixelSelector2.cc:243
    m2, n2, pot2 = ctx.select_make_maps(0, 2000 * 0.8)       # deterministic
    assert np.array_equal(m, m2) and n == n2
    # the potential adapts: asking for far fewer points raises it
    _, n3, pot3 = ctx.select_make_maps(0, 200.0, potential=3)
    assert pot3 > 3 and n3 < n
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("calib", ["c1", "c3", "c4"])
def test_selector_maps_bit_exact(gpu, oracle, calib):
    g = util.make_ctx(gpu, calib); o = util.make_ctx(oracle, calib)
    img = util.frames(calib)[0][0]
    g.make_images(0, img); o.make_images(0, img)
    pg = po = 3
    for density in (1500.0, 1500.0, 6000.0, 300.0, 50.0, 20000.0):     # potential carries over between calls
        mg, ng, pg = g.select_make_maps(0, density, potential=pg)
        mo, no, po = o.select_make_maps(0, density, potential=po)
        assert (ng, pg) == (no, po), (density, ng, no, pg, po)
        assert np.array_equal(mg, mo), (density, int((mg != mo).sum()))
    for pot in (1, 2, 5, 9):
        mg, ng, p1 = g.select_make_maps(0, 1000.0, potential=pot, recursions=0, th_factor=2.0)
        mo, no, p2 = o.select_make_maps(0, 1000.0, potential=pot, recursions=0, th_factor=2.0)
        assert (ng, p1) == (no, p2) and np.array_equal(mg, mo), pot
    g.close(); o.close()


@pytest.mark.gpu
def test_selector_low_texture_and_axis_aligned_gradients(gpu, oracle):
    """flat regions push the selection to pyramid levels 1 and 2; pure horizontal/vertical gradients make the 16-direction
    'has a winner' masks non-uniform (|g.dir| == 0 for some directions), which exercises the ordered n2 scan"""
    g = util.make_ctx(gpu, "c1"); o = util.make_ctx(oracle, "c1")
    rng = np.random.default_rng(3)
    img = np.full((480, 640), 100.0, dtype=np.float32)
    img[:, 200:] += 40.0                                  # one vertical edge: gradients exactly (gx, 0)
    img[300:, :] += 30.0                                  # one horizontal edge: (0, gy)
    img[100:140, 400:600] += np.floor(rng.uniform(0, 60, (40, 200))).astype(np.float32)
    blur = img.copy()
    for _ in range(6):                                    # wide soft ramps only visible at coarse levels
        blur[1:-1, 1:-1] = 0.2 * (blur[1:-1, 1:-1] + blur[:-2, 1:-1] + blur[2:, 1:-1] + blur[1:-1, :-2] + blur[1:-1, 2:])
    for im in (img, np.floor(blur).astype(np.float32)):
        g.make_images(0, im); o.make_images(0, im)
        for pot in (3, 1):
            mg, ng, p1 = g.select_make_maps(0, 2000.0, potential=pot)
            mo, no, p2 = o.select_make_maps(0, 2000.0, potential=pot)
            assert (ng, p1) == (no, p2), (ng, no)
            assert np.array_equal(mg, mo)
        assert {1.0} <= set(np.unique(mo))
    g.close(); o.close()


@pytest.mark.gpu
def test_selector_without_direction_distribution(gpu, oracle):
    g = util.make_ctx(gpu, "c1", select_direction_distribution=0); o = util.make_ctx(oracle, "c1", select_direction_distribution=0)
    img = util.frames("c1")[0][0]
    g.make_images(0, img); o.make_images(0, img)
    mg, ng, p1 = g.select_make_maps(0, 2000.0)
    mo, no, p2 = o.select_make_maps(0, 2000.0)
    assert (ng, p1) == (no, p2) and np.array_equal(mg, mo)
    g.close(); o.close()
