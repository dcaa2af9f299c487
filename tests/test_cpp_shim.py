"""include/ldso_shim.hpp (C++ facade with the reference's class names) gives the same results as the ctypes path.

CPU: the facade is compiled against the oracle (prefix macro) — checks that the header is valid C++ and drives the ABI
correctly.  GPU: compiled against libldso_gpu.so and run on the device."""
import os
import subprocess

import numpy as np
import pytest

from ldso_b200 import build
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
IDENT = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])


def _blob(path, calib):
    c = util.CALIB[calib]
    (img0, dep0, R0, t0), (img1, dep1, R1, t1) = util.frames(calib, 1001, (40, 42))
    u, v, idp, hdi = util.pick_points(img0, dep0, 600)
    K = np.array([[c["fx"], 0, c["cx"]], [0, c["fy"], c["cy"]], [0, 0, 1]], dtype=np.float32)
    Ki = np.linalg.inv(K.astype(np.float64)).astype(np.float32)
    R = (R1.T @ R0).astype(np.float32); t = (R1.T @ (t0 - t1)).astype(np.float32)
    tab = np.concatenate([(K @ R @ Ki).astype(np.float32).reshape(-1), (K @ t).astype(np.float32), [1.0, 0.0]]).astype(np.float32)
    with open(path, "wb") as f:
        np.array([c["w"], c["h"], len(u)], np.int32).tofile(f)
        np.array([c["fx"], c["fy"], c["cx"], c["cy"]], np.float32).tofile(f)
        for a in (img0, img1, u, v, idp, hdi, tab):
            np.ascontiguousarray(a, np.float32).tofile(f)
    return (img0, img1, u, v, idp, hdi, tab)


def _expected(api, calib, data):
    img0, img1, u, v, idp, hdi, tab = data
    ctx = util.make_ctx(api, calib)
    ctx.make_images(0, img0); ctx.make_images(1, img1)
    ctx.tracker_set_ref(0, 0, u, v, idp, hdi)
    r = ctx.tracker_track(0, 1, IDENT, [0.0, 0.0])
    su, sv, st, pot = ctx.select_points(0, 300.0, potential=3, border=5)
    z = np.zeros(0, np.float32)
    ctx.immature_set(np.zeros(0, np.int32), z, z, np.zeros((0, 8)), np.zeros((0, 8)), np.zeros((0, 3)), z, z, z, z, np.zeros(0, np.int32))
    ctx.immature_append(0, 0, su, sv)
    cnt = ctx.immature_trace(1, tab[:9].reshape(1, 9), tab[9:12].reshape(1, 3), tab[12:14].reshape(1, 2))
    g = ctx.immature_get()
    good = g["status"] == 0
    out = (r, len(su), pot, cnt, int(good.sum()), float((g["idepth_min"][good].astype(np.float64) + g["idepth_max"][good]).sum()))
    n = ctx.init_set_first(0)
    ri = ctx.init_track_frame(1)
    P = ctx.init_get_points(0)
    K1 = np.array([[0.5, 0, -0.25], [0, 0.5, -0.25], [0, 0, 1]], np.float32)[None]
    ctx.distmap_make(K1, np.zeros((1, 3), np.float32), np.zeros(len(u), np.int32), u, v, idp)
    extra = (n, ri, float(P["iR"].astype(np.float64).sum()), float(ctx.distmap_get().astype(np.float64).sum()))
    return out + (extra,)


def _run(exe, blob):
    out = subprocess.run([exe, blob], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    return {l.split()[0]: l.split()[1:] for l in out.stdout.strip().splitlines()}


def _check(got, exp):
    r, nsel, pot, cnt, ngood, dsum, (n_init, ri, irs, dmsum) = exp
    it = got["init"]
    assert [int(it[0]), int(it[1])] == [int(n_init[0]), int(n_init[1])] and int(it[2]) == int(ri["ready"]) and int(it[3]) == int(ri["snapped"])
    assert np.allclose([float(x) for x in it[4:7]], ri["pose"][9:], rtol=1e-4, atol=1e-7)
    assert abs(float(it[7]) - irs) <= 1e-4 * abs(irs)
    assert abs(float(got["distmap"][0]) - dmsum) <= 1e-6 * abs(dmsum)
    tr = [float(x) for x in got["track"]]
    assert int(tr[0]) == int(r["ok"])
    assert np.allclose(tr[1:13], r["pose"], rtol=0, atol=1e-12)
    assert np.allclose(tr[13:15], r["aff"], atol=1e-6)
    assert np.allclose(tr[15:20], r["last_residuals"], rtol=1e-6, equal_nan=True)
    assert [int(x) for x in got["select"]] == [nsel, pot]
    assert [int(x) for x in got["trace"]] == cnt.tolist()
    assert int(got["depth"][0]) == ngood and abs(float(got["depth"][1]) - dsum) <= 1e-5 * max(abs(dsum), 1)


def test_shim_compiles_and_matches_against_oracle(oracle, tmp_path):
    exe = str(tmp_path / "shim_smoke_cpu")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-DLDSO_SHIM_PREFIX=orc_",
                    "-DLDSO_SHIM_CTX=orc_ctx", "-DLDSO_SHIM_DECLARE_FOREIGN", os.path.join(ROOT, "tests", "cpp", "shim_smoke.cpp"),
                    "-o", exe, build.ORACLE_LIB, "-Wl,-rpath," + os.path.dirname(build.ORACLE_LIB)], check=True)
    blob = str(tmp_path / "in.bin")
    data = _blob(blob, "small")
    _check(_run(exe, blob), _expected(oracle, "small", data))


@pytest.mark.gpu
def test_shim_on_gpu(gpu, tmp_path):
    exe = str(tmp_path / "shim_smoke_gpu")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "shim_smoke.cpp"),
                    "-o", exe, build.GPU_LIB, "-Wl,-rpath," + os.path.dirname(build.GPU_LIB)], check=True)
    blob = str(tmp_path / "in.bin")
    data = _blob(blob, "c1")
    _check(_run(exe, blob), _expected(gpu, "c1", data))
