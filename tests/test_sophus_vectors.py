"""Pins the SE(3) algebra of include/ldso_smallmat.h (used by the tracker's LM step, FrameHessian state algebra, adjoints,
nullspaces) on the ONE set of vectors the reference holds for this path: the vendored Sophus unit tests
(thirdparty/Sophus/test/core/test_se3.cpp:34-50, tests.hpp:46-64, 208-254, 507-542).  The identities are the reference's own
(LieGroupTests: adjointTest, expLogTest, expMapTest, groupActionTest, product / inverse consistency), the expected values
come from dense 4x4 matrix algebra (numpy / scipy.linalg.expm), i.e. from outside the library under test."""
import ctypes as C
import json
import os

import numpy as np
import pytest
from scipy.linalg import expm

from ldso_b200 import build

HERE = os.path.dirname(os.path.abspath(__file__))
V = json.load(open(os.path.join(HERE, "golden", "sophus_se3_vectors.json")))
EPS = V["eps"]


@pytest.fixture(scope="module")
def lib():
    build.build_oracle()
    return C.CDLL(build.ORACLE_LIB)


def _d(a):
    return np.ascontiguousarray(a, np.float64).ctypes.data_as(C.POINTER(C.c_double))


def hat(x):
    u, w = x[:3], x[3:]
    M = np.zeros((4, 4))
    M[:3, :3] = [[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]]
    M[:3, 3] = u
    return M


def vee(M):
    return np.array([M[0, 3], M[1, 3], M[2, 3], M[2, 1], M[0, 2], M[1, 0]])


def mat(Rt):
    T = np.eye(4)
    T[:3, :3] = Rt[:9].reshape(3, 3)
    T[:3, 3] = Rt[9:]
    return T


def rt(T):
    return np.concatenate([T[:3, :3].reshape(-1), T[:3, 3]])


def call(lib, name, *ins, n_out):
    out = np.zeros(n_out)
    getattr(lib, name)(*[_d(np.asarray(i, np.float64)) for i in ins], _d(out))
    return out


def _poses(lib):
    """getTestSE3s() built through the library's own exp / product, as the Sophus test builds them"""
    poses = []
    for factors in V["poses"]:
        T = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)])
        for f in factors:
            val = lambda a: np.pi if a == "pi" else float(a)
            if f[0] == "so3exp":
                g = call(lib, "orc_se3_exp", [0, 0, 0] + list(f[1]), n_out=12)
                g[9:] = f[2]
            elif f[0] == "trans":
                g = np.concatenate([np.eye(3).reshape(-1), [val(a) for a in f[1:]]])
            else:
                w = [val(f[1]), 0, 0] if f[0] == "rotX" else [0, 0, val(f[1])]
                g = call(lib, "orc_se3_exp", [0, 0, 0] + w, n_out=12)
            T = call(lib, "orc_se3_mul", T, g, n_out=12)
        poses.append(T)
    return poses


def test_exp_equals_the_matrix_exponential_of_hat(lib):
    """expMapTest (tests.hpp:220-230): exp(x) == expm(hat(x)) to 10 kSmallEps"""
    for x in V["tangents"]:
        T = mat(call(lib, "orc_se3_exp", x, n_out=12))
        ref = expm(hat(np.array(x, float)))
        assert np.abs(T - ref).max() < 10 * EPS * max(1.0, np.abs(ref).max()), x


def test_exp_of_log_is_the_identity_map(lib):
    """expLogTest (tests.hpp:208-218) on getTestSE3s, incl. the rotations by pi and the near-identity ones"""
    for i, P in enumerate(_poses(lib)):
        lg = call(lib, "orc_se3_log", P, n_out=6)
        back = call(lib, "orc_se3_exp", lg, n_out=12)
        assert np.abs(mat(back) - mat(P)).max() < EPS * max(1.0, np.abs(P).max()), i
        # and the logarithm agrees with the dense one wherever that is unique (rotation angle < pi)
        if np.linalg.norm(lg[3:]) < 3.1:
            assert np.abs(expm(hat(lg)) - mat(P)).max() < 10 * EPS * max(1.0, np.abs(P).max())


def test_adjoint(lib):
    """adjointTest (tests.hpp:46-64): Ad x == vee(T hat(x) T^-1) to 10 kSmallEps"""
    for i, P in enumerate(_poses(lib)):
        T = mat(P)
        Ad = call(lib, "orc_se3_adj", P, n_out=36).reshape(6, 6)
        Ti = mat(call(lib, "orc_se3_inv", P, n_out=12))
        assert np.abs(Ti - np.linalg.inv(T)).max() < EPS * max(1.0, np.abs(T).max())
        for x in V["tangents"]:
            x = np.array(x, float)
            ad2 = vee(T @ hat(x) @ Ti)
            assert np.abs(Ad @ x - ad2).max() < 10 * EPS * max(1.0, np.abs(ad2).max()), (i, x)


def test_group_action_and_product(lib):
    """groupActionTest (tests.hpp:232-254): G * p == T p; product == matrix product"""
    poses = _poses(lib)
    for i, P in enumerate(poses):
        T = mat(P)
        for p in V["points"]:
            got = call(lib, "orc_se3_act", P, p, n_out=3)
            assert np.abs(got - (T @ np.array(p + [1.0]))[:3]).max() < EPS * max(1.0, np.abs(T).max())
        for Q in poses:
            PQ = mat(call(lib, "orc_se3_mul", P, Q, n_out=12))
            ref = T @ mat(Q)
            assert np.abs(PQ - ref).max() < EPS * max(1.0, np.abs(ref).max())


def test_raw_data_constructor(lib):
    """testRawDataAcces (test_se3.cpp:67-75): coefficients (x,y,z,w) = (0,1,0,0), translation (1,3,2)"""
    Rt = call(lib, "orc_se3_from_quat", [0, 0, 1, 0], [1, 3, 2], n_out=12)      # (w,x,y,z) = (0,0,1,0): rotation by pi about y
    assert np.abs(Rt[:9].reshape(3, 3) - np.diag([-1.0, 1.0, -1.0])).max() < EPS and np.array_equal(Rt[9:], [1, 3, 2])
