"""Host-side window algebra that feeds the BA kernels (the reference's H1 row: small dense fp64/fp32 math
that stays on the host): FrameFramePrecalc::Set, EnergyFunctional::setAdjointsF / setDeltaF.

Index convention everywhere: pair tables are [host + n * target].
"""
import numpy as np

SCALE_XI_ROT, SCALE_XI_TRANS, SCALE_A, SCALE_B = 1.0, 0.5, 10.0, 1000.0
PRECALC_STRIDE = 48


def hat(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=np.float64)


def se3_mul(a, b):
    return a[0] @ b[0], a[0] @ b[1] + a[1]


def se3_inv(a):
    return a[0].T, -a[0].T @ a[1]


def se3_adj(T):
    """6x6 adjoint [R, hat(t)R; 0, R] (thirdparty/Sophus/sophus/se3.hpp:100-110)"""
    R, t = T
    A = np.zeros((6, 6))
    A[:3, :3] = R
    A[:3, 3:] = hat(t) @ R
    A[3:, 3:] = R
    return A


def from_to_vec_exposure(expF, expT, g2F, g2T):
    """include/AffLight.h:27-35 (float arithmetic)"""
    expF, expT = np.float32(expF), np.float32(expT)
    if expF == 0 or expT == 0:
        expF = expT = np.float32(1)
    a = np.float32(np.exp(np.float32(g2T[0]) - np.float32(g2F[0]))) * expT / expF
    b = np.float32(g2T[1]) - np.float32(a) * np.float32(g2F[1])
    return np.float32(a), np.float32(b)


def from_to_vec_exposure_table(exposure, g2):
    """from_to_vec_exposure for every pair at once: a[t][h], b[t][h] (host = from, target = to), same float32 arithmetic"""
    e = np.asarray(exposure, np.float32)
    g = np.asarray(g2, np.float32).reshape(len(e), 2)
    eF = np.broadcast_to(e[None, :], (len(e), len(e))).copy()
    eT = np.broadcast_to(e[:, None], (len(e), len(e))).copy()
    z = (eF == 0) | (eT == 0)
    eF[z] = 1; eT[z] = 1
    a = np.exp(g[:, None, 0] - g[None, :, 0]).astype(np.float32) * eT / eF
    b = g[:, None, 1] - a * g[None, :, 1]
    return a.astype(np.float32), b.astype(np.float32)


def _pairs(w2c):
    """T_target * T_host^-1 for every pair, as stacks indexed [target][host] (float64)"""
    R = np.stack([T[0] for T in w2c]).astype(np.float64)
    t = np.stack([T[1] for T in w2c]).astype(np.float64)
    Ri = R.transpose(0, 2, 1)
    ti = -(Ri @ t[:, :, None])[:, :, 0]
    Rl = R[:, None] @ Ri[None, :]                       # [t][h]
    tl = (R[:, None] @ ti[None, :, :, None])[..., 0] + t[:, None]
    return Rl, tl


_K_CACHE = {}


def _k_matrices(fx, fy, cx, cy):
    key = (float(fx), float(fy), float(cx), float(cy))
    if key not in _K_CACHE:
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float32)
        _K_CACHE[key] = (K, np.linalg.inv(K.astype(np.float64)).astype(np.float32))
    return _K_CACHE[key]


def precalc_table(w2c_eval, w2c, calib, aff, aff0, exposure, pairs=None, ab=None):
    """FrameFramePrecalc::Set for every (host, target) pair (src/internal/FrameFramePrecalc.cc:6-35).
    w2c_eval / w2c: lists of (R, t) world-to-camera at the FEJ evaluation point / current state."""
    n = len(w2c)
    fx, fy, cx, cy = [np.float32(c) for c in calib]
    K, Ki = _k_matrices(fx, fy, cx, cy)
    R, tt = pairs if pairs is not None else _pairs(w2c)                    # pairs / ab: the caller's _pairs(w2c) and
    R0, t0 = (R, tt) if w2c_eval is w2c else _pairs(w2c_eval)                # from_to_vec_exposure_table(exposure, aff), reused
    R0f, t0f, Rf, tf = R0.astype(np.float32), t0.astype(np.float32), R.astype(np.float32), tt.astype(np.float32)
    KRKi = (K @ Rf @ Ki).astype(np.float32)
    Kt = (K @ tf[..., None])[..., 0].astype(np.float32)
    out = np.zeros((n, n, PRECALC_STRIDE), dtype=np.float32)      # [t][h] == flat index h + n * t
    out[..., 0:9] = R0f.reshape(n, n, 9); out[..., 9:12] = t0f
    out[..., 12:21] = KRKi.reshape(n, n, 9); out[..., 21:24] = Kt
    out[..., 27:36] = Rf.reshape(n, n, 9); out[..., 36:39] = tf
    out[..., 39] = np.linalg.norm(tt, axis=-1).astype(np.float32)
    a, b = ab if ab is not None else from_to_vec_exposure_table(exposure, aff)
    out[..., 24] = a; out[..., 25] = b
    out[..., 26] = np.asarray(aff0, np.float32).reshape(n, 2)[None, :, 1]
    return out.reshape(n * n, PRECALC_STRIDE)


def adjoints(w2c_eval, aff0, exposure, pairs=None, ab=None):
    """EnergyFunctional::setAdjointsF (src/internal/OptimizationBackend/EnergyFunctional.cc:551-609)"""
    n = len(w2c_eval)
    scale = np.array([SCALE_XI_TRANS] * 3 + [SCALE_XI_ROT] * 3 + [SCALE_A, SCALE_B])
    Rl, tl = pairs if pairs is not None else _pairs(w2c_eval)     # [t][h]
    H = np.zeros((n, n, 3, 3))
    H[..., 0, 1] = -tl[..., 2]; H[..., 0, 2] = tl[..., 1]; H[..., 1, 0] = tl[..., 2]
    H[..., 1, 2] = -tl[..., 0]; H[..., 2, 0] = -tl[..., 1]; H[..., 2, 1] = tl[..., 0]
    A = np.zeros((n, n, 6, 6))
    A[..., :3, :3] = Rl; A[..., :3, 3:] = H @ Rl; A[..., 3:, 3:] = Rl
    adH = np.zeros((n, n, 8, 8)); adT = np.zeros((n, n, 8, 8))
    adH[..., 6, 6] = 1; adH[..., 7, 7] = 1
    adT[..., np.arange(8), np.arange(8)] = 1
    adH[..., :6, :6] = -A.transpose(0, 1, 3, 2)
    a, _ = ab if ab is not None else from_to_vec_exposure_table(exposure, aff0)
    adT[..., 6, 6] = -a; adH[..., 6, 6] = a; adT[..., 7, 7] = -1; adH[..., 7, 7] = a
    adH *= scale[:, None]; adT *= scale[:, None]
    return adH.reshape(n * n, 8, 8), adT.reshape(n * n, 8, 8)


def ad_ht_delta(adH, adT, state_minus_zero):
    """EnergyFunctional::setDeltaF (EnergyFunctional.cc:523-549): adHTdeltaF[h + n t]"""
    n = len(state_minus_zero)
    adHF, adTF = adH.astype(np.float32), adT.astype(np.float32)
    out = np.zeros((n * n, 8), dtype=np.float32)
    for h in range(n):
        for t in range(n):
            i = h + n * t
            out[i] = state_minus_zero[h][:8].astype(np.float32) @ adHF[i] + state_minus_zero[t][:8].astype(np.float32) @ adTF[i]
    return out
