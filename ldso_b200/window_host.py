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


def precalc_table(w2c_eval, w2c, calib, aff, aff0, exposure):
    """FrameFramePrecalc::Set for every (host, target) pair (src/internal/FrameFramePrecalc.cc:6-35).
    w2c_eval / w2c: lists of (R, t) world-to-camera at the FEJ evaluation point / current state."""
    n = len(w2c)
    fx, fy, cx, cy = [np.float32(c) for c in calib]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float32)
    Ki = np.linalg.inv(K.astype(np.float64)).astype(np.float32)
    out = np.zeros((n * n, PRECALC_STRIDE), dtype=np.float32)
    for h in range(n):
        for t in range(n):
            l0 = se3_mul(w2c_eval[t], se3_inv(w2c_eval[h]))
            l = se3_mul(w2c[t], se3_inv(w2c[h]))
            R0, t0 = l0[0].astype(np.float32), l0[1].astype(np.float32)
            R, tt = l[0].astype(np.float32), l[1].astype(np.float32)
            KRKi = (K @ R @ Ki).astype(np.float32)
            Kt = (K @ tt).astype(np.float32)
            a, b = from_to_vec_exposure(exposure[h], exposure[t], aff[h], aff[t])
            row = out[h + n * t]
            row[0:9] = R0.reshape(-1); row[9:12] = t0
            row[12:21] = KRKi.reshape(-1); row[21:24] = Kt
            row[24] = a; row[25] = b; row[26] = np.float32(aff0[h][1])
            row[27:36] = R.reshape(-1); row[36:39] = tt; row[39] = np.float32(np.linalg.norm(l[1]))
    return out


def adjoints(w2c_eval, aff0, exposure):
    """EnergyFunctional::setAdjointsF (src/internal/OptimizationBackend/EnergyFunctional.cc:551-609)"""
    n = len(w2c_eval)
    adH = np.zeros((n * n, 8, 8)); adT = np.zeros((n * n, 8, 8))
    scale = np.array([SCALE_XI_TRANS] * 3 + [SCALE_XI_ROT] * 3 + [SCALE_A, SCALE_B])
    for h in range(n):
        for t in range(n):
            h2t = se3_mul(w2c_eval[t], se3_inv(w2c_eval[h]))
            AH = np.eye(8); AT = np.eye(8)
            AH[:6, :6] = -se3_adj(h2t).T
            a, _ = from_to_vec_exposure(exposure[h], exposure[t], aff0[h], aff0[t])
            AT[6, 6] = -a; AH[6, 6] = a; AT[7, 7] = -1; AH[7, 7] = a
            AH *= scale[:, None]; AT *= scale[:, None]
            adH[h + n * t] = AH; adT[h + n * t] = AT
    return adH, adT


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
