"""Synthetic textured-scene sequences with ground truth.

There is no dataset on the box, so BASELINE.json's configs are exercised on a procedural scene: a
closed box room (half-extents 4 x 2.5 x 6 m) whose six walls carry a smooth random albedo
    T(p) = clamp(128 + sum_k A_k sin(2 pi f_k . p_uv + phi_k) + 32 valuenoise(6 p_uv), 0, 255),
rendered by analytic ray/plane intersection for a pinhole camera, quantised to uint8 and cast to float
(what Undistort::undistort hands to makeImages when there is no photometric calibration,
reference src/frontend/Undistort.cc:198-206).  Everything is seeded and deterministic.
"""
import numpy as np

HALF = np.array([4.0, 2.5, 6.0])


class Scene:
    def __init__(self, seed=1001, n_sines=8):
        rng = np.random.Generator(np.random.PCG64(seed))
        self.seed = seed
        # per plane (6): sinusoid amplitudes, 2-D frequency vectors (cycles/m), phases, and a value-noise lattice
        self.A = rng.uniform(8, 24, size=(6, n_sines))
        f = rng.uniform(0.3, 6.0, size=(6, n_sines))
        ang = rng.uniform(0, 2 * np.pi, size=(6, n_sines))
        self.F = np.stack([f * np.cos(ang), f * np.sin(ang)], axis=-1)
        self.phi = rng.uniform(0, 2 * np.pi, size=(6, n_sines))
        self.noise = rng.uniform(-1, 1, size=(6, 160, 160))

    def albedo(self, plane, pu, pv):
        val = np.full(pu.shape, 128.0)
        for k in range(self.A.shape[1]):
            val += self.A[plane, k] * np.sin(2 * np.pi * (self.F[plane, k, 0] * pu + self.F[plane, k, 1] * pv) + self.phi[plane, k])
        # value noise, 6 cells / m, bilinear
        gx = (pu + 13.0) * 6.0
        gy = (pv + 13.0) * 6.0
        ix = np.floor(gx).astype(np.int64)
        iy = np.floor(gy).astype(np.int64)
        fx = gx - ix
        fy = gy - iy
        n = self.noise[plane]
        ix0 = np.mod(ix, 160); ix1 = np.mod(ix + 1, 160); iy0 = np.mod(iy, 160); iy1 = np.mod(iy + 1, 160)
        v = (n[iy0, ix0] * (1 - fx) * (1 - fy) + n[iy0, ix1] * fx * (1 - fy) + n[iy1, ix0] * (1 - fx) * fy + n[iy1, ix1] * fx * fy)
        val += 32.0 * v
        return np.clip(val, 0, 255)


def rot_yaw_pitch(yaw, pitch):
    cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    return Ry @ Rx


def trajectory(i, fps=30.0):
    """camera-to-world pose (R, t) of frame i; first 30 frames are a pure sideways translation."""
    t = i / fps
    if i < 30:
        return np.eye(3), np.array([0.25 * t, 0.0, -2.0])
    s = t - 1.0
    x = 0.25 + 0.6 * np.sin(0.5 * s)
    y = 0.08 * np.sin(1.1 * s)
    z = -2.0 + 1.5 * np.sin(0.17 * s)
    R = rot_yaw_pitch(0.08 * np.sin(0.35 * s), 0.03 * np.sin(0.6 * s))
    return R, np.array([x, y, z])


def render(scene, w, h, fx, fy, cx, cy, R, t):
    """returns (image float32 [h,w] with integer values 0..255, depth float32 [h,w] along the optical axis)"""
    u, v = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    dc = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], axis=-1)
    dw = dc @ R.T
    best = np.full((h, w), np.inf)
    img = np.zeros((h, w))
    for axis in range(3):
        for sgn, plane in ((+1, 2 * axis), (-1, 2 * axis + 1)):
            d = dw[..., axis]
            with np.errstate(divide="ignore", invalid="ignore"):
                lam = (sgn * HALF[axis] - t[axis]) / d
            ok = (lam > 1e-6) & (lam < best) & np.isfinite(lam)
            if not ok.any():
                continue
            p = t[None, None, :] + lam[..., None] * dw
            a0, a1 = [a for a in range(3) if a != axis]
            inside = ok & (np.abs(p[..., a0]) <= HALF[a0] + 1e-9) & (np.abs(p[..., a1]) <= HALF[a1] + 1e-9)
            if not inside.any():
                continue
            val = scene.albedo(plane, p[..., a0][inside], p[..., a1][inside])
            img[inside] = val
            best[inside] = lam[inside]
    img8 = np.floor(img + 0.5).clip(0, 255)
    return img8.astype(np.float32), best.astype(np.float32)


def rel_pose(R_ref, t_ref, R_new, t_new):
    """refToNew = inv(newToWorld) * refToWorld, as double[12] (row-major R | t)"""
    R = R_new.T @ R_ref
    t = R_new.T @ (t_ref - t_new)
    return np.concatenate([R.reshape(-1), t]).astype(np.float64)


def make_sequence(seed, n_frames, w=640, h=480, fx=384.0, fy=384.0, cx=319.5, cy=239.5, start=0, stride=1):
    sc = Scene(seed)
    frames, depths, poses = [], [], []
    for k in range(n_frames):
        R, t = trajectory(start + k * stride)
        img, dep = render(sc, w, h, fx, fy, cx, cy, R, t)
        frames.append(img); depths.append(dep); poses.append((R, t))
    return frames, depths, poses


def pick_points(img, depth, n, seed=7, border=6):
    """n high-gradient pixels spread over the image (one per grid cell), with ground-truth inverse depth."""
    h, w = img.shape
    gy, gx = np.gradient(img.astype(np.float64))
    g = gx * gx + gy * gy
    g[:border] = 0; g[-border:] = 0; g[:, :border] = 0; g[:, -border:] = 0
    cell = max(2, int(np.sqrt(w * h / float(n))))
    pts = []
    for y0 in range(0, h - cell + 1, cell):
        for x0 in range(0, w - cell + 1, cell):
            blk = g[y0:y0 + cell, x0:x0 + cell]
            k = int(np.argmax(blk))
            if blk.flat[k] > 4.0:
                pts.append((x0 + k % cell, y0 + k // cell))
    pts = np.array(pts[: n], dtype=np.float32)
    rng = np.random.default_rng(seed)
    u = pts[:, 0] + rng.uniform(-0.3, 0.3, len(pts)).astype(np.float32)
    v = pts[:, 1] + rng.uniform(-0.3, 0.3, len(pts)).astype(np.float32)
    d = depth[pts[:, 1].astype(int), pts[:, 0].astype(int)]
    idepth = (1.0 / d).astype(np.float32)
    hdi = rng.uniform(1e-4, 5e-3, len(pts)).astype(np.float32)
    return u.astype(np.float32), v.astype(np.float32), idepth, hdi


PATTERN = np.array([[0, -2], [-1, -1], [1, -1], [-2, 0], [0, 0], [2, 0], [-1, 1], [0, 2]], dtype=np.int64)   # src/Setting.cc:275


def make_window(seed=1001, frame_ids=(40, 46, 52, 58, 64, 70, 76), pts_per_frame=300, w=640, h=480, fx=384.0, fy=384.0,
                cx=319.5, cy=239.5, pose_noise=2e-4, idepth_noise=0.02, lin_frac=0.25, rng_seed=5, rendered=None):
    """A frozen sliding window with realistic distributions: n keyframes of the synthetic sequence, points hosted in
    every frame with ground-truth inverse depth (+noise), one residual per (point, other frame).  Returns a dict of the
    arrays the window C ABI takes.  Frame states are slightly perturbed from ground truth so that the system has a
    non-trivial gradient; the FEJ evaluation point differs slightly from the current state.
    rendered: optional list of (img, depth, (R, t)) already rendered for frame_ids (bench.py reuses its sequence)."""
    from . import window_host as wh
    rng = np.random.default_rng(rng_seed)
    n = len(frame_ids)
    imgs, deps, w2c_gt = [], [], []
    if rendered is None:
        sc = Scene(seed)
        rendered = []
        for i in frame_ids:
            R, t = trajectory(i)
            img, dep = render(sc, w, h, fx, fy, cx, cy, R, t)
            rendered.append((img, dep, (R, t)))
    for img, dep, (R, t) in rendered:
        imgs.append(img); deps.append(dep)
        w2c_gt.append((R.T, -R.T @ t))

    def perturb(T, s):
        r = rng.normal(0, s, 3); dt = rng.normal(0, s, 3)
        th = np.linalg.norm(r)
        K = wh.hat(r / th) if th > 0 else np.zeros((3, 3))
        dR = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
        return dR @ T[0], dR @ T[1] + dt

    w2c_eval = [perturb(T, pose_noise) for T in w2c_gt]
    w2c = [perturb(T, pose_noise * 0.3) for T in w2c_eval]
    aff0 = [(0.0, 0.0)] * n
    aff = [(float(rng.normal(0, 0.01)), float(rng.normal(0, 0.5))) for _ in range(n)]
    exposure = [1.0] * n
    calib = (fx, fy, cx, cy)
    precalc = wh.precalc_table(w2c_eval, w2c, calib, aff, aff0, exposure)
    adH, adT = wh.adjoints(w2c_eval, aff0, exposure)
    state_minus_zero = [np.concatenate([rng.normal(0, 1e-3, 6), [aff[i][0] / wh.SCALE_A, aff[i][1] / wh.SCALE_B]]) for i in range(n)]
    adHTdelta = wh.ad_ht_delta(adH, adT, state_minus_zero)
    U, V, ID, IDZ, HOST, COL, WGT, PRIOR = [], [], [], [], [], [], [], []
    res_begin, target, state, flags = [0], [], [], []
    for hidx in range(n):
        u, v, idp, _ = pick_points(imgs[hidx], deps[hidx], pts_per_frame, seed=hidx + 11, border=8)
        u = np.round(u); v = np.round(v)                      # DSO points sit on integer pixels
        gy, gx = np.gradient(imgs[hidx].astype(np.float64))
        for k in range(len(u)):
            xi, yi = int(u[k]), int(v[k])
            px = xi + PATTERN[:, 0]; py = yi + PATTERN[:, 1]
            col = imgs[hidx][py, px]
            g2 = gx[py, px] ** 2 + gy[py, px] ** 2
            wgt = np.sqrt(2500.0 / (2500.0 + g2))            # ImmaturePoint ctor, ImmaturePoint.cc:27-31
            idn = idp[k] * (1 + idepth_noise * rng.standard_normal())
            U.append(u[k]); V.append(v[k]); ID.append(idn); IDZ.append(idn * (1 + 1e-3 * rng.standard_normal()))
            HOST.append(hidx); COL.append(col); WGT.append(wgt)
            PRIOR.append(2500.0 if (hidx == 0 and k % 7 == 0) else 0.0)
            for t in range(n):
                if t == hidx:
                    continue
                target.append(t); state.append(0)
                lin = rng.uniform() < lin_frac and t != n - 1
                flags.append((1 | 2) if lin else 4)          # linearised residuals are active; fresh ones are new
            res_begin.append(len(target))
    nr = len(target)
    D = 8 * n + 4
    A = rng.normal(0, 1, (D, D))
    HM = (A @ A.T) * 1e2
    return dict(n=n, frame_ids=list(frame_ids), imgs=imgs, depths=deps, calib=np.array(calib, np.float32),
                precalc=precalc, frame_energy_th=np.full(n, 8 * 8 * 8.0, np.float32), adH=adH, adT=adT, adHTdelta=adHTdelta,
                c_delta=rng.normal(0, 1e-3, 4).astype(np.float32), c_prior=np.full(4, 5e9), 
                frame_prior=np.where(np.arange(8 * n) < 8, 1e10, 0.0).astype(np.float64).reshape(n, 8),
                frame_delta_prior=rng.normal(0, 1e-4, (n, 8)),
                u=np.array(U, np.float32), v=np.array(V, np.float32), idepth=np.array(ID, np.float32),
                idepth_zero=np.array(IDZ, np.float32), host=np.array(HOST, np.int32), color=np.array(COL, np.float32),
                weights=np.array(WGT, np.float32), prior=np.array(PRIOR, np.float32), res_begin=np.array(res_begin, np.int32),
                target=np.array(target, np.int32), state=np.array(state, np.int32), flags=np.array(flags, np.uint8),
                res_to_zero=rng.normal(0, 2.0, (nr, 8)).astype(np.float32), HM=HM, bM=rng.normal(0, 10, D),
                delta=rng.normal(0, 1e-3, D), w2c_gt=w2c_gt, w2c=w2c)


def replicate_window(W, B):
    """B copies of every point of a make_window() dict (and of its residuals), kept contiguous per host frame like the
    reference's allPoints walk: the residual population of B sequences in one window (BASELINE configs[4])"""
    npnt = len(W["u"])
    order = np.argsort(np.tile(W["host"], B), kind="stable")
    src = np.tile(np.arange(npnt), B)[order]                      # source point of every replicated point
    cnt = np.diff(W["res_begin"])[src]
    rb = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    ridx = np.concatenate([np.arange(W["res_begin"][p], W["res_begin"][p + 1]) for p in src])
    out = dict(W)
    for k in ("u", "v", "idepth", "idepth_zero", "host", "color", "weights", "prior"):
        out[k] = W[k][src]
    for k in ("target", "state", "flags", "res_to_zero"):
        out[k] = W[k][ridx]
    out["res_begin"] = rb
    return out
