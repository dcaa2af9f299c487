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
