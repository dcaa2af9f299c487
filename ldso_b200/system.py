"""ctypes binding of include/ldso_system.h: FullSystem::addActiveFrame as one C call per frame.

Prefix-agnostic like capi.Api: `System(api, ...)` drives ldso_gpu_system_* of libldso_gpu.so (the product) or, in the
tests and the CPU baseline, the checker library's mirror of the same entry points under its own symbol prefix — the same
C++ control flow (include/ldso_system.hpp) over the two implementations of the hot path."""
import ctypes as C

import numpy as np

from . import capi

ST = dict(initialized=0, keyframe=1, lost=2, init_failed=3, window=4, points=5, residuals=6, immature=7, tries=8, frame_id=9)
CT = dict(frames=0, keyframes=1, marg_frames=2, marg_points=3, lin_residuals=4, dropped_points=5, activated=6, ba_iterations=7,
          track_calls=8, res_in_m=9)
KF_LOG = ("rmse", "points", "residuals", "marg_points", "dropped_points", "lin_residuals", "activated", "marg_frames",
          "min_act_dist", "frame_energy_th", "ba_iterations", "hm_norm")
PHASES = ("trackNewCoarse+makeImages", "traceNewCoarse", "activatePointsMT", "optimize", "marginalisation",
          "makeNewTraces+setCoarseTrackingRef", "initializer", "other", "[makeImages call]", "[trackNewestCoarse calls]",
          "[motion hypotheses]", "[unused]")


class SystemConfig(C.Structure):
    _fields_ = [
        ("gpu", capi.Config),
        ("desired_immature_density", C.c_float), ("desired_point_density", C.c_float), ("min_points_remaining", C.c_float),
        ("max_log_aff_fac_in_window", C.c_float), ("min_frames", C.c_int), ("max_frames", C.c_int), ("min_frame_age", C.c_int),
        ("max_opt_iterations", C.c_int), ("min_opt_iterations", C.c_int), ("th_opt_iterations", C.c_float),
        ("max_shift_weight_t", C.c_float), ("max_shift_weight_r", C.c_float), ("max_shift_weight_rt", C.c_float),
        ("kf_global_weight", C.c_float), ("max_affine_weight", C.c_float), ("idepth_fix_prior_marg_fac", C.c_float),
        ("initial_rot_prior", C.c_float), ("initial_trans_prior", C.c_float), ("initial_aff_b_prior", C.c_float),
        ("initial_aff_a_prior", C.c_float), ("initial_calib_hessian", C.c_float), ("min_idepth_h_marg", C.c_float),
        ("marg_weight_fac", C.c_float), ("min_trace_quality", C.c_float), ("re_track_threshold", C.c_float),
        ("min_good_active_res_for_marg", C.c_int), ("min_good_res_for_marg", C.c_int), ("initializer_slack_factor", C.c_float),
        ("point_selection", C.c_int), ("npoints_counter_mode", C.c_int), ("reserved", C.c_int * 8),
    ]


class System:
    """One sequence: ldso_gpu_system (or the checker's mirror) + convenience read-backs."""

    def __init__(self, api, calib=None, gpu_kw=None, **kw):
        self.api = api
        f = api.fn
        f("system_default_config").restype = None
        f("system_create").argtypes = [C.POINTER(SystemConfig), C.POINTER(C.c_void_p)]
        f("system_destroy").argtypes = [C.c_void_p]
        f("system_destroy").restype = None
        f("system_last_error").argtypes = [C.c_void_p]
        f("system_last_error").restype = C.c_char_p
        f("system_ctx").argtypes = [C.c_void_p]
        f("system_ctx").restype = C.c_void_p
        f("system_add_active_frame").argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_float, C.c_void_p, C.c_void_p]
        for name in ("system_get_trajectory", "system_get_window", "system_get_counters", "system_get_prior",
                     "system_last_keyframe_log", "system_get_phase_seconds"):
            f(name).argtypes = None
        cfg = SystemConfig()
        f("system_default_config")(C.byref(cfg))
        if calib is not None:
            for k in ("w", "h", "fx", "fy", "cx", "cy"):
                setattr(cfg.gpu, k, calib[k])
        for k, v in (gpu_kw or {}).items():
            setattr(cfg.gpu, k, v)
        for k, v in kw.items():
            setattr(cfg, k, v)
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = f("system_create")(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            msg = f("system_last_error")(self.h).decode() if self.h else "?"
            if self.h:
                f("system_destroy")(self.h)
                self.h = None
            raise capi.LdsoError(f"{api.prefix}system_create failed rc={rc}: {msg}")
        self._pose = np.zeros(12, np.float64)
        self._status = np.zeros(16, np.int32)
        self._pose_p = self._pose.ctypes.data_as(C.c_void_p)
        self._status_p = self._status.ctypes.data_as(C.c_void_p)
        self._add = f("system_add_active_frame")
        self.n_fed = 0

    def close(self):
        if self.h:
            self.api.fn("system_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def ctx(self):
        """the underlying context as a capi.Ctx view (timers, profile, counters); owned by the system"""
        v = capi.Ctx.__new__(capi.Ctx)
        v.api, v.cfg, v.h = self.api, self.cfg.gpu, None
        v.h = C.c_void_p(self.api.fn("system_ctx")(self.h))
        v.levels = self.api.fn("pyr_levels")(v.h)
        v.close = lambda: None
        return v

    def add_frame_ptr(self, ptr, is_device=False, timestamp=0.0, exposure=1.0):
        """hot loop: raw pointer in, status array (shared, overwritten by the next call) out.  is_device: False / True, or an
        LDSO_IMAGE_* kind (2 = host 8-bit sensor image through the device-side Undistort, see include/ldso_system.h)"""
        rc = self._add(self.h, ptr, int(is_device), timestamp, exposure, self._pose_p, self._status_p)
        if rc < 0:
            raise capi.LdsoError(f"{self.api.prefix}system_add_active_frame rc={rc}: {self.api.fn('system_last_error')(self.h).decode()}")
        self.n_fed += 1
        return self._status

    def add_frame(self, img, timestamp=None, exposure=1.0):
        img = np.ascontiguousarray(img, dtype=np.float32)
        assert img.size == self.cfg.gpu.w * self.cfg.gpu.h
        st = self.add_frame_ptr(img.ctypes.data_as(C.c_void_p), False, float(self.n_fed if timestamp is None else timestamp), exposure)
        return {k: int(st[i]) for k, i in ST.items()}, self._pose.copy()

    def trajectory(self, relative_to_ref=True):
        n = self.n_fed + 1
        ids = np.zeros(n, np.int32); poses = np.zeros((n, 12)); kf = np.zeros(n, np.int32)
        k = self.api.fn("system_get_trajectory")(self.h, C.c_int(n), ids.ctypes.data_as(C.c_void_p), poses.ctypes.data_as(C.c_void_p),
                                                 kf.ctypes.data_as(C.c_void_p), C.c_int(1 if relative_to_ref else 0))
        return ids[:k].copy(), poses[:k].copy(), kf[:k].astype(bool)

    def centres(self, relative_to_ref=True):
        """camera centres (world frame) of every frame with a valid pose: ids, [n,3], is_keyframe"""
        ids, poses, kf = self.trajectory(relative_to_ref)
        R = poses[:, :9].reshape(-1, 3, 3); t = poses[:, 9:]
        return ids, -np.einsum("nji,nj->ni", R, t), kf

    def window(self):
        ids = np.zeros(16, np.int32)
        k = self.api.fn("system_get_window")(self.h, C.c_int(16), ids.ctypes.data_as(C.c_void_p))
        return ids[:k].tolist()

    def counters(self):
        out = np.zeros(16, np.int64)
        self.api.fn("system_get_counters")(self.h, out.ctypes.data_as(C.c_void_p))
        return {k: int(out[i]) for k, i in CT.items()}

    def prior(self):
        n = self.api.fn("system_get_prior")(self.h, None, None)
        D = 8 * n + 4
        HM = np.zeros((D, D)); bM = np.zeros(D)
        self.api.fn("system_get_prior")(self.h, HM.ctypes.data_as(C.c_void_p), bM.ctypes.data_as(C.c_void_p))
        return HM, bM

    def keyframe_log(self):
        out = np.zeros(12)
        self.api.fn("system_last_keyframe_log")(self.h, out.ctypes.data_as(C.c_void_p))
        return dict(zip(KF_LOG, out.tolist()))

    def phase_seconds(self):
        out = np.zeros(12)
        self.api.fn("system_get_phase_seconds")(self.h, out.ctypes.data_as(C.c_void_p))
        return dict(zip(PHASES, out.tolist()))


def run(api, calib, frames, raw8=False, **kw):
    """feed a list of images; returns (system, per-frame status dicts, per-keyframe logs).  raw8: hand the frames over as
    8-bit sensor images (LDSO_IMAGE_HOST_RAW8: Undistort in passthrough + makeImages inside the library)"""
    s = System(api, calib, **kw)
    if raw8:
        s.ctx().undistort_set(calib["w"], calib["h"], photometric_mode=0)
    sts, kfs = [], []
    for img in frames:
        if raw8:
            r8 = np.ascontiguousarray(img, dtype=np.uint8)
            assert np.array_equal(r8.astype(np.float32), img)
            st = s.add_frame_ptr(r8.ctypes.data_as(C.c_void_p), 2, float(s.n_fed))
            st = {k: int(st[i]) for k, i in ST.items()}
        else:
            st, _ = s.add_frame(img)
        sts.append(st)
        if st["keyframe"]:
            kfs.append(dict(s.keyframe_log(), frame=st["frame_id"], window=s.window()))
    return s, sts, kfs


def ate_rmse(est, gt):
    """translation RMSE after a similarity (Umeyama) alignment of the estimated camera centres to ground truth;
    returns (rmse, scale)"""
    mu_e, mu_g = est.mean(0), gt.mean(0)
    E, G = est - mu_e, gt - mu_g
    U, S, Vt = np.linalg.svd(G.T @ E / len(est))
    d = np.sign(np.linalg.det(U @ Vt))
    Dm = np.diag([1, 1, d])
    R = U @ Dm @ Vt
    s = np.trace(np.diag(S) @ Dm) / max((E ** 2).sum() / len(est), 1e-30)
    aligned = (s * (R @ E.T)).T + mu_g
    return float(np.sqrt(((aligned - gt) ** 2).sum(1).mean())), float(s)


class Batch:
    """several Systems stepped together by the library's worker threads (ldso_gpu_batch_*, BASELINE.json configs[4])"""

    def __init__(self, systems, n_threads=None):
        self.systems = list(systems)
        self.api = self.systems[0].api
        n = len(self.systems)
        f = self.api.fn
        f("batch_create").argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
        f("batch_destroy").argtypes = [C.c_void_p]
        f("batch_destroy").restype = None
        f("batch_add_active_frames").argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                 C.c_void_p, C.c_void_p]
        hs = (C.c_void_p * n)(*[s.h for s in self.systems])
        self.h = C.c_void_p()
        rc = f("batch_create")(n, hs, n if n_threads is None else n_threads, C.byref(self.h))
        if rc != 0:
            raise capi.LdsoError(f"{self.api.prefix}batch_create rc={rc}")
        self.poses = np.zeros((n, 12))
        self.status = np.zeros((n, 16), np.int32)
        self._ptrs = (C.c_void_p * n)()
        self._step = f("batch_add_active_frames")
        self._pp, self._sp = self.poses.ctypes.data_as(C.c_void_p), self.status.ctypes.data_as(C.c_void_p)

    def step_ptrs(self, ptrs, is_device=False, sync_after=False):
        """ptrs: one raw image pointer per system; returns the shared [n,16] status array (overwritten by the next call)"""
        for i, p in enumerate(ptrs):
            self._ptrs[i] = p
        rc = self._step(self.h, self._ptrs, int(is_device), None, None, 1 if sync_after else 0, self._pp, self._sp)
        if rc < 0:
            msgs = [s.api.fn("system_last_error")(s.h).decode() for s in self.systems]
            raise capi.LdsoError(f"{self.api.prefix}batch_add_active_frames rc={rc}: {[m for m in msgs if m]}")
        for s in self.systems:
            s.n_fed += 1
        return self.status

    def run_ptrs(self, ptrs, is_device=False, sync_after=True):
        """ptrs[system][frame]: raw image pointers; every system consumes its frames in order without waiting for the others
        between frames.  Returns status [n][n_frames][16] (a fresh array)."""
        n, nf = len(self.systems), len(ptrs[0])
        arr = (C.c_void_p * (n * nf))(*[p for row in ptrs for p in row])
        status = np.zeros((n, nf, 16), np.int32)
        f = self.api.fn("batch_run")
        f.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        rc = f(self.h, nf, arr, int(is_device), 1 if sync_after else 0, None, status.ctypes.data_as(C.c_void_p))
        if rc < 0:
            msgs = [s.api.fn("system_last_error")(s.h).decode() for s in self.systems]
            raise capi.LdsoError(f"{self.api.prefix}batch_run rc={rc}: {[m for m in msgs if m]}")
        for s in self.systems:
            s.n_fed += nf
        return status

    def close(self):
        if self.h:
            self.api.fn("batch_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
