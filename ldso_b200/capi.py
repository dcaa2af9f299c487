"""ctypes binding of the C ABI declared in include/ldso_gpu.h.

`load()` opens ldso_b200/lib/libldso_gpu.so and FAILS LOUDLY when it is missing: there is no CPU
fallback anywhere in this package.  The `Api` class is prefix-agnostic so that the test-suite can drive
a second library with the same signatures under another symbol prefix through the very same harness.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
GPU_LIB = os.path.join(ROOT, "lib", "libldso_gpu.so")

MAX_LEVELS = 6


class Config(C.Structure):
    _fields_ = [
        ("w", C.c_int), ("h", C.c_int),
        ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
        ("device", C.c_int), ("n_frame_slots", C.c_int),
        ("max_points", C.c_int), ("max_immature", C.c_int), ("max_residuals", C.c_int),
        ("huber_th", C.c_float), ("coarse_cutoff_th", C.c_float), ("gamma_weights_pixel_select", C.c_int),
        ("affine_opt_mode_a", C.c_float), ("affine_opt_mode_b", C.c_float),
        ("outlier_th", C.c_float), ("outlier_th_sum_component", C.c_float), ("overall_energy_th_weight", C.c_float),
        ("max_pix_search", C.c_float), ("trace_stepsize", C.c_float), ("trace_gn_iterations", C.c_int),
        ("trace_gn_threshold", C.c_float), ("trace_extra_slack_on_th", C.c_float), ("trace_slack_interval", C.c_float),
        ("trace_min_improvement", C.c_float), ("min_trace_test_radius", C.c_int),
        ("min_grad_hist_cut", C.c_float), ("min_grad_hist_add", C.c_float), ("grad_downweight_per_level", C.c_float),
        ("select_direction_distribution", C.c_int),
        ("idepth_fix_prior", C.c_float), ("min_idepth_h_act", C.c_float), ("gn_its_on_point_activation", C.c_int),
        ("frame_energy_th_n", C.c_float), ("frame_energy_th_fac_median", C.c_float),
        ("frame_energy_th_const_weight", C.c_float),
        ("tracker_solve_mode", C.c_int),
        ("tracker_lm_mode", C.c_int),
        ("reserved", C.c_int * 7),
    ]


def _ptr(a, ctype):
    """pointer to a numpy array's data for a ctypes call; from_buffer is ~8x cheaper than ndarray.ctypes (0.6 vs 4.8 us,
    15-20 pointers per frame), but needs a writable, non-empty buffer"""
    if a is None:
        return None
    if a.size and a.flags.writeable:
        return C.byref(ctype.from_buffer(a))
    return a.ctypes.data_as(C.POINTER(ctype))


def _fp(a):
    return _ptr(a, C.c_float)


def _dp(a):
    return _ptr(a, C.c_double)


def _ip(a):
    return _ptr(a, C.c_int)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class LdsoError(RuntimeError):
    pass


class Api:
    """One loaded library exporting the ABI of include/ldso_gpu.h under symbol prefix `prefix`."""

    def __init__(self, path, prefix):
        if not os.path.exists(path):
            raise LdsoError(f"{path} is missing — build it first (python -m ldso_b200.build); there is no fallback path")
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        f = self.fn
        f("default_config").restype = None
        f("create").argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
        f("destroy").argtypes = [C.c_void_p]
        f("destroy").restype = None
        f("last_error").argtypes = [C.c_void_p]
        f("last_error").restype = C.c_char_p
        f("launch_count").argtypes = [C.c_void_p]
        f("launch_count").restype = C.c_longlong

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def default_config(self, **kw):
        cfg = Config()
        self.fn("default_config")(C.byref(cfg))
        for k, v in kw.items():
            setattr(cfg, k, v)
        if os.environ.get("LDSO_LM_DBG"):           # measurement only: 3 = SM clocks per LM phase on stderr (profile_get)
            cfg.reserved[0] = int(os.environ["LDSO_LM_DBG"])
        return cfg

    def create(self, cfg):
        return Ctx(self, cfg)


class Ctx:
    """One sequence context.  Thin, one method per C entry point."""

    def __init__(self, api, cfg):
        self.api = api
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = api.fn("create")(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            msg = api.fn("last_error")(self.h).decode() if self.h else "?"
            if self.h:
                api.fn("destroy")(self.h)
                self.h = None
            raise LdsoError(f"{api.prefix}create failed rc={rc}: {msg}")
        self.levels = api.fn("pyr_levels")(self.h)
        self.imm_n = 0
        self.sizes = []
        for l in range(self.levels):
            w, h = C.c_int(), C.c_int()
            api.fn("level_size")(self.h, l, C.byref(w), C.byref(h))
            self.sizes.append((w.value, h.value))

    def close(self):
        if self.h:
            self.api.fn("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _call(self, name, *args, ok=(0,)):
        rc = self.api.fn(name)(self.h, *args)
        if rc not in ok:
            raise LdsoError(f"{self.api.prefix}{name} rc={rc}: {self.api.fn('last_error')(self.h).decode()}")
        return rc

    def sync(self):
        self._call("sync")

    def launch_count(self):
        return self.api.fn("launch_count")(self.h)

    def io_bytes(self):
        a, b = C.c_longlong(), C.c_longlong()
        self._call("io_bytes", C.byref(a), C.byref(b))
        return a.value, b.value

    def timer_start(self):
        self._call("timer_start")

    def timer_stop(self):
        ms = C.c_float()
        self._call("timer_stop", C.byref(ms))
        return ms.value

    def profile_enable(self, on=True):
        self._call("profile_enable", C.c_int(1 if on else 0))

    def profile_get(self, which=0):
        ms, n, units = C.c_double(), C.c_longlong(), C.c_double()
        self._call("profile_get", C.c_int(which), C.byref(ms), C.byref(n), C.byref(units))
        return ms.value, n.value, units.value

    def set_gamma(self, B):
        B = f32(B)
        assert B.size == 256
        self._call("set_gamma", _fp(B))

    # K1
    def make_images(self, slot, img):
        img = f32(img)
        assert img.size == self.cfg.w * self.cfg.h
        self._call("frame_make_images", C.c_int(slot), _fp(img))

    def make_images_dev(self, slot, dev_ptr):
        self._call("frame_make_images_dev", C.c_int(slot), C.c_void_p(dev_ptr))

    def download(self, slot, lvl, what):
        w, h = self.sizes[lvl]
        out = np.empty((h, w, 3) if what == 0 else (h, w), dtype=np.float32)
        self._call("frame_download", C.c_int(slot), C.c_int(lvl), C.c_int(what), _fp(out))
        return out

    # K2
    def tracker_make_k(self, trk, fx, fy, cx, cy):
        self._call("tracker_make_k", C.c_int(trk), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy))

    def tracker_set_ref(self, trk, ref_slot, u, v, idepth, hdi, exposure=1.0, aff_a=0.0, aff_b=0.0):
        u, v, idepth, hdi = f32(u), f32(v), f32(idepth), f32(hdi)
        self._call("tracker_set_ref", C.c_int(trk), C.c_int(ref_slot), C.c_int(u.size), _fp(u), _fp(v), _fp(idepth), _fp(hdi),
                   C.c_float(exposure), C.c_float(aff_a), C.c_float(aff_b))

    def tracker_get_pc(self, trk, lvl):
        w, h = self.sizes[lvl]
        n = C.c_int()
        arrs = [np.empty(w * h, dtype=np.float32) for _ in range(4)]
        self._call("tracker_get_pc", C.c_int(trk), C.c_int(lvl), C.byref(n), *[_fp(a) for a in arrs])
        return [a[: n.value].copy() for a in arrs]

    # K3 / K4
    def tracker_calc_res(self, trk, new_slot, lvl, pose12, aff_a, aff_b, cutoff, exposure=1.0):
        pose12 = np.ascontiguousarray(pose12, dtype=np.float64)
        out = np.empty(6, dtype=np.float32)
        self._call("tracker_calc_res", C.c_int(trk), C.c_int(new_slot), C.c_float(exposure), C.c_int(lvl), _dp(pose12),
                   C.c_float(aff_a), C.c_float(aff_b), C.c_float(cutoff), _fp(out))
        return out

    def tracker_calc_res_noalloc(self, trk, new_slot, lvl, pose12, aff_a, aff_b, cutoff, exposure=1.0):
        """allocation-free variant (scratch in a file-scope device array of the library); returns (out6, sums52)"""
        pose12 = np.ascontiguousarray(pose12, dtype=np.float64)
        out = np.empty(6, dtype=np.float32)
        sums = np.empty(52, dtype=np.float64)
        self._call("tracker_calc_res_noalloc", C.c_int(trk), C.c_int(new_slot), C.c_float(exposure), C.c_int(lvl), _dp(pose12),
                   C.c_float(aff_a), C.c_float(aff_b), C.c_float(cutoff), _fp(out), _dp(sums))
        return out, sums

    def tracker_calc_gs(self, trk, lvl, pose12, aff_a, aff_b):
        pose12 = np.ascontiguousarray(pose12, dtype=np.float64)
        H = np.empty((8, 8), dtype=np.float64)
        b = np.empty(8, dtype=np.float64)
        self._call("tracker_calc_gs", C.c_int(trk), C.c_int(lvl), _dp(pose12), C.c_float(aff_a), C.c_float(aff_b), _dp(H), _dp(b))
        return H, b

    def tracker_track(self, trk, new_slot, pose12, aff, coarsest_lvl=None, min_res=None, exposure=1.0):
        pose = np.array(pose12, dtype=np.float64).copy()
        aff = np.array(aff, dtype=np.float32).copy()
        if coarsest_lvl is None:
            coarsest_lvl = self.levels - 1
        mr = f32(min_res if min_res is not None else [np.nan] * 5)
        last = np.empty(5, dtype=np.float32)
        flow = np.empty(3, dtype=np.float32)
        iters = np.zeros(2, dtype=np.int32)
        rc = self._call("tracker_track", C.c_int(trk), C.c_int(new_slot), C.c_float(exposure), _dp(pose), _fp(aff),
                        C.c_int(coarsest_lvl), _fp(mr), _fp(last), _fp(flow), _ip(iters), ok=(0, 1))
        return dict(ok=(rc == 0), pose=pose, aff=aff, last_residuals=last, flow=flow, iters=iters)


    # A2
    def select_make_maps(self, slot, density, potential=3, recursions=1, th_factor=1.0):
        w, h = self.sizes[0]
        out = np.zeros((h, w), dtype=np.float32)
        pot = C.c_int(potential)
        n = C.c_int()
        self._call("select_make_maps", C.c_int(slot), C.c_float(density), C.c_int(recursions), C.c_float(th_factor), C.byref(pot),
                   _fp(out), C.byref(n))
        return out, n.value, pot.value

    def select_points(self, slot, density, potential=3, recursions=1, th_factor=1.0, border=5, max_points=None):
        """makeMaps + row-major list of the selected pixels inside the border: (u, v, type, potential)"""
        w, h = self.sizes[0]
        cap = max_points if max_points is not None else (w * h) // 2
        u = np.empty(cap, np.float32); v = np.empty(cap, np.float32); t = np.empty(cap, np.float32)
        pot = C.c_int(potential)
        n = C.c_int()
        self._call("select_points", C.c_int(slot), C.c_float(density), C.c_int(recursions), C.c_float(th_factor), C.byref(pot),
                   C.c_int(border), C.c_int(cap), _fp(u), _fp(v), _fp(t), C.byref(n))
        return u[: n.value].copy(), v[: n.value].copy(), t[: n.value].copy(), pot.value

    # ---- K5 / K6: immature points ----
    def immature_create(self, slot, u, v):
        """ImmaturePoint constructor: (color[n,8], weights[n,8], gradH[n,3], energyTH[n])"""
        u, v = f32(u), f32(v)
        n = len(u)
        col = np.zeros((n, 8), np.float32); wgt = np.zeros((n, 8), np.float32)
        gh = np.zeros((n, 3), np.float32); eth = np.zeros(n, np.float32)
        self._call("immature_create", C.c_int(slot), C.c_int(n), _fp(u), _fp(v), _fp(col), _fp(wgt), _fp(gh), _fp(eth))
        return col, wgt, gh, eth

    def immature_set(self, host, u, v, color, weights, grad_h, energy_th, idepth_min, idepth_max, quality, status):
        host = np.ascontiguousarray(host, dtype=np.int32); status = np.ascontiguousarray(status, dtype=np.int32)
        a = [f32(x) for x in (u, v, color, weights, grad_h, energy_th, idepth_min, idepth_max, quality)]
        self.imm_n = len(host)
        self._call("immature_set", C.c_int(self.imm_n), _ip(host), *[_fp(x) for x in a], _ip(status))

    def immature_trace(self, new_slot, KRKi, Kt, aff):
        KRKi, Kt, aff = f32(KRKi), f32(Kt), f32(aff)
        cnt = np.zeros(6, np.int32)
        self._call("immature_trace", C.c_int(new_slot), C.c_int(len(Kt)), _fp(KRKi), _fp(Kt), _fp(aff), _ip(cnt))
        return cnt

    def immature_trace_async(self, new_slot, KRKi, Kt, aff):
        """the same trace queued without waiting; counts come from immature_trace_counts()"""
        KRKi, Kt, aff = f32(KRKi), f32(Kt), f32(aff)
        self._call("immature_trace_async", C.c_int(new_slot), C.c_int(len(Kt)), _fp(KRKi), _fp(Kt), _fp(aff))

    def immature_trace_counts(self):
        cnt = np.zeros(6, np.int32)
        self._call("immature_trace_counts", _ip(cnt))
        return cnt

    def immature_get(self, fields=None):
        """read back pool state; fields: subset of (idepth_min, idepth_max, quality, status, uv, interval), default all"""
        n = self.imm_n
        names = ("idepth_min", "idepth_max", "quality", "status", "uv", "interval")
        want = names if fields is None else fields
        out = {}
        for k in names:
            if k in want:
                out[k] = np.zeros((n, 2) if k == "uv" else n, np.int32 if k == "status" else np.float32)
        g = lambda k: (_ip(out[k]) if k == "status" else _fp(out[k])) if k in out else None
        self._call("immature_get", g("idepth_min"), g("idepth_max"), g("quality"), g("status"), g("uv"), g("interval"))
        return out

    def immature_compact(self, drop_host=-1, drop_oob=True, remove=None):
        rm = np.ascontiguousarray(remove if remove is not None else [], dtype=np.int32)
        n = C.c_int()
        self._call("immature_compact", C.c_int(drop_host), C.c_int(1 if drop_oob else 0), C.c_int(len(rm)), _ip(rm), C.byref(n))
        self.imm_n = n.value
        return n.value

    def immature_append(self, slot, host_idx, u, v, types=None):
        u, v = f32(u), f32(v)
        n = C.c_int()
        first = self.imm_n
        self._call("immature_append", C.c_int(slot), C.c_int(host_idx), C.c_int(len(u)), _fp(u), _fp(v), C.byref(n))
        self.imm_n = n.value
        if types is not None and len(u):
            self.immature_set_types(first, types)
        return n.value

    # ---- N2: CoarseDistanceMap + activation walk ----
    def immature_set_types(self, first, types):
        t = f32(types)
        self._call("immature_set_types", C.c_int(first), C.c_int(len(t)), _fp(t))

    def distmap_make(self, KRKi, Kt, host, u, v, idepth):
        KRKi, Kt = f32(KRKi).reshape(-1, 9), f32(Kt).reshape(-1, 3)
        host = np.ascontiguousarray(host, dtype=np.int32)
        u, v, idepth = f32(u), f32(v), f32(idepth)
        self._call("distmap_make", C.c_int(len(KRKi)), _fp(KRKi), _fp(Kt), C.c_int(len(host)), _ip(host), _fp(u), _fp(v), _fp(idepth))

    def distmap_get(self):
        w, h = self.sizes[1]
        d = np.zeros((h, w), np.float32)
        self._call("distmap_get", _fp(d))
        return d

    def immature_select_activation(self, KRKi, Kt, host_skip, host_flagged, min_act_dist, min_trace_quality=3.0):
        """returns (action[n_pool] uint8: 0 keep / 1 delete / 2 activate, (n_accepted, n_deleted))"""
        KRKi, Kt = f32(KRKi).reshape(-1, 9), f32(Kt).reshape(-1, 3)
        sk = np.ascontiguousarray(host_skip, dtype=np.uint8); fl = np.ascontiguousarray(host_flagged, dtype=np.uint8)
        assert len(sk) == len(KRKi) == len(fl) == len(Kt)
        act = np.zeros(max(self.imm_n, 1), np.uint8)
        cnt = np.zeros(2, np.int32)
        self._call("immature_select_activation", C.c_int(len(KRKi)), _fp(KRKi), _fp(Kt), sk.ctypes.data_as(C.c_void_p),
                   fl.ctypes.data_as(C.c_void_p), C.c_float(min_act_dist), C.c_float(min_trace_quality),
                   act.ctypes.data_as(C.c_void_p), _ip(cnt))
        return act[:self.imm_n], (int(cnt[0]), int(cnt[1]))

    # ---- N3: FeatureDetector::DetectCorners ----
    def detect_corners(self, slot, n_features, max_out=None):
        w, h = self.sizes[0]
        max_out = max_out or 4 * n_features + 64
        x, y, sc = (np.zeros(max_out, np.float32) for _ in range(3))
        fl = np.zeros(max_out, np.uint8)
        n, nc = C.c_int(), C.c_int()
        self._call("detect_corners", C.c_int(slot), C.c_int(n_features), C.c_int(max_out), _fp(x), _fp(y), _fp(sc),
                   fl.ctypes.data_as(C.c_void_p), C.byref(n), C.byref(nc))
        k = n.value
        return x[:k], y[:k], sc[:k], fl[:k].astype(bool), nc.value

    # ---- N4: undistortion in front of makeImages ----
    def undistort_set(self, w_org, h_org, G=None, vignette_inv=None, remap_x=None, remap_y=None, photometric_mode=2):
        G = None if G is None else f32(G)
        vig = None if vignette_inv is None else f32(vignette_inv)
        rx = None if remap_x is None else f32(remap_x)
        ry = None if remap_y is None else f32(remap_y)
        self._call("undistort_set", C.c_int(w_org), C.c_int(h_org), C.c_int(0 if G is None else len(G)), _fp(G), _fp(vig), _fp(rx),
                   _fp(ry), C.c_int(photometric_mode))

    def make_images_raw(self, slot, raw, exposure_time=1.0, factor=1.0):
        raw = np.ascontiguousarray(raw)
        assert raw.dtype in (np.uint8, np.uint16)
        self._call("frame_make_images_raw", C.c_int(slot), raw.ctypes.data_as(C.c_void_p), C.c_int(raw.dtype.itemsize),
                   C.c_float(exposure_time), C.c_float(factor))

    def undistort_get_image(self):
        w, h = self.sizes[0]
        out = np.zeros((h, w), np.float32)
        self._call("undistort_get_image", _fp(out))
        return out

    # ---- N1: CoarseInitializer ----
    def init_set_first(self, slot):
        n = np.zeros(MAX_LEVELS, np.int32)
        self._call("init_set_first", C.c_int(slot), _ip(n))
        self.init_n = n[:self.levels].copy()
        return self.init_n

    def init_track_frame(self, slot, exposure_first=1.0, exposure_new=1.0):
        pose = np.zeros(12, np.float64); aff = np.zeros(2, np.float32); res = np.zeros(3, np.float32)
        snapped, ready = C.c_int(), C.c_int()
        iters = np.zeros(MAX_LEVELS, np.int32)
        self._call("init_track_frame", C.c_int(slot), C.c_float(exposure_first), C.c_float(exposure_new), _dp(pose), _fp(aff),
                   C.byref(snapped), C.byref(ready), _fp(res), _ip(iters))
        return dict(pose=pose, aff=aff, snapped=bool(snapped.value), ready=bool(ready.value), res=res, iters=iters[:self.levels].copy())

    def init_get_points(self, lvl):
        n = int(self.init_n[lvl])
        m = max(n, 1)
        a = {k: np.zeros(m, np.float32) for k in ("u", "v", "idepth", "iR", "type", "lastHessian", "energy")}
        good = np.zeros(m, np.uint8)
        nn = C.c_int()
        self._call("init_get_points", C.c_int(lvl), C.byref(nn), _fp(a["u"]), _fp(a["v"]), _fp(a["idepth"]), _fp(a["iR"]),
                   good.ctypes.data_as(C.c_void_p), _fp(a["type"]), _fp(a["lastHessian"]), _fp(a["energy"]))
        assert nn.value == n
        out = {k: x[:n] for k, x in a.items()}
        out["isGood"] = good[:n].astype(bool)
        return out

    def init_get_graph(self, lvl):
        n = int(self.init_n[lvl])
        nb = np.zeros((max(n, 1), 10), np.int32); par = np.zeros(max(n, 1), np.int32)
        self._call("init_get_graph", C.c_int(lvl), _ip(nb), _ip(par))
        return nb[:n], par[:n]

    def init_calc_res_and_gs(self, slot, lvl, pose12, aff=(0.0, 0.0)):
        pose = np.ascontiguousarray(pose12, dtype=np.float64)
        H = np.zeros(64, np.float32); b = np.zeros(8, np.float32); Hsc = np.zeros(64, np.float32); bsc = np.zeros(8, np.float32)
        res = np.zeros(3, np.float32)
        self._call("init_calc_res_and_gs", C.c_int(slot), C.c_int(lvl), _dp(pose), C.c_float(aff[0]), C.c_float(aff[1]),
                   _fp(H), _fp(b), _fp(Hsc), _fp(bsc), _fp(res))
        return H.reshape(8, 8), b, Hsc.reshape(8, 8), bsc, res

    def immature_activate(self, cand, min_obs=1):
        cand = np.ascontiguousarray(cand, dtype=np.int32)
        n = len(cand)
        idepth = np.zeros(n, np.float32); ok = np.zeros(n, np.int32); st = np.zeros((n, self.win_n), np.int32)
        self._call("immature_activate", C.c_int(n), _ip(cand), C.c_int(min_obs), _fp(idepth), _ip(ok), _ip(st))
        return idepth, ok, st

    # ---- windowed BA (K7-K10, H1) ----
    def window_set_frames(self, W, slots=None):
        n = W["n"]
        slots = np.ascontiguousarray(slots if slots is not None else np.arange(n), dtype=np.int32)
        self.win_n = n
        pc, eth, cal = f32(W["precalc"]), f32(W["frame_energy_th"]), f32(W["calib"])
        self._call("window_set_frames", C.c_int(n), _ip(slots), _fp(pc), _fp(eth), _fp(cal))

    def window_load(self, W, slots=None):
        """upload a window description dict (see synth.make_window); frames must already be in their slots"""
        n = W["n"]
        slots = np.ascontiguousarray(slots if slots is not None else np.arange(n), dtype=np.int32)
        self.win_n, self.win_np, self.win_nr = n, len(W["u"]), len(W["target"])
        self.window_set_frames(W, slots)
        a = [f32(W[k]) for k in ("u", "v", "idepth", "idepth_zero")]
        host = np.ascontiguousarray(W["host"], dtype=np.int32)
        col, wgt, pri = f32(W["color"]), f32(W["weights"]), f32(W["prior"])
        rb = np.ascontiguousarray(W["res_begin"], dtype=np.int32)
        self._call("window_set_points", C.c_int(self.win_np), _fp(a[0]), _fp(a[1]), _fp(a[2]), _fp(a[3]), _ip(host), _fp(col),
                   _fp(wgt), _fp(pri), _ip(rb))
        tg = np.ascontiguousarray(W["target"], dtype=np.int32)
        st = np.ascontiguousarray(W["state"], dtype=np.int32)
        fl = np.ascontiguousarray(W["flags"], dtype=np.uint8)
        rtz = f32(W["res_to_zero"])
        self._call("window_set_residuals", C.c_int(self.win_nr), _ip(tg), _ip(st), fl.ctypes.data_as(C.c_void_p), _fp(rtz))
        adH = np.ascontiguousarray(W["adH"], dtype=np.float64); adT = np.ascontiguousarray(W["adT"], dtype=np.float64)
        dl, cd = f32(W["adHTdelta"]), f32(W["c_delta"])
        cp = np.ascontiguousarray(W["c_prior"], dtype=np.float64)
        fp_ = np.ascontiguousarray(W["frame_prior"], dtype=np.float64)
        fd = np.ascontiguousarray(W["frame_delta_prior"], dtype=np.float64)
        self._call("window_set_adjoints", _dp(adH), _dp(adT), _fp(dl), _fp(cd), _dp(cp), _dp(fp_), _dp(fd))

    def window_set_adjoints_only(self, adH, adT, W):
        """setAdjointsF after a pose step; the delta / prior terms of W are re-sent unchanged"""
        adH = np.ascontiguousarray(adH, dtype=np.float64); adT = np.ascontiguousarray(adT, dtype=np.float64)
        dl, cd = f32(W["adHTdelta"]), f32(W["c_delta"])
        cp = np.ascontiguousarray(W["c_prior"], dtype=np.float64)
        fp_ = np.ascontiguousarray(W["frame_prior"], dtype=np.float64)
        fd = np.ascontiguousarray(W["frame_delta_prior"], dtype=np.float64)
        self._call("window_set_adjoints", _dp(adH), _dp(adT), _fp(dl), _fp(cd), _dp(cp), _dp(fp_), _dp(fd))

    def window_set_idepth(self, idepth, idepth_zero):
        a, b = f32(idepth), f32(idepth_zero)
        self._call("window_set_idepth", _fp(a), _fp(b))

    def window_linearize(self, fix=False):
        stats = np.zeros(3, dtype=np.float64)
        th = C.c_float()
        self._call("window_linearize", C.c_int(1 if fix else 0), _dp(stats), C.byref(th))
        return stats, th.value

    def window_apply_res(self):
        self._call("window_apply_res")

    def window_reset_oob(self):
        self._call("window_reset_oob")

    def window_fix_linearization(self, mask):
        m = np.ascontiguousarray(mask, dtype=np.uint8)
        self._call("window_fix_linearization", m.ctypes.data_as(C.c_void_p))

    def window_accumulate(self):
        D = 8 * self.win_n + 4
        Hs = [np.zeros((D, D)) for _ in range(3)]
        bs = [np.zeros(D) for _ in range(3)]
        cnt = np.zeros(2, dtype=np.int32)
        self._call("window_accumulate", _dp(Hs[0]), _dp(bs[0]), _dp(Hs[1]), _dp(bs[1]), _dp(Hs[2]), _dp(bs[2]), _ip(cnt))
        return dict(HA=Hs[0], bA=bs[0], HL=Hs[1], bL=bs[1], Hsc=Hs[2], bsc=bs[2], counts=cnt)

    def window_marginalize_points(self, marg, fac=600.0 * 600.0):
        """EnergyFunctional::marginalizePointsF: returns (H, b, n_res) with H = M - Msc of the masked points"""
        D = 8 * self.win_n + 4
        m = np.ascontiguousarray(marg, dtype=np.uint8)
        assert len(m) == self.win_np
        H, b, nres = np.zeros((D, D)), np.zeros(D), C.c_int(0)
        self._call("window_marginalize_points", m.ctypes.data_as(C.c_void_p), C.c_float(fac), _dp(H), _dp(b), C.byref(nres))
        return H, b, nres.value

    def window_marginalize_frame(self, n, idx, HM, bM, prior, delta_prior):
        """EnergyFunctional::marginalizeFrame (dense part): returns (HM', bM') of dimension 8(n-1)+4"""
        c64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        HM, bM, prior, delta_prior = c64(HM), c64(bM), c64(prior), c64(delta_prior)
        D = 8 * n + 4
        assert HM.shape == (D, D) and bM.shape == (D,)
        Ho, bo = np.zeros((D - 8, D - 8)), np.zeros(D - 8)
        self._call("window_marginalize_frame", C.c_int(n), C.c_int(idx), _dp(HM), _dp(bM), _dp(prior), _dp(delta_prior), _dp(Ho), _dp(bo))
        return Ho, bo

    def window_calc_energy(self, HM, bM, delta):
        """(calcLEnergyF_MT, calcMEnergyF)"""
        c64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        HM, bM, delta = c64(HM), c64(bM), c64(delta)
        e = np.zeros(2)
        self._call("window_calc_energy", _dp(HM), _dp(bM), _dp(delta), _dp(e))
        return float(e[0]), float(e[1])

    def window_solve(self, iteration, lam, acc, HM, bM, delta, N=None):
        D = 8 * self.win_n + 4
        x = np.zeros(D)
        c64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        HM, bM, delta = c64(HM), c64(bM), c64(delta)
        Nn = c64(N) if N is not None else None
        self._call("window_solve", C.c_int(iteration), C.c_double(lam), _dp(acc["HA"]), _dp(acc["bA"]), _dp(acc["HL"]),
                   _dp(acc["bL"]), _dp(acc["Hsc"]), _dp(acc["bsc"]), _dp(HM), _dp(bM), _dp(delta), _dp(Nn),
                   C.c_int(0 if N is None else Nn.shape[1]), _dp(x))
        return x

    def window_resubstitute(self, x, lam=0.0):
        x = np.ascontiguousarray(x, dtype=np.float64)
        step = np.zeros(self.win_np, dtype=np.float32)
        self._call("window_resubstitute", _dp(x), C.c_float(lam), _fp(step))
        return step

    def window_residual_floats(self, what):
        k = {0: 72, 1: 8, 2: 8, 3: 3}.get(what, 1)
        out = np.zeros((self.win_nr, k) if k > 1 else self.win_nr, dtype=np.float32)
        self._call("window_get_residual_floats", C.c_int(what), _fp(out))
        return out

    def window_residual_ints(self, what):
        out = np.zeros(self.win_nr, dtype=np.int32)
        self._call("window_get_residual_ints", C.c_int(what), _ip(out))
        return out

    def window_point_floats(self, what):
        k = 4 if what in (2, 5) else 1
        out = np.zeros((self.win_np, k) if k > 1 else self.win_np, dtype=np.float32)
        self._call("window_get_point_floats", C.c_int(what), _fp(out))
        return out


_gpu_api = None


def load():
    """The product library.  Raises if it has not been built."""
    global _gpu_api
    if _gpu_api is None:
        _gpu_api = Api(GPU_LIB, "ldso_gpu_")
    return _gpu_api
