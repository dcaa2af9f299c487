"""Closed-loop direct sparse odometry on top of the C ABI — TEST / BENCH HARNESS, not product code.

A compact restatement of the control flow of the reference's FullSystem (src/frontend/FullSystem.cc: addActiveFrame
:78-187, trackNewCoarse :218-468, makeKeyFrame :485-699, optimize :844-1007, activatePointsMT :1197-1337,
flagPointsForRemoval :1356-1424, makeNewTraces :1426-1482) that drives EVERY stage of the hot path through one
prefix-agnostic binding, so the same loop runs the CUDA library and the CPU oracle and their trajectories can be
compared with each other and with ground truth (the ATE parity gate).

Deliberate simplifications (host-side orchestration is outside SURVEY section 8a; they are the same for both libraries):
  * bootstrap either through CoarseInitializer (N1; add_frame without gt_depth) or from ground-truth depth of the first
    frame (add_frame(gt_depth=...), kept for the tests that need a metric trajectory);
  * no marginalisation prior: points and frames that leave the window are dropped, HM = bM = 0, the gauge is held by a
    prior on the oldest keyframe; consequently Jacobians are evaluated at the current state (no FEJ, no linearised
    residuals) and steps are always accepted (setting_forceAceptStep = true in the reference, Setting.cc);
  * the immature pool lives on the device (append / trace / select_activation / activate / compact); the host keeps only
    the static fields it needs to turn an activated point into a PointHessian (u, v, color, weights)."""
import time

import numpy as np

from . import window_host as wh

IN, OOB, OUTLIER = 0, 1, 2
GOOD, T_OOB, T_OUTLIER, SKIPPED, BADCOND, UNINIT = range(6)


def se3_exp(xi):
    """xi = [upsilon, omega] -> (R, t) (Sophus SE3::exp)"""
    ups, om = np.asarray(xi[:3], float), np.asarray(xi[3:], float)
    th = np.linalg.norm(om)
    O = wh.hat(om)
    if th < 1e-10:
        R = np.eye(3) + O
        V = np.eye(3) + 0.5 * O
    else:
        A, B, C = np.sin(th) / th, (1 - np.cos(th)) / th ** 2, (th - np.sin(th)) / th ** 3
        R = np.eye(3) + A * O + B * O @ O
        V = np.eye(3) + B * O + C * O @ O
    return R, V @ ups


_LIBC = []


def _libc():
    """glibc rand(): the reference sub-selects the initializer's points with it (FullSystem.cc:1486, :1516)"""
    if not _LIBC:
        import ctypes
        _LIBC.append(ctypes.CDLL("libc.so.6"))
    return _LIBC[0]


def odometry_from_rt12(p):
    return p[:9].reshape(3, 3).copy(), p[9:].copy()


def rt12(T):
    return np.concatenate([T[0].reshape(-1), T[1]])


def from_rt12(p):
    return p[:9].reshape(3, 3).copy(), p[9:].copy()


class Odometry:
    def __init__(self, api, calib, desired_points=2000, desired_immature=1500, max_frames=7, ba_iterations=6, seed=0,
                 point_selection=0, **cfg_kw):
        self.c = dict(calib)
        cfg = api.default_config(w=calib["w"], h=calib["h"], fx=calib["fx"], fy=calib["fy"], cx=calib["cx"], cy=calib["cy"],
                                 affine_opt_mode_a=0.0, affine_opt_mode_b=0.0, n_frame_slots=max_frames + 3,
                                 max_points=4 * desired_points, max_residuals=4 * desired_points * max_frames, **cfg_kw)
        self.ctx = api.create(cfg)
        self.desired_points, self.desired_immature, self.max_frames, self.ba_its = desired_points, desired_immature, max_frames, ba_iterations
        self.point_selection = point_selection       # setting_pointSelection: 0 = DSO PixelSelector, 1 = LDSO corner detector (N3)
        self.K = np.array([[calib["fx"], 0, calib["cx"]], [0, calib["fy"], calib["cy"]], [0, 0, 1]], dtype=np.float32)
        self.Ki = np.linalg.inv(self.K.astype(np.float64)).astype(np.float32)
        self.frames = []                 # keyframes in the window: dict(id, slot, w2c=(R,t), aff=(a,b))
        self.free_slots = list(range(cfg.n_frame_slots))
        # active points (PointHessian)
        self.P = dict(host=np.zeros(0, np.int64), u=np.zeros(0, np.float32), v=np.zeros(0, np.float32), idepth=np.zeros(0, np.float32),
                      color=np.zeros((0, 8), np.float32), weights=np.zeros((0, 8), np.float32), hdi=np.zeros(0, np.float32))
        self.res = np.zeros((0, 0), np.int8)       # [point][frame position]: -1 no residual, else ResState
        # immature points (host-managed pool; host = frame id)
        self.I = None
        self.potential = 3
        self.min_act_dist = 2.0
        self.poses = {}                  # frame id -> world-to-camera (R, t) of every tracked frame
        self.last_kf_pose_rel = None     # newest keyframe -> last frame
        self.prev_rel = None
        self.ref_id = None
        self.first_rmse = -1.0
        self.log = []
        self.phase = {}
        self.n_frames_seen = 0
        self.init_first = None           # first frame of the CoarseInitializer while the system is not initialised

    # ---- small helpers -------------------------------------------------------------------------------
    def pos_of(self, fid):
        return [f["id"] for f in self.frames].index(fid)

    def positions(self, host_ids):
        """window position of each host frame id (keyframe ids are increasing along self.frames)"""
        ids = np.array([f["id"] for f in self.frames], np.int64)
        return np.searchsorted(ids, np.asarray(host_ids, np.int64))

    def _window_stacks(self):
        """stacked R, t and affine of the window frames, rebuilt only when a frame's pose / affine object changed"""
        Rs = [f["w2c"][0] for f in self.frames]; ts = [f["w2c"][1] for f in self.frames]; affs = [f["aff"] for f in self.frames]
        c = getattr(self, "_stack_cache", None)
        if (c is None or len(c[0]) != len(Rs) or any(a is not b for a, b in zip(c[0], Rs)) or any(a is not b for a, b in zip(c[1], ts))
                or c[2] != affs):
            c = self._stack_cache = (Rs, ts, affs, np.stack(Rs), np.stack(ts), np.array(affs, np.float32))   # holds the keyed objects alive
        return c[3], c[4], c[5]

    def rel_to(self, w2c_new):
        """(R, t) of every window frame -> the frame with pose w2c_new, stacked"""
        Rh, th, _ = self._window_stacks()
        R = w2c_new[0] @ Rh.transpose(0, 2, 1)
        t = w2c_new[1] - np.einsum("nij,nj->ni", R, th)
        return R, t

    def host_tables(self, w2c_new, aff_new=(0.0, 0.0)):
        """hostToFrame KRKi / Kt / affine per keyframe of the window (FullSystem.cc:1171-1176)"""
        R, t = self.rel_to(w2c_new)
        KRKi = (self.K @ R.astype(np.float32)) @ self.Ki
        Kt = t.astype(np.float32) @ self.K.T
        fa = self._window_stacks()[2]
        a = np.exp(np.float32(aff_new[0]) - fa[:, 0]).astype(np.float32)          # AffLight::fromToVecExposure, exposures 1
        b = (np.float32(aff_new[1]) - a * fa[:, 1]).astype(np.float32)
        return np.ascontiguousarray(KRKi, np.float32), np.ascontiguousarray(Kt, np.float32), np.stack([a, b], 1)

    def make_new_traces(self, f):
        """FullSystem::makeNewTraces (:1426-1482): selector + ImmaturePoint constructor on keyframe f, appended to the device pool"""
        if self.point_selection == 1:
            # FullSystem.cc:1428-1438: every feature of FeatureDetector::DetectCorners becomes an immature point of type 1
            u, v, _, corner, _ = self.ctx.detect_corners(f["slot"], self.desired_immature)
            typ = np.ones(len(u), np.float32)
            f["n_corners"] = int(corner.sum())
        else:
            u, v, typ, self.potential = self.ctx.select_points(f["slot"], self.desired_immature, potential=self.potential, border=5)
        col, wgt, gh, eth = self.ctx.immature_create(f["slot"], u, v)
        ok = np.isfinite(eth)
        k = int(ok.sum())
        new = dict(host=np.full(k, f["id"], np.int64), u=u[ok], v=v[ok], color=col[ok], weights=wgt[ok], type=typ[ok])
        self.I = new if self.I is None else {k2: np.concatenate([self.I[k2], new[k2]]) for k2 in new}
        self.ctx.immature_append(f["slot"], len(self.frames) - 1, new["u"], new["v"], types=new["type"])
        return k

    def fetch_trace_counts(self):
        """status counts of the trace queued by the previous frame (waits only if it is still running)"""
        if getattr(self, "_trace_log", None) is not None:
            self._trace_log["trace"] = self.ctx.immature_trace_counts().tolist()
            self._trace_log = None

    # ---- per frame (FullSystem::addActiveFrame, :78-187) ---------------------------------------------
    def tick(self, name, t0):
        t1 = time.perf_counter()
        self.phase[name] = self.phase.get(name, 0.0) + (t1 - t0)
        return t1

    def add_frame(self, img, gt_depth=None, gt_w2c=None, dev_ptr=None):
        """img: host float image; dev_ptr: optional device pointer of the same image (skips the upload)"""
        fid = self.n_frames_seen
        self.n_frames_seen += 1
        slot = self.free_slots.pop()
        t = time.perf_counter()
        if dev_ptr is None:
            self.ctx.make_images(slot, img)
        else:
            self.ctx.make_images_dev(slot, dev_ptr)
        t = self.tick("K1 makeImages", t)
        if not self.frames:
            w2c = gt_w2c if gt_w2c is not None else (np.eye(3), np.zeros(3))
            if gt_depth is None:
                return self.initializer_step(fid, slot, w2c)
            f = dict(id=fid, slot=slot, w2c=(w2c[0].copy(), w2c[1].copy()), aff=(0.0, 0.0))
            self.bootstrap(f, gt_depth)
            self.poses[fid] = f["w2c"]
            return dict(kf=True, ok=True)
        ref = self.frames[-1]
        ok, rel, aff, tr = self.track(slot)
        self.fetch_trace_counts()
        t = self.tick("K3/K4 trackNewestCoarse", t)
        if not ok:
            self.free_slots.append(slot)
            self.log.append(dict(frame=fid, lost=True))
            return dict(kf=False, ok=False)
        w2c = wh.se3_mul(rel, ref["w2c"])
        self.poses[fid] = w2c
        # keyframe decision (:160-176, setting_keyframesPerSecond = 0); the shift weights are tied to 640+480 in Setting.cc
        wsum = self.c["w"] + self.c["h"]
        a_rel, _ = wh.from_to_vec_exposure(1.0, 1.0, ref["aff"], aff)
        score = (0.04 * 1120 * np.sqrt(tr["flow"][0]) + 0.0 * np.sqrt(tr["flow"][1]) + 0.02 * 1120 * np.sqrt(tr["flow"][2])) / wsum \
            + 2.0 * abs(np.log(a_rel))
        res0 = float(tr["last_residuals"][0])
        if self.first_rmse < 0:
            self.first_rmse = res0
        need_kf = score > 1 or 2 * self.first_rmse < res0
        # traceNewCoarse (:1157-1195)
        # The trace is queued without waiting: nothing reads its results before the next keyframe's activation, so its
        # status counts (log only) are fetched after the NEXT frame's tracker has returned, when the stream has drained.
        traced = self.I is not None and len(self.I["host"]) > 0
        if traced:
            KRKi, Kt, affs = self.host_tables(w2c, aff)
            self.ctx.immature_trace_async(slot, KRKi, Kt, affs)
        t = self.tick("K5 traceNewCoarse", t)
        self.log.append(dict(frame=fid, kf=bool(need_kf), score=float(score), res=res0, iters=tr["iters"].tolist(), trace=None))
        self._trace_log = self.log[-1] if traced else None
        if need_kf:
            self.make_keyframe(dict(id=fid, slot=slot, w2c=w2c, aff=(float(aff[0]), float(aff[1]))))
            t = self.tick("KF makeKeyFrame total (Python orchestration included)", t)
        else:
            self.free_slots.insert(0, slot)      # oldest-first reuse: the queued trace may still be reading this slot
        return dict(kf=bool(need_kf), ok=True)

    def track(self, slot):
        """FullSystem::trackNewCoarse (:218-468): motion hypotheses, first acceptable one wins"""
        ref = self.frames[-1]
        last = self.n_frames_seen - 2                                         # the two previous frames, normally consecutive ids
        ids = [last - 1, last] if last in self.poses and last - 1 in self.poses else sorted(self.poses)
        T_ref = ref["w2c"]
        T_last = self.poses[ids[-1]]
        last_rel = wh.se3_mul(T_last, wh.se3_inv(T_ref))                     # lastF_2_slast

        def tries():                                                          # built on demand: the first one usually wins
            if len(ids) >= 2:
                motion = wh.se3_mul(T_last, wh.se3_inv(self.poses[ids[-2]]))   # sprelast -> slast, assumed to repeat
                yield wh.se3_mul(motion, last_rel)                            # constant motion
                yield wh.se3_mul(motion, wh.se3_mul(motion, last_rel))        # double motion
            yield last_rel                                                    # zero motion
            yield (np.eye(3), np.zeros(3))                                    # zero motion from the keyframe
        aff0 = self.last_aff if hasattr(self, "last_aff") else ref["aff"]
        best = None
        for T in tries():
            r = self.ctx.tracker_track(0, slot, rt12(T), [aff0[0], aff0[1]])
            if r["ok"] and np.isfinite(r["last_residuals"][0]):
                if best is None or r["last_residuals"][0] < best[3]["last_residuals"][0]:
                    best = (True, from_rt12(r["pose"]), r["aff"], r)
                if r["last_residuals"][0] < 1.5 * max(self.first_rmse, 1.0) or self.first_rmse < 0:
                    break
        if best is None:
            return False, None, None, None
        self.last_aff = (float(best[2][0]), float(best[2][1]))
        return best

    # ---- bootstrap through CoarseInitializer (FullSystem::addActiveFrame :92-126, initializeFromInitializer :1484-1665) ----
    def initializer_step(self, fid, slot, w2c_first):
        """frames before the system is initialised: setFirst on the first one, trackFrame on the following ones; when the
        initializer reports ready the first frame becomes keyframe 0 with the initializer's points and the current frame
        goes through makeKeyFrame"""
        t = time.perf_counter()
        if self.init_first is None:
            self.init_first = dict(id=fid, slot=slot, w2c=(w2c_first[0].copy(), w2c_first[1].copy()), aff=(0.0, 0.0))
            self.ctx.init_set_first(slot)
            self.tick("N1 CoarseInitializer", t)
            return dict(kf=False, ok=True, init=True)
        r = self.ctx.init_track_frame(slot)
        self.tick("N1 CoarseInitializer", t)
        if not r["ready"]:
            self.free_slots.append(slot)
            self.log.append(dict(frame=fid, init=True, snapped=r["snapped"]))
            return dict(kf=False, ok=True, init=True)
        first = self.init_first
        P0 = self.ctx.init_get_points(0)
        npts = len(P0["u"])
        rescale = 1.0 / ((1e-5 + float(np.sum(P0["iR"], dtype=np.float32))) / (1e-5 + npts))
        keep = self.desired_points / float(npts)
        libc = _libc()
        libc.srand(549812361)                                              # FullSystem.cc:1486
        sel = np.array([libc.rand() / float(2147483647) <= keep for _ in range(npts)], bool)
        u, v = (P0["u"][sel] + np.float32(0.5)).astype(np.float32), (P0["v"][sel] + np.float32(0.5)).astype(np.float32)
        col, wgt, gh, eth = self.ctx.immature_create(first["slot"], u, v)
        ok = np.isfinite(eth)
        idp = (P0["iR"][sel][ok] * np.float32(rescale)).astype(np.float32)
        n = int(ok.sum())
        self.frames.append(first)
        self.poses[first["id"]] = first["w2c"]
        self.P = dict(host=np.full(n, first["id"], np.int64), u=u[ok], v=v[ok], idepth=idp, color=col[ok], weights=wgt[ok],
                      hdi=np.full(n, 1e-3, np.float32))
        self.res = np.full((n, 1), -1, np.int8)
        R, tt = odometry_from_rt12(r["pose"])
        rel = (R, tt / rescale)                                            # firstToNew.translation() /= rescaleFactor (:1547)
        w2c = wh.se3_mul(rel, first["w2c"])
        self.poses[fid] = w2c
        self.init_first = None
        self.log.append(dict(frame=fid, kf=True, init_done=True, rescale=float(rescale), init_points=n))
        self.make_keyframe(dict(id=fid, slot=slot, w2c=w2c, aff=(float(r["aff"][0]), float(r["aff"][1]))))
        self.tick("KF makeKeyFrame total (Python orchestration included)", time.perf_counter())
        return dict(kf=True, ok=True, init=True)

    # ---- bootstrap: the first keyframe with ground-truth depth (stands in for CoarseInitializer) ----
    def bootstrap(self, f, gt_depth):
        self.frames.append(f)
        u, v, typ, self.potential = self.ctx.select_points(f["slot"], self.desired_points, potential=self.potential, border=5)
        col, wgt, gh, eth = self.ctx.immature_create(f["slot"], u, v)
        ok = np.isfinite(eth)
        u, v, col, wgt = u[ok], v[ok], col[ok], wgt[ok]
        idp = (1.0 / gt_depth[v.astype(int), u.astype(int)]).astype(np.float32)
        n = len(u)
        self.P = dict(host=np.full(n, f["id"], np.int64), u=u, v=v, idepth=idp, color=col, weights=wgt, hdi=np.full(n, 1e-3, np.float32))
        self.res = np.full((n, 1), -1, np.int8)
        self.make_new_traces(f)
        self.ctx.tracker_set_ref(0, f["slot"], u, v, idp, self.P["hdi"], 1.0, 0.0, 0.0)
        self.ref_id = f["id"]
        self.first_rmse = -1.0

    # ---- keyframe (FullSystem::makeKeyFrame, :485-699) -------------------------------------------------
    def make_keyframe(self, f):
        t = time.perf_counter()
        drop = self.flag_frames_for_marginalization(f)
        self.frames.append(f)
        npts = len(self.P["host"])
        self.res = np.concatenate([self.res, np.full((npts, 1), IN, np.int8)], axis=1)     # new residuals into f (:549-560)
        t = self.tick("KF  pool sync + frame selection", t)
        n_act = self.activate_points(drop)
        t = self.tick("KF  activatePointsMT (K6 + host candidate spacing)", t)
        rmse = self.optimize()
        t = self.tick("KF  optimize (K7-K10, H1 + host pose tables)", t)
        st_newest = self.res[:, -1].copy()
        st_prev = self.res[:, -2].copy() if self.res.shape[1] >= 2 else st_newest
        self.res[self.res != IN] = -1                                      # linearizeAll(true): non-IN residuals are dropped
        # flagPointsForRemoval (:1356-1424) without the marginalisation branch: isOOB points and points of the leaving frame go
        host_pos = self.positions(self.P["host"])
        kill = (self.P["idepth"] < 0) | ((self.res >= 0).sum(1) == 0)
        kill |= (st_newest == OOB) | ((st_newest == OUTLIER) & (st_prev == OUTLIER))
        for d in drop:
            kill |= host_pos == d
        self.keep_points(~kill)
        for d in sorted(drop, reverse=True):
            fr = self.frames.pop(d)
            self.free_slots.append(fr["slot"])
            self.res = np.delete(self.res, d, axis=1)
            if self.I is not None:
                m = self.I["host"] != fr["id"]
                self.I = {k: a[m] for k, a in self.I.items()}
                self.ctx.immature_compact(drop_host=d, drop_oob=False)
        self.keep_points((self.res >= 0).sum(1) > 0)
        # setCoarseTrackingRef on the new keyframe (:614-621, CoarseTracker.cc:347-538)
        t = self.tick("KF  point / frame removal", t)
        self.set_tracking_ref()
        t = self.tick("KF  setCoarseTrackingRef (K2)", t)
        n_new = self.make_new_traces(f)
        t = self.tick("KF  makeNewTraces (A2 + K6 ctor)", t)
        self.log[-1].update(dict(points=int(len(self.P["host"])), activated=int(n_act), immature=int(len(self.I["host"])),
                                 window=[fr["id"] for fr in self.frames], rmse=float(rmse), new_immature=int(n_new)))

    def keep_points(self, m):
        self.P = {k: a[m] for k, a in self.P.items()}
        self.res = self.res[m]

    def flag_frames_for_marginalization(self, newf):
        """FullSystem::flagFramesForMarginalization (:763-842): positions (in self.frames, before newf is added) to drop"""
        n = len(self.frames)
        if n < self.max_frames:
            return []
        latest = self.frames[-1]
        best, best_score = None, None
        for i, fr in enumerate(self.frames[:-1]):                         # the newest keyframe is never dropped (minFrameAge)
            c_i = -fr["w2c"][0].T @ fr["w2c"][1]
            dist_latest = np.linalg.norm(c_i - (-latest["w2c"][0].T @ latest["w2c"][1]))
            s = 0.0
            for j, fo in enumerate(self.frames):
                if j == i:
                    continue
                s += 1.0 / (1e-5 + np.linalg.norm(c_i - (-fo["w2c"][0].T @ fo["w2c"][1])))
            s *= -np.sqrt(dist_latest)
            if best is None or s < best_score:
                best, best_score = i, s
        return [best]

    def level1_tables(self):
        """KRKi = K[1] R Ki[0], Kt = K[1] t of every window frame -> newest keyframe (FullSystem.cc:1236-1238)"""
        K1 = self.K.copy(); K1[0, 0] *= 0.5; K1[1, 1] *= 0.5; K1[0, 2] = (K1[0, 2] + 0.5) / 2 - 0.5; K1[1, 2] = (K1[1, 2] + 0.5) / 2 - 0.5
        R, t = self.rel_to(self.frames[-1]["w2c"])
        KRKi = (K1 @ R.astype(np.float32)) @ self.Ki
        Kt = t.astype(np.float32) @ K1.T
        return np.ascontiguousarray(KRKi, np.float32), np.ascontiguousarray(Kt, np.float32)

    def activate_points(self, flagged=()):
        """FullSystem::activatePointsMT (:1197-1337): distance map, ordered candidate walk (N2) and activation (K6), all on
        the device-resident pool"""
        I = self.I
        if I is None or len(I["host"]) == 0:
            return 0
        npts = len(self.P["host"])
        r = npts / float(self.desired_points)
        for lim, d in ((0.66, -0.8), (0.8, -0.5), (0.9, -0.2), (1.0, -0.1)):
            if r < lim:
                self.min_act_dist += d
                break
        else:
            for lim, d in ((1.5, 0.8), (1.3, 0.5), (1.15, 0.2), (1.0, 0.1)):
                if r > lim:
                    self.min_act_dist += d
                    break
        self.min_act_dist = float(np.clip(self.min_act_dist, 0.0, 4.0))
        n = len(self.frames)
        KRKi, Kt = self.level1_tables()
        self.ctx.distmap_make(KRKi, Kt, self.positions(self.P["host"]), self.P["u"], self.P["v"], self.P["idepth"])
        skip = np.zeros(n, np.uint8); skip[-1] = 1
        flg = np.zeros(n, np.uint8); flg[list(flagged)] = 1
        action, _ = self.ctx.immature_select_activation(KRKi, Kt, skip, flg, self.min_act_dist)
        cand = np.nonzero(action == 2)[0].astype(np.int32)
        n_ok = 0
        if len(cand):
            self.window_frames_only()
            idp, okk, rst = self.ctx.immature_activate(cand, min_obs=1)
            good = okk == 1
            n_ok = int(good.sum())
            if n_ok:
                g = cand[good]
                newP = dict(host=I["host"][g], u=I["u"][g], v=I["v"][g], idepth=idp[good], color=I["color"][g],
                            weights=I["weights"][g], hdi=np.full(n_ok, 1e-3, np.float32))
                self.P = {k: np.concatenate([self.P[k], newP[k]]) for k in self.P}
                rs = rst[good].astype(np.int8)
                rs[rs != IN] = -1                                          # only IN residuals are created (:1130-1150)
                self.res = np.concatenate([self.res, rs], axis=0)
        gone = np.nonzero(action != 0)[0]                                  # deleted, activated or discarded (:1301-1318)
        if len(gone):
            self.ctx.immature_compact(drop_host=-1, drop_oob=False, remove=gone)
            self.I = {k: a[action == 0] for k, a in I.items()}
        return n_ok

    # ---- window BA (FullSystem::optimize, :844-1007) ---------------------------------------------------
    def frame_tables(self):
        n = len(self.frames)
        w2c = [f["w2c"] for f in self.frames]
        aff = [f["aff"] for f in self.frames]
        calib = (self.c["fx"], self.c["fy"], self.c["cx"], self.c["cy"])
        pairs, ab = wh._pairs(w2c), wh.from_to_vec_exposure_table([1.0] * n, aff)   # shared by both tables
        precalc = wh.precalc_table(w2c, w2c, calib, aff, aff, [1.0] * n, pairs=pairs, ab=ab)   # evaluation point = current state (no FEJ)
        adH, adT = wh.adjoints(w2c, aff, [1.0] * n, pairs=pairs, ab=ab)
        eth = np.array([f.get("eth", 8 * 8 * 8.0) for f in self.frames], np.float32)
        return dict(n=n, precalc=precalc, frame_energy_th=eth, calib=np.array(calib, np.float32), adH=adH, adT=adT,
                    slots=np.array([f["slot"] for f in self.frames], np.int32))

    def window_frames_only(self):
        T = self.frame_tables()
        self.ctx.window_set_frames(T, T["slots"])
        return T

    def optimize(self):
        ctx, n = self.ctx, len(self.frames)
        host_pos = self.positions(self.P["host"])
        order = np.argsort(host_pos, kind="stable")                          # allPoints order: contiguous per host frame
        self.P = {k: a[order] for k, a in self.P.items()}
        self.res = self.res[order]
        host_pos = host_pos[order]
        pidx, target = np.nonzero(self.res >= 0)
        counts = np.bincount(pidx, minlength=len(host_pos))
        res_begin = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
        nr = len(target)
        D = 8 * n + 4
        T = self.frame_tables()
        fprior = np.zeros((n, 8)); fprior[0, :] = 1e10                       # gauge: the oldest keyframe stays put
        W = dict(T, u=self.P["u"], v=self.P["v"], idepth=self.P["idepth"], idepth_zero=self.P["idepth"], host=host_pos.astype(np.int32),
                 color=self.P["color"], weights=self.P["weights"], prior=np.zeros(len(host_pos), np.float32), res_begin=res_begin,
                 target=target.astype(np.int32), state=self.res[pidx, target].astype(np.int32),
                 flags=np.where(target == n - 1, 4, 0).astype(np.uint8) | 4, res_to_zero=np.zeros((nr, 8), np.float32),
                 adHTdelta=np.zeros((n * n, 8), np.float32), c_delta=np.zeros(4, np.float32), c_prior=np.full(4, 5e9),
                 frame_prior=fprior, frame_delta_prior=np.zeros((n, 8)))
        ctx.window_load(W, T["slots"])
        HM, bM, delta = np.zeros((D, D)), np.zeros(D), np.zeros(D)
        st, _ = ctx.window_linearize(False)
        ctx.window_apply_res()
        idepth = self.P["idepth"].copy()
        lam = 1e-1
        sc = np.array([wh.SCALE_XI_TRANS] * 3 + [wh.SCALE_XI_ROT] * 3 + [wh.SCALE_A, wh.SCALE_B])
        for it in range(self.ba_its):
            acc = ctx.window_accumulate()
            x = ctx.window_solve(it, lam, acc, HM, bM, delta)
            pstep = ctx.window_resubstitute(x, lam)
            if not (np.all(np.isfinite(x)) and np.all(np.isfinite(pstep))):
                break
            for h, f in enumerate(self.frames):                             # doStepFromBackup (:1826-1932), step = -x
                s = -x[4 + 8 * h: 12 + 8 * h] * sc
                f["w2c"] = wh.se3_mul(se3_exp(s[:6]), f["w2c"])
                f["aff"] = (f["aff"][0] + s[6], f["aff"][1] + s[7])
            idepth = (idepth + pstep).astype(np.float32)
            T = self.frame_tables()
            ctx.window_set_frames(T, T["slots"])
            ctx.window_set_adjoints_only(T["adH"], T["adT"], W)
            ctx.window_set_idepth(idepth, idepth)
            st, _ = ctx.window_linearize(False)
            ctx.window_apply_res()
            lam = max(lam * 0.25, 1e-4)
            if np.abs(x[4:]).max() < 1e-4 and it >= 1:
                break
        st, th = ctx.window_linearize(True)
        ctx.window_apply_res()
        self.frames[-1]["eth"] = th
        states = ctx.window_residual_ints(0)
        self.P["idepth"] = idepth
        self.P["hdi"] = ctx.window_point_floats(6)
        self.res[pidx, target] = states.astype(np.int8)
        self._centers = (pidx, target, ctx.window_residual_floats(3), states)
        self._hdi_at_opt = self.P["hdi"].copy()
        for f in self.frames:
            self.poses[f["id"]] = f["w2c"]
        return np.sqrt(st[0] / max(8 * st[2], 1))

    def set_tracking_ref(self):
        """CoarseTracker::setCoarseTrackingRef / makeCoarseDepthL0 (CoarseTracker.cc:347-538)"""
        pidx, target, cen, states = self._centers
        newest = self.frames[-1]
        # _centers was taken before points / frames were dropped: the newest frame was the last column then
        m = (target == target.max()) & (states == IN)
        hdi_all = self._hdi_at_opt
        self.ctx.tracker_set_ref(0, newest["slot"], cen[m, 0], cen[m, 1], cen[m, 2], hdi_all[pidx[m]], 1.0, newest["aff"][0], newest["aff"][1])
        self.ref_id = newest["id"]
        self.first_rmse = -1.0
        self.last_aff = newest["aff"]


def run(api, calib_name, frames, depths, poses_c2w, calib, n_frames=None, bootstrap="gt", **kw):
    """feed a synthetic sequence; returns (odometry object, estimated camera centres, ground-truth centres, frame ids)"""
    od = Odometry(api, calib, **kw)
    n = len(frames) if n_frames is None else n_frames
    R0, t0 = poses_c2w[0]
    for i in range(n):
        w2c0 = (R0.T, -R0.T @ t0)
        od.add_frame(frames[i], gt_depth=depths[i] if (i == 0 and bootstrap == "gt") else None, gt_w2c=w2c0 if i == 0 else None)
    ids = sorted(od.poses)
    est = np.array([-od.poses[i][0].T @ od.poses[i][1] for i in ids])
    gt = np.array([poses_c2w[i][1] for i in ids])
    return od, est, gt, ids


def ate_rmse(est, gt):
    """translation RMSE after a similarity (Umeyama) alignment of the estimated camera centres to ground truth"""
    mu_e, mu_g = est.mean(0), gt.mean(0)
    E, G = est - mu_e, gt - mu_g
    U, S, Vt = np.linalg.svd(G.T @ E / len(est))
    d = np.sign(np.linalg.det(U @ Vt))
    Dm = np.diag([1, 1, d])
    R = U @ Dm @ Vt
    s = np.trace(np.diag(S) @ Dm) / max((E ** 2).sum() / len(est), 1e-30)
    aligned = (s * (R @ E.T)).T + mu_g
    return float(np.sqrt(((aligned - gt) ** 2).sum(1).mean())), float(s)
