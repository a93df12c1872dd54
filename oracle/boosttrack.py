"""Numpy restatement of the reference BoostTrack tracker -- TEST INFRASTRUCTURE ONLY.

Follows boxmot/trackers/boosttrack/boosttrack.py (convert_bbox_to_z :19-28, convert_x_to_bbox
:31-42, KalmanBoxTracker :45-121, BoostTrack.update :221-336, filter_outputs :338-346,
get_mh_dist_matrix :356-369, duo_confidence_boost :372-412, dlo_confidence_boost :414-456),
boosttrack/kalmanfilter.py:29-157 (8-state constant-velocity filter, ConstantNoise) and
boosttrack/assoc.py (shape_similarity :9-34, MhDist_similarity :37-47, iou_batch :50-66,
soft_biou_batch :69-103, match :106-114, linear_assignment :117-153, associate :156-200).
ReID and CMC are side inputs: ``update(dets, embs=..., warp=...)``.

Parity status: pinned differentially against the reference run through oracle/ref_shim.py
(tests/golden/ref_boosttrack_*.npz, tests/test_oracle.py).

Track ids are per tracker instance here; the reference draws them from a class-level counter
that is never reset, oracle/ref_shim.py resets it per tracker.
"""
import numpy as np
import scipy.linalg

from oracle import lap

MH_LIMIT = 13.2767


def bbox_to_z(b):
    w = b[2] - b[0]
    h = b[3] - b[1]
    return np.array([b[0] + w / 2.0, b[1] + h / 2.0, h, w / float(h + 1e-6)])


def x_to_bbox(x):
    h, r = x[2], x[3]
    w = 0 if r <= 0 else r * h
    return np.array([x[0] - w / 2.0, x[1] - h / 2.0, x[0] + w / 2.0, x[1] + h / 2.0])


_F = np.eye(8)
for _i in range(4):
    _F[_i, _i + 4] = 1
_H = np.eye(4, 8)
_Q = np.eye(8)
_Q[4:, 4:] *= 0.01
_R = np.diag([1, 1, 10, 0.01])


class _KF:
    """kalmanfilter.py:29-157."""

    def __init__(self, z):
        self.x = np.zeros(8)
        self.x[:4] = z
        p = np.eye(8)
        p[4:, 4:] *= 1000.0
        p *= 10.0
        self.P = p

    def predict(self):
        self.x = _F @ self.x
        self.P = np.linalg.multi_dot((_F, self.P, _F.T)) + _Q

    def update(self, z):
        pm = _H @ self.x
        pc = np.linalg.multi_dot((_H, self.P, _H.T)) + _R
        cf = scipy.linalg.cho_factor(pc, lower=True, check_finite=False)
        K = scipy.linalg.cho_solve(cf, (self.P @ _H.T).T, check_finite=False).T
        self.x = self.x + (z - pm) @ K.T
        self.P = self.P - np.linalg.multi_dot((K, pc, K.T))


class _Track:
    def __init__(self, det7, tid, emb):
        self.tsu = 0
        self.id = tid
        self.kf = _KF(bbox_to_z(det7[:4]))
        self.conf, self.cls, self.det_ind = det7[4], det7[5], det7[6]
        self.emb = emb
        self.hit_streak = 0
        self.age = 0

    def confidence(self, coef=0.9):
        n = 7
        if self.age < n:
            return coef ** (n - self.age)
        return coef ** (self.tsu - 1)

    def camera_update(self, warp):
        wm = np.vstack([np.asarray(warp, dtype=float).reshape(2, 3), [0.0, 0.0, 1.0]])
        x1, y1, x2, y2 = x_to_bbox(self.kf.x)
        x1_, y1_, _ = wm @ np.array([x1, y1, 1.0])
        x2_, y2_, _ = wm @ np.array([x2, y2, 1.0])
        w, h = x2_ - x1_, y2_ - y1_
        self.kf.x[:4] = [x1_ + w / 2, y1_ + h / 2, h, w / h]

    def predict(self):
        self.kf.predict()
        self.age += 1
        if self.tsu > 0:
            self.hit_streak = 0
        self.tsu += 1
        return x_to_bbox(self.kf.x)

    def update(self, det7):
        self.tsu = 0
        self.hit_streak += 1
        self.kf.update(bbox_to_z(det7))
        self.conf, self.cls, self.det_ind = det7[4], det7[5], det7[6]

    def update_emb(self, emb, alpha):
        self.emb = alpha * self.emb + (1 - alpha) * emb
        self.emb = self.emb / np.linalg.norm(self.emb)


def iou_batch(a, b):
    b = np.expand_dims(b, 0)
    a = np.expand_dims(a, 1)
    w = np.maximum(0.0, np.minimum(a[..., 2], b[..., 2]) - np.maximum(a[..., 0], b[..., 0]))
    h = np.maximum(0.0, np.minimum(a[..., 3], b[..., 3]) - np.maximum(a[..., 1], b[..., 1]))
    wh = w * h
    with np.errstate(divide="ignore", invalid="ignore"):
        return wh / ((a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1])
                     + (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1]) - wh)


def soft_biou_batch(a, b):
    b = np.expand_dims(b, 0)
    a = np.expand_dims(a, 1)
    k1, k2 = 0.25, 0.5
    c = b[..., 4]
    a_x1 = a[..., 0] - (a[..., 2] - a[..., 0]) * (1 - c) * k1
    b_x1 = b[..., 0] - (b[..., 2] - b[..., 0]) * (1 - c) * k2
    a_y1 = a[..., 1] - (a[..., 3] - a[..., 1]) * (1 - c) * k1
    b_y1 = b[..., 1] - (b[..., 3] - b[..., 1]) * (1 - c) * k2
    a_x2 = a[..., 2] + (a[..., 2] - a[..., 0]) * (1 - c) * k1
    b_x2 = b[..., 2] + (b[..., 2] - b[..., 0]) * (1 - c) * k2
    a_y2 = a[..., 3] + (a[..., 3] - a[..., 1]) * (1 - c) * k1
    b_y2 = b[..., 3] + (b[..., 3] - b[..., 1]) * (1 - c) * k2
    w = np.maximum(0.0, np.minimum(a_x2, b_x2) - np.maximum(a_x1, b_x1))
    h = np.maximum(0.0, np.minimum(a_y2, b_y2) - np.maximum(a_y1, b_y1))
    wh = w * h
    return wh / ((a_x2 - a_x1) * (a_y2 - a_y1) + (b_x2 - b_x1) * (b_y2 - b_y1) - wh)


def shape_similarity(d, t, corr):
    if d.size == 0 or t.size == 0:
        return np.zeros((0, 0))
    dw = (d[:, 2] - d[:, 0]).reshape((-1, 1))
    dh = (d[:, 3] - d[:, 1]).reshape((-1, 1))
    tw = (t[:, 2] - t[:, 0]).reshape((1, -1))
    th = (t[:, 3] - t[:, 1]).reshape((1, -1))
    den2 = np.maximum(dh, th) if corr else np.maximum(dw, tw)     # v1 divides both terms by the widths
    return np.exp(-(np.abs(dw - tw) / np.maximum(dw, tw) + np.abs(dh - th) / den2))


def mh_similarity(mh, temp=1.0):
    mh = mh.copy()
    mask = mh > MH_LIMIT
    mh[mask] = MH_LIMIT
    mh = MH_LIMIT - mh
    e = np.exp(mh / temp)
    return np.where(mask, 0, e / e.sum(0).reshape((1, -1)))


class BoostTrackOracle:
    def __init__(self, max_age=60, min_hits=3, det_thresh=0.6, iou_threshold=0.3, use_ecc=True,
                 min_box_area=10, aspect_ratio_thresh=1.6, cmc_method="ecc", lambda_iou=0.5,
                 lambda_mhd=0.25, lambda_shape=0.25, use_dlo_boost=True, use_duo_boost=True,
                 dlo_boost_coef=0.65, s_sim_corr=False, use_rich_s=False, use_sb=False, use_vt=False,
                 with_reid=False, per_class=False, **_kw):
        assert not per_class, "oracle covers per_class=False"
        self.max_age, self.min_hits, self.det_thresh, self.thr = max_age, min_hits, det_thresh, iou_threshold
        self.use_ecc, self.min_box_area, self.ar_thresh = use_ecc, min_box_area, aspect_ratio_thresh
        self.l_iou, self.l_mhd, self.l_shape = lambda_iou, lambda_mhd, lambda_shape
        self.dlo, self.duo, self.dlo_coef, self.s_corr = use_dlo_boost, use_duo_boost, dlo_boost_coef, s_sim_corr
        self.rich_s, self.sb, self.vt, self.with_reid = use_rich_s, use_sb, use_vt, with_reid
        self.frame_count = 0
        self.tracks = []
        self.next_id = 0
        self.n_updates = 0
        self.counters = {}

    # -- helpers over the current track list -------------------------------------------
    def _mh(self, dets):
        if not self.tracks:
            return np.zeros((0, 0))
        z = np.array([bbox_to_z(d) for d in dets]).reshape(len(dets), 4)
        x = np.array([t.kf.x[:4] for t in self.tracks])
        si = np.array([np.reciprocal(np.diag(t.kf.P[:4, :4])) for t in self.tracks])
        return ((z.reshape((-1, 1, 4)) - x.reshape((1, -1, 4))) ** 2 * si.reshape((1, -1, 4))).sum(axis=2)

    def _trk_boxes(self):
        return np.array([np.r_[x_to_bbox(t.kf.x), t.confidence()] for t in self.tracks]).reshape(-1, 5)

    def _dlo(self, dets):
        if len(dets) == 0 or not self.tracks:
            return dets
        trk = self._trk_boxes()
        sb = soft_biou_batch(dets, trk)
        if self.rich_s:
            S = (mh_similarity(self._mh(dets), 1) + shape_similarity(dets, trk, self.s_corr) + sb) / 3
        else:
            S = iou_batch(dets, trk)
        if not self.sb and not self.vt:
            dets[:, 4] = np.maximum(dets[:, 4], S.max(1) * self.dlo_coef)
            return dets
        if self.sb:
            dets[:, 4] = np.maximum(dets[:, 4], 0.65 * dets[:, 4] + (1 - 0.65) * S.max(1) ** 1.5)
        if self.vt:
            lim = np.maximum(0.95 - np.array([t.tsu - 1 for t in self.tracks]), 0.8)
            hit = (S > lim).max(1)
            sc = dets[:, 4].copy()
            sc[hit] = np.maximum(sc[hit], self.det_thresh + 1e-5)
            dets[:, 4] = sc
        return dets

    def _duo(self, dets):
        if len(dets) == 0:
            return dets
        mh = self._mh(dets)
        if mh.size == 0:
            return dets
        cand = np.where((mh.min(1) > MH_LIMIT) & (dets[:, 4] < self.det_thresh))[0]
        if len(cand) == 0:
            return dets
        bd = iou_batch(dets[cand], dets[cand]) - np.eye(len(cand))
        keep = []
        for i, ci in enumerate(cand):
            rivals = np.where(bd[i] > 0.3)[0]
            if len(rivals) == 0 or dets[ci, 4] == max(dets[cand[rivals], 4].max(), dets[ci, 4]):
                keep.append(ci)
        dets[keep, 4] = self.det_thresh + 1e-4
        return dets

    def update(self, dets, img=None, embs=None, warp=None):
        dets = (np.empty((0, 6)) if dets is None or len(dets) == 0
                else np.asarray(dets).astype(np.float32).astype(np.float64))
        dets = np.hstack([dets, np.arange(len(dets)).reshape(-1, 1)])
        self.frame_count += 1
        if self.use_ecc and warp is not None:
            for t in self.tracks:
                t.camera_update(warp)
        self.n_updates += len(self.tracks)
        trks, confs = [], []
        for t in self.tracks:
            pos = t.predict()
            confs.append(t.confidence())
            trks.append(np.r_[pos, confs[-1]])
        trks = np.vstack(trks) if trks else np.empty((0, 5))
        if self.dlo:
            dets = self._dlo(dets)
        if self.duo:
            dets = self._duo(dets)
        keep = dets[:, 4] >= self.det_thresh
        dets = dets[keep]
        scores = dets[:, 4]
        dets_embs = np.ones((len(dets), 1))
        if self.with_reid and embs is not None and len(keep):
            dets_embs = np.asarray(embs)[keep]
        emb_cost = None
        if self.with_reid and self.tracks:
            te = np.array([t.emb for t in self.tracks])
            emb_cost = (np.empty((0, len(te))) if len(dets) == 0
                        else dets_embs.reshape(len(dets), -1) @ te.reshape(len(te), -1).T)

        matched, um_d = [], list(range(len(dets)))
        if self.tracks:
            iou = iou_batch(dets, trks)
            cost = iou.copy()
            conf = scores.reshape((-1, 1)) * np.array(confs).reshape((1, -1))
            conf[iou < self.thr] = 0
            cost += self.l_iou * conf * iou
            mh = self._mh(dets)
            if mh.size > 0:
                cost += self.l_mhd * mh_similarity(mh)
                cost += self.l_shape * conf * shape_similarity(dets, trks, self.s_corr)
            if emb_cost is not None:
                cost += (1 + self.l_iou + self.l_shape + self.l_mhd) * 1.5 * emb_cost
            if cost.size == 0:
                pairs = np.empty((0, 2), dtype=int)
            else:
                a = (cost > self.thr).astype(np.int32)
                if a.sum(1).max() == 1 and a.sum(0).max() == 1:
                    pairs = np.stack(np.where(a), axis=1)
                else:
                    _, x, _ = lap.lapjv(-cost, extend_cost=True)
                    pairs = np.array([[r, c] for r, c in enumerate(x) if c >= 0], dtype=int).reshape(-1, 2)
                    self.counters["lap"] = self.counters.get("lap", 0) + 1
            um_d = [d for d in range(len(dets)) if d not in pairs[:, 0]]
            for d, t in pairs:
                ok = iou[d, t] >= self.thr or (emb_cost is not None and iou[d, t] >= self.thr / 2
                                               and emb_cost[d, t] >= 0.75)
                if ok:
                    matched.append((d, t))
                else:
                    um_d.append(d)
        if len(dets):
            trust = (dets[:, 4] - self.det_thresh) / (1 - self.det_thresh)
            alpha = 0.95 + (1 - 0.95) * (1 - trust)
        for d, t in matched:
            self.tracks[t].update(dets[d])
            self.tracks[t].update_emb(dets_embs[d], alpha[d])
        for d in um_d:
            if dets[d, 4] >= self.det_thresh:
                self.next_id += 1
                self.tracks.append(_Track(dets[d], self.next_id, dets_embs[d]))
        rows = []
        for t in self.tracks:
            if t.tsu < 1 and (t.hit_streak >= self.min_hits or self.frame_count <= self.min_hits):
                rows.append(np.r_[x_to_bbox(t.kf.x), t.id, t.conf, t.cls, t.det_ind])
        self.tracks = [t for t in self.tracks if t.tsu <= self.max_age]
        if not rows:
            return np.empty((0, 8))
        rows = np.vstack(rows)
        w, h = rows[:, 2] - rows[:, 0], rows[:, 3] - rows[:, 1]
        return rows[(w / h <= self.ar_thresh) & (w * h > self.min_box_area)]
