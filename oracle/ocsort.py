"""Numpy restatement of the reference OC-SORT tracker -- TEST INFRASTRUCTURE ONLY.

Follows boxmot/trackers/ocsort/ocsort.py (KalmanBoxTracker :56-192, OcSort.update
:246-439) with the association of boxmot/utils/association.py:377-533 (+R0-R3, see
oracle/ref_shim.py).  Track ids are per tracker instance: the first track
gets internal id 0 and is emitted as id+1 (ocsort.py:430).

Parity status: pinned differentially against the reference run through
oracle/ref_shim.py on MOT17-mini det.txt and synthetic streams
(tests/golden/ocsort_*.npz, tests/test_oracle_vs_reference.py).

The per-track observation dict of the reference is collapsed to what is ever read:
a delta_t-deep ring of (age, box) plus the last observation (ocsort.py:17-28,146-152).
"""
import numpy as np

from oracle import assoc
from oracle.kalman import XYSRFilter, xyxy_to_z, x_to_xyxy

_PLACEHOLDER = np.array([-1.0, -1.0, -1.0, -1.0, -1.0])


def _direction(b1, b2):
    """ocsort.py:48-53 speed_direction -> [dy, dx] / (norm + 1e-6)."""
    cx1, cy1 = (b1[0] + b1[2]) / 2.0, (b1[1] + b1[3]) / 2.0
    cx2, cy2 = (b2[0] + b2[2]) / 2.0, (b2[1] + b2[3]) / 2.0
    v = np.array([cy2 - cy1, cx2 - cx1])
    return v / (np.sqrt((cy2 - cy1) ** 2 + (cx2 - cx1) ** 2) + 1e-6)


class _Track:
    def __init__(self, det5, cls, det_ind, tid, delta_t, q_xy, q_s):
        self.kf = XYSRFilter(xyxy_to_z(det5), q_xy, q_s)
        self.id = tid
        self.tsu = 0
        self.hits = 0
        self.hit_streak = 0
        self.age = 0
        self.conf = det5[4]
        self.cls = cls
        self.det_ind = det_ind
        self.last_obs = _PLACEHOLDER.copy()
        self.obs = {}            # age -> box5, pruned to the last delta_t ages
        self.velocity = None
        self.delta_t = delta_t

    def lookback(self):
        """k_previous_obs (ocsort.py:17-28) at the current age."""
        if not self.obs:
            return _PLACEHOLDER
        for dt in range(self.delta_t, 0, -1):
            if self.age - dt in self.obs:
                return self.obs[self.age - dt]
        return self.last_obs

    def predict(self):
        kf = self.kf
        if kf.x[6] + kf.x[2] <= 0:
            kf.x[6] *= 0.0
        kf.predict()
        self.age += 1
        if self.tsu > 0:
            self.hit_streak = 0
        self.tsu += 1
        return x_to_xyxy(kf.x)

    def update(self, det5, cls, det_ind):
        self.det_ind = det_ind
        if det5 is None:
            self.kf.update(None)
            return
        self.conf = det5[4]
        self.cls = cls
        if self.last_obs.sum() >= 0:
            prev = None
            for dt in range(self.delta_t, 0, -1):
                if self.age - dt in self.obs:
                    prev = self.obs[self.age - dt]
                    break
            if prev is None:
                prev = self.last_obs
            self.velocity = _direction(prev, det5)
        self.last_obs = det5
        self.obs[self.age] = det5
        for k in [k for k in self.obs if k < self.age - self.delta_t]:
            del self.obs[k]
        self.tsu = 0
        self.hits += 1
        self.hit_streak += 1
        self.kf.update(xyxy_to_z(det5))


class OcSortOracle:
    def __init__(self, min_conf=0.1, det_thresh=0.2, max_age=30, min_hits=3,
                 asso_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.2,
                 use_byte=False, Q_xy_scaling=0.01, Q_s_scaling=0.0001, per_class=False,
                 img_wh=(1920, 1080)):
        assert not per_class, "oracle covers per_class=False"
        self.min_conf, self.det_thresh = min_conf, det_thresh
        self.max_age, self.min_hits = max_age, min_hits
        self.thr, self.delta_t, self.inertia = asso_threshold, delta_t, inertia
        self.use_byte = use_byte
        self.q_xy, self.q_s = Q_xy_scaling, Q_s_scaling
        self.iou_fn = assoc.asso_func(asso_func, *img_wh)
        self.frame_count = 0
        self.tracks = []
        self.next_id = 0
        self.n_updates = 0       # live tracks entering predict (the bench's unit)
        self.lap_calls = 0

    def update(self, dets, img=None, embs=None):
        # basetracker.py:122-128: ndarray has a ``.data`` memoryview, so the reference
        # round-trips EVERY detection array through float32 before use.
        dets = (np.empty((0, 6)) if dets is None or len(dets) == 0
                else np.asarray(dets).astype(np.float32).astype(np.float64))
        self.frame_count += 1
        dets = np.hstack([dets, np.arange(len(dets)).reshape(-1, 1)])
        conf = dets[:, 4]
        second = dets[(conf > self.min_conf) & (conf < self.det_thresh)]
        dets = dets[conf > self.det_thresh]

        self.n_updates += len(self.tracks)
        boxes = np.zeros((len(self.tracks), 5))
        bad = []
        for t, trk in enumerate(self.tracks):
            boxes[t, :4] = trk.predict()
            if np.any(np.isnan(boxes[t])):
                bad.append(t)
        boxes = np.delete(boxes, bad, axis=0)
        for t in reversed(bad):
            self.tracks.pop(t)
        T = len(self.tracks)
        vel = np.array([trk.velocity if trk.velocity is not None else (0.0, 0.0)
                        for trk in self.tracks]).reshape(T, 2)
        last = np.array([trk.last_obs for trk in self.tracks]).reshape(T, 5)
        kobs = np.array([trk.lookback() for trk in self.tracks]).reshape(T, 5)

        matched, um_d, um_t = assoc.ocm_associate(
            dets[:, :5], boxes, self.iou_fn, self.thr, vel, kobs, self.inertia)
        for d, t in matched:
            self.tracks[t].update(dets[d, :5], dets[d, 5], dets[d, 6])

        if self.use_byte and len(second) > 0 and um_t.shape[0] > 0:
            iou_left = np.array(self.iou_fn(second, boxes[um_t]))
            if iou_left.max() > self.thr:
                taken = []
                for r, c in assoc.lap_pairs_unlimited(-iou_left):
                    if iou_left[r, c] < self.thr:
                        continue
                    t = um_t[c]
                    self.tracks[t].update(second[r, :5], second[r, 5], second[r, 6])
                    taken.append(t)
                um_t = np.setdiff1d(um_t, np.array(taken))

        if um_d.shape[0] > 0 and um_t.shape[0] > 0:
            iou_left = np.array(self.iou_fn(dets[um_d], last[um_t]))
            if iou_left.max() > self.thr:
                gone_d, gone_t = [], []
                for r, c in assoc.lap_pairs_unlimited(-iou_left):
                    if iou_left[r, c] < self.thr:
                        continue
                    d, t = um_d[r], um_t[c]
                    self.tracks[t].update(dets[d, :5], dets[d, 5], dets[d, 6])
                    gone_d.append(d)
                    gone_t.append(t)
                um_d = np.setdiff1d(um_d, np.array(gone_d))
                um_t = np.setdiff1d(um_t, np.array(gone_t))

        for t in um_t:
            self.tracks[t].update(None, None, None)
        for d in um_d:
            self.tracks.append(_Track(dets[d, :5], dets[d, 5], dets[d, 6], self.next_id,
                                      self.delta_t, self.q_xy, self.q_s))
            self.next_id += 1

        rows = []
        i = len(self.tracks)
        for trk in reversed(self.tracks):
            box = x_to_xyxy(trk.kf.x) if trk.last_obs.sum() < 0 else trk.last_obs[:4]
            if trk.tsu < 1 and (trk.hit_streak >= self.min_hits
                                or self.frame_count <= self.min_hits):
                rows.append(np.r_[box, trk.id + 1, trk.conf, trk.cls, trk.det_ind])
            i -= 1
            if trk.tsu > self.max_age:
                self.tracks.pop(i)
        return np.array(rows) if rows else np.empty((0, 8))
