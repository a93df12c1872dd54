"""Numpy restatement of the reference ByteTrack tracker -- TEST INFRASTRUCTURE ONLY.

Follows boxmot/trackers/bytetrack/bytetrack.py (STrack :14-116, ByteTrack.update :158-302,
joint/sub/remove_duplicate_stracks :308-346) over KalmanFilterXYAH
(boxmot/motion/kalman_filters/aabb/xyah_kf.py, base_kalman_filter.py:43-155; restated in
oracle/kalman.py cv8_*), iou_distance / fuse_score / linear_assignment
(boxmot/utils/matching.py:184-227, :488-560, :30-108; oracle/assoc.py) and the box
conversions of boxmot/utils/ops.py:10-103.  Ids are per tracker instance.

Parity status: pinned differentially against the reference run through oracle/ref_shim.py
(tests/golden/ref_bytetrack_*.npz, tests/test_oracle.py).

Reference behaviours kept on purpose:
  * a lost track that ages out is only flagged Removed in the frame it ages out
    (bytetrack.py:273-276 runs after nothing else reads the list, and :284 subtracts the
    removed list as it was BEFORE this frame): it stays in the pool for one more frame and
    can be re-found there; from then on its id is in `removed_stracks` for good, so the next
    time it is lost it vanishes at once (:283-284);
  * unconfirmed tracks are never predicted (only the pool is, :200-202);
  * detections are matched with the box rebuilt from (xc, yc, w, h) (:105-116), not the raw box.
"""
import numpy as np

from oracle import assoc
from oracle.kalman import cv8_initiate, cv8_multi_predict, cv8_update

NEW, TRACKED, LOST, REMOVED = 0, 1, 2, 3


def _xywh_to_xyxy(v):
    """ops.py:26-43."""
    return np.array([v[0] - v[2] / 2, v[1] - v[3] / 2, v[0] + v[2] / 2, v[1] + v[3] / 2])


class _STrack:
    def __init__(self, det7):
        b = det7[0:4]
        xywh = np.array([(b[0] + b[2]) / 2, (b[1] + b[3]) / 2, b[2] - b[0], b[3] - b[1]])   # ops.py:10-23
        tl = np.array([xywh[0] - xywh[2] / 2, xywh[1] - xywh[3] / 2, xywh[2], xywh[3]])     # ops.py:46-60
        self.xywh = xywh
        self.xyah = np.array([tl[0] + tl[2] / 2, tl[1] + tl[3] / 2, tl[2] / tl[3], tl[3]])  # ops.py:90-103
        self.conf, self.cls, self.det_ind = det7[4], det7[5], det7[6]
        self.mean = self.cov = None
        self.is_activated = False
        self.state = NEW
        self.id = 0
        self.frame_id = self.start_frame = 0
        self.in_removed = False          # id present in ByteTrack.removed_stracks

    @property
    def xyxy(self):
        """bytetrack.py:105-116."""
        if self.mean is None:
            v = self.xywh.copy()
        else:
            v = self.mean[:4].copy()
            v[2] *= v[3]
        return _xywh_to_xyxy(v)

    def take(self, det):
        self.conf, self.cls, self.det_ind = det.conf, det.cls, det.det_ind


class ByteTrackOracle:
    def __init__(self, min_conf=0.1, track_thresh=0.45, match_thresh=0.8, track_buffer=25, frame_rate=30):
        self.min_conf, self.track_thresh, self.match_thresh = min_conf, track_thresh, match_thresh
        self.det_thresh = track_thresh
        self.max_time_lost = int(frame_rate / 30.0 * track_buffer)
        self.active, self.lost = [], []
        self.frame_count = 0
        self.next_id = 0
        self.n_updates = 0               # pool tracks entering predict (bench unit)

    def _kf_update(self, t, det):
        t.mean, t.cov = cv8_update("xyah", t.mean, t.cov, det.xyah)

    def _matched(self, t, det, refind):
        """STrack.update (:82-103) / re_activate (:68-80) -- identical but for the lists."""
        t.frame_id = self.frame_count
        self._kf_update(t, det)
        was_tracked = t.state == TRACKED
        t.state, t.is_activated = TRACKED, True
        t.take(det)
        if not was_tracked:
            refind.append(t)

    def update(self, dets):
        # every ndarray goes through float32 on the way in (basetracker.py:122-128)
        dets = np.asarray(dets, dtype=np.float32).astype(np.float64).reshape(-1, 6)
        dets = np.hstack([dets, np.arange(len(dets)).reshape(-1, 1)])
        self.frame_count += 1
        conf = dets[:, 4]
        high = [_STrack(d) for d in dets[conf > self.track_thresh]]
        second = [_STrack(d) for d in dets[(conf > self.min_conf) & (conf < self.track_thresh)]]
        unconfirmed = [t for t in self.active if not t.is_activated]
        tracked = [t for t in self.active if t.is_activated]
        new_tracks, refind, new_lost, removed = [], [], [], []

        # ---- first association: tracked + lost vs high-score detections (:197-219)
        pool = tracked + [t for t in self.lost if all(t is not a for a in tracked)]
        self.n_updates += len(pool)
        for t in pool:                                   # multi_predict (:41-54)
            m = t.mean.copy()
            if t.state != TRACKED:
                m[7] = 0
            mm, cc = cv8_multi_predict("xyah", m[None], t.cov[None])      # STrack.multi_predict
            t.mean, t.cov = mm[0], cc[0]
        d = assoc.iou_distance([t.xyxy for t in pool], [x.xyxy for x in high])
        d = assoc.fuse_score(d, [x.conf for x in high])
        m, u_trk, u_det = assoc.lap_thresholded(d, self.match_thresh)
        for it, idt in m:
            self._matched(pool[it], high[idt], refind)

        # ---- second association: still-tracked leftovers vs low-score detections (:221-250)
        r_tracked = [pool[i] for i in u_trk if pool[i].state == TRACKED]
        d = assoc.iou_distance([t.xyxy for t in r_tracked], [x.xyxy for x in second])
        m, u_trk2, _ = assoc.lap_thresholded(d, 0.5)
        for it, idt in m:
            self._matched(r_tracked[it], second[idt], refind)
        for it in u_trk2:
            t = r_tracked[it]
            if t.state != LOST:
                t.state = LOST
                new_lost.append(t)

        # ---- unconfirmed tracks vs the remaining high-score detections (:252-264)
        rest = [high[i] for i in u_det]
        d = assoc.iou_distance([t.xyxy for t in unconfirmed], [x.xyxy for x in rest])
        d = assoc.fuse_score(d, [x.conf for x in rest])
        m, u_unc, u_det = assoc.lap_thresholded(d, 0.7)
        for it, idt in m:
            self._matched(unconfirmed[it], rest[idt], refind)
        for it in u_unc:
            unconfirmed[it].state = REMOVED
            removed.append(unconfirmed[it])

        # ---- births (:266-272)
        for i in u_det:
            t = rest[i]
            if t.conf < self.det_thresh:
                continue
            self.next_id += 1
            t.id = self.next_id
            t.mean, t.cov = cv8_initiate("xyah", t.xyah)
            t.state = TRACKED
            t.is_activated = self.frame_count == 1
            t.frame_id = t.start_frame = self.frame_count
            new_tracks.append(t)

        # ---- list bookkeeping (:273-290)
        for t in self.lost:
            if self.frame_count - t.frame_id > self.max_time_lost:
                t.state = REMOVED
                removed.append(t)
        active = [t for t in self.active if t.state == TRACKED] + new_tracks
        active += [t for t in refind if all(t is not a for a in active)]
        lost = [t for t in self.lost if all(t is not a for a in active)] + new_lost
        lost = [t for t in lost if not t.in_removed]
        for t in removed:
            t.in_removed = True
        self.active, self.lost = self._remove_duplicates(active, lost)

        rows = [np.r_[t.xyxy, t.id, t.conf, t.cls, t.det_ind] for t in self.active if t.is_activated]
        return np.asarray(rows) if rows else np.empty((0, 8))

    @staticmethod
    def _remove_duplicates(a, b):
        """bytetrack.py:332-346."""
        pd = assoc.iou_distance([t.xyxy for t in a], [t.xyxy for t in b])
        dupa, dupb = set(), set()
        for p, q in zip(*np.where(pd < 0.15)):
            if a[p].frame_id - a[p].start_frame > b[q].frame_id - b[q].start_frame:
                dupb.add(q)
            else:
                dupa.add(p)
        return ([t for i, t in enumerate(a) if i not in dupa], [t for i, t in enumerate(b) if i not in dupb])
