"""Numpy restatement of the reference BoT-SORT tracker -- TEST INFRASTRUCTURE ONLY.

Follows boxmot/trackers/botsort/botsort.py (BotSort.update :94-166 and its stages :168-411),
botsort_track.py (STrack :10-159: feature EMA :40-49, class voting :51-64, multi_predict
:76-90, multi_gmc :92-104), botsort_utils.py (joint/sub/remove_duplicate_stracks :10-81) over
KalmanFilterXYWH (oracle/kalman.py cv8_*), iou_distance / fuse_score / embedding_distance /
linear_assignment (boxmot/utils/matching.py:184-227, :488-560, :230-321, :30-108).
ReID and CMC are side inputs: embeddings come in through ``update(dets, embs)`` and the warp
is a 2x3 matrix (identity by default), exactly how oracle/ref_shim.py drives the reference.

Parity status: pinned differentially against the reference run through oracle/ref_shim.py
(tests/golden/ref_botsort_*.npz, tests/test_oracle.py).

Feature arithmetic is float32 as in the reference (features arrive as float32 and
``0.9 * f32_array`` stays float32 under NumPy 2 promotion); a new detection's feature ends up
normalised twice because ``smooth_feat`` aliases ``curr_feat`` in STrack.__init__
(botsort_track.py:42-48).
"""
import numpy as np
from scipy.spatial.distance import cdist

from oracle import assoc
from oracle.bytetrack import LOST, NEW, REMOVED, TRACKED, _xywh_to_xyxy
from oracle.kalman import cv8_initiate, cv8_multi_predict, cv8_update


class _STrack:
    def __init__(self, det7, feat=None):
        b = det7[0:4]
        self.xywh = np.array([(b[0] + b[2]) / 2, (b[1] + b[3]) / 2, b[2] - b[0], b[3] - b[1]])   # ops.py:10-23
        self.conf, self.cls, self.det_ind = det7[4], det7[5], det7[6]
        self.mean = self.cov = None
        self.is_activated = False
        self.state = NEW
        self.id = 0
        self.frame_id = self.start_frame = 0
        self.in_removed = False
        self.cls_hist = []
        self.smooth_feat = self.curr_feat = None
        self.vote(self.cls, self.conf)
        if feat is not None:
            self.update_features(feat)

    def update_features(self, feat):
        """botsort_track.py:40-49 (in-place operations kept: they alias on the first call)."""
        feat /= np.linalg.norm(feat)
        self.curr_feat = feat
        if self.smooth_feat is None:
            self.smooth_feat = feat
        else:
            self.smooth_feat = 0.9 * self.smooth_feat + (1 - 0.9) * feat
        self.smooth_feat /= np.linalg.norm(self.smooth_feat)

    def vote(self, cls, conf):
        """update_cls (botsort_track.py:51-64)."""
        best, found = 0, False
        for c in self.cls_hist:
            if cls == c[0]:
                c[1] += conf
                found = True
            if c[1] > best:
                best = c[1]
                self.cls = c[0]
        if not found:
            self.cls_hist.append([cls, conf])
            self.cls = cls

    @property
    def xyxy(self):
        return _xywh_to_xyxy(self.mean[:4] if self.mean is not None else self.xywh)


def embedding_distance(tracks, dets):
    """enhanced_embedding_distance with its defaults (matching.py:230-321)."""
    if not len(tracks) or not len(dets):
        return np.zeros((len(tracks), len(dets)), dtype=np.float32)
    d = np.asarray([x.curr_feat for x in dets], dtype=np.float32)
    t = np.asarray([x.smooth_feat for x in tracks], dtype=np.float32)
    d = d / (np.linalg.norm(d, axis=1, keepdims=True) + 1e-8)
    t = t / (np.linalg.norm(t, axis=1, keepdims=True) + 1e-8)
    return np.maximum(0.0, cdist(t, d, "cosine"))


class BotSortOracle:
    def __init__(self, track_high_thresh=0.5, track_low_thresh=0.1, new_track_thresh=0.6, track_buffer=30,
                 match_thresh=0.8, proximity_thresh=0.5, appearance_thresh=0.25, frame_rate=30,
                 fuse_first_associate=False, with_reid=True):
        self.high, self.low, self.new_thresh = track_high_thresh, track_low_thresh, new_track_thresh
        self.match_thresh, self.prox, self.app = match_thresh, proximity_thresh, appearance_thresh
        self.fuse_first, self.with_reid = fuse_first_associate, with_reid
        self.max_time_lost = int(frame_rate / 30.0 * track_buffer)
        self.active, self.lost = [], []
        self.frame_count = 0
        self.next_id = 0
        self.n_updates = 0

    def _matched(self, t, det, refind):
        """STrack.update (:131-149) / re_activate (:118-129)."""
        t.frame_id = self.frame_count
        t.mean, t.cov = cv8_update("xywh", t.mean, t.cov, det.xywh)
        if det.curr_feat is not None:
            t.update_features(det.curr_feat)
        was_tracked = t.state == TRACKED
        t.state, t.is_activated = TRACKED, True
        t.conf, t.cls, t.det_ind = det.conf, det.cls, det.det_ind
        t.vote(det.cls, det.conf)
        if not was_tracked:
            refind.append(t)

    def _fused(self, tracks, dets, fuse):
        iou_d = assoc.iou_distance([t.xyxy for t in tracks], [x.xyxy for x in dets])
        far = iou_d > self.prox
        if fuse:
            iou_d = assoc.fuse_score(iou_d, [x.conf for x in dets])
        if not self.with_reid:
            return iou_d
        emb = embedding_distance(tracks, dets) / 2.0
        emb[emb > self.app] = 1.0
        emb[far] = 1.0
        return np.minimum(iou_d, emb)

    @staticmethod
    def _gmc(tracks, warp):
        """multi_gmc (botsort_track.py:92-104)."""
        R8 = np.kron(np.eye(4), warp[:2, :2])
        for t in tracks:
            m = R8.dot(t.mean)
            m[:2] += warp[:2, 2]
            t.mean, t.cov = m, R8.dot(t.cov).dot(R8.T)

    def update(self, dets, embs=None, warp=None):
        dets = np.asarray(dets, dtype=np.float32).astype(np.float64).reshape(-1, 6)
        dets = np.hstack([dets, np.arange(len(dets)).reshape(-1, 1)])
        self.frame_count += 1
        conf = dets[:, 4]
        first = conf > self.high
        feats = [None] * int(first.sum())
        if self.with_reid:
            feats = list(np.array(embs, dtype=np.float32)[first])
        high = [_STrack(d, f) for d, f in zip(dets[first], feats)]
        second = [_STrack(d) for d in dets[(conf > self.low) & (conf < self.high)]]
        unconfirmed = [t for t in self.active if not t.is_activated]
        tracked = [t for t in self.active if t.is_activated]
        new_tracks, refind, new_lost, removed = [], [], [], []

        pool = tracked + [t for t in self.lost if all(t is not a for a in tracked)]
        self.n_updates += len(pool)
        for t in pool:                                   # multi_predict (:76-90)
            m = t.mean.copy()
            if t.state != TRACKED:
                m[6:8] = 0
            mm, cc = cv8_multi_predict("xywh", m[None], t.cov[None])      # STrack.multi_predict
            t.mean, t.cov = mm[0], cc[0]
        warp = np.eye(2, 3) if warp is None else np.asarray(warp, dtype=float)
        self._gmc(pool, warp)
        self._gmc(unconfirmed, warp)

        # ---- first association (:203-246)
        m, u_trk, u_det = assoc.lap_thresholded(self._fused(pool, high, self.fuse_first), self.match_thresh)
        for it, idt in m:
            self._matched(pool[it], high[idt], refind)
        # ---- second association (:248-291)
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
        # ---- unconfirmed (:293-344)
        rest = [high[i] for i in u_det]
        m, u_unc, u_det = assoc.lap_thresholded(self._fused(unconfirmed, rest, True), 0.7)
        for it, idt in m:
            self._matched(unconfirmed[it], rest[idt], refind)
        for it in u_unc:
            unconfirmed[it].state = REMOVED
            removed.append(unconfirmed[it])
        # ---- births (:346-353)
        for i in u_det:
            t = rest[i]
            if t.conf < self.new_thresh:
                continue
            self.next_id += 1
            t.id = self.next_id
            t.mean, t.cov = cv8_initiate("xywh", t.xywh)
            t.state = TRACKED
            t.is_activated = self.frame_count == 1
            t.frame_id = t.start_frame = self.frame_count
            new_tracks.append(t)
        # ---- bookkeeping (:385-411), identical to ByteTrack's
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
        pd = assoc.iou_distance([t.xyxy for t in a], [t.xyxy for t in b])
        dupa, dupb = set(), set()
        for p, q in zip(*np.where(pd < 0.15)):
            if a[p].frame_id - a[p].start_frame > b[q].frame_id - b[q].start_frame:
                dupb.add(q)
            else:
                dupa.add(p)
        return ([t for i, t in enumerate(a) if i not in dupa], [t for i, t in enumerate(b) if i not in dupb])
