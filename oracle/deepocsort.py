"""Numpy restatement of the reference DeepOCSORT tracker -- TEST INFRASTRUCTURE ONLY.

Follows boxmot/trackers/deepocsort/deepocsort.py (KalmanBoxTracker :51-235, DeepOcSort.update
:316-498), the association of boxmot/utils/association.py:377-533 with the embedding term
(`emb_cost[iou <= 0] = 0`, `compute_aw_max_metric` :320-374) and the CMC correction
`KalmanFilterXYSR.apply_affine_correction` (xysr_kf.py:111-135).  ReID and CMC are side inputs:
embeddings come in through ``update(dets, embs=...)`` and the warp as ``warp=`` (2x3).

Parity status: pinned differentially against the reference run through oracle/ref_shim.py
(tests/golden/ref_deepocsort_*.npz, tests/test_oracle.py).

Reference behaviours kept on purpose:
  * ids start at 1 and are emitted as they are (deepocsort.py:301,489) -- the same numbers as
    OC-SORT's ``id + 1``;
  * `last_observation`, `observations[age]` and the detection row are ONE numpy array in the
    reference (deepocsort.py:166-168 store views of the same row), so a CMC warp is applied
    to the newest observation once through `last_observation` (if its sum is > 0) and once more
    through the `observations` window (:198-207) -- i.e. TWICE when both conditions hold;
  * the track embedding starts as the raw detection embedding (not normalised) and is
    normalised by every `update_emb` (:184-186); the OCR round ignores embeddings (:381-392).
"""
import numpy as np

from oracle import assoc
from oracle.kalman import x_to_xyxy
from oracle.ocsort import _Track as _OcTrack


def aw_max_metric(emb, w_emb, bottom):
    """compute_aw_max_metric (association.py:320-374)."""
    w = np.full_like(emb, w_emb)
    for axis in (0, 1):
        lines = emb if axis == 0 else emb.T
        for i, line in enumerate(lines):
            pos = line[line > 0]
            if pos.size < 2:
                continue
            top = np.sort(pos)[::-1]
            weight = 1 - max(top[1] / top[0] - bottom, 0) / (1 - bottom)
            if axis == 0:
                w[i] *= weight
            else:
                w[:, i] *= weight
    return w * emb


def deep_associate(dets, trks, iou_fn, iou_thr, velocities, prev_obs, inertia, emb_cost, w_emb,
                   aw_off, aw_param, counters=None):
    """enhanced_associate (association.py:377-533 + R1-R3) with the embedding term."""
    D, T = len(dets), len(trks)
    if T == 0:
        return np.empty((0, 2), dtype=int), np.arange(D), np.empty((0,), dtype=int)
    prev = prev_obs
    cxd, cyd = (dets[:, 0] + dets[:, 2]) / 2.0, (dets[:, 1] + dets[:, 3]) / 2.0
    cxp, cyp = (prev[:, 0] + prev[:, 2]) / 2.0, (prev[:, 1] + prev[:, 3]) / 2.0
    dx = cxd[None, :] - cxp[:, None]
    dy = cyd[None, :] - cyp[:, None]
    norm = np.sqrt(dx ** 2 + dy ** 2) + 1e-6
    dx, dy = dx / norm, dy / norm
    cosv = np.clip(velocities[:, 1][:, None] * dx + velocities[:, 0][:, None] * dy, -1, 1)
    ang = (np.pi / 2.0 - np.abs(np.arccos(cosv))) / np.pi
    valid = np.where(prev[:, 4] < 0, 0.0, 1.0)[:, None]
    motion = ((valid * ang) * inertia).T
    iou = iou_fn(dets, trks)
    total = iou + motion
    if emb_cost is not None:
        emb = emb_cost.copy()
        emb[iou <= 0] = 0
        emb = aw_max_metric(emb, w_emb, aw_param) if not aw_off else emb * w_emb
        total = total + emb
    if min(D, T) > 0:
        a = (iou > iou_thr).astype(np.int32)
        if a.sum(1).max() == 1 and a.sum(0).max() == 1:
            pairs = np.stack(np.where(a), axis=1)
        else:
            pairs = assoc.lap_pairs_unlimited(-total)
            if counters is not None:
                counters["lap"] = counters.get("lap", 0) + 1
    else:
        pairs = np.empty((0, 2), dtype=int)
    um_d = [d for d in range(D) if d not in pairs[:, 0]]
    um_t = [t for t in range(T) if t not in pairs[:, 1]]
    keep = []
    for d, t in pairs:
        if iou[d, t] < iou_thr:
            um_d.append(d)
            um_t.append(t)
        else:
            keep.append((d, t))
    m = np.asarray(keep, dtype=int) if keep else np.empty((0, 2), dtype=int)
    return m, np.asarray(um_d, dtype=int), np.asarray(um_t, dtype=int)


def _warp_box(box, m, t):
    """deepocsort.py:195-197: both corners through x -> m x + t."""
    ps = m @ box[:4].reshape(2, 2).T + t
    box[:4] = ps.T.reshape(-1)


class _Track(_OcTrack):
    def __init__(self, det5, cls, det_ind, tid, delta_t, q_xy, q_s, emb):
        super().__init__(det5, cls, det_ind, tid, delta_t, q_xy, q_s)
        self.emb = emb

    def update(self, det5, cls, det_ind):
        if det5 is None:
            # deepocsort.py:179-181: update(None) does not touch det_ind
            self.kf.update(None)
            return
        super().update(det5, cls, det_ind)
        # the reference stores ONE array as last_observation and observations[age]
        self.obs[self.age] = self.last_obs

    def update_emb(self, emb, alpha):
        self.emb = alpha * self.emb + (1 - alpha) * emb
        self.emb = self.emb / np.linalg.norm(self.emb)

    def affine(self, m, t):
        """deepocsort.py:191-208 + xysr_kf.py:111-135."""
        if self.last_obs.sum() > 0:
            _warp_box(self.last_obs, m, t)
        for dt in range(self.delta_t, -1, -1):
            if self.age - dt in self.obs:
                _warp_box(self.obs[self.age - dt], m, t)
        # xysr_kf.py:111-135.  The frozen copy's `last_measurement` is warped there too, but
        # unfreeze() interpolates from the LIVE observation history (:184-192), which is not:
        # the ORU anchor stays unwarped.
        kf = self.kf
        states = [(kf.x, kf.P)]
        if not kf.observed and kf.has_saved:
            states.append((kf.saved_x, kf.saved_P))
        for x, P in states:
            x[:2] = (m @ x[:2].reshape(2, 1) + t).ravel()
            x[4:6] = m @ x[4:6]
            P[:2, :2] = m @ P[:2, :2] @ m.T
            P[4:6, 4:6] = m @ P[4:6, 4:6] @ m.T


class DeepOcSortOracle:
    def __init__(self, det_thresh=0.3, max_age=30, min_hits=3, iou_threshold=0.3, delta_t=3,
                 asso_func="iou", inertia=0.2, w_association_emb=0.5, alpha_fixed_emb=0.95,
                 aw_param=0.5, embedding_off=False, cmc_off=False, aw_off=False,
                 Q_xy_scaling=0.01, Q_s_scaling=0.0001, per_class=False, img_wh=(1920, 1080), **_kw):
        assert not per_class, "oracle covers per_class=False"
        self.det_thresh, self.max_age, self.min_hits = det_thresh, max_age, min_hits
        self.thr, self.delta_t, self.inertia = iou_threshold, delta_t, inertia
        self.w_emb, self.alpha_fixed, self.aw_param = w_association_emb, alpha_fixed_emb, aw_param
        self.embedding_off, self.cmc_off, self.aw_off = embedding_off, cmc_off, aw_off
        self.q_xy, self.q_s = Q_xy_scaling, Q_s_scaling
        self.iou_fn = assoc.asso_func(asso_func, *img_wh)
        self.frame_count = 0
        self.tracks = []
        self.next_id = 1
        self.n_updates = 0
        self.counters = {}

    def update(self, dets, img=None, embs=None, warp=None):
        dets = (np.empty((0, 6)) if dets is None or len(dets) == 0
                else np.asarray(dets).astype(np.float32).astype(np.float64))
        self.frame_count += 1
        dets = np.hstack([dets, np.arange(len(dets)).reshape(-1, 1)])
        keep = dets[:, 4] > self.det_thresh
        dets = dets[keep]
        if self.embedding_off or len(dets) == 0:
            dets_embs = np.ones((len(dets), 1))
        else:
            dets_embs = np.asarray(embs)[keep]
        if not self.cmc_off and warp is not None:
            w = np.asarray(warp, dtype=float).reshape(2, 3)
            m, t = w[:, :2], w[:, 2].reshape(2, 1)
            for trk in self.tracks:
                trk.affine(m, t)
        trust = (dets[:, 4] - self.det_thresh) / (1 - self.det_thresh)
        af = self.alpha_fixed
        alpha = af + (1 - af) * (1 - trust)

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
        trk_embs = np.vstack([trk.emb for trk in self.tracks]) if T else np.empty((0,))
        vel = np.array([trk.velocity if trk.velocity is not None else (0.0, 0.0)
                        for trk in self.tracks]).reshape(T, 2)
        last = np.array([trk.last_obs for trk in self.tracks]).reshape(T, 5)
        kobs = np.array([trk.lookback() for trk in self.tracks]).reshape(T, 5)

        if self.embedding_off or len(dets) == 0 or T == 0:
            emb_cost = None
        else:
            emb_cost = dets_embs @ trk_embs.T
        matched, um_d, um_t = deep_associate(
            dets[:, :5], boxes, self.iou_fn, self.thr, vel, kobs, self.inertia, emb_cost,
            self.w_emb, self.aw_off, self.aw_param, self.counters)
        for d, t in matched:
            self.tracks[t].update(dets[d, :5].copy(), dets[d, 5], dets[d, 6])
            self.tracks[t].update_emb(dets_embs[d], alpha[d])

        if um_d.shape[0] > 0 and um_t.shape[0] > 0:
            iou_left = np.array(self.iou_fn(dets[um_d], last[um_t]))
            if iou_left.max() > self.thr:
                gone_d, gone_t = [], []
                for r, c in assoc.lap_pairs_unlimited(-iou_left):
                    if iou_left[r, c] < self.thr:
                        continue
                    d, t = um_d[r], um_t[c]
                    self.tracks[t].update(dets[d, :5].copy(), dets[d, 5], dets[d, 6])
                    self.tracks[t].update_emb(dets_embs[d], alpha[d])
                    gone_d.append(d)
                    gone_t.append(t)
                um_d = np.setdiff1d(um_d, np.array(gone_d))
                um_t = np.setdiff1d(um_t, np.array(gone_t))

        for t in um_t:
            self.tracks[t].update(None, None, None)
        for d in um_d:
            self.tracks.append(_Track(dets[d, :5].copy(), dets[d, 5], dets[d, 6], self.next_id,
                                      self.delta_t, self.q_xy, self.q_s, dets_embs[d]))
            self.next_id += 1

        rows = []
        i = len(self.tracks)
        for trk in reversed(self.tracks):
            box = x_to_xyxy(trk.kf.x) if trk.last_obs.sum() < 0 else trk.last_obs[:4]
            if trk.tsu < 1 and (trk.hit_streak >= self.min_hits
                                or self.frame_count <= self.min_hits):
                rows.append(np.r_[box, trk.id, trk.conf, trk.cls, trk.det_ind])
            i -= 1
            if trk.tsu > self.max_age:
                self.tracks.pop(i)
        return np.array(rows) if rows else np.empty((0, 8))
