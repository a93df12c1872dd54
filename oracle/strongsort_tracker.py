"""Numpy restatement of the fork's StrongSORT END TO END -- TEST INFRASTRUCTURE ONLY.

Restates, in the reference's only runnable configuration (R4 = unmatched_tracks_3
initialised, M1 = tracks born Confirmed as under GITHUB_ACTIONS=true, M2 = YAML keys filtered to
the constructor, R5 = handle_occlusions=False and crowd_detection=False):

  StrongSort.update / _compute_detection_quality / _format_outputs
                     boxmot/trackers/strongsort/strongsort.py:115-205, :301-327, :329-359
  Tracker.predict / update / _enhanced_match / _attempt_id_recovery / _initiate_track
                     boxmot/trackers/strongsort/sort/tracker.py:102-111, :118-181, :183-281,
                     :300-351, :395-410
  matching_cascade / min_cost_matching
                     sort/linear_assignment.py:90-171, :14-87 (scipy linear_sum_assignment itself)
  NearestNeighborDistanceMetric.partial_fit / distance
                     sort/linear_assignment.py:538-593, :595-618
  Track (predict / update / mark_missed / camera_update, feature ring of 10)
                     sort/track.py:78-317 -- through the flat-table restatements of
                     oracle/strongsort.py (TS_* columns), which are pinned on the reference's own
                     Track objects (tests/golden/ref_strongsort_units.npz)

The per-pair cost pieces come from oracle/strongsort.py.  Every np.linalg.norm / np.dot of a
feature vector that can decide something at the last ulp (partial_fit's quality ranking, the EMA
chain that produces the ranked vectors) goes through oracle.lap.ddot / norm, i.e. OpenBLAS'
SkylakeX summation order (oracle/numerics.c), so the oracle returns the same result on every host.

Parity status: pinned against the imported reference on tests/golden/ref_strongsort_*.npz
(oracle/gen_golden.py --strongsort; tests/test_oracle.py).
"""
import numpy as np
from scipy.optimize import linear_sum_assignment

from oracle import kalman as K
from oracle import lap
from oracle import strongsort as so

INFTY_COST = so.INFTY_COST
MAX_FEATURES = 10            # track.py:109
LOST_BUFFER = 30             # tracker.py:100
RECOVERY_SIM = 0.7           # tracker.py:331


class _Feature:
    """One stored feature vector: the data, its partial_fit quality (norm) and an identity."""
    __slots__ = ("v", "q")

    def __init__(self, v):
        self.v = v
        self.q = lap.norm(v)


class _Track:
    def __init__(self, det, tid, born_confirmed):
        # Track.__init__ (track.py:78-142) + Tracker._initiate_track (tracker.py:395-410)
        self.id = tid
        self.mean, self.cov = K.cv8_initiate("xyah", det["xyah"])
        a = np.zeros(so.TS_FIELDS)
        a[so.TRK_AGE] = 1
        a[so.TRK_HITS] = 1
        a[so.TRK_MOTION] = 1.0
        a[so.TRK_APPEAR] = 1.0
        a[so.TRK_QUALITY] = det["quality"]
        a[so.TRK_HI_STREAK] = 1 if det["conf"] > 0.7 else 0
        a[so.TRK_CONF] = det["conf"]
        a[so.TRK_CLS] = det["cls"]
        a[so.TS_STATE] = so.CONFIRMED if born_confirmed else so.TENTATIVE
        a[so.TS_CONFIRMED] = 1
        a[so.TS_NPOS] = 1
        a[so.TS_PX:so.TS_PY + 1] = det["xyah"][:2]
        a[so.TS_NCONF] = 1
        a[so.TS_CONF0] = det["conf"]
        a[so.TS_DETIND] = det["det_ind"]
        self.a = a
        f = det["feat"]
        self.features = [_Feature(f / (lap.norm(f) + 1e-8))]          # track.py:113-115
        self.lost_frame = 0

    state = property(lambda self: int(self.a[so.TS_STATE]))

    def tlwh(self):
        r = self.mean[:4].copy()
        r[2] *= r[3]
        r[:2] -= r[2:] / 2
        return r

    def tlbr(self):
        r = self.tlwh()
        r[2:] = r[:2] + r[2:]
        return r

    def update(self, det, n_init, ema_alpha):
        """Track.update (track.py:221-301): the feature step restated with the BLAS summation
        order, the rest through oracle.strongsort.track_update."""
        f = det["feat"]
        new = f / (lap.norm(f) + 1e-8)
        last = self.features[-1].v
        sim = lap.ddot(new, last) / (lap.norm(new) * lap.norm(last) + 1e-8)
        ac = 0.9 * self.a[so.TRK_APPEAR] + 0.1 * sim
        conf = det["conf"]
        cf = 1.0 if conf > 0.7 else (0.5 if conf > 0.3 else 0.2)
        af = 1.0 if sim > 0.7 else (0.7 if sim > 0.4 else 0.4)
        alpha = float(np.clip(ema_alpha * cf * af, 0.1, 0.95))
        smooth = alpha * last + (1 - alpha) * new
        smooth = smooth / (lap.norm(smooth) + 1e-8)
        self.features.append(_Feature(smooth))
        self.features = self.features[-MAX_FEATURES:]
        self.a[so.TRK_APPEAR] = ac
        self.mean, self.cov = so.track_update(self.mean, self.cov, self.a, det["xyah"], conf, det["cls"],
                                              det["det_ind"], n_init)


class StrongSortOracle:
    def __init__(self, min_conf=0.1, max_cos_dist=0.15, max_iou_dist=0.7, max_age=50, n_init=2, nn_budget=150,
                 mc_lambda=0.995, ema_alpha=0.9, conf_thresh_high=0.7, conf_thresh_low=0.3,
                 id_preservation_weight=0.1, born_confirmed=True, **_ignored):
        self.min_conf, self.max_cos_dist, self.max_iou_dist = min_conf, max_cos_dist, max_iou_dist
        self.max_age, self.n_init, self.budget = max_age, n_init, nn_budget
        self.mc_lambda, self.ema_alpha = mc_lambda, ema_alpha
        self.hi, self.lo, self.idw = conf_thresh_high, conf_thresh_low, id_preservation_weight
        self.born_confirmed = born_confirmed
        self.tracks, self.lost = [], []
        self.bank = {}                 # id -> list of [feature, seq] (metric.samples / sample_qualities)
        self.next_id, self.frame_count, self.seq = 1, 0, 0
        self.n_history = 0             # len(matching_history), capped at 100 (tracker.py:388-393)
        self.n_updates = 0             # tracks entering predict (the bench's unit of work)

    # ---- NearestNeighborDistanceMetric -----------------------------------------------------
    def _distance(self, feats, track_list):
        """metric.distance (linear_assignment.py:595-618) for the given tracks."""
        out = np.zeros((len(track_list), len(feats)))
        b = feats / (np.linalg.norm(feats, axis=1, keepdims=True) + 1e-8)
        for i, t in enumerate(track_list):
            ent = self.bank.get(t.id, [])
            if not ent:
                out[i] = INFTY_COST
                continue
            x = np.array([e[0].v for e in ent])
            a = x / (np.linalg.norm(x, axis=1, keepdims=True) + 1e-8)
            out[i] = (1.0 - np.clip(a @ b.T, -1.0, 1.0)).min(axis=0)
        return out

    def _partial_fit(self):
        """tracker.py:165-178 + linear_assignment.py:538-593.  Within one quality value the list
        order is insertion order (the sorts are stable), so "keep the best `budget`" drops, of the
        lowest quality value, the sample inserted last."""
        conf = [t for t in self.tracks if t.state == so.CONFIRMED]
        if not any(t.features for t in conf):
            return
        for t in conf:
            ent = self.bank.setdefault(t.id, [])
            for f in t.features:
                self.seq += 1
                ent.append([f, self.seq])
                if self.budget is not None and len(ent) > self.budget:
                    ent.sort(key=lambda e: e[0].q, reverse=True)
                    del ent[self.budget:]
        active = {t.id for t in conf}
        keep = min(self.budget // 4, 5) if self.budget else 5
        for tid, ent in self.bank.items():
            if tid not in active and len(ent) > keep:
                ent.sort(key=lambda e: e[0].q, reverse=True)
                del ent[keep:]

    # ---- matching --------------------------------------------------------------------------
    def _gated_cost(self, dets, tis, dis):
        trks = [self.tracks[i] for i in tis]
        feats = np.array([dets[i]["feat"] for i in dis])
        cost = self._distance(feats, trks)
        means = np.array([t.mean for t in trks])
        covs = np.array([t.cov for t in trks])
        tab = np.array([t.a for t in trks])
        meas = np.array([dets[i]["xyah"] for i in dis])
        cost = so.gated_metric(cost, means, covs, tab, meas, self.mc_lambda, self.idw)
        return cost, tab

    def _min_cost_matching(self, cost, tab, dets, tis, dis, max_distance):
        """linear_assignment.py:14-87 from the metric's matrix on."""
        q = np.array([dets[i]["quality"] for i in dis])
        c = np.array([dets[i]["conf"] for i in dis])
        k = np.array([dets[i]["cls"] for i in dis])
        cost = so.enhance_cost_matrix(cost, tab, q, c, k)
        cost[cost > max_distance] = max_distance + 1e-5
        rows, cols = linear_sum_assignment(cost)
        matches, um_t, um_d = [], [], []
        for col, d in enumerate(dis):
            if col not in cols:
                um_d.append(d)
        for row, t in enumerate(tis):
            if row not in rows:
                um_t.append(t)
        for r, c_ in zip(rows, cols):
            if cost[r, c_] > max_distance:
                um_t.append(tis[r])
                um_d.append(dis[c_])
            else:
                matches.append((tis[r], dis[c_]))
        return matches, um_t, um_d

    def _cascade(self, dets, tis, dis, max_distance):
        """matching_cascade (linear_assignment.py:90-171) with gated_metric (tracker.py:186-206)."""
        unmatched = list(dis)
        levels = {}
        for t in tis:
            levels.setdefault(int(self.tracks[t].a[so.TRK_TSU]), []).append(t)
        matches = []
        for age in sorted(levels):
            if age > self.max_age:
                break
            lvl = sorted(levels[age], key=lambda t: -(self.tracks[t].a[so.TRK_QUALITY] +
                                                      self.tracks[t].a[so.TS_STABILITY]))
            if not unmatched or not lvl:
                continue
            cost, tab = self._gated_cost(dets, lvl, unmatched)
            m, _, unmatched = self._min_cost_matching(cost, tab, dets, lvl, unmatched, max_distance)
            matches += m
        return matches, unmatched

    def _enhanced_match(self, dets):
        """tracker.py:183-281 (R4: unmatched_tracks_3 = [] when the IoU stage is skipped)."""
        confirmed = [i for i, t in enumerate(self.tracks) if t.state == so.CONFIRMED]
        non_tentative = [i for i, t in enumerate(self.tracks) if t.state != so.TENTATIVE]
        high = [i for i, d in enumerate(dets) if d["conf"] >= self.hi]
        med = [i for i, d in enumerate(dets) if self.lo <= d["conf"] < self.hi]
        low = {i for i, d in enumerate(dets) if d["conf"] < self.lo}
        all_m, um_t, um_d, um_t3 = [], list(confirmed), list(range(len(dets))), []
        if high and confirmed:
            m, _ = self._cascade(dets, confirmed, high, self.max_cos_dist * 0.8)
            all_m += m
            um_t = [t for t in um_t if t not in [x[0] for x in m]]
            um_d = [d for d in um_d if d not in [x[1] for x in m]]
        rem_t = [t for t in um_t if t in confirmed]
        rem_med = [d for d in med if d in um_d]
        if rem_med and rem_t:
            m, _ = self._cascade(dets, rem_t, rem_med, self.max_cos_dist)
            all_m += m
            um_t = [t for t in um_t if t not in [x[0] for x in m]]
            um_d = [d for d in um_d if d not in [x[1] for x in m]]
        cand = non_tentative + [k for k in um_t if self.tracks[k].a[so.TRK_TSU] == 1]
        rem_d = [d for d in um_d if d not in low]
        if rem_d and cand:
            trks = [self.tracks[i] for i in cand]
            cost = so.iou_cost(np.array([t.mean for t in trks]), np.array([t.a[so.TRK_TSU] for t in trks]),
                               np.array([dets[i]["tlwh"] for i in rem_d]))
            m, um_t3, _ = self._min_cost_matching(cost, np.array([t.a for t in trks]), dets, cand, rem_d,
                                                  self.max_iou_dist)
            all_m += m
            um_t = [t for t in um_t if t not in [x[0] for x in m]]
            um_d = [d for d in um_d if d not in [x[1] for x in m]]
        final_um_t = [t for t in um_t if t not in cand] + list(um_t3)
        return all_m, final_um_t, um_d

    def _attempt_id_recovery(self, um_d, dets):
        """tracker.py:300-351: at most one lost track is recovered per frame."""
        if not self.lost or not um_d:
            return []
        df = np.array([dets[i]["feat"] for i in um_d])
        for lt in self.lost:
            if not lt.features:
                continue
            tf = lt.features[-1].v.reshape(1, -1)
            sim = np.dot(df, tf.T).flatten()
            sim = sim / (np.linalg.norm(df, axis=1) * np.linalg.norm(tf))
            b = int(np.argmax(sim))
            if sim[b] > RECOVERY_SIM:
                d = um_d[b]
                lt.a[so.TS_STATE] = so.TENTATIVE            # "state = 1": TrackState.Tentative (track.py:21)
                lt.a[so.TRK_TSU] = 0
                lt.update(dets[d], self.n_init, self.ema_alpha)
                self.tracks.append(lt)
                self.lost.remove(lt)
                um_d.remove(d)
                return [(len(self.tracks) - 1, d)]
        return []

    # ---- frame -----------------------------------------------------------------------------
    def _tracker_update(self, dets, frame_id):
        matches, um_t, um_d = self._enhanced_match(dets)
        for t, d in matches:
            self.tracks[t].update(dets[d], self.n_init, self.ema_alpha)
        for t in um_t:
            so.track_mark_missed(self.tracks[t].a, self.max_age)
        rec = self._attempt_id_recovery(um_d, dets)
        for d in [i for i in um_d if i not in [m[1] for m in rec]]:
            self.tracks.append(_Track(dets[d], self.next_id, self.born_confirmed))
            self.next_id += 1
        active = []
        for t in self.tracks:
            if t.state == so.DELETED:
                now = frame_id if frame_id else self.n_history
                if len(self.lost) < LOST_BUFFER:
                    t.lost_frame = now
                    self.lost.append(t)
                self.lost = [x for x in self.lost if now - x.lost_frame < self.max_age]
            else:
                active.append(t)
        self.tracks = active
        self._partial_fit()
        self.n_history = min(self.n_history + 1, 100)

    def update(self, dets, embs, warp=None):
        """dets (N, 6) [x1, y1, x2, y2, conf, cls]; embs (N, E); warp 2x3 or None (identity).
        Returns (M, 8) rows [x1, y1, x2, y2, id, conf, cls, det_ind] (strongsort.py:329-359, first 8
        of the fork's 10 columns)."""
        dets = np.asarray(dets, dtype=float).reshape(-1, 6)
        self.frame_count += 1
        idx = np.nonzero(dets[:, 4] >= self.min_conf)[0]
        self.n_updates += len(self.tracks)
        if len(idx) == 0:
            for t in self.tracks:
                t.mean, t.cov = so.track_predict(t.mean, t.cov, t.a)
            self._tracker_update([], None)
            return self._outputs()
        if self.tracks:
            # strongsort.py:168-172: camera_update runs with whatever the CMC returns -- the identity
            # is not a no-op (mean[:4] is rebuilt from the corners, motion consistency is updated)
            w = np.eye(2, 3) if warp is None else warp
            for t in self.tracks:
                t.mean = so.track_camera_update(t.mean, t.a, w)
        dl = []
        for i in idx:
            x1, y1, x2, y2, conf, cls = dets[i]
            tlwh = np.array([x1, y1, x2 - x1, y2 - y1])                  # ops.xyxy2tlwh
            xyah = tlwh.copy()
            xyah[:2] += xyah[2:] / 2
            xyah[2] /= xyah[3]
            f = np.asarray(embs[i], dtype=float).copy()
            q = 0.7 * conf + 0.3 * min(lap.norm(f) / 10.0, 1.0)          # strongsort.py:301-327
            if tlwh[3] > 0:
                q = 0.9 * q + 0.1 * max(0.1, 1.0 - abs(tlwh[2] / tlwh[3] - 0.5) / 2.0)
            dl.append(dict(tlwh=tlwh, xyah=xyah, conf=conf, cls=cls, det_ind=float(i), feat=f, quality=q))
        dl.sort(key=lambda d: d["quality"], reverse=True)
        for t in self.tracks:
            t.mean, t.cov = so.track_predict(t.mean, t.cov, t.a)
        self._tracker_update(dl, self.frame_count)
        return self._outputs()

    def _outputs(self):
        rows = [np.r_[t.tlbr(), t.id, t.a[so.TRK_CONF], t.a[so.TRK_CLS], t.a[so.TS_DETIND]]
                for t in self.tracks if t.state == so.CONFIRMED and t.a[so.TRK_TSU] < 1]
        return np.array(rows) if rows else np.empty((0, 8))
