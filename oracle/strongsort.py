"""Numpy restatement of StrongSORT's function-level association pieces -- TEST INFRASTRUCTURE ONLY.

Sources restated (paths under /root/reference/boxmot/trackers/strongsort/sort/):
  nn_cosine_distance   linear_assignment.py:595-618 NearestNeighborDistanceMetric.distance with
                       _nn_cosine_distance :468-497 and _cosine_distance :382-413 (INFTY_COST :11)
  min_cost_matching    linear_assignment.py:14-87 on a given cost matrix (the fork's
                       _enhance_cost_matrix :251-273 is the identity for neutral quality /
                       confidence attributes, which is how the fixture drives the reference)

  gated_metric         tracker.py:186-206 after metric.distance: gate_cost_matrix
                       (linear_assignment.py:174-248 with _compute_enhanced_motion_cost :288-307,
                       _compute_adaptive_lambda :310-329, _apply_track_specific_adjustments :332-352)
                       and Tracker._apply_id_preservation_weighting (tracker.py:283-298)
  enhance_cost_matrix  linear_assignment.py:251-273 (_compute_quality_factor :355-364,
                       _compute_confidence_factor :367-379)
  feature_update       track.py:238-268 with _calculate_adaptive_ema_alpha :319-335 (the feature EMA
                       of Track.update: similarity to the last stored feature, appearance
                       consistency, adaptive alpha, renormalised blend)
  track_predict / track_update / track_mark_missed
                       track.py:195-219, :221-301 (without the feature step), :303-317 on a flat
                       per-track table (TS_* columns)
  iou_cost             iou_matching.py:10-87 (tlwh IoU, INFTY_COST rows for time_since_update > 1)

Parity status: pinned against the reference's own functions imported through oracle/ref_shim.py
(tests/golden/ref_strongsort_units.npz, oracle/gen_golden.py --strongsort-units).  The fork's
end-to-end StrongSORT (tracker.py / track.py / occlusion_handler.py) is NOT restated: as shipped
it never confirms a track outside CI and create_tracker raises (M2).
"""
import numpy as np
from scipy.optimize import linear_sum_assignment

INFTY_COST = 1e5


def nn_cosine_distance(banks, feats):
    """banks: list of (n_t, E) arrays (n_t may be 0); feats (D, E).  Returns (T, D)."""
    feats = np.asarray(feats)
    out = np.zeros((len(banks), len(feats)))
    for i, x in enumerate(banks):
        if len(x) == 0:
            out[i] = INFTY_COST
            continue
        a = np.asarray(x) / (np.linalg.norm(x, axis=1, keepdims=True) + 1e-8)
        b = feats / (np.linalg.norm(feats, axis=1, keepdims=True) + 1e-8)
        out[i] = (1.0 - np.clip(a @ b.T, -1.0, 1.0)).min(axis=0)
    return out


def min_cost_matching(cost, max_distance):
    """Returns (matches [(row, col)], unmatched rows, unmatched cols) in the reference's order."""
    cost = np.array(cost, dtype=float)
    nr, nc = cost.shape
    if nr == 0 or nc == 0:
        return [], list(range(nr)), list(range(nc))
    cost[cost > max_distance] = max_distance + 1e-5
    rows, cols = linear_sum_assignment(cost)
    um_c = [c for c in range(nc) if c not in cols]
    um_r = [r for r in range(nr) if r not in rows]
    m = []
    for r, c in zip(rows, cols):
        if cost[r, c] > max_distance:
            um_r.append(r)
            um_c.append(c)
        else:
            m.append((r, c))
    return m, um_r, um_c


# columns of the per-track attribute table the units take (one row per track)
TRK_AGE, TRK_HITS, TRK_TSU, TRK_MOTION, TRK_APPEAR, TRK_QUALITY, TRK_HI_STREAK, TRK_LO_STREAK, TRK_HAS_VEL, \
    TRK_VX, TRK_VY, TRK_CONF, TRK_CLS, TRK_FIELDS = range(14)
CHI2INV95_4 = 9.4877


def _project(mean, cov):
    """KalmanFilterXYAH.project with confidence 0 (base_kalman_filter.py:86-109, xyah_kf.py)."""
    std = np.array([1.0 / 20 * mean[3], 1.0 / 20 * mean[3], 1e-1, 1.0 / 20 * mean[3]])
    h = np.eye(4, 8)
    return h @ mean, np.linalg.multi_dot((h, cov, h.T)) + np.diag(np.square(std))


def gating_distance(mean, cov, meas):
    """base_kalman_filter.py:166-194, metric "maha"."""
    import scipy.linalg
    m, c = _project(mean, cov)
    d = meas - m
    z = scipy.linalg.solve_triangular(np.linalg.cholesky(c), d.T, lower=True, check_finite=False,
                                      overwrite_b=True)
    return np.sum(z * z, axis=0)


def gated_metric(cost, means, covs, trk, meas, mc_lambda, id_weight):
    """cost (T, D) appearance distances; means (T, 8), covs (T, 8, 8); trk (T, TRK_FIELDS);
    meas (D, 4) detections as xyah.  Returns the (T, D) matrix matching_cascade hands to
    min_cost_matching."""
    out = np.array(cost, dtype=float)
    for r in range(len(out)):
        a = trk[r]
        gd = gating_distance(means[r], covs[r], meas)
        out[r, gd > CHI2INV95_4] = INFTY_COST
        motion = gd.copy()
        motion *= 2.0 - a[TRK_MOTION]
        if a[TRK_HAS_VEL]:
            pred = means[r, :2] + a[TRK_VX:TRK_VY + 1]
            for i in range(len(meas)):
                motion[i] *= 1.0 + min(np.linalg.norm(meas[i, :2] - pred) / 50.0, 1.0)
        lam = mc_lambda + (1 - mc_lambda) * min(a[TRK_AGE] / 10.0, 1.0) * 0.1
        lam = lam * (0.8 + 0.4 * a[TRK_MOTION])
        if a[TRK_APPEAR] < 0.5:
            lam = min(lam * 1.2, 0.99)
        lam = np.clip(lam, 0.1, 0.99)
        row = lam * out[r] + (1 - lam) * motion
        if a[TRK_QUALITY] > 0.8:
            row = row * 0.95
        if a[TRK_HITS] > 10 and a[TRK_TSU] == 0:
            row = row * 0.98
        if a[TRK_HI_STREAK] > 3:
            row = row * 0.97
        if a[TRK_LO_STREAK] > 2:
            row = row * 1.05
        if id_weight > 0:
            row = row * (1.0 - id_weight * min(a[TRK_HITS] / 10.0, 1.0))
        out[r] = row
    return out


def enhance_cost_matrix(cost, trk, det_quality, det_conf, det_cls):
    out = np.array(cost, dtype=float)
    for r in range(out.shape[0]):
        for c in range(out.shape[1]):
            q = np.clip(1.0 - ((trk[r, TRK_QUALITY] + det_quality[c]) / 2.0 - 0.5) * 0.2, 0.8, 1.2)
            out[r, c] *= q
            if trk[r, TRK_CLS] == det_cls[c]:
                out[r, c] *= 0.9
            if trk[r, TRK_CONF] > 0.7 and det_conf[c] > 0.7:
                out[r, c] *= 0.9
            elif trk[r, TRK_CONF] < 0.3 or det_conf[c] < 0.3:
                out[r, c] *= 1.1
            else:
                out[r, c] *= 1.0
    return out


def iou_cost(means, tsu, det_tlwh):
    out = np.zeros((len(means), len(det_tlwh)))
    for r, m in enumerate(means):
        if tsu[r] > 1:
            out[r] = INFTY_COST
            continue
        b = m[:4].copy()
        b[2] *= b[3]
        b[:2] -= b[2:] / 2
        tl = np.maximum(b[:2], det_tlwh[:, :2])
        br = np.minimum(b[:2] + b[2:], det_tlwh[:, :2] + det_tlwh[:, 2:])
        wh = np.maximum(0.0, br - tl)
        inter = wh.prod(axis=1)
        out[r] = 1.0 - inter / (b[2:].prod() + det_tlwh[:, 2:].prod(axis=1) - inter)
    return out


def feature_update(last, feat, conf, appearance_consistency, base_alpha=0.9):
    """One Track.update feature step.  last: the track's newest stored feature or None.
    Returns (feature to append, appearance_consistency, similarity or nan)."""
    feat = np.asarray(feat, dtype=float)
    new = feat / (np.linalg.norm(feat) + 1e-8)
    if last is None:
        return new, appearance_consistency, float("nan")
    last = np.asarray(last, dtype=float)
    sim = np.dot(new, last) / (np.linalg.norm(new) * np.linalg.norm(last) + 1e-8)
    ac = 0.9 * appearance_consistency + 0.1 * sim
    cf = 1.0 if conf > 0.7 else (0.5 if conf > 0.3 else 0.2)
    af = 1.0 if sim > 0.7 else (0.7 if sim > 0.4 else 0.4)
    alpha = float(np.clip(base_alpha * cf * af, 0.1, 0.95))
    smooth = alpha * last + (1 - alpha) * new
    smooth = smooth / (np.linalg.norm(smooth) + 1e-8)
    return smooth, ac, sim


# ---- Track state as a flat table (one row per track), restating track.py:195-317 -------------
# columns 0..12 are the TRK_* columns above (TRK_HAS_VEL holds len(velocity_history), capped at 10)
(TS_STATE, TS_STABILITY, TS_CONFIRMED, TS_MISSED, TS_NPOS, TS_PX, TS_PY, TS_NCONF, TS_CONF0) = range(13, 22)
TS_CONF_CAP = 20
TS_DETIND = TS_CONF0 + TS_CONF_CAP
TS_FIELDS = TS_DETIND + 1
TENTATIVE, CONFIRMED, DELETED = 1, 2, 3


def _motion_consistency(a, prev, cur):
    """track.py:379-399 (a[TRK_VX:VY] is velocity_history[-1])."""
    if a[TRK_HAS_VEL] >= 2:
        pv = a[TRK_VX:TRK_VY + 1]
        actual = cur - prev
        pm = (prev + pv) - prev
        if np.linalg.norm(pm) > 0:
            err = np.linalg.norm(actual - pm)
            cons = max(0, 1.0 - err / max(np.linalg.norm(pm) * 0.5, 10.0))
        else:
            cons = 1.0 if np.linalg.norm(actual) < 5.0 else 0.5
        a[TRK_MOTION] = 0.8 * a[TRK_MOTION] + 0.2 * cons


def track_predict(mean, cov, a):
    """Track.predict (track.py:195-219) on (mean, cov, row a); returns the new mean / cov, edits a."""
    from oracle import kalman as K
    mean, cov = K.cv8_predict("xyah", mean, cov)
    a[TRK_AGE] += 1
    a[TRK_TSU] += 1
    a[TRK_VX:TRK_VY + 1] = mean[4:6]
    a[TRK_HAS_VEL] = min(a[TRK_HAS_VEL] + 1, 10)
    prev = a[TS_PX:TS_PY + 1].copy()
    had = a[TS_NPOS]
    a[TS_PX:TS_PY + 1] = mean[:2]
    a[TS_NPOS] = min(a[TS_NPOS] + 1, 10)
    if had >= 1:
        _motion_consistency(a, prev, mean[:2].copy())
    return mean, cov


def track_update(mean, cov, a, xyah, conf, cls, det_ind, n_init):
    """Track.update without the feature step (track.py:221-301; run feature_update first and store
    its appearance consistency in a[TRK_APPEAR], as the reference's order of statements does)."""
    from oracle import kalman as K
    a[TRK_CONF], a[TRK_CLS], a[TS_DETIND] = conf, cls, det_ind
    mean, cov = K.cv8_update("xyah", mean, cov, xyah, conf)
    n = int(a[TS_NCONF])
    hist = list(a[TS_CONF0:TS_CONF0 + n]) + [conf]
    if len(hist) > TS_CONF_CAP:
        hist.pop(0)
    a[TS_NCONF] = len(hist)
    a[TS_CONF0:TS_CONF0 + len(hist)] = hist
    if conf > 0.7:
        a[TRK_HI_STREAK] += 1
        a[TRK_LO_STREAK] = 0
    elif conf < 0.3:
        a[TRK_LO_STREAK] += 1
        a[TRK_HI_STREAK] = 0
    else:
        a[TRK_LO_STREAK] = a[TRK_HI_STREAK] = 0
    a[TRK_HITS] += 1
    a[TS_CONFIRMED] += 1
    a[TRK_TSU] = 0
    cq = conf
    if len(hist) > 1:
        cq = 0.7 * cq + 0.3 * np.mean(hist) * (1.0 - np.std(hist))
    q = cq + min(a[TRK_HITS] / 20.0, 0.2) + max(0, (a[TRK_APPEAR] - 0.5) * 0.2) + max(0, (a[TRK_MOTION] - 0.5) * 0.1)
    a[TRK_QUALITY] = np.clip(q, 0.0, 1.0)
    cs = 1.0 - min(np.std(hist), 1.0) if len(hist) > 3 else 0.5
    hit_rate = a[TS_CONFIRMED] / max(a[TRK_AGE], 1)
    cons = 0.4 * a[TRK_APPEAR] + 0.3 * a[TRK_MOTION] + 0.3 * cs
    a[TS_STABILITY] = np.clip(0.5 * hit_rate + 0.5 * cons, 0.0, 1.0)
    if a[TS_STATE] == TENTATIVE and (a[TRK_HITS] >= n_init or (a[TRK_HITS] >= 1 and a[TRK_QUALITY] > 0.8)):
        a[TS_STATE] = CONFIRMED
    return mean, cov


def track_mark_missed(a, max_age):
    """Track.mark_missed (track.py:303-317)."""
    a[TS_MISSED] += 1
    thr = max_age
    if a[TRK_QUALITY] > 0.8:
        thr = int(max_age * 1.5)
    elif a[TRK_QUALITY] < 0.3:
        thr = int(max_age * 0.5)
    if a[TS_STATE] == TENTATIVE:
        a[TS_STATE] = DELETED
    elif a[TRK_TSU] > thr:
        a[TS_STATE] = DELETED


def track_camera_update(mean, a, warp):
    """Track.camera_update (track.py:173-189): both corners of to_tlbr() through the 2x3 warp, the
    mean's (x, y, a, h) rebuilt from them, motion consistency against the previous centre."""
    w = np.asarray(warp, dtype=float).reshape(2, 3)
    ret = mean[:4].copy()
    ret[2] *= ret[3]
    ret[:2] -= ret[2:] / 2
    ret[2:] = ret[:2] + ret[2:]
    x1, y1, x2, y2 = ret
    x1_, y1_ = w @ np.array([x1, y1, 1.0])
    x2_, y2_ = w @ np.array([x2, y2, 1.0])
    bw, bh = x2_ - x1_, y2_ - y1_
    cx, cy = x1_ + bw / 2, y1_ + bh / 2
    prev = mean[:2].copy()
    mean = mean.copy()
    mean[:4] = [cx, cy, bw / bh, bh]
    _motion_consistency(a, prev, np.array([cx, cy]))
    return mean


def matching_cascade(means, covs, trk, banks, det_xyah, det_attr, feats, track_indices, detection_indices,
                     max_distance, cascade_depth, mc_lambda=0.995, id_weight=0.1):
    """linear_assignment.py:90-171 with tracker.py:186-206's gated_metric as the distance, composed
    of the restatements above.  banks: list of (n_t, E) arrays by track index."""
    track_indices, unmatched = list(track_indices), list(detection_indices)
    levels = {}
    for t in track_indices:
        levels.setdefault(int(trk[t, TRK_TSU]), []).append(t)
    matches = []
    for age in sorted(levels):
        if age > cascade_depth:
            break
        order = sorted(levels[age], key=lambda t: -(trk[t, TRK_QUALITY] + trk[t, TS_STABILITY]))
        if not order or not unmatched:
            continue
        cost = nn_cosine_distance([banks[t] for t in order], feats[unmatched])
        cost = gated_metric(cost, means[order], covs[order], trk[order], det_xyah[unmatched], mc_lambda, id_weight)
        cost = enhance_cost_matrix(cost, trk[order], det_attr[unmatched, 0], det_attr[unmatched, 1],
                                   det_attr[unmatched, 2])
        m, _, um_c = min_cost_matching(cost, max_distance)
        matches += [(order[r], unmatched[c]) for r, c in m]
        unmatched = [unmatched[c] for c in um_c]
    matched_t = {t for t, _ in matches}
    return matches, [t for t in track_indices if t not in matched_t], unmatched
