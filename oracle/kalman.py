"""Numpy restatement of the reference Kalman filters -- TEST INFRASTRUCTURE ONLY.

Parity status: pinned differentially against the reference's own classes imported
through oracle/ref_shim.py (tests/test_oracle_vs_reference.py, fixtures under
tests/golden/); the reference ships no golden vectors for these functions.

Sources restated (paths under /root/reference/):
  XYSR  boxmot/motion/kalman_filters/aabb/xysr_kf.py   predict :137-175, update :211-291,
        freeze/unfreeze (OC-SORT ORU) :177-209, constants set at
        boxmot/trackers/ocsort/ocsort.py:83-113
  XYAH  boxmot/motion/kalman_filters/aabb/base_kalman_filter.py :43-53 initiate,
        :62-77 predict, :86-109 project, :111-127 multi_predict, :129-155 update,
        :166-194 gating_distance; noise models xyah_kf.py:17-79
  XYWH  same base class; noise models xywh_kf.py:17-66
"""
import numpy as np
import scipy.linalg

# --------------------------------------------------------------------------- XYSR

XYSR_F = np.eye(7)
XYSR_F[0, 4] = XYSR_F[1, 5] = XYSR_F[2, 6] = 1.0
XYSR_H = np.eye(4, 7)


def xysr_constants(q_xy=0.01, q_s=0.0001):
    """(P0, Q, R) as KalmanBoxTracker.__init__ builds them (ocsort.py:104-111)."""
    R = np.eye(4)
    R[2:, 2:] *= 10.0
    P0 = np.eye(7)
    P0[4:, 4:] *= 1000.0
    P0 *= 10.0
    Q = np.eye(7)
    Q[4:6, 4:6] *= q_xy
    Q[-1, -1] *= q_s
    return P0, Q, R


def xysr_predict(x, P, Q):
    """xysr_kf.py:164-171 with B=None, alpha=1: x <- F x ; P <- F P F' + Q."""
    x = XYSR_F @ x
    P = (XYSR_F @ P) @ XYSR_F.T + Q
    return x, P


def xysr_update(x, P, z, R):
    """xysr_kf.py:264-283: explicit inverse of S, Joseph-form covariance."""
    H = XYSR_H
    y = z - H @ x
    PHT = P @ H.T
    S = H @ PHT + R
    K = PHT @ np.linalg.inv(S)
    x = x + K @ y
    I_KH = np.eye(7) - K @ H
    P = (I_KH @ P) @ I_KH.T + (K @ R) @ K.T
    return x, P


def xyxy_to_z(b):
    """R0: [x1,y1,x2,y2] -> [xc, yc, s, r] with r = w/(h+1e-6)."""
    w = b[2] - b[0]
    h = b[3] - b[1]
    return np.array([b[0] + w / 2.0, b[1] + h / 2.0, w * h, w / (h + 1e-6)])


def x_to_xyxy(x):
    """ocsort.py:31-45 convert_x_to_bbox."""
    w = np.sqrt(x[2] * x[3])
    h = x[2] / w
    return np.array([x[0] - w / 2.0, x[1] - h / 2.0, x[0] + w / 2.0, x[1] + h / 2.0])


class XYSRFilter:
    """State of one KalmanFilterXYSR with the observation history collapsed to what
    the filter ever reads back: the last real measurement, the number of None
    updates since, and the state frozen at the first None (xysr_kf.py:177-209,
    :231-254).  x is a flat (7,) vector here."""

    def __init__(self, z0, q_xy, q_s):
        self.P, self.Q, self.R = xysr_constants(q_xy, q_s)
        self.x = np.zeros(7)
        self.x[:4] = z0
        self.observed = False
        self.has_saved = False
        self.saved_x = None
        self.saved_P = None
        self.last_z = None      # last non-None measurement (4,)
        self.n_none = 0         # None-updates since last_z

    def predict(self):
        self.x, self.P = xysr_predict(self.x, self.P, self.Q)

    def update(self, z):
        if z is None:
            if self.observed:
                # freeze: snapshot taken at the first miss (state is the prior of
                # that frame, since predict() ran before this call)
                self.saved_x = self.x.copy()
                self.saved_P = self.P.copy()
                self.has_saved = True
            self.observed = False
            self.n_none += 1
            return
        if not self.observed and self.has_saved:
            self._replay_virtual_trajectory(z)
        self.observed = True
        self.x, self.P = xysr_update(self.x, self.P, z, self.R)
        self.last_z = np.array(z, dtype=float)
        self.n_none = 0

    def _replay_virtual_trajectory(self, z):
        """unfreeze(): restore the frozen state and re-run update(+predict) along a
        straight line (in x, y, w, h) from the last measurement to the new one."""
        self.x = self.saved_x.copy()
        self.P = self.saved_P.copy()
        x1, y1, s1, r1 = self.last_z
        w1, h1 = np.sqrt(s1 * r1), np.sqrt(s1 / r1)
        x2, y2, s2, r2 = z
        w2, h2 = np.sqrt(s2 * r2), np.sqrt(s2 / r2)
        gap = self.n_none + 1
        dx, dy = (x2 - x1) / gap, (y2 - y1) / gap
        dw, dh = (w2 - w1) / gap, (h2 - h1) / gap
        for i in range(gap):
            xx, yy = x1 + (i + 1) * dx, y1 + (i + 1) * dy
            ww, hh = w1 + (i + 1) * dw, h1 + (i + 1) * dh
            zi = np.array([xx, yy, ww * hh, ww / float(hh)])
            self.x, self.P = xysr_update(self.x, self.P, zi, self.R)
            if i != gap - 1:
                self.predict()


# --------------------------------------------------------------------- XYAH / XYWH

_W_POS = 1.0 / 20
_W_VEL = 1.0 / 160
_MOTION = np.eye(8)
for _i in range(4):
    _MOTION[_i, 4 + _i] = 1.0
_UPDATE = np.eye(4, 8)

chi2inv95 = {1: 3.8415, 2: 5.9915, 3: 7.8147, 4: 9.4877, 5: 11.070,
             6: 12.592, 7: 14.067, 8: 15.507, 9: 16.919}


def _stds(kind, m, which):
    """Standard deviations of the XYAH / XYWH noise models.
    which: 'init' | 'proc' | 'meas'; m is the measurement (init) or the mean."""
    if kind == "xyah":
        h = m[3]
        if which == "init":
            return np.array([2 * _W_POS * h, 2 * _W_POS * h, 1e-2, 2 * _W_POS * h,
                             10 * _W_VEL * h, 10 * _W_VEL * h, 1e-5, 10 * _W_VEL * h])
        if which == "proc":
            return np.array([_W_POS * h, _W_POS * h, 1e-2, _W_POS * h,
                             _W_VEL * h, _W_VEL * h, 1e-5, _W_VEL * h])
        return np.array([_W_POS * h, _W_POS * h, 1e-1, _W_POS * h])
    w, h = m[2], m[3]
    if which == "init":
        return np.array([2 * _W_POS * w, 2 * _W_POS * h, 2 * _W_POS * w, 2 * _W_POS * h,
                         10 * _W_VEL * w, 10 * _W_VEL * h, 10 * _W_VEL * w, 10 * _W_VEL * h])
    if which == "proc":
        return np.array([_W_POS * w, _W_POS * h, _W_POS * w, _W_POS * h,
                         _W_VEL * w, _W_VEL * h, _W_VEL * w, _W_VEL * h])
    return np.array([_W_POS * w, _W_POS * h, _W_POS * w, _W_POS * h])


def cv8_initiate(kind, z):
    mean = np.r_[z, np.zeros(4)]
    return mean, np.diag(np.square(_stds(kind, z, "init")))


def cv8_predict(kind, mean, cov):
    q = np.diag(np.square(_stds(kind, mean, "proc")))
    mean = mean @ _MOTION.T
    cov = np.linalg.multi_dot((_MOTION, cov, _MOTION.T)) + q
    return mean, cov


def cv8_multi_predict(kind, means, covs):
    """base_kalman_filter.py:111-127: np.dot(np.dot(F, P), F^T) + Q -- NOT predict()'s multi_dot, which
    evaluates F (P F^T): the two differ in the last bit once P is no longer bit-symmetric (update's
    P - K (S K^T) is not)."""
    out_m = np.empty_like(means)
    out_c = np.empty_like(covs)
    for i in range(len(means)):
        q = np.diag(np.square(_stds(kind, means[i], "proc")))
        out_m[i] = means[i] @ _MOTION.T
        out_c[i] = np.dot(np.dot(_MOTION, covs[i]), _MOTION.T) + q
    return out_m, out_c


def cv8_project(kind, mean, cov, confidence=0.0):
    std = (1 - confidence) * _stds(kind, mean, "meas")
    return _UPDATE @ mean, np.linalg.multi_dot((_UPDATE, cov, _UPDATE.T)) + np.diag(np.square(std))


def cv8_update(kind, mean, cov, z, confidence=0.0):
    pm, pc = cv8_project(kind, mean, cov, confidence)
    cf = scipy.linalg.cho_factor(pc, lower=True, check_finite=False)
    K = scipy.linalg.cho_solve(cf, (cov @ _UPDATE.T).T, check_finite=False).T
    new_mean = mean + (z - pm) @ K.T
    new_cov = cov - np.linalg.multi_dot((K, pc, K.T))
    return new_mean, new_cov


def cv8_gating_distance(kind, mean, cov, measurements, only_position=False, metric="maha"):
    pm, pc = cv8_project(kind, mean, cov)
    if only_position:
        pm, pc = pm[:2], pc[:2, :2]
        measurements = measurements[:, :2]
    d = measurements - pm
    if metric == "gaussian":
        return np.sum(d * d, axis=1)
    L = np.linalg.cholesky(pc)
    zz = scipy.linalg.solve_triangular(L, d.T, lower=True, check_finite=False)
    return np.sum(zz * zz, axis=0)
