"""Numpy restatement of StrongSORT's function-level association pieces -- TEST INFRASTRUCTURE ONLY.

Sources restated (paths under /root/reference/boxmot/trackers/strongsort/sort/):
  nn_cosine_distance   linear_assignment.py:595-618 NearestNeighborDistanceMetric.distance with
                       _nn_cosine_distance :468-497 and _cosine_distance :382-413 (INFTY_COST :11)
  min_cost_matching    linear_assignment.py:14-87 on a given cost matrix (the fork's
                       _enhance_cost_matrix :251-273 is the identity for neutral quality /
                       confidence attributes, which is how the fixture drives the reference)

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
