"""Numpy restatement of the reference cost / assignment helpers -- TEST INFRASTRUCTURE ONLY.

Parity status: pinned differentially against the reference functions imported through
oracle/ref_shim.py (tests/test_oracle_vs_reference.py); the reference has no golden
vectors for them.  ``lapjv`` tie-breaking: parity unpinned (oracle/lapjv.c header).

Sources restated (paths under /root/reference/boxmot/):
  iou family           utils/iou.py:50-67 (iou), :79-127 (hmiou), :129-169 (giou),
                       :171-184 (centroid), :199-264 (ciou), :266-307 (diou)
  OCM association      utils/association.py:10-20, :377-533 with repairs R1-R3
  legacy LAP wrapper   utils/association.py:105-114
  thresholded LAP      utils/matching.py:30-108 (method 'lap')
  iou_distance         utils/matching.py:184-227
  fuse_score           utils/matching.py:488-560 (fork semantics)
"""
import numpy as np

from oracle import lap


def iou_batch(a, b):
    a = np.asarray(a, dtype=float)[:, None, :]
    b = np.asarray(b, dtype=float)[None, :, :]
    iw = np.maximum(0.0, np.minimum(a[..., 2], b[..., 2]) - np.maximum(a[..., 0], b[..., 0]))
    ih = np.maximum(0.0, np.minimum(a[..., 3], b[..., 3]) - np.maximum(a[..., 1], b[..., 1]))
    inter = iw * ih
    area_a = (a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1])
    area_b = (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1])
    with np.errstate(divide="ignore", invalid="ignore"):
        return inter / (area_a + area_b - inter)


def hmiou_batch(a, b):
    a = np.asarray(a, dtype=float)[:, None, :]
    b = np.asarray(b, dtype=float)[None, :, :]
    ih = np.maximum(0.0, np.minimum(a[..., 3], b[..., 3]) - np.maximum(a[..., 1], b[..., 1]))
    uh = np.maximum(1e-10, np.maximum(a[..., 3], b[..., 3]) - np.minimum(a[..., 1], b[..., 1]))
    o = ih / uh
    iw = np.maximum(0.0, np.minimum(a[..., 2], b[..., 2]) - np.maximum(a[..., 0], b[..., 0]))
    inter = iw * ih
    area_a = (a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1])
    area_b = (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1])
    return inter / (area_a + area_b - inter + 1e-10) * o


def giou_batch(a, b):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    iou = iou_batch(a, b)
    A = a[:, None, :]
    B = b[None, :, :]
    wc = np.maximum(A[..., 2], B[..., 2]) - np.minimum(A[..., 0], B[..., 0])
    hc = np.maximum(A[..., 3], B[..., 3]) - np.minimum(A[..., 1], B[..., 1])
    iw = np.maximum(0.0, np.minimum(A[..., 2], B[..., 2]) - np.maximum(A[..., 0], B[..., 0]))
    ih = np.maximum(0.0, np.minimum(A[..., 3], B[..., 3]) - np.maximum(A[..., 1], B[..., 1]))
    inter = iw * ih
    union = ((A[..., 2] - A[..., 0]) * (A[..., 3] - A[..., 1])
             + (B[..., 2] - B[..., 0]) * (B[..., 3] - B[..., 1]) - inter)
    area_c = wc * hc
    g = iou - (area_c - union) / area_c
    return (g + 1.0) / 2.0


def diou_batch(a, b):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    iou = iou_batch(a, b)
    A = a[:, None, :]
    B = b[None, :, :]
    cxa, cya = (A[..., 0] + A[..., 2]) / 2.0, (A[..., 1] + A[..., 3]) / 2.0
    cxb, cyb = (B[..., 0] + B[..., 2]) / 2.0, (B[..., 1] + B[..., 3]) / 2.0
    inner = (cxa - cxb) ** 2 + (cya - cyb) ** 2
    ow = np.maximum(A[..., 2], B[..., 2]) - np.minimum(A[..., 0], B[..., 0])
    oh = np.maximum(A[..., 3], B[..., 3]) - np.minimum(A[..., 1], B[..., 1])
    outer = ow ** 2 + oh ** 2
    return (iou - inner / outer + 1.0) / 2.0


def ciou_batch(a, b):
    """iou.py:199-264 (fork: eps=1e-7 in iou / outer diagonal / heights / alpha, and the
    aspect term is ADDED: iou - rho + alpha*v)."""
    eps = 1e-7
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    A = a[:, None, :]
    B = b[None, :, :]
    iw = np.maximum(0.0, np.minimum(A[..., 2], B[..., 2]) - np.maximum(A[..., 0], B[..., 0]))
    ih = np.maximum(0.0, np.minimum(A[..., 3], B[..., 3]) - np.maximum(A[..., 1], B[..., 1]))
    inter = iw * ih
    area_a = (A[..., 2] - A[..., 0]) * (A[..., 3] - A[..., 1])
    area_b = (B[..., 2] - B[..., 0]) * (B[..., 3] - B[..., 1])
    iou = inter / (area_a + area_b - inter + eps)
    cxa, cya = (A[..., 0] + A[..., 2]) / 2.0, (A[..., 1] + A[..., 3]) / 2.0
    cxb, cyb = (B[..., 0] + B[..., 2]) / 2.0, (B[..., 1] + B[..., 3]) / 2.0
    inner = (cxa - cxb) ** 2 + (cya - cyb) ** 2
    ow = np.maximum(A[..., 2], B[..., 2]) - np.minimum(A[..., 0], B[..., 0])
    oh = np.maximum(A[..., 3], B[..., 3]) - np.minimum(A[..., 1], B[..., 1])
    outer = ow ** 2 + oh ** 2 + eps
    w1 = A[..., 2] - A[..., 0]
    h1 = A[..., 3] - A[..., 1] + eps
    w2 = B[..., 2] - B[..., 0]
    h2 = B[..., 3] - B[..., 1] + eps
    arct = np.arctan(w2 / h2) - np.arctan(w1 / h1)
    v = (4 / (np.pi ** 2)) * (arct ** 2)
    alpha = v / ((1 - iou) + v + eps)
    return (iou - inner / outer + alpha * v + 1) / 2.0


def centroid_batch(a, b, w, h):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    ca = np.stack(((a[:, 0] + a[:, 2]) / 2, (a[:, 1] + a[:, 3]) / 2), axis=-1)[:, None, :]
    cb = np.stack(((b[:, 0] + b[:, 2]) / 2, (b[:, 1] + b[:, 3]) / 2), axis=-1)[None, :, :]
    dist = np.sqrt(np.sum((ca - cb) ** 2, axis=-1))
    return 1 - dist / np.sqrt(w ** 2 + h ** 2)


def asso_func(name, w=1920, h=1080):
    table = {"iou": iou_batch, "hmiou": hmiou_batch, "giou": giou_batch,
             "diou": diou_batch, "ciou": ciou_batch}
    if name == "centroid":
        return lambda a, b: centroid_batch(a, b, w, h)
    return table[name]


# ------------------------------------------------------------------ assignment wrappers

def lap_pairs_unlimited(cost):
    """association.py:105-114: lapjv(cost, extend_cost=True) -> (row, col) pairs in row order."""
    _, x, _ = lap.lapjv(cost, extend_cost=True)
    rows = np.nonzero(x >= 0)[0]
    return np.stack([rows, x[rows]], axis=1) if rows.size else np.empty((0, 2), dtype=int)


def lap_thresholded(cost, thresh):
    """matching.py:30-108 (method 'lap'): matches with cost <= thresh, rows/cols left out."""
    if cost.size == 0:
        return (np.empty((0, 2), dtype=int), np.arange(cost.shape[0]), np.arange(cost.shape[1]))
    _, x, y = lap.lapjv(cost, extend_cost=True, cost_limit=thresh)
    m = [[i, j] for i, j in enumerate(x) if j >= 0 and cost[i, j] <= thresh]
    m = np.asarray(m, dtype=int) if m else np.empty((0, 2), dtype=int)
    return m, np.nonzero(x < 0)[0], np.nonzero(y < 0)[0]


def iou_distance(a_boxes, b_boxes):
    if len(a_boxes) == 0 or len(b_boxes) == 0:
        return np.zeros((len(a_boxes), len(b_boxes)), dtype=np.float32)
    return 1 - iou_batch(np.asarray(a_boxes), np.asarray(b_boxes))


def fuse_score(cost, det_confs):
    """matching.py:520-544 (confidence_threshold=0.5, confidence_weighting=True)."""
    if cost.size == 0:
        return cost
    c = np.repeat(np.asarray(det_confs, dtype=float)[None, :], cost.shape[0], axis=0)
    sim = 1 - cost
    weight = np.where(c > 0.7, c * 1.2, c)
    fused = 1 - sim * weight * (c >= 0.5)
    return np.where(c < 0.5, fused * 2.0, fused)


# ------------------------------------------------------------------ OCM association

def ocm_associate(dets, trks, iou_fn, iou_thr, velocities, prev_obs, inertia):
    """enhanced_associate (association.py:377-533 + R1-R3), emb_cost=None path.

    dets (D,5) trks (T,5) velocities (T,2)=[vy,vx] prev_obs (T,5).
    Returns matches (M,2)[det,trk], unmatched_dets, unmatched_trks (order as the
    reference builds them: never-assigned ascending, then IoU-rejected in row order)."""
    D, T = len(dets), len(trks)
    if T == 0:
        return np.empty((0, 2), dtype=int), np.arange(D), np.empty((0,), dtype=int)
    prev = prev_obs
    cxd, cyd = (dets[:, 0] + dets[:, 2]) / 2.0, (dets[:, 1] + dets[:, 3]) / 2.0
    cxp, cyp = (prev[:, 0] + prev[:, 2]) / 2.0, (prev[:, 1] + prev[:, 3]) / 2.0
    dx = cxd[None, :] - cxp[:, None]
    dy = cyd[None, :] - cyp[:, None]
    norm = np.sqrt(dx ** 2 + dy ** 2) + 1e-6
    dx, dy = dx / norm, dy / norm                      # (T, D)
    cosv = velocities[:, 1][:, None] * dx + velocities[:, 0][:, None] * dy
    cosv = np.clip(cosv, -1, 1)
    ang = (np.pi / 2.0 - np.abs(np.arccos(cosv))) / np.pi
    valid = np.where(prev[:, 4] < 0, 0.0, 1.0)[:, None]
    motion = ((valid * ang) * inertia).T              # (D, T)
    iou = iou_fn(dets, trks)
    total = iou + motion
    if min(D, T) > 0:
        a = (iou > iou_thr).astype(np.int32)
        if a.sum(1).max() == 1 and a.sum(0).max() == 1:
            pairs = np.stack(np.where(a), axis=1)
        else:
            pairs = lap_pairs_unlimited(-total)
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
