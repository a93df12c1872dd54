"""StrongSORT's matching cascade composed from the C-ABI units (function-level; the Tracker-level
control flow of boxmot/trackers/strongsort/sort/tracker.py is not built).

``matching_cascade`` follows boxmot/trackers/strongsort/sort/linear_assignment.py:90-171 with the
``gated_metric`` of tracker.py:186-206 as the distance: per time_since_update level (ascending,
tracks of a level ordered by -(quality + stability), :276-285) the appearance distance
(bm_nn_cosine_distance, tcgen05), its gating / motion blend / id-preservation weighting
(bm_strongsort_gated_metric), _enhance_cost_matrix (bm_strongsort_enhance_cost) and the thresholded
exact assignment (bm_min_cost_matching_batch) all run on the device; the host only slices the level's
rows and columns.  No CPU fallback: every step needs the CUDA library.

Difference kept on purpose: detections left unmatched by a level are handed to the next level in
ascending index order; the reference hands them over in scipy's order (never-assigned columns, then
rejected assignments), which only matters for ties between equal costs.
"""
import numpy as np
import torch

from boxmot_b200 import _lib

TRK_COST_FIELDS = 13          # leading columns of the per-track table read by the cost units
TS_TSU, TS_QUALITY, TS_STABILITY = 2, 5, 14


def _dev(a, dtype, device):
    return torch.as_tensor(np.ascontiguousarray(a, dtype=dtype)).to(device)


def gated_min_cost_matching(means, covs, trk, bank, nb, det_xyah, det_attr, feats, track_indices, detection_indices,
                            max_distance, mc_lambda, id_weight, device="cuda:0"):
    """min_cost_matching(gated_metric, ...) for one list of tracks / detections.  Returns
    (matches [(track index, detection index)], unmatched detection indices ascending)."""
    ti, di = list(track_indices), list(detection_indices)
    if not ti or not di:
        return [], di
    L = _lib.lib()
    T, D, E, cap = len(ti), len(di), feats.shape[1], bank.shape[1]
    with torch.cuda.device(device):
        d_bank, d_nb = _dev(bank[ti], np.float32, device), _dev(nb[ti], np.int32, device)
        d_feat = _dev(feats[di], np.float32, device)
        d_nn = torch.empty((T, D), dtype=torch.float32, device=device)
        d_work = torch.empty((max(cap, D) + T * cap + D + 64,), dtype=torch.float32, device=device)
        _lib.check(L.bm_nn_cosine_distance(d_bank.data_ptr(), d_nb.data_ptr(), T, cap, d_feat.data_ptr(), D, E,
                                           d_nn.data_ptr(), d_work.data_ptr(), None), "bm_nn_cosine_distance")
        d_cost = d_nn.double().contiguous()
        d_mean, d_cov = _dev(means[ti], np.float64, device), _dev(covs[ti], np.float64, device)
        d_trk = _dev(trk[ti][:, :TRK_COST_FIELDS], np.float64, device)
        d_meas = _dev(det_xyah[di], np.float64, device)
        _lib.check(L.bm_strongsort_gated_metric(d_cost.data_ptr(), d_mean.data_ptr(), d_cov.data_ptr(), d_trk.data_ptr(),
                                                T, d_meas.data_ptr(), D, float(mc_lambda), float(id_weight), None),
                   "bm_strongsort_gated_metric")
        d_det = _dev(det_attr[di], np.float64, device)
        _lib.check(L.bm_strongsort_enhance_cost(d_cost.data_ptr(), d_trk.data_ptr(), T, d_det.data_ptr(), D, None),
                   "bm_strongsort_enhance_cost")
        d_nr, d_nc = _dev([T], np.int32, device), _dev([D], np.int32, device)
        d_x = torch.empty((1, T), dtype=torch.int32, device=device)
        d_y = torch.empty((1, D), dtype=torch.int32, device=device)
        _lib.check(L.bm_min_cost_matching_batch(d_cost.data_ptr(), 1, T, D, d_nr.data_ptr(), d_nc.data_ptr(),
                                                float(max_distance), d_x.data_ptr(), d_y.data_ptr(), None),
                   "bm_min_cost_matching_batch")
        x, y = d_x.cpu().numpy()[0], d_y.cpu().numpy()[0]
    matches = [(ti[r], di[c]) for r, c in enumerate(x) if c >= 0]
    return matches, [di[c] for c in range(D) if y[c] < 0]


def matching_cascade(means, covs, trk, bank, nb, det_xyah, det_attr, feats, track_indices, detection_indices,
                     max_distance, cascade_depth, mc_lambda=0.995, id_weight=0.1, device="cuda:0"):
    """Returns (matches, unmatched track indices, unmatched detection indices)."""
    track_indices, unmatched = list(track_indices), sorted(detection_indices)
    levels = {}
    for t in track_indices:
        levels.setdefault(int(trk[t, TS_TSU]), []).append(t)
    matches = []
    for age in sorted(levels):
        if age > cascade_depth:
            break
        order = sorted(levels[age], key=lambda t: -(trk[t, TS_QUALITY] + trk[t, TS_STABILITY]))
        m, unmatched = gated_min_cost_matching(means, covs, trk, bank, nb, det_xyah, det_attr, feats, order, unmatched,
                                               max_distance, mc_lambda, id_weight, device)
        matches += m
    matched_t = {t for t, _ in matches}
    return matches, [t for t in track_indices if t not in matched_t], unmatched
