"""TEST INFRASTRUCTURE (oracle): the reference's per_class_decorator (boxmot/trackers/basetracker.py:154-201) over
the numpy oracles, for the trackers whose classes it couples.  For every class id 0 .. nr_classes-1 the decorator
makes per_class_active_tracks[cls] the tracker's active_tracks (:183), resets frame_count to the frame's starting
value (:186), calls update with that class's rows (det_ind = index within them) and stores active_tracks back (:192);
frame_count ends at start + 1 (:198).  Nothing else is swapped: ByteTrack / BoT-SORT keep ONE lost list
(bytetrack.py:143-144, botsort.py:68-69), BoostTrack keeps ONE self.trackers (boosttrack.py:185, swap=False)."""
import numpy as np


def per_class_update(oracle, dets, embs=None, nr_classes=80, swap=True, **kw):
    dets = np.asarray(dets, dtype=float).reshape(-1, 6)
    if swap and not hasattr(oracle, "_per_class_active"):
        oracle._per_class_active = {c: [] for c in range(nr_classes)}
    fc = oracle.frame_count
    out = []
    for c in range(nr_classes):
        m = dets[:, 5] == c
        if swap:
            oracle.active = oracle._per_class_active[c]
        oracle.frame_count = fc
        r = oracle.update(dets[m], **kw) if embs is None else oracle.update(dets[m], embs=embs[m], **kw)
        if swap:
            oracle._per_class_active[c] = oracle.active
        r = np.asarray(r, dtype=float)
        if r.size:
            out.append(r.reshape(-1, 8))
    oracle.frame_count = fc + 1
    return np.vstack(out) if out else np.empty((0, 8))
