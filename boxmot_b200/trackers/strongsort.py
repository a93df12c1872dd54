"""StrongSORT (the fork's "enhanced" StrongSORT) behind the reference plugin API
(boxmot/trackers/strongsort/strongsort.py:18-359).

Like the reference's class this one is NOT a BaseTracker: its ``update`` does its own input checks
(:119-134, AssertionError) and has no float32 round trip; the engine takes float32 detections, so
the rows are rounded once here.  Output = the first 8 of the fork's 10 columns.

Scope: the fork only runs with its occlusion handler and crowd detection off (R5: the
handler raises TypeError on the first mutual occlusion), with the unbound variable of
``_enhanced_match`` repaired (R4) and -- for any track ever to be confirmed -- with tracks born
Confirmed, which is what the reference does in its own CI mode (M1, track.py:98-105).  That is the
behaviour reproduced; ``handle_occlusions`` / ``crowd_detection`` are accepted and have no effect.

ReID and CMC are side inputs of the association path: ``embs`` is required (the reference runs its
ReID model when it is None, :175-178), the camera-motion warp that ``self.cmc.apply`` returns
(:169-172) is passed as ``update(..., warp=H)`` (2x3; identity when omitted)."""
import warnings

import numpy as np

from boxmot_b200.engine import StrongSortEngine


class StrongSort(object):
    def __init__(self, reid_weights=None, device=None, half=None, per_class=False, min_conf=0.1, max_cos_dist=0.15,
                 max_iou_dist=0.7, max_age=50, n_init=2, nn_budget=150, mc_lambda=0.995, ema_alpha=0.9,
                 conf_thresh_high=0.7, conf_thresh_low=0.3, id_preservation_weight=0.1, adaptive_matching=True,
                 appearance_weight=0.6, motion_weight=0.4, occlusion_threshold=0.3, handle_occlusions=True,
                 crowd_detection=True, born_confirmed=True, track_cap=1024, det_cap=256, pool_cap=None,
                 cuda_device="cuda:0"):
        self.per_class = per_class
        self.min_conf = min_conf
        self.conf_thresh_high, self.conf_thresh_low = conf_thresh_high, conf_thresh_low
        self.id_preservation_weight = id_preservation_weight
        self.adaptive_matching, self.appearance_weight, self.motion_weight = adaptive_matching, appearance_weight, motion_weight
        self.handle_occlusions, self.crowd_detection = handle_occlusions, crowd_detection
        self.occlusion_threshold = occlusion_threshold
        self._kw = dict(min_conf=min_conf, max_cos_dist=max_cos_dist, max_iou_dist=max_iou_dist, max_age=max_age,
                        n_init=n_init, nn_budget=nn_budget, mc_lambda=mc_lambda, ema_alpha=ema_alpha,
                        conf_thresh_high=conf_thresh_high, conf_thresh_low=conf_thresh_low,
                        id_preservation_weight=id_preservation_weight, born_confirmed=born_confirmed)
        self._caps = (track_cap, det_cap, pool_cap)
        self._device = cuda_device
        self._engine = None
        self._skipped = 0
        self.frame_count = 0
        if per_class:
            # the reference's decorator reads attributes this class never defines (basetracker.py:176-190)
            raise AttributeError("'StrongSort' object has no attribute 'nr_classes'")

    def update(self, dets, img, embs=None, warp=None):
        if dets is None or len(dets) == 0:               # per_class_decorator (basetracker.py:162-163)
            dets = np.empty((0, 6))
        assert isinstance(dets, np.ndarray), \
            f"Unsupported 'dets' input format '{type(dets)}', valid format is np.ndarray"
        assert isinstance(img, np.ndarray), \
            f"Unsupported 'img' input format '{type(img)}', valid format is np.ndarray"
        assert len(dets.shape) == 2, "Unsupported 'dets' dimensions, valid number of dimensions is two"
        assert dets.shape[1] == 6, "Unsupported 'dets' 2nd dimension lenght, valid lenghts is 6"
        if embs is not None:
            assert dets.shape[0] == embs.shape[0], "Missmatch between detections and embeddings sizes"
        if embs is None and len(dets):
            raise RuntimeError("StrongSort needs `embs`: the ReID backend is outside the association path")
        if self._engine is None:
            e = int(embs.shape[1]) if embs is not None and embs.ndim == 2 else 0
            if e == 0:
                self.frame_count += 1                    # nothing to track yet; replayed once the size is known
                self._skipped += 1
                return np.array([])
            self._engine = StrongSortEngine(1, track_cap=self._caps[0], det_cap=self._caps[1],
                                            pool_cap=self._caps[2], device=self._device, emb_dim=e, **self._kw)
            for _ in range(self._skipped):
                self._engine.step_host([np.empty((0, 6))], [np.empty((0, e))])
        self.frame_count += 1
        n = len(dets)
        rows = self._engine.step_host([np.asarray(dets, dtype=np.float32)],
                                      [np.asarray(embs, dtype=np.float32) if n else np.empty((0, self._engine.E))],
                                      None if warp is None else [np.asarray(warp, dtype=np.float64)])[0]
        self._engine.status()
        return rows if len(rows) else np.array([])

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
        self.frame_count = 0
        self._skipped = 0
