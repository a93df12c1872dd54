"""BoostTrack(++) behind the reference plugin API (boxmot/trackers/boosttrack/boosttrack.py:124-336).

ReID features and the camera-motion warp are SIDE INPUTS of the association path: pass ``embs``
to ``update`` when ``with_reid`` (the reference computes them with its ReID backend when ``embs``
is None, :264-267 -- no model runs here); ``camera_update`` is applied when ``update(..., warp=H)``
provides the 2x3 matrix that ``self.cmc.apply`` returns (:243-246)."""
import numpy as np

from boxmot_b200.engine import BoostTrackEngine
from boxmot_b200.trackers.basetracker import BaseTracker


class BoostTrack(BaseTracker):
    def __init__(self, reid_weights=None, device=None, half=None, max_age=60, min_hits=3, det_thresh=0.6,
                 iou_threshold=0.3, use_ecc=True, min_box_area=10, aspect_ratio_thresh=1.6, cmc_method="ecc",
                 lambda_iou=0.5, lambda_mhd=0.25, lambda_shape=0.25, use_dlo_boost=True, use_duo_boost=True,
                 dlo_boost_coef=0.65, s_sim_corr=False, use_rich_s=False, use_sb=False, use_vt=False,
                 with_reid=False, per_class=False, track_cap=512, det_cap=512, cuda_device="cuda:0"):
        super().__init__(per_class=bool(per_class))
        self.max_age, self.min_hits, self.det_thresh, self.iou_threshold = max_age, min_hits, det_thresh, iou_threshold
        self.use_ecc, self.min_box_area, self.aspect_ratio_thresh = use_ecc, min_box_area, aspect_ratio_thresh
        self.cmc_method = cmc_method
        self.lambda_iou, self.lambda_mhd, self.lambda_shape = lambda_iou, lambda_mhd, lambda_shape
        self.use_dlo_boost, self.use_duo_boost, self.dlo_boost_coef = use_dlo_boost, use_duo_boost, dlo_boost_coef
        self.s_sim_corr, self.use_rich_s, self.use_sb, self.use_vt = s_sim_corr, use_rich_s, use_sb, use_vt
        self.with_reid = with_reid
        self._kw = dict(max_age=max_age, min_hits=min_hits, det_thresh=det_thresh, iou_threshold=iou_threshold,
                        use_ecc=use_ecc, min_box_area=min_box_area, aspect_ratio_thresh=aspect_ratio_thresh,
                        lambda_iou=lambda_iou, lambda_mhd=lambda_mhd, lambda_shape=lambda_shape,
                        use_dlo_boost=use_dlo_boost, use_duo_boost=use_duo_boost, dlo_boost_coef=dlo_boost_coef,
                        s_sim_corr=s_sim_corr, use_rich_s=use_rich_s, use_sb=use_sb, use_vt=use_vt,
                        with_reid=with_reid)
        self._caps = (track_cap, det_cap)
        self._device = cuda_device
        self._engine = None
        self._skipped = 0

    def update(self, dets, img=None, embs=None, warp=None):
        dets = self._prepare(dets, img)
        self.check_inputs(dets, img, embs)
        if self.with_reid and embs is None and len(dets):
            raise RuntimeError("with_reid=True needs `embs`: the ReID backend is outside the association path")
        if self._engine is None:
            e = int(embs.shape[1]) if (self.with_reid and embs is not None) else 0
            if self.with_reid and e == 0:
                self.frame_count += 1          # nothing to associate yet; the engine starts with the first embeddings
                self._skipped += 1
                return np.empty((0, 8))
            self._engine = BoostTrackEngine(1, track_cap=self._caps[0], det_cap=self._caps[1], device=self._device,
                                            emb_dim=max(e, 1), **self._kw)
            for _ in range(self._skipped):     # empty frames seen before: advance the frame counter
                self._engine.step_host([np.empty((0, 6))], [np.empty((0, max(e, 1)))])
        self.frame_count += 1
        if self.per_class:
            # per_class_decorator (basetracker.py:154-201) over a tracker whose tracks live in self.trackers
            # (boosttrack.py:185), which the decorator does not swap: every class call predicts, matches and ages the
            # ONE pool with that class's detections at the same frame_count -- a track survives a frame only if fewer
            # than max_age class calls miss it.  Reproduced as nr_classes steps of the one stream.
            C = self.nr_classes
            sel = [(dets[:, 5] == c) if dets.size else np.zeros((0,), bool) for c in range(C)]
            warps = None
            if warp is not None:
                w = np.asarray(warp, dtype=np.float64)
                warps = w if w.ndim == 3 else np.concatenate([w[None], np.tile(np.eye(2, 3), (C - 1, 1, 1))])
            rows = self._engine.step_classes_host([dets[m] for m in sel],
                                                  [embs[m] for m in sel] if self.with_reid else None, warps)
            rows = [r for r in rows if len(r)]
            self._engine.status()
            return np.vstack(rows) if rows else np.empty((0, 8))
        rows = self._engine.step_host([dets], [embs] if self.with_reid else None,
                                      None if warp is None else [np.asarray(warp, dtype=np.float64)])[0]
        self._engine.status()
        return rows if len(rows) else np.empty((0, 8))

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
        self.frame_count = 0
