"""DeepOCSORT behind the reference plugin API (boxmot/trackers/deepocsort/deepocsort.py:238-498).

ReID features and the camera-motion warp are SIDE INPUTS of the association path: pass ``embs``
to ``update`` (the reference computes them with its ReID backend when ``embs`` is None,
deepocsort.py:346-351 -- no model runs here); the CMC correction is applied when
``update(..., warp=H)`` provides the 2x3 matrix that ``self.cmc.apply`` returns (:354-357)."""
import numpy as np

from boxmot_b200.engine import DeepOcSortEngine
from boxmot_b200.trackers.basetracker import BaseTracker


class DeepOcSort(BaseTracker):
    def __init__(self, reid_weights=None, device=None, half=None, per_class=False, det_thresh=0.3, max_age=30,
                 min_hits=3, iou_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.2, w_association_emb=0.5,
                 alpha_fixed_emb=0.95, aw_param=0.5, embedding_off=False, cmc_off=False, aw_off=False,
                 Q_xy_scaling=0.01, Q_s_scaling=0.0001, track_cap=512, det_cap=512, cuda_device="cuda:0",
                 **kwargs):
        # unknown keys (e.g. the YAML's `iou_thresh`) are swallowed like the reference's **kwargs
        super().__init__(max_age=max_age, per_class=bool(per_class), asso_func=asso_func)
        self.max_age, self.min_hits, self.iou_threshold, self.det_thresh = max_age, min_hits, iou_threshold, det_thresh
        self.delta_t, self.asso_func, self.inertia = delta_t, asso_func, inertia
        self.w_association_emb, self.alpha_fixed_emb, self.aw_param = w_association_emb, alpha_fixed_emb, aw_param
        self.Q_xy_scaling, self.Q_s_scaling = Q_xy_scaling, Q_s_scaling
        self.embedding_off, self.cmc_off, self.aw_off = embedding_off, cmc_off, aw_off
        self._kw = dict(det_thresh=det_thresh, max_age=max_age, min_hits=min_hits, iou_threshold=iou_threshold,
                        delta_t=delta_t, asso_func=asso_func, inertia=inertia, w_association_emb=w_association_emb,
                        alpha_fixed_emb=alpha_fixed_emb, aw_param=aw_param, embedding_off=embedding_off,
                        cmc_off=cmc_off, aw_off=aw_off, Q_xy_scaling=Q_xy_scaling, Q_s_scaling=Q_s_scaling)
        self._caps = (track_cap, det_cap)
        self._device = cuda_device
        self._engine = None
        self._skipped = 0

    def update(self, dets, img=None, embs=None, warp=None):
        dets = self._prepare(dets, img)
        self.check_inputs(dets, img, embs)
        if not self.embedding_off and embs is None and len(dets):
            raise RuntimeError("DeepOCSORT needs `embs` unless embedding_off: the ReID backend is outside the "
                               "association path")
        if self._engine is None:
            e = int(embs.shape[1]) if (not self.embedding_off and embs is not None) else 0
            if not self.embedding_off and e == 0:
                self.frame_count += 1          # nothing to associate yet; the engine starts with the first embeddings
                self._skipped += 1
                return np.asarray([])
            h, w = (img.shape[0], img.shape[1]) if img is not None else (1080, 1920)
            n = self.nr_classes if self.per_class else 1
            self._created = np.zeros((n,), dtype=np.int64)      # tracks created so far, per class
            self._global_ids = [[] for _ in range(n)]           # class-local id -> tracker-wide id
            self._count = 1                                      # KalmanBoxTracker.count (deepocsort.py:301)
            self._engine = DeepOcSortEngine(n, track_cap=self._caps[0], det_cap=self._caps[1], device=self._device,
                                            emb_dim=max(e, 1), img_wh=(w, h), **self._kw)
            for _ in range(self._skipped):     # empty frames seen before: advance the frame counter
                self._engine.step_host([np.empty((0, 6))] * n, [np.empty((0, max(e, 1)))] * n)
        self.frame_count += 1
        if self.per_class:
            return self._update_per_class(dets, embs, warp)
        rows = self._engine.step_host([dets], None if self.embedding_off else [embs],
                                      None if warp is None else [np.asarray(warp, dtype=np.float64)])[0]
        self._engine.status()
        return rows if len(rows) else np.asarray([])

    def _update_per_class(self, dets, embs, warp):
        """per_class_decorator (basetracker.py:155-201): update runs once per class id on the rows of
        that class (det_ind relative to the subset, the frame's warp applied to every class list);
        new ids come from ONE counter while the classes are walked in order (deepocsort.py:56-58)."""
        n = self.nr_classes
        idx = [np.where(dets[:, 5] == c)[0] if dets.size else np.empty((0,), dtype=int) for c in range(n)]
        per_dets = [dets[i] for i in idx]
        per_embs = None if self.embedding_off else [
            embs[i] if embs is not None and len(embs) else np.empty((0, self._engine.E)) for i in idx]
        warps = None if warp is None else [np.asarray(warp, dtype=np.float64)] * n
        rows = self._engine.step_host(per_dets, per_embs, warps)
        self._engine.status()
        _, created = self._engine.stream_info()
        for c in range(n):
            for _ in range(int(created[c] - self._created[c])):
                self._global_ids[c].append(self._count)
                self._count += 1
        self._created[:] = created
        out = []
        for c, r in enumerate(rows):
            if len(r):
                r = r.copy()
                r[:, 4] = [self._global_ids[c][int(i) - 1] for i in r[:, 4]]
                out.append(r)
        return np.vstack(out) if out else np.empty((0, 8))

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
            self._created[:] = 0
            self._global_ids = [[] for _ in self._global_ids]
            self._count = 1
        self.frame_count = 0
