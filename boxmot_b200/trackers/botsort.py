"""BoT-SORT behind the reference plugin API (boxmot/trackers/botsort/botsort.py:26-411).

ReID features and the camera-motion warp are SIDE INPUTS of the association path: pass
``embs`` to ``update`` (the reference computes them with its ReID backend when ``embs`` is
None, botsort.py:105-108 -- no model runs here), and the warp is the identity unless
``update(..., warp=H)`` provides a 2x3 matrix (what ``self.cmc.apply`` returns, :218)."""
import numpy as np

from boxmot_b200.engine import BotSortEngine
from boxmot_b200.trackers.basetracker import BaseTracker


class BotSort(BaseTracker):
    def __init__(self, reid_weights=None, device=None, half=None, per_class=False, track_high_thresh=0.5,
                 track_low_thresh=0.1, new_track_thresh=0.6, track_buffer=30, match_thresh=0.8,
                 proximity_thresh=0.5, appearance_thresh=0.25, cmc_method="ecc", frame_rate=30,
                 fuse_first_associate=False, with_reid=True, track_cap=512, det_cap=512, cuda_device="cuda:0"):
        super().__init__(per_class=bool(per_class))
        self.lost_stracks = []
        self.removed_stracks = []
        self.track_high_thresh = track_high_thresh
        self.track_low_thresh = track_low_thresh
        self.new_track_thresh = new_track_thresh
        self.match_thresh = match_thresh
        self.buffer_size = int(frame_rate / 30.0 * track_buffer)
        self.max_time_lost = self.buffer_size
        self.proximity_thresh = proximity_thresh
        self.appearance_thresh = appearance_thresh
        self.with_reid = with_reid
        self.fuse_first_associate = fuse_first_associate
        self._kw = dict(track_high_thresh=track_high_thresh, track_low_thresh=track_low_thresh,
                        new_track_thresh=new_track_thresh, track_buffer=track_buffer, match_thresh=match_thresh,
                        proximity_thresh=proximity_thresh, appearance_thresh=appearance_thresh,
                        frame_rate=frame_rate, fuse_first_associate=fuse_first_associate, with_reid=with_reid)
        self._caps = (track_cap, det_cap)
        self._device = cuda_device
        self._engine = None

    def update(self, dets, img=None, embs=None, warp=None):
        dets = self._prepare(dets, img)
        self.check_inputs(dets, img, embs)
        if self.with_reid and embs is None and len(dets):
            raise RuntimeError("with_reid=True needs `embs`: the ReID backend is outside the association path")
        if self._engine is None:
            e = int(embs.shape[1]) if (self.with_reid and embs is not None) else 0
            if self.with_reid and e == 0:
                self.frame_count += 1          # nothing to associate yet; the engine starts with the first embeddings
                self._skipped = getattr(self, "_skipped", 0) + 1
                return np.asarray([])
            self._engine = BotSortEngine(1, track_cap=self._caps[0], det_cap=self._caps[1], device=self._device,
                                         emb_dim=max(e, 1), **self._kw)
            for _ in range(getattr(self, "_skipped", 0)):     # empty frames seen before: advance the frame counter
                self._engine.step_host([np.empty((0, 6))], [np.empty((0, max(e, 1)))])
        self.frame_count += 1
        if self.per_class:
            return self._update_per_class(dets, embs, warp)
        rows = self._engine.step_host([dets], [embs] if self.with_reid else None,
                                      None if warp is None else [np.asarray(warp, dtype=np.float64)])[0]
        self._engine.status()
        return rows if len(rows) else np.asarray([])

    def _update_per_class(self, dets, embs, warp):
        """per_class_decorator (basetracker.py:154-201): one update per class id on that class's active_tracks with
        the lost / removed lists shared by all classes (botsort.py:68-69).  ``warp``: None, an array [nr_classes, 2, 3]
        (what ``self.cmc.apply`` returned in each class call), or one 2x3 matrix, which is then the FIRST class call's
        warp and the identity afterwards -- what the reference's stateful CMC yields when asked nr_classes times about
        the same image."""
        C = self.nr_classes
        sel = [(dets[:, 5] == c) if dets.size else np.zeros((0,), bool) for c in range(C)]
        per_cls = [dets[m] for m in sel]
        per_emb = [embs[m] for m in sel] if self.with_reid else None
        warps = None
        if warp is not None:
            w = np.asarray(warp, dtype=np.float64)
            warps = w if w.ndim == 3 else np.concatenate([w[None], np.tile(np.eye(2, 3), (C - 1, 1, 1))])
        rows = [r for r in self._engine.step_classes_host(per_cls, per_emb, warps) if len(r)]
        self._engine.status()
        return np.vstack(rows) if rows else np.empty((0, 8))

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
        self.frame_count = 0
